"""NVLink P2P collectives: (1) all ranks emulated in one launch on a single GPU; (2) real multi-process run."""
import json
import os
import subprocess
import sys

import pytest
import torch

from rt1b200.ops import reference as ref

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
DEV = "cuda"


@pytest.mark.parametrize("world", [2, 4, 8])
def test_emulated_allreduce_adamw_matches_mean_gradient_adam(world):
    from rt1b200.parallel.symm import EmulatedComm
    torch.manual_seed(world)
    n, off = 64 * 300, 64 * 5                      # a bucket that starts inside the flat buffers
    total = off + n + 64
    comm = EmulatedComm(world, ctas_per_rank=4)
    master0 = torch.randn(total, device=DEV)
    masters = [master0.clone() for _ in range(world)]
    ms = [torch.zeros(total, device=DEV) for _ in range(world)]
    vs = [torch.zeros(total, device=DEV) for _ in range(world)]
    params = [master0.to(torch.bfloat16) for _ in range(world)]
    want_p, want_m, want_v = master0.clone(), torch.zeros(total, device=DEV), torch.zeros(total, device=DEV)
    for step in range(1, 4):
        grads = [(torch.randn(total, device=DEV) * 0.1).to(torch.bfloat16) for _ in range(world)]
        g_sum = torch.stack([g.float() for g in grads]).sum(0)
        h = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, bc1=1 - 0.9 ** step, bc2=1 - 0.999 ** step)
        keep = [g.clone() for g in grads]
        comm.allreduce_adamw(grads, params, masters, ms, vs, off, off + n, h)
        torch.cuda.synchronize()
        ref.adamw_step(want_p[off:off + n], g_sum[off:off + n], want_m[off:off + n], want_v[off:off + n], 1e-3, 0.9, 0.999,
                       1e-8, 0.0, step, grad_scale=1.0 / world)
        # every rank's bf16 parameters agree with the oracle, inside the bucket only
        for r in range(world):
            # bf16 rounding of two fp32 values one ulp apart may differ: allow one bf16 ulp
            torch.testing.assert_close(params[r][off:off + n].float(), want_p[off:off + n].to(torch.bfloat16).float(),
                                       rtol=8e-3, atol=1e-6, msg=lambda m: f"rank {r} step {step}: {m}")
            assert torch.equal(params[r][:off], master0[:off].to(torch.bfloat16))
            assert float(grads[r][off:off + n].abs().max()) == 0.0           # bucket cleared
            assert torch.equal(grads[r][:off], keep[r][:off])                # outside untouched
        # fp32 state: each element updated exactly once, by its owner
        chunk = ((n + world - 1) // world + 7) & ~7
        for r in range(world):
            a, b = min(n, chunk * r), min(n, chunk * r + chunk)
            torch.testing.assert_close(masters[r][off + a:off + b], want_p[off + a:off + b], rtol=1e-6, atol=1e-7)
            torch.testing.assert_close(ms[r][off + a:off + b], want_m[off + a:off + b], rtol=1e-5, atol=1e-8)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("two_shot", [True, False])
def test_emulated_allreduce(dtype, two_shot):
    from rt1b200.parallel.symm import EmulatedComm
    world, n = 4, 8 * 1000
    comm = EmulatedComm(world)
    bufs = [torch.zeros(2 * n, dtype=dtype, device=DEV) for _ in range(world)]
    src = [(torch.randn(n, device=DEV)).to(dtype) for _ in range(world)]
    for b, s in zip(bufs, src):
        b[:n].copy_(s)
    want = torch.stack([s.float() for s in src]).sum(0) * 0.25
    comm.allreduce_(bufs, n, scale=0.25, two_shot=two_shot)
    torch.cuda.synchronize()
    for b in bufs:
        got = b[:n] if two_shot else b[n:]
        torch.testing.assert_close(got.float(), want.to(dtype).float(), rtol=1e-2 if dtype == torch.bfloat16 else 1e-6,
                                   atol=1e-2 if dtype == torch.bfloat16 else 1e-6)


def _launch(nproc, port, extra=(), timeout=900):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "_p2p_worker.py"), *extra]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0 and "P2P_WORKER_OK" in r.stdout, r.stdout[-4000:] + r.stderr[-4000:]
    return r.stdout


@pytest.mark.multigpu
def test_p2p_world2_real_processes():
    _launch(2, 29621)


@pytest.mark.multigpu
def test_p2p_all_visible_gpus():
    n = torch.cuda.device_count()
    if n < 4:
        pytest.skip("needs >= 4 GPUs")
    _launch(min(n, 8), 29622)
