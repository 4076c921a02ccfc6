"""world_size=2 data-parallel plumbing (BASELINE.json config 1, second half): gloo on CPU here, and the same
worker over NCCL / the P2P engine on a multi-GPU box."""
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def _launch(nproc, backend, port, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "_dp_worker.py"), backend]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    assert r.returncode == 0 and "DP_WORKER_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_gloo_world2_matches_mean_gradient_adam():
    _launch(2, "cpu", 29611)


@pytest.mark.gpu
@pytest.mark.multigpu
def test_nccl_world2():
    _launch(2, "torchdist", 29612)


@pytest.mark.gpu
@pytest.mark.multigpu
def test_p2p_world2():
    _launch(2, "p2p", 29613)
