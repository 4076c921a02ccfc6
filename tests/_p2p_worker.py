"""Multi-process worker: symmetric memory, P2P all-reduce and the fused all-reduce+AdamW vs NCCL + reference Adam,
then a full RT-1 training step through the engine with backend='p2p'.  Optionally a bandwidth sweep (--sweep)."""
import json
import os
import sys
import time
from argparse import Namespace

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rt1b200.ops import reference as ref  # noqa: E402
from rt1b200.parallel.symm import SymmetricComm  # noqa: E402


def timed(fn, iters, world, dev):
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    comm = SymmetricComm()
    if rank == 0:
        print(f"[p2p] allocator={comm.allocator} ctas={comm.ctas_per_rank}", flush=True)

    # 1. all-reduce, every schedule, both dtypes
    for dtype in (torch.bfloat16, torch.float32):
        for n in (8 * 16, 8 * 12345, 8 * 1024 * 128):
            buf = comm.alloc_tensor(2 * n, dtype)
            algos = ["two_shot", "one_shot"] + (["nvls"] if buf._rt1_mc_ptr else [])
            for algo in algos:
                torch.manual_seed(7 + rank)
                x = torch.randn(n, device=dev).to(dtype)
                buf[:n].copy_(x)
                want = x.float().clone()
                dist.all_reduce(want)
                torch.cuda.synchronize(); dist.barrier()
                out = comm.allreduce_(buf, scale=1.0, n=n, algo=algo)
                torch.cuda.synchronize()
                tol = 3e-2 if dtype == torch.bfloat16 else 1e-5
                torch.testing.assert_close(out.float(), want.to(dtype).float(), rtol=tol, atol=tol,
                                           msg=lambda m: f"{algo} {dtype} n={n}: {m}")
    if rank == 0:
        print(f"[p2p] all-reduce ok; multicast={'yes' if buf._rt1_mc_ptr else 'no'}", flush=True)
    s = comm.allreduce_small(torch.tensor([float(rank + 1), 2.0], device=dev))
    assert s.tolist() == [world * (world + 1) / 2, 2.0 * world], s

    # 2. fused all-reduce + AdamW on a flat buffer vs NCCL + reference Adam
    n = 64 * 4096
    grad = comm.alloc_tensor(n, torch.bfloat16)
    param = comm.alloc_tensor(n, torch.bfloat16)
    comm.register_grad_buffer(grad); comm.register_param_buffer(param)
    torch.manual_seed(0)
    master = torch.randn(n, device=dev)
    param.copy_(master)
    m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    want_p, want_m, want_v = master.clone(), m.clone(), v.clone()
    flat = Namespace(master=master, exp_avg=m, exp_avg_sq=v)
    default_algo = comm.algo
    for step in range(1, 7):
        # alternate the P2P and (where a multicast mapping exists) the NVLS schedule on the same state
        comm.algo = "nvls" if (step % 2 == 0 and comm.has_multicast) else "p2p"
        torch.manual_seed(100 * step + rank)
        g = (torch.randn(n, device=dev) * 0.1).to(torch.bfloat16)
        grad.copy_(g)
        gs = g.float().clone(); dist.all_reduce(gs)
        if comm.algo == "nvls":
            gs = gs.to(torch.bfloat16).float()       # the switch returns the fp32-accumulated sum rounded to bf16
        h = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, bc1=1 - 0.9 ** step, bc2=1 - 0.999 ** step)
        for lo, hi, last in ((n // 2, n, False), (0, n // 2, True)):            # two chained buckets
            comm.allreduce_adamw(lo, hi, flat, h, final=last)
        torch.cuda.synchronize()
        ref.adamw_step(want_p, gs, want_m, want_v, 1e-3, 0.9, 0.999, 1e-8, 0.0, step, grad_scale=1.0 / world)
        # NVLS: the switch's bf16 rounding of the sum can differ by one ulp from ours; near-zero parameters then
        # differ by a fraction of the lr-sized update, hence the absolute floor
        torch.testing.assert_close(param.float(), want_p.to(torch.bfloat16).float(), rtol=8e-3,
                                   atol=1e-4 if comm.has_multicast else 1e-6,
                                   msg=lambda m_: f"step {step} algo {comm.algo}: {m_}")
        assert float(grad.abs().max()) == 0.0
    comm.algo = default_algo
    dist.barrier()
    if rank == 0:
        print(f"[p2p] fused all-reduce+AdamW ok (nvls tested: {comm.has_multicast})", flush=True)

    # 3. a real RT-1 training step through the engine
    from rt1b200.data.dataset import SyntheticBatchSource
    from rt1b200.models.factory import RT1Config
    from rt1b200.train.trainer import RT1Module, Trainer
    args = Namespace(height=64, width=96, seq_len=3, lr=1e-3, milestones=[], exp_name="p2p", log_dir="/tmp/rt1_p2p/logs",
                     ckpt_dir="/tmp/rt1_p2p/ckpt", log_every_n_steps=1000, ckpt_every_n_epochs=1, max_epochs=1)
    cfg = RT1Config(height=64, width=96, seq_len=3, num_layers=2, layer_size=128, num_heads=8, dropout_rate=0.0)
    torch.manual_seed(1234 + rank)
    tr = Trainer(RT1Module(args, cfg), args, device=dev, backend="p2p", param_dtype=torch.bfloat16, loggers=[], verbose=False)
    assert tr.engine.backend == "p2p"
    src = SyntheticBatchSource(2, 3, 64, 96, pool=2, seed=10 + rank, image_dtype=torch.uint8, pin=False)
    losses = [float(tr.train_step(next(src))) for _ in range(3)]
    torch.cuda.synchronize()
    assert all(l == l for l in losses), losses
    # parameters identical on every rank after the fused exchange
    p = tr.engine.flat.param.float()
    ref_p = p.clone(); dist.broadcast(ref_p, 0)
    assert torch.equal(p, ref_p), "bf16 parameters diverged across ranks"
    tr.engine.gather_sharded_state()
    mast = tr.engine.flat.master.clone(); dist.broadcast(mast, 0)
    assert torch.equal(tr.engine.flat.master, mast)
    torch.testing.assert_close(tr.engine.flat.master.to(torch.bfloat16).float(), p, rtol=0, atol=0)

    # 4. optional sweep: bus bandwidth of P2P two-shot / one-shot vs NCCL, and the fused step's roofline fraction
    if "--sweep" in sys.argv:
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from bench import ClockSampler
        sampler = ClockSampler(local)
        if rank == 0:
            sampler.start()
        rows = []
        sizes = [1 << k for k in range(10, 29)]          # 1 KB .. 256 MB
        big = comm.alloc_tensor((256 << 20) // 2 * 2, torch.bfloat16)
        nccl_buf = torch.zeros((256 << 20) // 2, dtype=torch.bfloat16, device=dev)
        for nbytes in sizes:
            n_el = max(8, nbytes // 2)
            iters = 50 if nbytes <= (1 << 22) else 10
            row = {"bytes": n_el * 2, "world": world}
            for name, fn in (("p2p_two_shot_us", lambda: comm.allreduce_(big, n=n_el, algo="two_shot")),
                             ("p2p_one_shot_us", (lambda: comm.allreduce_(big, n=n_el, algo="one_shot")) if nbytes <= (8 << 20) else None),
                             ("nvls_us", (lambda: comm.allreduce_(big, n=n_el, algo="nvls")) if big._rt1_mc_ptr else None),
                             ("nccl_us", lambda: dist.all_reduce(nccl_buf[:n_el]))):
                if fn is None:
                    continue
                for _ in range(3):
                    fn()
                row[name] = timed(fn, iters, world, dev) * 1e3
            best = min(row.get("p2p_two_shot_us", 1e30), row.get("p2p_one_shot_us", 1e30), row.get("nvls_us", 1e30))
            row["p2p_busbw_gbs"] = 2 * (world - 1) / world * row["bytes"] / (best * 1e-6) / 1e9
            row["nccl_busbw_gbs"] = 2 * (world - 1) / world * row["bytes"] / (row["nccl_us"] * 1e-6) / 1e9
            rows.append(row)
        # fused all-reduce + AdamW over an RT-1 sized flat buffer (35.3M elements) and the scaled variant (143M)
        fused = []
        for n_par in (35_324_352, 142_976_000):
            n_par = n_par // 64 * 64
            g2 = comm.alloc_tensor(n_par, torch.bfloat16); p2 = comm.alloc_tensor(n_par, torch.bfloat16)
            comm.register_grad_buffer(g2); comm.register_param_buffer(p2)
            fl = Namespace(master=torch.zeros(n_par, device=dev), exp_avg=torch.zeros(n_par, device=dev),
                           exp_avg_sq=torch.zeros(n_par, device=dev))
            h = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, bc1=0.1, bc2=0.001)
            import math
            nb_overlap = min(4, max(1, math.ceil(n_par * 2 / 2 ** 20 / 48.0)))      # the engine's bucket policy with overlap
            bounds = []

            def set_buckets(nb):
                bucket = (n_par // nb + 63) // 64 * 64
                bounds[:] = [(a, min(n_par, a + bucket)) for a in range(0, n_par, bucket)]

            def fused_step():
                for k, (lo, hi) in enumerate(reversed(bounds)):
                    comm.allreduce_adamw(lo, hi, fl, h, final=(k == len(bounds) - 1))

            def nccl_step():
                dist.all_reduce(g2)
                from rt1b200.ops import kernels
                kernels.fused_adamw(p2, fl.master, g2, fl.exp_avg, fl.exp_avg_sq, h, grad_scale=1.0 / world, zero_grad=True)
            for _ in range(3):
                nccl_step()
            t_n = timed(nccl_step, 20, world, dev)
            for nb in sorted({1, nb_overlap}):
                set_buckets(nb)
                for algo in (["p2p", "nvls"] if comm.has_multicast else ["p2p"]):
                    for ctas in (96, 148):
                        comm.algo, comm.ctas_per_rank = algo, ctas
                        for _ in range(3):
                            fused_step()
                        t_f = timed(fused_step, 20, world, dev)
                        # bytes per direction per GPU: P2P pulls (W-1)/W of the grads and pushes (W-1)/W of the params;
                        # NVLS sends N up for the in-switch reduce + N/W of parameters, receives N/W + N
                        nbytes = n_par * 2
                        link_bytes = 2 * (world - 1) / world * nbytes if algo == "p2p" else nbytes * (1 + 1 / world)
                        t_link900, t_link770 = link_bytes / 900e9 * 1e3, link_bytes / 770e9 * 1e3
                        hbm_bytes = n_par * (2 + 2) + n_par / world * (2 * world + 24 + 2)   # zero+incoming params, shard state
                        t_hbm = hbm_bytes / 6551e9 * 1e3
                        fused.append({"params": n_par, "world": world, "algo": algo, "ctas": ctas, "buckets": len(bounds),
                                      "fused_ms": t_f, "nccl_plus_fused_adam_ms": t_n, "roof_link_ms_at_900GBs": t_link900,
                                      "roof_link_ms_at_770GBs_measured_peer_copy": t_link770, "roof_hbm_ms": t_hbm,
                                      "roofline_fraction_nominal_link_900": max(t_link900, t_hbm) / t_f,
                                      "roofline_fraction_measured_link_770": max(t_link770, t_hbm) / t_f})
            comm.algo, comm.ctas_per_rank = "auto", 128
        if rank == 0:
            clocks = sampler.stop()
            os.makedirs("gpurun_out", exist_ok=True)
            with open(f"gpurun_out/p2p_sweep_w{world}.json", "w") as f:
                json.dump({"allreduce": rows, "fused_allreduce_adamw": fused, "clocks": clocks,
                           "timing": "CUDA events around 10-50 back-to-back calls, max over ranks"}, f, indent=1)
            for f_ in fused:
                print(json.dumps(f_))
    dist.barrier()
    if rank == 0:
        print("P2P_WORKER_OK", flush=True)
    tr.engine.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
