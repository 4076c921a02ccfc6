#!/usr/bin/env python
"""Headline benchmark: RT-1 Language-Table training throughput (samples/s), 1 process per GPU.

    python bench.py --gpus N --steps K --warmup W            # this framework (bf16, fused P2P reduce+AdamW)
    python bench.py --impl reference --gpus N ...            # the UNMODIFIED reference (baseline/_ref) under
                                                             # torch DDP(find_unused_parameters=True) + Adam, fp32

Config (BASELINE.json config 2): RT-1 = FiLM-EfficientNet-B3 + TokenLearner(8) + 8-layer decoder, 6-frame
history, 256x456 frames, 3 action tokens x 256 bins, per-GPU batch 8 (weak scaling: global batch = 8 N),
synthetic Language-Table-shaped episodes, random-init weights.  One "step" = forward + backward + gradient
all-reduce + optimizer update, nothing skipped.

Timing: W >= 3 warm-up steps; exactly K steps between CUDA events, barrier + synchronize on both sides, MAX
over ranks; `nvidia-smi` clocks sampled during the timed region.  Two numbers:
  value      — device-resident inputs (a rotating pool of distinct batches), no host traffic in the loop;
  e2e.value  — through the public API (`Trainer.train_step(host_batch)` + `Trainer.prefetch(next_host_batch)`, the
               loop `Trainer.fit` runs): every step copies that step's batch from pinned host memory (on a copy
               stream, overlapping the previous step's compute) and reads the step's loss back to the host.
Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

REPO = os.path.dirname(os.path.abspath(__file__))
PER_GPU_BATCH = 8
SEQ_LEN, HEIGHT, WIDTH = 6, 256, 456
METRIC = "rt1_train_samples_per_sec"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", choices=["ours", "reference"], default="ours")
    p.add_argument("--batch", type=int, default=PER_GPU_BATCH, help="per-GPU batch (default 8 = reference recipe)")
    p.add_argument("--model", choices=["rt1", "scaled"], default="rt1")
    p.add_argument("--backend", default="auto")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--reps", type=int, default=3,
                   help="back-to-back repetitions of the --steps window; the MEDIAN window is reported (all are listed)")
    p.add_argument("--ref-precision", choices=["fp32", "bf16"], default="fp32",
                   help="reference arm only: fp32 = the reference's stock precision (what the driver runs); bf16 = the same "
                        "unmodified training_step under torch.autocast(bfloat16) with channels_last convolutions")
    p.add_argument("--cnn", choices=["kernels", "cudnn"], default="kernels",
                   help="our arm only: 'cudnn' runs OUR model with the image tokenizer on stock PyTorch bf16 channels-last ops "
                        "(cuDNN/cuBLAS) — the library number every encoder kernel family has to beat")
    return p.parse_args()


# ------------------------------------------------------------------------------------------------ utilities
class ClockSampler:
    """`nvidia-smi -lms 200` for the duration of the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.path = gpu_index, None, f"/tmp/rt1_clocks_{os.getpid()}.csv"

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        try:
            os.remove(self.path)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", device_id=dev)
    if world != args.gpus and rank == 0:
        print(f"[bench] note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)
    return rank, world, local, dev


def barrier_sync(world):
    import torch
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()


def max_over_ranks(ms: float, world: int, dev) -> float:
    import torch
    if world == 1:
        return ms
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def timed_loop(step_fn, steps: int, world: int, dev):
    """Exactly `steps` calls between two CUDA events; barrier + synchronize on both sides; max over ranks."""
    import torch
    barrier_sync(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step_fn(i)
    e1.record()
    barrier_sync(world)
    return max_over_ranks(e0.elapsed_time(e1), world, dev)


def timed_reps(step_fn, steps: int, reps: int, world: int, dev):
    """`reps` windows of exactly `steps` steps each; returns (median window ms, all windows)."""
    wins = [timed_loop(step_fn, steps, world, dev) for _ in range(max(1, reps))]
    return statistics.median(wins), wins


def params_equal_across_ranks(flat_param, world: int, dev) -> bool:
    """After training: every rank must hold bit-identical bf16 parameters (the fused reduce+update writes them to all peers)."""
    import torch
    if world == 1:
        return True
    v = flat_param.view(torch.int16).to(torch.int64)
    idx = torch.arange(1, v.numel() + 1, device=dev, dtype=torch.int64)
    digest = torch.stack([v.sum(), (v * (idx % 65521)).sum(), (v ^ (idx & 0x7FFF)).sum()])
    lo, hi = digest.clone(), digest.clone()
    torch.distributed.all_reduce(lo, op=torch.distributed.ReduceOp.MIN)
    torch.distributed.all_reduce(hi, op=torch.distributed.ReduceOp.MAX)
    return bool((lo == hi).all().item())


def make_batches(batch, pool, seed, uint8: bool, pin: bool):
    from rt1b200.data.dataset import SyntheticBatchSource
    import torch
    return SyntheticBatchSource(batch, SEQ_LEN, HEIGHT, WIDTH, pool=pool, seed=seed,
                                image_dtype=torch.uint8 if uint8 else torch.float32, pin=pin)


# --------------------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    from argparse import Namespace
    from rt1b200.data.dataset import to_device, tree_leaves
    from rt1b200.models.factory import RT1Config
    from rt1b200.ops import _lib
    from rt1b200.train.trainer import RT1Module, Trainer

    if args.cnn == "cudnn":
        os.environ["RT1B200_CNN_BENCH_CUDNN"] = "1"      # read by rt1b200.ops.cnn (bench/tests only)
    rank, world, local, dev = dist_setup(args)
    torch.manual_seed(0)
    targs = Namespace(height=HEIGHT, width=WIDTH, seq_len=SEQ_LEN, lr=5e-4, milestones=[50, 75, 90], exp_name="bench",
                      log_dir="/tmp/rt1_bench/logs", ckpt_dir="/tmp/rt1_bench/ckpt", log_every_n_steps=10 ** 9,
                      ckpt_every_n_epochs=1, max_epochs=1)
    cfg = RT1Config.language_table() if args.model == "rt1" else RT1Config.scaled()
    module = RT1Module(targs, cfg)
    trainer = Trainer(module, targs, device=dev, backend=args.backend, param_dtype=torch.bfloat16, loggers=[],
                      verbose=False)
    src = make_batches(args.batch, 4, 100 + rank, uint8=True, pin=True)
    dev_pool = [to_device(b, dev, non_blocking=False) for b in src.pool]
    h2d = sum(t.numel() * t.element_size() for t in tree_leaves(src.pool[0]))

    def dev_step(i):
        trainer.train_step(dev_pool[i % len(dev_pool)])

    for i in range(max(args.warmup, 3)):
        dev_step(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = _lib.launch_count()
    ms, windows = timed_reps(dev_step, args.steps, args.reps, world, dev)
    launches = (_lib.launch_count() - launches0) // max(1, args.reps)
    clocks = sampler.stop() if rank == 0 else {}

    e2e = None
    if not args.no_e2e:
        host_loss = []

        def e2e_step(i):
            # every step: this step's batch goes pinned host → device inside the API (staged on the trainer's copy
            # stream by the previous iteration's prefetch, like a data loader with a one-batch look-ahead — the copy
            # itself still happens inside the timed region, once per step), and the step's loss is read back to the host
            trainer.train_step(src.pool[i % len(src.pool)])
            trainer.prefetch(src.pool[(i + 1) % len(src.pool)])
            host_loss.append(trainer.read_loss())       # D2H of THIS step's loss (pinned, side stream; waits for its forward)

        for i in range(2):
            e2e_step(i)
        sampler2 = ClockSampler(local)
        if rank == 0:
            sampler2.start()
        ms_e2e, windows_e2e = timed_reps(e2e_step, args.steps, args.reps, world, dev)
        clocks_e2e = sampler2.stop() if rank == 0 else {}
        e2e = {"value": args.batch * world * args.steps / (ms_e2e / 1e3), "unit": "samples/s",
               "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
               "last_loss": host_loss[-1], "windows_ms": [round(w, 3) for w in windows_e2e], "clocks": clocks_e2e}
    equal = params_equal_across_ranks(trainer.engine.flat.param, world, dev)
    if world > 1:
        torch.distributed.barrier()
    result = None
    if rank == 0:
        result = {
            "metric": METRIC, "value": args.batch * world * args.steps / (ms / 1e3), "unit": "samples/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "impl": "ours",
            "config": {"model": "RT-1 FiLM-EfficientNet-B3 + TokenLearner(8) + 8L decoder (35.3M params)"
                       if args.model == "rt1" else "RT-1 scaled (24L decoder, d_model 1024)",
                       "global_batch": args.batch * world, "per_gpu_batch": args.batch, "seq_len": SEQ_LEN,
                       "frames": f"{HEIGHT}x{WIDTH}", "tokens_per_sample": 66,
                       "parallelism": f"dp{world}", "backend": trainer.engine.backend,
                       "l2": "working set >> L2: rotating pool of 4 distinct input batches + GBs of activations per step",
                       "optimizer": "Adam(lr 5e-4) as fused reduce+AdamW, fp32 master weights"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "windows_ms": [round(w, 3) for w in windows], "reps": len(windows), "timing": "median of `reps` windows of `steps` steps",
            "params_equal_across_ranks": equal, "cnn": args.cnn,
            "kernels_loaded": _lib.available(), "native_lib": os.path.relpath(_lib.LIB_PATH, REPO),
        }
    if world > 1:
        torch.distributed.destroy_process_group()
    return result


# --------------------------------------------------------------------------------------------- reference arm
def run_reference(args):
    sys.path.insert(0, os.path.join(REPO, "baseline"))
    import refenv
    if not refenv.available() and not refenv.try_install():
        return {"impl": "reference", "unavailable": refenv.why_unavailable()}
    refenv.activate()          # reference + shims first on sys.path, this repository's packages removed
    import torch
    from argparse import Namespace
    rank, world, local, dev = dist_setup(args)
    torch.manual_seed(0)
    import distribute_train as ref_train               # baseline/_ref/distribute_train.py (unmodified)
    targs = Namespace(height=HEIGHT, width=WIDTH, seq_len=SEQ_LEN, lr=5e-4, milestones=[50, 75, 90])
    module = ref_train.RT1_Lightning(targs)
    if args.model == "scaled":
        # BASELINE config 3 (24 layers, d_model 1024): the reference hard-codes its hyper-parameters inside RT1_Lightning
        # (distribute_train.py:42-55), so the scaled variant is built through the reference's own TransformerNetwork
        # constructor with the two changed arguments and swapped into the unmodified Lightning module
        from pytorch_robotics_transformer.transformer_network import TransformerNetwork as RefNet
        module.model = RefNet(input_tensor_space=module.state_space, output_tensor_space=module.action_space, vocab_size=256,
                              token_embedding_size=512, num_layers=24, layer_size=128, num_heads=8, feed_forward_size=1024,
                              dropout_rate=0.1, time_sequence_length=SEQ_LEN, crop_size=236, use_token_learner=True)
    module = module.to(dev)
    module.train()
    # what Lightning's DDPStrategy(find_unused_parameters=True) + configure_optimizers() set up
    opt = torch.optim.Adam(module.parameters(), lr=targs.lr)

    g = torch.Generator().manual_seed(100 + rank)
    B = args.batch

    def make_batch():
        img = torch.rand(B, SEQ_LEN, 3, HEIGHT, WIDTH, generator=g)
        emb = torch.randn(B, 1, 512, generator=g).expand(-1, SEQ_LEN, -1).contiguous()
        term = torch.zeros(B, SEQ_LEN, dtype=torch.int64)
        act = (torch.rand(B, SEQ_LEN, 2, generator=g) - 0.5) * 0.2
        return {"action_label": {"terminate_episode": term, "action": act},
                "train_observation": {"image": img, "natural_language_embedding": emb}}

    def pin(b):
        return {k: {kk: vv.pin_memory() for kk, vv in v.items()} for k, v in b.items()}

    def to_dev(b):
        return {k: {kk: vv.to(dev, non_blocking=True) for kk, vv in v.items()} for k, v in b.items()}

    host_pool = [pin(make_batch()) for _ in range(4)]
    dev_pool = [to_dev(b) for b in host_pool]
    h2d = sum(vv.numel() * vv.element_size() for v in host_pool[0].values() for vv in v.values())

    class _StepWrapper(torch.nn.Module):
        """DDP hooks fire on ``forward``; route it to the reference's own ``training_step``."""

        def __init__(self, m):
            super().__init__()
            self.m = m

        def forward(self, batch):
            return self.m.training_step(batch, 0)

    if world > 1:
        ddp = torch.nn.parallel.DistributedDataParallel(_StepWrapper(module), device_ids=[local],
                                                        find_unused_parameters=True)
        run = lambda b: ddp(b)            # noqa: E731
    else:
        run = lambda b: module.training_step(b, 0)   # noqa: E731

    bf16 = args.ref_precision == "bf16"
    if bf16:
        # same unmodified training_step; what `Trainer(precision="bf16-mixed")` + channels_last would do in Lightning:
        # fp32 parameters and Adam state, autocast to bf16 around forward/loss, NHWC convolutions
        module.to(memory_format=torch.channels_last)
        run_plain = run

        def run(b):                                  # noqa: F811
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return run_plain(b)

    def train_step(batch):
        loss = run(batch)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss.detach()

    def dev_step(i):
        train_step(dev_pool[i % 4])

    for i in range(max(args.warmup, 3)):
        dev_step(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, windows = timed_reps(dev_step, args.steps, args.reps, world, dev)
    clocks = sampler.stop() if rank == 0 else {}
    e2e = None
    if not args.no_e2e:
        losses = []

        def e2e_step(i):
            losses.append(float(train_step(to_dev(host_pool[i % 4]))))

        for i in range(2):
            e2e_step(i)
        ms_e2e, windows_e2e = timed_reps(e2e_step, args.steps, args.reps, world, dev)
        e2e = {"value": B * world * args.steps / (ms_e2e / 1e3), "unit": "samples/s", "ms_per_step": ms_e2e / args.steps,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "last_loss": losses[-1],
               "windows_ms": [round(w, 3) for w in windows_e2e]}
    result = None
    if rank == 0:
        result = {
            "metric": METRIC, "value": B * world * args.steps / (ms / 1e3), "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16-autocast" if bf16 else "fp32",
            "data": "synthetic", "impl": "reference",
            "config": {"model": "reference RT1_Lightning (unmodified, baseline/_ref)"
                       + (" with TransformerNetwork(num_layers=24, feed_forward_size=1024)" if args.model == "scaled" else ""),
                       "global_batch": B * world,
                       "per_gpu_batch": B, "seq_len": SEQ_LEN, "frames": f"{HEIGHT}x{WIDTH}",
                       "parallelism": f"dp{world}" + (" torch DDP(find_unused_parameters=True) over NCCL" if world > 1 else ""),
                       "optimizer": "torch.optim.Adam(lr 5e-4)",
                       "precision": "bf16 autocast + channels_last (NOT the reference's stock path; like-for-like precision arm)"
                       if bf16 else "fp32 (Lightning default 32-true)",
                       "l2": "working set >> L2"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": 0,
            "windows_ms": [round(w, 3) for w in windows], "reps": len(windows), "timing": "median of `reps` windows of `steps` steps",
        }
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    return result


def main():
    args = parse()
    if args.impl == "reference":
        try:
            res = run_reference(args)
        except Exception as e:  # the reference arm must never take the driver down
            if int(os.environ.get("RANK", "0")) == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {str(e)[:300]}"}))
            return 0
    else:
        res = run_ours(args)
    if res is not None:
        print(json.dumps(res), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
