#!/usr/bin/env python
"""Training input pipeline on disk-backed episodes (reference-format .npy files): loader ceilings and end-to-end training rate.

    python tools/bench_input_pipeline.py [--episodes 48] [--workers 15] [--steps 40] [--out gpurun_out/input_pipeline.json]

Measures, on the same synthetic Language-Table-shaped episode files (180 x 320 frames, batch 8, T = 6):
  loader_reference   samples/s of the UNMODIFIED reference ``EmbodiedIntelligenceDataset`` + PIL transform under a DataLoader
                     (its own input-pipeline ceiling: SURVEY §6.3 step 4);
  loader_pil         our dataset with the same PIL transform (decoded-episode cache instead of un-pickling per sample);
  loader_raw         our dataset shipping raw uint8 frames + crop boxes (what ``distribute_train.py`` does on a GPU);
  train_e2e          ``Trainer.fit``-style loop on loader_raw: H2D + GPU crop/resize + the training step, samples/s.
"""
import argparse
import json
import os
import sys
import time

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from bench import ClockSampler  # noqa: E402


def loader_rate(loader, n_batches, batch):
    it = iter(loader)
    next(it)                                     # worker start-up
    t0 = time.time()
    n = 0
    for _ in range(n_batches):
        try:
            next(it)
        except StopIteration:
            break
        n += 1
    return n * batch / (time.time() - t0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--episodes", type=int, default=48)
    ap.add_argument("--workers", type=int, default=15)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--out", default="gpurun_out/input_pipeline.json")
    ap.add_argument("--train-only", action="store_true", help="skip the loader-only measurements (implied under torchrun)")
    a = ap.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:                                        # one process per GPU: every rank reads its own shuffled stream
        torch.cuda.set_device(local)
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
        a.train_only = True
    from argparse import Namespace
    from torch.utils.data import DataLoader
    from rt1b200.data.dataset import DecodeAndRandomResizedCrop, EmbodiedIntelligenceDataset, collate_fn
    from rt1b200.data.episodes import write_synthetic_dataset
    from rt1b200.data.gpu_transforms import GpuCropResize, RawFramesWithBoxes
    root = "/tmp/rt1_input_bench"
    if rank == 0 and not os.path.isdir(os.path.join(root, "train")):
        write_synthetic_dataset(root, a.episodes, 0, 0, min_len=30, max_len=40, height=180, width=320)
    if world > 1:
        torch.distributed.barrier()
    ids = list(range(a.episodes))
    res = {"batch": a.batch, "seq_len": 6, "workers": a.workers, "episodes": a.episodes, "frame": "180x320 uint8 -> 256x456",
           "n_gpus": world}

    def mk(ds):
        return DataLoader(ds, batch_size=a.batch, shuffle=True, collate_fn=collate_fn, num_workers=a.workers, pin_memory=True,
                          drop_last=True, persistent_workers=a.workers > 0)

    n_b = 60
    raw_ds = EmbodiedIntelligenceDataset(os.path.join(root, "train"), ids, 6, RawFramesWithBoxes(0.95))
    if not a.train_only:
        res["loader_pil_samples_per_s"] = loader_rate(mk(EmbodiedIntelligenceDataset(
            os.path.join(root, "train"), ids, 6, DecodeAndRandomResizedCrop(0.95, (456, 256)))), n_b, a.batch)
        res["loader_raw_samples_per_s"] = loader_rate(mk(raw_ds), n_b, a.batch)
    # the unmodified reference dataset + transform (needs the reference installation and its import shims)
    try:
        if a.train_only:
            raise RuntimeError("skipped (--train-only)")
        sys.path.insert(0, os.path.join(REPO, "baseline"))
        import refenv
        if refenv.available() or refenv.try_install():
            import subprocess
            code = ("import sys,time,os;sys.path.insert(0,%r);import refenv;refenv.activate();import load_np_dataset as L;"
                    "from torch.utils.data import DataLoader;"
                    "ds=L.EmbodiedIntelligenceDataset(%r,list(range(%d)),6,L.DecodeAndRandomResizedCrop(0.95,(456,256)));"
                    "dl=DataLoader(ds,batch_size=%d,shuffle=True,collate_fn=L.collate_fn,num_workers=%d,drop_last=True);"
                    "it=iter(dl);next(it);t0=time.time();n=0\n"
                    "for _ in range(%d):\n  next(it);n+=1\n"
                    "print('RATE',n*%d/(time.time()-t0))") % (os.path.join(REPO, "baseline"), os.path.join(root, "train"), a.episodes,
                                                              a.batch, a.workers, n_b, a.batch)
            r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
            for line in r.stdout.splitlines():
                if line.startswith("RATE"):
                    res["loader_reference_samples_per_s"] = float(line.split()[1])
            if "loader_reference_samples_per_s" not in res:
                res["loader_reference_error"] = (r.stderr or r.stdout)[-300:]
    except Exception as e:                                   # noqa: BLE001
        res["loader_reference_error"] = f"{type(e).__name__}: {e}"

    if torch.cuda.is_available():
        from rt1b200.train.trainer import RT1Module, Trainer
        dev = torch.device("cuda", local)
        torch.cuda.set_device(dev)
        args = Namespace(height=256, width=456, seq_len=6, lr=5e-4, milestones=[], exp_name="inp", log_dir="/tmp/rt1_inp/logs",
                         ckpt_dir="/tmp/rt1_inp/ckpt", log_every_n_steps=10 ** 9, ckpt_every_n_epochs=1, max_epochs=1)
        tr = Trainer(RT1Module(args), args, device=dev, param_dtype=torch.bfloat16, loggers=[], verbose=False)
        tr.batch_transform = GpuCropResize(256, 456)
        loader = mk(raw_ds)
        it = iter(loader)
        nxt = next(it)
        for _ in range(5):                                   # warm-up incl. graph capture
            batch, nxt = nxt, next(it)
            tr.train_step(batch)
            tr.prefetch(nxt)
        torch.cuda.synchronize()
        sampler = ClockSampler(local)
        sampler.start()
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        losses = []
        for _ in range(a.steps):
            batch = nxt
            tr.train_step(batch)
            nxt = next(it, None)
            if nxt is None:
                it = iter(loader)
                nxt = next(it)
            tr.prefetch(nxt)
            losses.append(tr.read_loss())
        e1.record()
        torch.cuda.synchronize()
        wall = time.time() - t0
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:                                    # the job's rate is set by its slowest rank
            torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        ms = float(ms.item())
        res["train_e2e_samples_per_s"] = world * a.steps * a.batch / (ms / 1e3)
        res["train_e2e_wall_samples_per_s"] = world * a.steps * a.batch / wall
        res["train_e2e_ms_per_step"] = ms / a.steps
        res["h2d_bytes_per_step"] = sum(t.numel() * t.element_size() for v in batch.values() for t in v.values())
        res["last_loss"] = losses[-1]
        res["clocks"] = sampler.stop()
    if rank == 0:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        json.dump(res, open(a.out, "w"), indent=1)
        print(json.dumps(res))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
