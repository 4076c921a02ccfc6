#!/usr/bin/env python
"""RT-1 distributed training CLI — same flags as the reference's ``distribute_train.py`` (``:270-295``).

    python distribute_train.py --gpus 0,1,2,3 --lr 5e-4 --exp_name exp_rt1 \
        --dataset_dir /data/pytorch_b2b_sim_np --ckpt_dir ./exp/ckpt

Differences from the reference, all additive:
  * the default action is *training* (the reference's ``__main__`` calls ``debug(args)``; use ``--mode debug``
    for that smoke run);
  * ``--gpus a,b,c`` with more than one id re-launches this script once per GPU (one process per GPU,
    rendezvous on 127.0.0.1) — what Lightning's DDP launcher did; under ``torchrun`` it uses the given ranks;
  * ``--synthetic`` trains on generated Language-Table-shaped episodes (no dataset/USE offline),
    ``--resume PATH`` restores a checkpoint (model, Adam state, scheduler, counters), ``--max_steps`` bounds a
    run, ``--backend {auto,p2p,torchdist,local}`` picks the gradient-reduction path, ``--precision {bf16,fp32}``.
"""
from __future__ import annotations

import os
import subprocess
import sys
from argparse import ArgumentParser

import torch

from rt1b200.train.trainer import RT1Module, Trainer  # noqa: F401  (RT1Module re-exported for checkpoints)

RT1_Lightning = RT1Module   # name used by code written against the reference (language_table/eval/main_rt1.py:24)


def print_keys(obj, prefix: str = "") -> None:
    """Print the dotted paths of the leaves of a nested dict (the reference's checkpoint-inspection helper, ``:184-189``)."""
    if isinstance(obj, dict):
        for key, value in obj.items():
            print_keys(value, prefix + str(key) + ".")
    else:
        print(prefix[:-1])


def build_parser() -> ArgumentParser:
    p = ArgumentParser()
    p.add_argument("--device", type=str, default="gpu")
    p.add_argument("--gpus", type=str, default="0,1,2,3")
    p.add_argument("--max_epochs", type=int, default=100)
    p.add_argument("--batch_size", type=int, default=8)
    p.add_argument("--num_workers", type=int, default=15)
    p.add_argument("--milestones", type=int, nargs="+", default=[50, 75, 90])
    p.add_argument("--lr", type=float, default=5e-4)
    p.add_argument("--exp_name", type=str, default="exp_rt1")
    p.add_argument("--log_dir", type=str, default="./exp/logs")
    p.add_argument("--ckpt_dir", type=str, default="./exp/ckpt")
    p.add_argument("--log_every_n_steps", type=int, default=500)
    p.add_argument("--ckpt_every_n_epochs", type=int, default=1)
    p.add_argument("--dataset_dir", type=str, default="/gemini/data-2/data/pytorch_b2b_sim_np")
    p.add_argument("--train_episode", type=int, default=7800)
    p.add_argument("--test_episode", type=int, default=50)
    p.add_argument("--eval_episode", type=int, default=50)
    p.add_argument("--random_crop_factor", type=float, default=0.95)
    p.add_argument("--height", type=int, default=256)
    p.add_argument("--width", type=int, default=456)
    p.add_argument("--seq_len", type=int, default=6)
    # additions
    p.add_argument("--mode", choices=["train", "debug"], default="train")
    p.add_argument("--synthetic", action="store_true", help="generate a synthetic dataset under --dataset_dir if absent")
    p.add_argument("--resume", type=str, default=None)
    p.add_argument("--max_steps", type=int, default=None)
    p.add_argument("--backend", choices=["auto", "p2p", "torchdist", "local"], default="auto")
    p.add_argument("--precision", choices=["bf16", "fp32"], default=None)
    p.add_argument("--broadcast_buffers", action="store_true",
                   help="re-broadcast rank 0's BatchNorm statistics after every step, like torch DDP's broadcast_buffers=True; "
                        "without it they are broadcast before every evaluation pass, which gives the same evaluation and "
                        "checkpoints (no rank reads another rank's running statistics while training)")
    p.add_argument("--cpu_input", dest="gpu_input", action="store_false",
                   help="crop + resize frames in the loader workers with PIL like the reference (default on a GPU: ship raw uint8 "
                        "frames and crop/resize with a kernel)")
    return p


def _pick_device(args) -> torch.device:
    if args.device in ("gpu", "cuda") and torch.cuda.is_available():
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        return torch.device("cuda", local)
    return torch.device("cpu")


def _init_distributed(device: torch.device):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1 and not torch.distributed.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        # NCCL is used for bootstrap/broadcast plumbing only; gradient traffic goes through the P2P engine
        torch.distributed.init_process_group("nccl" if device.type == "cuda" else "gloo",
                                             device_id=device if device.type == "cuda" else None)
    return world


def _maybe_relaunch(args) -> bool:
    """``--gpus 0,1,2,3`` outside torchrun → spawn one process per listed GPU."""
    ids = [g for g in args.gpus.split(",") if g.strip() != ""]
    if "RANK" in os.environ or len(ids) <= 1 or args.device not in ("gpu", "cuda") or not torch.cuda.is_available():
        return False
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=",".join(ids))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={len(ids)}",
           "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29533"),
           os.path.abspath(__file__)] + sys.argv[1:]
    sys.exit(subprocess.call(cmd, env=env))


def make_loaders(args, rank: int, world: int):
    from torch.utils.data import DataLoader
    from rt1b200.data.dataset import (DecodeAndRandomResizedCrop, DistributedWindowSampler,
                                      EmbodiedIntelligenceDataset, collate_fn)
    from rt1b200.data.episodes import write_synthetic_dataset
    if args.synthetic and not os.path.isdir(os.path.join(args.dataset_dir, "train")):
        if rank == 0:
            write_synthetic_dataset(args.dataset_dir, args.train_episode, args.eval_episode, args.test_episode)
        if world > 1:
            torch.distributed.barrier()
    if getattr(args, "gpu_input", True) and args.device in ("gpu", "cuda") and torch.cuda.is_available():
        # workers ship raw uint8 frames + one crop box per frame; crop + bilinear resize run as a kernel on the GPU
        from rt1b200.data.gpu_transforms import RawFramesWithBoxes
        transform = RawFramesWithBoxes(args.random_crop_factor)
    else:
        transform = DecodeAndRandomResizedCrop(args.random_crop_factor, (args.width, args.height))
    out = []
    for split, n, shuffle in (("train", args.train_episode, True), ("test", args.test_episode, False),
                              ("val", args.eval_episode, False)):
        ds = EmbodiedIntelligenceDataset(os.path.join(args.dataset_dir, split), list(range(n)), args.seq_len,
                                         transform=transform)
        sampler = DistributedWindowSampler(len(ds), rank, world, shuffle=shuffle, drop_last=shuffle)
        out.append(DataLoader(ds, batch_size=args.batch_size, sampler=sampler, collate_fn=collate_fn,
                              num_workers=args.num_workers, pin_memory=torch.cuda.is_available(),
                              drop_last=shuffle, persistent_workers=args.num_workers > 0))
    return out


def train(args):
    _maybe_relaunch(args)
    if "RANK" not in os.environ and args.device in ("gpu", "cuda"):
        os.environ["CUDA_VISIBLE_DEVICES"] = args.gpus
    device = _pick_device(args)
    world = _init_distributed(device)
    rank = int(os.environ.get("RANK", "0"))
    if rank == 0:
        print("Loading datasets ...", flush=True)
    train_loader, test_loader, eval_loader = make_loaders(args, rank, world)
    if rank == 0:
        print("Building RT1 model ...", flush=True)
    module = RT1Module(args)
    dtype = {None: None, "bf16": torch.bfloat16, "fp32": torch.float32}[args.precision]
    trainer = Trainer(module, args, device=device, backend=args.backend, param_dtype=dtype)
    if device.type == "cuda" and getattr(args, "gpu_input", True):
        from rt1b200.data.gpu_transforms import GpuCropResize
        trainer.batch_transform = GpuCropResize(args.height, args.width)
    if rank == 0:
        print("Start training ...", flush=True)
    trainer.fit(train_loader, eval_loader, resume=args.resume, max_steps=args.max_steps)
    if rank == 0:
        print("Start final testing ...", flush=True)
    trainer.test(test_loader)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    return trainer


def debug(args):
    device = _pick_device(args)
    agent = RT1Module(args).to(device)
    print("=========Training Debug=========")
    agent.train()
    pred_action, action_loss, mean_loss = agent.debug(training=True, device=device)
    print(f"pred_action:{pred_action['action'].shape}")
    print(f"is_terminate:{pred_action['terminate_episode'].shape}")
    print(f"action_loss:{action_loss.shape}, mean_loss:{mean_loss.shape}")
    print("=========Inference Debug=========")
    agent.eval()
    pred_action = agent.debug(training=False, device=device)
    print(f"pred_action:{pred_action['action'].shape}")
    print(f"is_terminate:{pred_action['terminate_episode'].shape}")
    return agent


def main(argv=None):
    args = build_parser().parse_args(argv)
    if args.mode == "debug":
        return debug(args)
    return train(args)


if __name__ == "__main__":
    main()
