"""One RT-1 training step bracketed by cudaProfilerStart/Stop (for `ncu --profile-from-start off`).
Usage: python tools/profile_step.py [--impl ours|reference] [--batch 8]"""
import argparse
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--warmup", type=int, default=4)
a = ap.parse_args()
from argparse import Namespace  # noqa: E402
from rt1b200.data.dataset import SyntheticBatchSource, to_device  # noqa: E402
from rt1b200.models.factory import RT1Config  # noqa: E402
from rt1b200.train.trainer import RT1Module, Trainer  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
args = Namespace(height=256, width=456, seq_len=6, lr=5e-4, milestones=[], exp_name="prof", log_dir="/tmp/rt1_prof/logs",
                 ckpt_dir="/tmp/rt1_prof/ckpt", log_every_n_steps=10 ** 9, ckpt_every_n_epochs=1, max_epochs=1)
tr = Trainer(RT1Module(args, RT1Config.language_table()), args, device=dev, param_dtype=torch.bfloat16, loggers=[], verbose=False)
src = SyntheticBatchSource(a.batch, 6, 256, 456, pool=2, image_dtype=torch.uint8, pin=False)
pool = [to_device(b, dev) for b in src.pool]
for i in range(a.warmup):
    tr.train_step(pool[i % 2])
torch.cuda.synchronize()
torch.cuda.profiler.start()
tr.train_step(pool[0])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled one step")
