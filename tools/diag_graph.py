"""Localise a failing CUDA-graph capture of the training step: capture (a) the forward alone, (b) forward + autograd.grad."""
import gc
import os
import sys
from argparse import Namespace

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from rt1b200.data.dataset import SyntheticBatchSource, to_device  # noqa: E402
from rt1b200.models.factory import RT1Config  # noqa: E402
from rt1b200.train.trainer import RT1Module, Trainer  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
H, W, T = 64, 96, 3
args = Namespace(height=H, width=W, seq_len=T, lr=1e-3, milestones=[], exp_name="d", log_dir="/tmp/d/logs", ckpt_dir="/tmp/d/ckpt",
                 log_every_n_steps=10 ** 9, ckpt_every_n_epochs=1, max_epochs=1)
cfg = RT1Config.language_table(height=H, width=W, seq_len=T, num_layers=2, dropout_rate=0.1)
tr = Trainer(RT1Module(args, cfg), args, device=dev, param_dtype=torch.bfloat16, loggers=[], verbose=False, graph=False)
src = SyntheticBatchSource(2, T, H, W, pool=1, image_dtype=torch.uint8, pin=False)
batch = to_device(src.pool[0], dev)
tr.module.train()
model = tr.module.model
model._device_rng = True
tr.engine.remove_ready_hooks()
stream = torch.cuda.Stream()


def body(mode):
    if mode == "fwd_nograd":
        with torch.no_grad():
            return tr.module.training_step(batch, 0)
    loss = tr.module.training_step(batch, 0)
    if mode == "fwd":
        return loss
    grads = torch.autograd.grad(loss, tr.engine.flat.params, allow_unused=True)
    return loss


for mode in ("fwd_nograd", "fwd", "fwd_bwd"):
    model._loss = None
    model._aux_info = {}
    gc.collect()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        body(mode)
    torch.cuda.current_stream().wait_stream(stream)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g, stream=stream):
            out = body(mode)
        g.replay()
        torch.cuda.synchronize()
        print(f"[diag_graph] {mode}: captured and replayed, loss {float(out):.6f}", flush=True)
    except Exception as e:
        print(f"[diag_graph] {mode}: FAILED {type(e).__name__}: {str(e)[:200]}", flush=True)
        break
