"""Trainer on the GPU: staged input copies, early loss read-back, the weight-gradient lane against the in-line schedule."""
import os
from argparse import Namespace

import pytest
import torch

pytestmark = pytest.mark.gpu


def _trainer(seed=0, hw=(64, 96)):
    from rt1b200.models.factory import RT1Config
    from rt1b200.train.trainer import RT1Module, Trainer
    torch.manual_seed(seed)
    (H, W), T = hw, 3
    args = Namespace(height=H, width=W, seq_len=T, lr=1e-3, milestones=[], exp_name="t", log_dir="/tmp/rt1_t/logs",
                     ckpt_dir="/tmp/rt1_t/ckpt", log_every_n_steps=10 ** 9, ckpt_every_n_epochs=1, max_epochs=1)
    cfg = RT1Config.language_table(height=H, width=W, seq_len=T, num_layers=2, dropout_rate=0.0)
    return Trainer(RT1Module(args, cfg), args, device=torch.device("cuda", 0), param_dtype=torch.bfloat16, loggers=[], verbose=False)


def test_prefetch_and_read_loss_match_inline_path():
    from rt1b200.data.dataset import SyntheticBatchSource
    src = SyntheticBatchSource(2, 3, 64, 96, pool=3, seed=1, image_dtype=torch.uint8, pin=True)
    a, b = _trainer(), _trainer()
    la, lb = [], []
    for i in range(3):
        la.append(float(a.train_step(src.pool[i])))                   # copy in line, loss read on the main stream
        b.train_step(src.pool[i])
        if i + 1 < 3:
            b.prefetch(src.pool[i + 1])                               # staged on the copy stream
        lb.append(b.read_loss())                                      # pinned read-back after the forward
    torch.cuda.synchronize()
    assert b._prefetched is None
    torch.testing.assert_close(torch.tensor(la), torch.tensor(lb), rtol=2e-2, atol=1e-4)


def test_wgrad_lane_matches_inline_gradients():
    """Same model, same batch: weight gradients computed on the side stream equal the in-line ones — up to the run-to-run
    noise of the in-line schedule itself (BatchNorm statistics are accumulated with fp32 atomics and, at this tiny test
    size, a 26-block BatchNorm backward amplifies the last-bit differences), which is measured alongside."""
    from rt1b200.data.dataset import SyntheticBatchSource, to_device
    from rt1b200.ops import kernels
    src = SyntheticBatchSource(4, 3, 96, 160, pool=1, seed=2, image_dtype=torch.uint8, pin=False)

    def grads_with(lane: str):
        os.environ["RT1B200_WGRAD_STREAM"] = lane
        try:
            kernels._LANES.clear()
            tr = _trainer(seed=3, hw=(96, 160))
            batch = to_device(src.pool[0], tr.device)
            tr.module.train()
            tr.engine.zero_grad()
            loss = tr.module.training_step(batch, 0)
            loss.backward()
            torch.cuda.synchronize()
            return tr.engine.flat.grad.float().clone()
        finally:
            os.environ.pop("RT1B200_WGRAD_STREAM", None)
            kernels._LANES.clear()

    inline_a, inline_b, lane = grads_with("0"), grads_with("0"), grads_with("1")
    assert torch.isfinite(lane).all()
    denom = inline_a.norm().clamp_min(1e-6)
    noise = float((inline_a - inline_b).norm() / denom)
    diff = float((lane - inline_a).norm() / denom)
    assert diff < max(3.0 * noise, 2e-2), (diff, noise)
