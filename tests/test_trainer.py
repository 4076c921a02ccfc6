"""Trainer / CLI / checkpoint / data tests on CPU, including the 2-process gloo data-parallel plumbing test."""
import os
import subprocess
import sys
from argparse import Namespace

import numpy as np
import pytest
import torch

from rt1b200.data.dataset import (DecodeAndRandomResizedCrop, DistributedWindowSampler, EmbodiedIntelligenceDataset,
                                  SyntheticBatchSource, collate_fn)
from rt1b200.data.episodes import load_episode, write_synthetic_dataset
from rt1b200.models.factory import RT1Config
from rt1b200.train.trainer import MultiStepLR, RT1Module, Trainer, load_policy_from_checkpoint

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _args(tmp, **kw):
    d = dict(height=64, width=96, seq_len=3, lr=1e-3, milestones=[1, 2], exp_name="t", log_dir=str(tmp / "logs"),
             ckpt_dir=str(tmp / "ckpt"), log_every_n_steps=1, ckpt_every_n_epochs=1, max_epochs=1, batch_size=2)
    d.update(kw)
    return Namespace(**d)


def _tiny_module(args):
    cfg = RT1Config(height=args.height, width=args.width, seq_len=args.seq_len, num_layers=1, layer_size=16,
                    num_heads=2, dropout_rate=0.0)
    return RT1Module(args, cfg)


def test_dataset_windows_and_collate(tmp_path):
    root = write_synthetic_dataset(str(tmp_path / "ds"), 2, 1, 1, min_len=5, max_len=7, height=60, width=80)
    ep = load_episode(os.path.join(root, "train", "episode_0.npy"))
    assert {"rgb", "instruction", "action", "is_terminal", "is_first"} <= set(ep[0])
    T = 3
    ds = EmbodiedIntelligenceDataset(os.path.join(root, "train"), [0, 1], T,
                                     transform=DecodeAndRandomResizedCrop(0.95, (96, 64)))
    n0 = len(load_episode(os.path.join(root, "train", "episode_0.npy")))
    n1 = len(load_episode(os.path.join(root, "train", "episode_1.npy")))
    assert len(ds) == n0 + n1          # (n + T-1) - T + 1 windows per episode
    first = ds[0]
    # front padding: the first window is T copies of step 0
    assert torch.equal(first["action_label"]["action"][0], first["action_label"]["action"][1])
    assert first["train_observation"]["image"].shape == (T, 3, 64, 96)
    assert first["train_observation"]["image"].dtype == torch.float32
    assert float(first["train_observation"]["image"].max()) <= 1.0
    last = ds[n0 - 1]
    assert int(last["action_label"]["terminate_episode"][-1]) == 1
    batch = collate_fn([ds[0], ds[1]])
    assert batch["train_observation"]["image"].shape == (2, T, 3, 64, 96)
    assert batch["action_label"]["terminate_episode"].dtype == torch.int64
    assert batch["train_observation"]["natural_language_embedding"].shape == (2, T, 512)
    # the reference's "is my dataset ready" check: python load_np_dataset.py (README, "Convert Dataset")
    import load_np_dataset
    out = load_np_dataset.main(["--data_dir", os.path.join(root, "train"), "--window_length", str(T), "--batch_size", "4",
                                "--width", "96", "--height", "64", "--save_image", str(tmp_path / "frame.png")])
    assert out["train_observation"]["image"].shape == (4, T, 3, 64, 96) and out["action_label"]["action"].shape == (4, T, 2)
    assert open(tmp_path / "frame.png", "rb").read(4) == b"\x89PNG"


def test_dataset_matches_reference_loader(tmp_path, reference):
    import importlib
    sys.path.insert(0, os.path.join(REPO, "baseline", "_ref"))
    try:
        ref_ds_mod = importlib.import_module("load_np_dataset")
    finally:
        sys.path.pop(0)
    root = write_synthetic_dataset(str(tmp_path / "ds"), 2, 0, 0, min_len=4, max_len=6, height=60, width=80)
    ours = EmbodiedIntelligenceDataset(os.path.join(root, "train"), [0, 1], 3, transform=None)
    theirs = ref_ds_mod.EmbodiedIntelligenceDataset(os.path.join(root, "train"), [0, 1], 3, transform=None)
    assert len(ours) == len(theirs)
    for i in range(len(ours)):
        a, b = ours[i], theirs[i]
        for grp in a:
            for k in a[grp]:
                torch.testing.assert_close(a[grp][k], b[grp][k])


def test_sampler_partitions():
    seen = []
    for r in range(3):
        s = DistributedWindowSampler(10, r, 3, shuffle=True, seed=1)
        seen.append(list(s))
    assert all(len(x) == 4 for x in seen)
    assert set(sum(seen, [])) == set(range(10))


def test_multistep_lr():
    s = MultiStepLR(1.0, [2, 4], 0.1)
    lrs = []
    for _ in range(6):
        lrs.append(s.get_lr()); s.step()
    assert lrs == pytest.approx([1, 1, 0.1, 0.1, 0.01, 0.01])


def test_fit_checkpoint_resume_and_reference_load(tmp_path, reference):
    torch.manual_seed(0)
    args = _args(tmp_path)
    module = _tiny_module(args)
    tr = Trainer(module, args, device="cpu", loggers=[])
    src = SyntheticBatchSource(2, 3, 64, 96, pool=2, image_dtype=torch.float32, pin=False)
    losses = [float(tr.train_step(next(src))) for _ in range(4)]
    assert all(np.isfinite(losses))
    tr.epoch = 1
    path = tr.save_checkpoint({"eval_loss": 0.5, "train_loss_epoch": losses[-1]})
    assert os.path.basename(path).startswith("epoch=0-eval_loss=0.500000-train_loss_epoch=")
    assert os.path.exists(os.path.join(args.ckpt_dir, "t", "last.ckpt"))
    ck = torch.load(path, weights_only=False)
    assert all(k.startswith("model.") for k in ck["state_dict"])
    assert {"optimizer_states", "lr_schedulers", "epoch", "global_step"} <= set(ck)
    # a fresh trainer resumes bit-identically
    module2 = _tiny_module(args)
    tr2 = Trainer(module2, args, device="cpu", loggers=[])
    tr2.load_checkpoint(path)
    assert tr2.global_step == 4 and tr2.engine.step_count == 4
    b = next(src)
    torch.manual_seed(3); l1 = float(tr.train_step(b))
    torch.manual_seed(3); l2 = float(tr2.train_step(b))
    assert l1 == pytest.approx(l2, rel=1e-6)
    # Adam state is torch.optim.Adam compatible
    osd = ck["optimizer_states"][0]
    opt = torch.optim.Adam(module2.model.parameters(), lr=1e-3)
    groups = dict(osd["param_groups"][0]); groups.pop("param_names")
    opt.load_state_dict({"state": osd["state"], "param_groups": [groups]})
    # the checkpoint loads strictly into the *reference* model class
    import gym.spaces as gs  # shim
    from collections import OrderedDict
    ss = gs.Dict({"image": gs.Box(0.0, 1.0, (3, 64, 96), np.float32),
                  "natural_language_embedding": gs.Box(-np.inf, np.inf, [512], np.float32)})
    asp = gs.Dict(OrderedDict([("terminate_episode", gs.Discrete(2)), ("action", gs.Box(-0.1, 0.1, (2,), np.float32))]))
    ref = reference.transformer_network.TransformerNetwork(
        input_tensor_space=ss, output_tensor_space=asp, vocab_size=256, token_embedding_size=512, num_layers=1,
        layer_size=16, num_heads=2, feed_forward_size=512, dropout_rate=0.0, time_sequence_length=3)
    ref.load_state_dict({k[len("model."):]: v for k, v in ck["state_dict"].items()}, strict=True)
    pol = load_policy_from_checkpoint(path)
    assert not pol.training and pol.cfg.num_layers == 1


def test_open_loop_action_accuracy_tool(tmp_path):
    """tools/eval_action_accuracy.py: a checkpoint of the full model against held-out episodes, with the mode / marginal baselines."""
    import importlib.util
    args = _args(tmp_path, height=64, width=96, seq_len=3)
    tr = Trainer(RT1Module(args), args, device="cpu", loggers=[])
    tr.epoch = 1
    path = tr.save_checkpoint({"eval_loss": 0.5, "train_loss_epoch": 0.5})
    root = write_synthetic_dataset(str(tmp_path / "ds"), 1, 1, 0, min_len=5, max_len=6, height=60, width=80)
    spec = importlib.util.spec_from_file_location("eval_action_accuracy", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "eval_action_accuracy.py"))
    tool = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(tool)
    out = tool.main(["--ckpt", path, "--data_dir", os.path.join(root, "val"), "--height", "64", "--width", "96", "--seq_len", "3",
                     "--batch_size", "2", "--max_batches", "2", "--device", "cpu"])
    assert out["steps"] == 2 * 2 * 3 and 0.0 <= out["model"]["token_accuracy"] <= 1.0
    # an untrained policy is near the uniform cross-entropy; the label marginal of a dozen steps is far below it
    assert 3.0 < out["model"]["ce_per_token"] < 7.0 and out["marginal"]["ce_per_token"] < out["model"]["ce_per_token"]
    assert out["mode"]["token_accuracy_within_2"] >= out["mode"]["token_accuracy"] > 0 and out["action_range_m"][0] < 0 < out["action_range_m"][1]
