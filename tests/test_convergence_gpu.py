"""Training-trajectory parity: the bf16 sm_100a kernel path against the UNMODIFIED fp32 reference model (baseline/_ref) over
60 optimisation steps — same initial weights, same batches, same (centred) crop, dropout and stochastic depth off.

No single-step test pins the bf16 path to the fp32 model over TRAINING: a 26-block BatchNorm network in bf16 differs from
fp32 in every step by rounding, and what matters is that the differences do not accumulate into a different optimisation
trajectory."""
from argparse import Namespace
from collections import OrderedDict

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_bf16_kernel_training_tracks_fp32_reference(reference, monkeypatch):
    from rt1b200.data.dataset import SyntheticBatchSource, to_device
    from rt1b200.models import RT1Config, default_spaces
    from rt1b200.train.trainer import RT1Module, Trainer
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    H, W, T, B, STEPS, LR = 64, 96, 3, 4, 60, 5e-4
    cfg = RT1Config.language_table(height=H, width=W, seq_len=T, num_layers=2, dropout_rate=0.0)
    ss, _ = default_spaces(cfg)
    gs = reference.spaces
    rss = gs.Dict({k: gs.Box(low=v.low, high=v.high, shape=v.shape, dtype=np.float32) for k, v in ss.items()})
    rasp = gs.Dict(OrderedDict([("terminate_episode", gs.Discrete(2)),
                                ("action", gs.Box(low=cfg.action_low, high=cfg.action_high, shape=(2,), dtype=np.float32))]))
    ref = reference.transformer_network.TransformerNetwork(
        input_tensor_space=rss, output_tensor_space=rasp, vocab_size=cfg.vocab_size, token_embedding_size=512,
        num_layers=cfg.num_layers, layer_size=cfg.layer_size, num_heads=cfg.num_heads, feed_forward_size=cfg.feed_forward_size,
        dropout_rate=0.0, time_sequence_length=T, crop_size=236, use_token_learner=True)
    args = Namespace(height=H, width=W, seq_len=T, lr=LR, milestones=[], exp_name="conv", log_dir="/tmp/rt1_conv/logs",
                     ckpt_dir="/tmp/rt1_conv/ckpt", log_every_n_steps=10 ** 9, ckpt_every_n_epochs=1, max_epochs=1)
    module = RT1Module(args, cfg)
    module.model.load_state_dict(ref.state_dict(), strict=True)                  # identical initial weights
    for m in list(module.modules()) + list(ref.modules()):                       # stochastic depth off on both sides
        if hasattr(m, "drop_rate"):
            m.drop_rate = 0.0
        if type(m).__name__ == "StochasticDepth":
            m.p = 0.0
    module.model._crop_shift_override = (int(H * 0.07), int(W * 0.07))           # centred window on our side ...
    real_randint = torch.randint

    def centred_randint(low, high=None, size=None, **kw):                        # ... and in the reference's pad-shift crop
        if size == [] and not kw:
            return torch.tensor((high if high is not None else low) // 2)
        return real_randint(low, high, size, **kw) if high is not None else real_randint(low, size, **kw)

    tr = Trainer(module, args, device=dev, param_dtype=torch.bfloat16, loggers=[], verbose=False, graph=True)
    # A second copy of the fp32 reference whose INITIAL weights are rounded to bf16: how far a rounding-sized perturbation
    # moves the fp32 model's own trajectory on these batches is the yardstick for "the bf16 path tracks it".
    import copy
    ref_b = copy.deepcopy(ref)
    with torch.no_grad():
        for p_ in ref_b.parameters():
            p_.copy_(p_.to(torch.bfloat16).float())
    src = SyntheticBatchSource(B, T, H, W, pool=6, seed=11, image_dtype=torch.float32, pin=False)

    def run_reference(net):
        net = net.to(dev).train()
        opt = torch.optim.Adam(net.parameters(), lr=LR)
        state = {"context_image_tokens": torch.zeros(B, T, 8, 512, device=dev),
                 "action_tokens": torch.zeros(B, T, 3, dtype=torch.long, device=dev), "seq_idx": torch.zeros(B, dtype=torch.long, device=dev)}
        out = []
        for i in range(STEPS):
            batch = to_device(src.pool[i % len(src.pool)], dev)
            monkeypatch.setattr(torch, "randint", centred_randint)
            try:
                net.set_actions(batch["action_label"])
                net(batch["train_observation"], state)
                loss = net.get_actor_loss().mean()
            finally:
                monkeypatch.setattr(torch, "randint", real_randint)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
            out.append(float(loss))
        return np.array(out)

    ours = np.array([float(tr.train_step(to_device(src.pool[i % len(src.pool)], dev))) for i in range(STEPS)])
    theirs = run_reference(ref)
    theirs_b = run_reference(ref_b)
    assert np.isfinite(ours).all() and np.isfinite(theirs).all()
    assert theirs[-10:].mean() < 0.7 * theirs[:5].mean(), "the reference itself did not learn these batches"
    # identical start (same weights, same batch): bf16 rounding only
    assert abs(ours[0] - theirs[0]) <= 0.02 * theirs[0], (ours[0], theirs[0])
    # The curves stay together while the loss DESCENDS (10-step windowed means, from step 10, while the reference is above 15 % of
    # its starting loss).  Tolerance: 5 %, widened to twice the spread the fp32 reference shows against ITSELF when only its
    # initial weights are rounded to bf16.  Once these six batches are fitted (loss down 10x and more) the fp32 model's own two
    # runs differ by 25 % (measured) — a 26-block BatchNorm network at batch 4 x 3 frames under Adam at 5e-4 amplifies
    # rounding-sized perturbations — so there only a factor-of-two sanity bound applies.
    windows = [(a, ours[a:a + 10].mean(), theirs[a:a + 10].mean(), theirs_b[a:a + 10].mean()) for a in range(10, STEPS - 5, 5)]
    descent = [w for w in windows if w[2] >= 0.15 * theirs[0]]
    print(f"[convergence] ours/ref per window {[round(float(o / r), 3) for _, o, r, _ in windows]}; "
          f"fp32'/fp32 {[round(float(rb / r), 3) for _, _, r, rb in windows]}; descent windows {len(descent)}")
    assert len(descent) >= 2, "the comparison needs at least two windows in the descent phase"
    spread = max(abs(rb - r) / r for _, _, r, rb in descent)
    tol = max(0.05, 2.0 * spread)
    if tol > 0.35:          # measured spreads: 5-9 %; beyond this the fp32 model's own two runs disagree too much to judge anything
        pytest.skip(f"the fp32 reference is chaotic already in the descent phase on this box (spread {spread:.3f})")
    for a, o, r, _ in descent:
        assert abs(o - r) <= tol * r, (a, o, r, tol)
    for a, o, r, _ in windows:
        assert 0.5 * r <= o <= 2.0 * r, (a, o, r)
