"""Parity of rt1b200.models with the unmodified reference (state-dict layout, forward, loss, inference)."""
from collections import OrderedDict

import numpy as np
import pytest
import torch

from rt1b200 import spaces
from rt1b200.models import RT1Config, TransformerNetwork, build_rt1, default_spaces


def _small_cfg():
    return RT1Config(height=64, width=96, seq_len=3, num_layers=2, layer_size=16, num_heads=4, dropout_rate=0.0)


def _build_pair(reference, cfg):
    ss, asp = default_spaces(cfg)
    kw = dict(vocab_size=cfg.vocab_size, token_embedding_size=512, num_layers=cfg.num_layers,
              layer_size=cfg.layer_size, num_heads=cfg.num_heads, feed_forward_size=cfg.feed_forward_size,
              dropout_rate=cfg.dropout_rate, time_sequence_length=cfg.seq_len, crop_size=236,
              use_token_learner=True)
    gs = reference.spaces   # the reference type-checks against its own (shimmed) gym.spaces classes
    rss = gs.Dict({k: gs.Box(low=v.low, high=v.high, shape=v.shape, dtype=np.float32) for k, v in ss.items()})
    rasp = gs.Dict(OrderedDict([("terminate_episode", gs.Discrete(2)),
                                ("action", gs.Box(low=cfg.action_low, high=cfg.action_high, shape=(2,), dtype=np.float32))]))
    ref = reference.transformer_network.TransformerNetwork(input_tensor_space=rss, output_tensor_space=rasp, **kw)
    ours = TransformerNetwork(input_tensor_space=ss, output_tensor_space=asp, **kw)
    return ref, ours


def _perturb_film(model, seed=0):
    # FiLM projections are zero-initialised; give them values so the parity test exercises them
    g = torch.Generator().manual_seed(seed)
    for name, p in model.named_parameters():
        if "_projection_" in name:
            with torch.no_grad():
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)


def test_full_config_state_dict_layout(reference):
    cfg = RT1Config.language_table()
    ref, ours = _build_pair(reference, cfg)
    rsd, osd = ref.state_dict(), ours.state_dict()
    assert list(rsd.keys()) == list(osd.keys())
    assert len(osd) == 806
    for k in rsd:
        assert rsd[k].shape == osd[k].shape, k
        assert rsd[k].dtype == osd[k].dtype, k
    assert sum(p.numel() for p in ours.parameters()) == 35_324_320
    assert len(list(ours.parameters())) == 572
    ours.load_state_dict(rsd, strict=True)
    ref.load_state_dict(osd, strict=True)


def test_masks_match_reference(reference):
    ref, ours = _build_pair(reference, _small_cfg())
    assert torch.equal(ref._default_attention_mask.to(torch.uint8), ours._default_attention_mask)
    assert ref._action_tokens_mask == ours._action_tokens_mask
    assert ref._all_num_tokens == ours._all_num_tokens


def test_train_forward_and_loss_parity(reference, cuda_literal_to_cpu):
    cfg = _small_cfg()
    ref, ours = _build_pair(reference, cfg)
    _perturb_film(ref)
    ours.load_state_dict(ref.state_dict())
    b, t = 2, cfg.seq_len
    g = torch.Generator().manual_seed(1)
    obs = {"image": torch.rand(b, t, 3, cfg.height, cfg.width, generator=g),
           "natural_language_embedding": torch.randn(b, t, 512, generator=g)}
    acts = {"terminate_episode": torch.randint(0, 2, (b, t), generator=g),
            "action": (torch.rand(b, t, 2, generator=g) - 0.5) * 0.25}
    state = {"context_image_tokens": torch.zeros(b, t, 8, 512), "action_tokens": torch.zeros(b, t, 3, dtype=torch.long),
             "seq_idx": torch.zeros(b, dtype=torch.long)}
    ref.train(); ours.train()
    ref.set_actions(acts); ours.set_actions(acts)
    torch.manual_seed(7)
    ra, _ = ref(obs, state)
    rl = ref.get_actor_loss()
    rl.mean().backward()
    torch.manual_seed(7)
    oa, _ = ours(obs, state)
    ol = ours.get_actor_loss()
    ol.mean().backward()
    assert ol.shape == (b, t)
    torch.testing.assert_close(ol, rl, rtol=1e-4, atol=1e-7)
    for k in ra:
        torch.testing.assert_close(oa[k], ra[k])
    assert torch.equal(ours.get_aux_info()["action_labels"], ref.get_aux_info()["action_labels"])
    rg = dict(ref.named_parameters()); og = dict(ours.named_parameters())
    n_checked = 0
    for k, p in og.items():
        if rg[k].grad is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
            continue
        torch.testing.assert_close(p.grad, rg[k].grad, rtol=2e-3, atol=1e-6, msg=k)
        n_checked += 1
    assert n_checked > 400
    # the action-token embedding is evaluated-then-zeroed upstream: no gradient either way
    assert og["_action_token_emb.weight"].grad is None
    # BN running statistics advanced identically
    for k, v in ours.state_dict().items():
        if "running_" in k or "num_batches" in k:
            torch.testing.assert_close(v, ref.state_dict()[k], rtol=1e-4, atol=1e-6, msg=k)


def test_inference_rollout_parity(reference, cuda_literal_to_cpu):
    cfg = _small_cfg()
    ref, ours = _build_pair(reference, cfg)
    _perturb_film(ref, 3)
    ours.load_state_dict(ref.state_dict())
    ref.eval(); ours.eval()
    T = cfg.seq_len
    rstate = {"context_image_tokens": torch.zeros(1, T, 8, 512), "action_tokens": torch.zeros(1, T, 3, dtype=torch.long),
              "seq_idx": torch.zeros(1, dtype=torch.long)}
    ostate = {k: v.clone() for k, v in rstate.items()}
    g = torch.Generator().manual_seed(5)
    emb = torch.randn(1, 512, generator=g)
    with torch.no_grad():
        for step in range(T + 3):                      # runs past the window so the roll is exercised
            img = torch.rand(1, 3, cfg.height, cfg.width, generator=g)
            obs = {"image": img, "natural_language_embedding": emb}
            torch.manual_seed(100 + step)
            ra, rstate = ref(obs, rstate)
            torch.manual_seed(100 + step)
            oa, ostate = ours(obs, ostate)
            for k in ra:
                torch.testing.assert_close(oa[k], ra[k], msg=f"step {step} {k}")
            torch.testing.assert_close(ostate["context_image_tokens"], rstate["context_image_tokens"],
                                       rtol=1e-4, atol=1e-5)
            assert int(ostate["seq_idx"]) == int(rstate["seq_idx"])
            assert ostate["seq_idx"].dim() == 0
            torch.testing.assert_close(ours.get_aux_info()["action_predictions_logits"],
                                       ref.get_aux_info()["action_predictions_logits"], rtol=1e-4, atol=1e-4)
            # both expect the caller to re-wrap seq_idx to shape [1] (reference policy.py:83)
            rstate["seq_idx"] = rstate["seq_idx"].reshape(1)
            ostate["seq_idx"] = ostate["seq_idx"].reshape(1)


def test_batched_policy_step_matches_single(reference):
    cfg = _small_cfg()
    torch.manual_seed(0)
    m = build_rt1(cfg, eval_center_crop=True).eval()
    _perturb_film(m, 1)
    B, T = 3, cfg.seq_len
    g = torch.Generator().manual_seed(2)
    emb = torch.randn(B, 512, generator=g)
    state = m.initial_state(B)
    # desynchronise the environments: env i has already seen i frames
    singles = [m.initial_state(1) for _ in range(B)]
    frames = torch.rand(T + 2, B, 3, cfg.height, cfg.width, generator=g)
    for i in range(B):
        for s in range(i):
            _, singles[i] = m.policy_step({"image": frames[s, i:i + 1], "natural_language_embedding": emb[i:i + 1]}, singles[i])
    state = {k: torch.cat([s[k] for s in singles], 0) for k in state}
    assert state["seq_idx"].tolist() == [0, 1, 2]
    for s in range(2, T + 2):
        obs = {"image": frames[s], "natural_language_embedding": emb}
        act, state = m.policy_step(obs, state)
        for i in range(B):
            a1, singles[i] = m.policy_step({"image": frames[s, i:i + 1], "natural_language_embedding": emb[i:i + 1]}, singles[i])
            torch.testing.assert_close(act["action"][i:i + 1], a1["action"])
            torch.testing.assert_close(state["context_image_tokens"][i:i + 1], singles[i]["context_image_tokens"],
                                       rtol=1e-4, atol=1e-5)
