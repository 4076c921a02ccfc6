"""PyTorch restatement of the vendored LAVA baseline (reference ``language_table/train``, ``common/clip_tokenizer.py``,
``eval/main.py``).  Mirrors the reference's ``resnet_v1_test.py`` (exact parameter counts) and
``clip_tokenizer_test.py`` (BPE behaviour, on a tiny merge table since the real one is not shipped), and adds what the
reference never tested: network shapes, a BC training run that learns, normalisation statistics, the CLI."""
import numpy as np
import pytest
import torch

from language_table.common import clip_tokenizer
from language_table.train import main as train_main
from language_table.train import normalization
from language_table.train import train as train_lib
from language_table.train.configs import language_table_resnet_sim_local, language_table_sim_local
from language_table.train.networks import lava, pixel, resnet_v1


@pytest.mark.parametrize("name,count", [("ResNet18", 11_689_512), ("ResNet34", 21_797_672), ("ResNet50", 25_557_032),
                                        ("ResNet101", 44_549_160), ("ResNet152", 60_192_808), ("ResNet200", 64_673_832)])
def test_resnet_parameter_counts(name, count):
    model = getattr(resnet_v1, name)(num_classes=1000)
    assert sum(p.numel() for p in model.parameters()) == count


def test_multiscale_resnet_pyramid():
    feats = resnet_v1.MultiscaleResNet(stage_sizes=(3, 4))(torch.rand(1, 3, 224, 224))
    assert [tuple(f.shape[1:]) for f in feats] == [(64, 112, 112), (64, 56, 56), (256, 56, 56), (512, 28, 28)]


def _small(cfg):
    cfg.model.dense_resnet_width, cfg.model.lava_d_model, cfg.model.lava_num_layers = 64, 32, 2
    cfg.data_target_height, cfg.data_target_width = 64, 96
    cfg.synthetic_data, cfg.per_device_batch_size = True, 4
    cfg.log_loss_every_steps, cfg.checkpoint_every_steps = 10 ** 6, 5
    return cfg


@pytest.mark.parametrize("encoder", ["conv_maxpool", "resnet"])
def test_sequence_lava_shapes_and_frozen_resnet(encoder):
    cfg = _small(language_table_sim_local.get_config() if encoder == "conv_maxpool" else language_table_resnet_sim_local.get_config())
    model = train_lib.create_model(cfg)
    obs = {"rgb": torch.rand(2, 4, 64, 96, 3), "clip_embedding": torch.randn(2, 4, 512)}
    out = model(obs)
    assert out.shape == (2, 2)
    out.sum().backward()
    if encoder == "resnet":
        assert all(p.grad is None for p in model.encoder.image_encoder.resnet.parameters())
        model.train()
        assert not model.encoder.image_encoder.resnet.training          # frozen tower keeps inference-mode BatchNorm


def test_pixel_lang_mse_shapes():
    model = pixel.PixelLangMSE(2, 64, 2, sequence_length=4)
    assert model({"rgb": torch.rand(3, 4, 64, 96, 3), "clip_embedding": torch.randn(3, 4, 512)}).shape == (3, 2)


def test_positional_codes():
    pe = lava.positional_encoding2d(32, 5, 7)
    assert pe.shape == (1, 35, 32) and float(pe.abs().max()) <= 1.0
    s = lava.sinusoid_1d(6, 16)
    assert s.shape == (1, 6, 16) and torch.allclose(s[0, 0, 0::2], torch.zeros(8)) and torch.allclose(s[0, 0, 1::2], torch.ones(8))


def test_chan_statistics_match_numpy():
    rng = np.random.default_rng(0)
    data = rng.normal(2.0, 3.0, size=(1000, 4))
    st = normalization.ChanRunningStatistics()
    for chunk in np.array_split(data, 7):
        st.update_running_statistics(chunk)
    assert st.n == 1000
    np.testing.assert_allclose(st.mean, data.mean(0), rtol=1e-10)
    np.testing.assert_allclose(st.std, data.std(0), rtol=1e-10)
    norm, denorm = normalization.StdNormalizer(st.mean, st.std), normalization.StdDenormalizer(st.mean, st.std)
    np.testing.assert_allclose(denorm(norm(data[:5].astype(np.float32))), data[:5], rtol=1e-4)
    mm = normalization.MinMaxNormalizer(data.min(0), data.max(0))
    assert float(np.abs(mm(data)).max()) <= 1.0 + 1e-5
    np.testing.assert_allclose(normalization.MinMaxDenormalizer(data.min(0), data.max(0))(mm(data[:5])), data[:5], rtol=1e-4, atol=1e-4)


def test_bc_training_learns_checkpoints_and_resumes(tmp_path):
    cfg = _small(language_table_sim_local.get_config())
    cfg.image_photometric_distortions = False
    cfg.num_train_steps = 12
    out = train_lib.train(cfg, str(tmp_path), device="cpu", verbose=False)
    assert out["steps"] == 12 and np.isfinite(out["losses"]).all()
    assert np.mean(out["losses"][-4:]) < np.mean(out["losses"][:4])          # the BC loss goes down
    assert train_lib.latest_checkpoint(str(tmp_path)).endswith("ckpt_12.pt")
    cfg.num_train_steps = 14
    again = train_lib.train(cfg, str(tmp_path), device="cpu", verbose=False)   # restores step 12, runs 2 more
    assert again["steps"] == 14 and len(again["losses"]) == 2
    # the evaluation-side policy adapter loads the checkpoint and acts inside the dataset's action range
    from language_table.train.policy import BCJaxPyPolicy
    pol = BCJaxPyPolicy.from_checkpoint(train_lib.create_model(cfg), train_lib.latest_checkpoint(str(tmp_path)), device="cpu")
    act = pol.action({"rgb_sequence": np.random.rand(4, 64, 96, 3).astype(np.float32),
                      "instruction_tokenized_use": np.random.randn(4, 512).astype(np.float32)})
    assert act.shape == (2,) and np.all(act >= pol.amin - 1e-6) and np.all(act <= pol.amax + 1e-6)


def test_train_cli(tmp_path):
    import os
    cfg_path = os.path.join(os.path.dirname(train_main.__file__), "configs", "language_table_sim_local.py")
    cfg = train_main.load_config(cfg_path)
    assert cfg.model_name == "sequence_lav_mse" and cfg.sequence_length == 4 and cfg.per_device_batch_size == 4
    # ml_collections-style overrides on the command line: --config.<path>=<literal>
    train_main.apply_config_overrides(cfg, ["--config.agent.learning_rate=3e-4", "--config.model.lava_num_layers", "2",
                                            "--config.data_skip_normalize_keys=['rgb']", "--config.model.lava_image_encoder=resnet"])
    assert cfg.agent.learning_rate == 3e-4 and cfg.model.lava_num_layers == 2 and cfg.data_skip_normalize_keys == ["rgb"]
    assert cfg.model.lava_image_encoder == "resnet"
    for bad in (["--config.agent.learning_rat=1"], ["--config.nosuch.x=1"], ["--lr", "1"], ["--config.seed"]):
        with pytest.raises(SystemExit):
            train_main.apply_config_overrides(cfg, bad)
    # the whole entry point, two steps of a shrunken model
    rc = train_main.main(["--config", cfg_path, "--workdir", str(tmp_path / "cli"), "--synthetic_data", "--num_train_steps", "2",
                          "--config.model.dense_resnet_width=64", "--config.model.lava_d_model=32", "--config.model.lava_num_layers=1",
                          "--config.data_target_height=36", "--config.data_target_width=64", "--config.checkpoint_every_steps=2"])
    assert rc == 0 and os.path.exists(tmp_path / "cli" / "checkpoints" / "ckpt_2.pt")
    import json
    assert json.load(open(tmp_path / "cli" / "hparams.json"))["model"]["lava_d_model"] == 32


def test_clip_bpe_tokenizer():
    # tiny merge table: "pu"+"sh</w>" etc. — the real table is CLIP's 48 894 merges
    merges = [("p", "u"), ("s", "h</w>"), ("pu", "sh</w>"), ("t", "h"), ("th", "e</w>")]
    enc, ranks = clip_tokenizer.create_vocab(merges=merges)
    assert len(enc) == 512 + len(merges) + 2
    tok = clip_tokenizer.ClipTokenizer(enc, ranks)
    ids = tok.tokenize("  Push   the RED block! ")
    assert ids[0] == enc["push</w>"] and ids[1] == enc["the</w>"]
    assert ids[-1] == enc["!</w>"]
    # offsets point into the cleaned text, in bytes, piece by piece (multi-byte characters count their bytes)
    ids2, begin, end = tok.tokenize_with_offsets("  Push   thé RED block! ")
    clean = tok.clean("  Push   thé RED block! ").encode()
    assert clean == "push thé red block!".encode() and len(ids2) == len(begin) == len(end)
    assert (begin[0], end[0]) == (0, 4) and clean[begin[-1]:end[-1]] == b"!"
    assert b"".join(clean[b:e] for b, e in zip(begin, end)) == clean.replace(b" ", b"")
    assert all(b0 < e0 <= b1 for (b0, e0), b1 in zip(zip(begin, end), begin[1:] + [len(clean)]))
    out = clip_tokenizer.tokenize_text(["push the block", "push"], tok, vocab_size=len(enc))
    assert out.shape == (2, 77) and out[0, 0] == len(enc) - 2 and out[1, 2] == len(enc) - 1 and out[1, 3] == 0


def test_dataset_example_and_plan_plot(tmp_path):
    from language_table.environments.oracles import plot
    from language_table.examples import dataset_example
    eps = dataset_example.main(["--episodes", "1"])
    assert len(eps) == 1 and "rgb" in eps[0][0]
    img = plot.plot_plan([(0.2, -0.2), (0.4, 0.0), (0.5, 0.2)], [(0.35, -0.05, 0.03)], (0.2, -0.2), (0.5, 0.2),
                         filename=str(tmp_path / "plan.ppm"))
    assert img.shape == (400, 400, 3) and (img == (220, 30, 30)).all(-1).any()
    assert (tmp_path / "plan.ppm").stat().st_size > 400 * 400 * 3


def test_lava_closed_loop_eval_entry(tmp_path):
    """``language_table/eval/main.py``: five task families, oracle-validated resets, results.json."""
    from language_table.eval import main as eval_main
    cfg = _small(language_table_sim_local.get_config())
    cfg.num_train_steps, cfg.checkpoint_every_steps = 2, 2
    train_lib.train(cfg, str(tmp_path / "run"), device="cpu", verbose=False)
    res = eval_main.evaluate_checkpoint(train_lib.latest_checkpoint(str(tmp_path / "run")), str(tmp_path / "eval"), cfg,
                                        num_evals_per_reward=1, max_episode_steps=2, device="cpu")
    assert list(res) == ["blocktoblock", "blocktoabsolutelocation", "blocktoblockrelativelocation", "blocktorelativelocation", "separate"]
    assert all(v["episodes"] == 1 for v in res.values()) and (tmp_path / "eval" / "results.json").exists()


def test_clip_text_tower_architecture_pooling_and_checkpoint_names():
    """The CLIP text tower behind ``lang_encoder="clip"`` (reference lava.py:423-432): ViT-B/32's text transformer."""
    from language_table.train.networks.clip_text import ClipTextEncoder, load_clip_text_weights
    torch.manual_seed(0)
    full = ClipTextEncoder()
    assert sum(p.numel() for p in full.parameters()) == 63_428_096          # CLIP ViT-B/32 text tower
    enc = ClipTextEncoder(vocab_size=300, features=64, num_heads=4, num_layers=2, out_features=32, context_length=16).eval()
    sot, eot = 298, 299
    toks = torch.zeros(3, 16, dtype=torch.long)
    toks[0, :5] = torch.tensor([sot, 5, 6, 7, eot])
    toks[1, :5] = torch.tensor([sot, 5, 6, 7, eot])
    toks[1, 5:9] = torch.tensor([1, 2, 3, 4])            # padding garbage AFTER the end-of-text token
    toks[2, :4] = torch.tensor([sot, 5, 9, eot])
    with torch.no_grad():
        out = enc(toks)
    assert out.shape == (3, 32)
    # causal attention + pooling at the end-of-text position: what follows it cannot change the embedding
    torch.testing.assert_close(out[0], out[1])
    assert not torch.allclose(out[0], out[2])
    # public CLIP checkpoints name the blocks transformer.resblocks.<i>.* and carry a visual half
    other = ClipTextEncoder(vocab_size=300, features=64, num_heads=4, num_layers=2, out_features=32, context_length=16)
    ck = {("transformer." + k if k.startswith("resblocks.") else k): v.clone() for k, v in other.state_dict().items()}
    ck["visual.conv1.weight"] = torch.zeros(1)
    ck["logit_scale"] = torch.zeros(())
    load_clip_text_weights(enc, ck)
    with torch.no_grad():
        torch.testing.assert_close(enc(toks), other.eval()(toks))
    with pytest.raises(KeyError):
        load_clip_text_weights(enc, {k: v for k, v in ck.items() if not k.startswith("ln_final")})
    frozen = ClipTextEncoder(vocab_size=300, features=64, num_heads=4, num_layers=1, out_features=32, context_length=16).freeze()
    assert not any(p.requires_grad for p in frozen.parameters()) and not frozen.training


def test_sequence_lava_with_in_graph_clip_tower():
    from language_table.train.networks import lava
    from language_table.train.networks.clip_text import ClipTextEncoder
    tower = ClipTextEncoder(vocab_size=300, features=64, num_heads=4, num_layers=1, out_features=512, context_length=16).freeze()
    enc = lava.SequenceLAVAEncoder(lang_encoder="clip", sequence_length=4, text_encoder=tower)
    toks = torch.zeros(2, 4, 16, dtype=torch.long)
    toks[:, :, 0], toks[:, :, 1], toks[:, :, 2] = 298, 7, 299
    out = enc({"rgb": torch.rand(2, 4, 64, 96, 3), "instruction_tokenized_clip": toks})
    assert out.shape[0] == 2
    out.sum().backward()
    assert all(p.grad is None for p in tower.parameters())               # frozen, like the reference's


def test_normalization_handoff_between_processes(tmp_path):
    """Process 0 computes and writes the normalisation statistics, the others wait for the file (reference
    input_pipeline_rlds.py:191-233): same numbers everywhere, also after a restart."""
    import threading
    from language_table.train import input_pipeline_rlds as ip
    cfg = _small(language_table_sim_local.get_config())
    cfg.synthetic_data = True
    path = str(tmp_path / "norm_info")
    got = {}
    waiter = threading.Thread(target=lambda: got.update(r1=ip.create_datasets(cfg, seed=0, normalization_path=path, process_index=1)[1]))
    waiter.start()                                      # rank 1 starts first and has to wait
    _, info0 = ip.create_datasets(cfg, seed=0, normalization_path=path, process_index=0)
    waiter.join(timeout=60)
    assert not waiter.is_alive() and "r1" in got
    a0, a1 = info0["action_statistics"], got["r1"]["action_statistics"]
    assert np.allclose(a0["mean"], a1["mean"]) and np.allclose(a0["std"], a1["std"])
    # a restarted process 0 re-loads the file instead of recomputing: poison the file and see the poison come back
    info0["action_min"] = np.asarray(info0["action_min"]) - 123.0
    ip._save_norm_info(path, info0)
    _, again = ip.create_datasets(cfg, seed=0, normalization_path=path, process_index=0)
    assert np.allclose(again["action_min"], info0["action_min"])


def test_clip_checkpoint_conversion_script(tmp_path):
    from language_table.train import download_clip_flax_ckpt as conv
    from language_table.train.networks.clip_text import ClipTextEncoder
    src = ClipTextEncoder(vocab_size=300, features=64, num_heads=1, num_layers=2, out_features=32, context_length=16)
    ck = {("transformer." + k if k.startswith("resblocks.") else k): v for k, v in src.state_dict().items()}
    ck["visual.proj"] = torch.zeros(2)
    torch.save(ck, tmp_path / "clip.pt")
    out = conv.convert(str(tmp_path / "clip.pt"), str(tmp_path / "ckpt"))
    back = torch.load(out)
    assert set(back) == set(src.state_dict()) and all(torch.equal(back[k], v) for k, v in src.state_dict().items())


def test_lava_train_state_metrics_and_multi_step(tmp_path):
    """TrainState / TrainMetrics / flattened_traversal (reference train/bc.py:32-47, :237), pretrained-checkpoint key
    replacement (bc.py:92-108), multi_train_step / merge_batch_stats / the run's side files (train/train.py)."""
    import csv
    import os
    from language_table.train import bc
    from language_table.train.configs import Hyper, apply_overrides
    m = bc.TrainMetrics.single_from_model_output(loss=1.0, learning_rate=1e-3).merge(
        bc.TrainMetrics.single_from_model_output(loss=3.0, learning_rate=5e-4))
    out = m.compute()
    assert out["loss"] == 2.0 and abs(out["loss_std"] - 1.0) < 1e-12 and out["learning_rate"] == 5e-4
    assert bc.TrainMetrics.empty().merge(m).compute() == out
    labels = bc.flattened_traversal(lambda path, v: "/".join(path) + ":" + str(v))({"a": {"b": 1, "c": {"d": 2}}, "e": 3})
    assert labels == {"a": {"b": "a/b:1", "c": {"d": "a/c/d:2"}}, "e": "e:3"}

    cfg = _small(language_table_sim_local.get_config())
    cfg.num_train_steps, cfg.checkpoint_every_steps, cfg.log_loss_every_steps, cfg.num_steps_per_train_iter = 6, 3, 2, 2
    res = train_lib.train(cfg, str(tmp_path / "run"), device="cpu", verbose=False)
    agent = res["agent"]
    assert res["steps"] == 6 and len(res["losses"]) == 6
    rows = list(csv.DictReader(open(tmp_path / "run" / "metrics.csv")))
    assert [int(r["step"]) for r in rows] == [2, 4, 6] and all(float(r["loss_std"]) >= 0 for r in rows)
    assert abs(float(rows[0]["loss"]) - sum(res["losses"][:2]) / 2) < 1e-6
    overview = open(tmp_path / "run" / "parameters.txt").read()
    assert "trainable" in overview and f"total: {sum(p.numel() for p in agent.model.parameters()):,}" in overview
    assert os.path.exists(tmp_path / "run" / "hparams.json")
    assert sorted(os.listdir(tmp_path / "run" / "checkpoints")) == ["ckpt_3.pt", "ckpt_6.pt"]

    state, initial = agent.create_train_state()
    assert isinstance(state, bc.TrainState) and state.step == 6 and initial.count == 0
    assert set(state.params) == {n for n, _ in agent.model.named_parameters()} and "observation_statistics" in state.norm_info
    assert state.replace(step=7).step == 7 and state.step == 6
    flat = [v for v in str(agent.frozen_labels()).replace("'", " ").split() if v in ("adam", "zero")]
    assert len(flat) == len(state.params) and "adam" in flat
    assert train_lib.merge_batch_stats(agent) is agent                       # one process: nothing to average

    # a second agent picks up one sub-tree of the first run's checkpoint under another prefix
    ck = train_lib.latest_checkpoint(str(tmp_path / "run"))
    sd = torch.load(ck, weights_only=False)["params"]
    prefix = next(k for k in sd if sd[k].dim() >= 2).split(".")[0]
    renamed = {"old." + k: v for k, v in sd.items() if k.startswith(prefix + ".")}
    torch.save(renamed, tmp_path / "part.pt")
    model2 = train_lib.create_model(cfg)
    agent2 = bc.BCAgent(model2, cfg.sequence_length, 1e-3, {}, {}, None, None,
                        pretrained_checkpoints=[(str(tmp_path / "part.pt"), [("old." + prefix, prefix)])])
    assert len(agent2.loaded_pretrained) == len(renamed)
    sd2 = model2.state_dict()
    assert all(torch.equal(sd2[k[4:]], v) for k, v in renamed.items())

    sweep = language_table_sim_local.get_hyper(Hyper)
    assert sweep == [{"config.trial": 0}] and apply_overrides(cfg, {"config.agent.learning_rate": 0.5}).agent.learning_rate == 0.5


def test_lava_two_process_training(tmp_path):
    """pmap / pmean of the reference as two gloo processes: tests/_lava_worker.py."""
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29641", os.path.join(here, "_lava_worker.py"), str(tmp_path / "run")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, OMP_NUM_THREADS="2"))
    assert r.returncode == 0 and "LAVA_WORKER_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_lava_image_transforms():
    """DecodeAndRandomResizedCrop / PhotometricDistortions (reference train/input_pipeline_rlds.py:326-452)."""
    import colorsys
    import io
    from PIL import Image
    from language_table.train import input_pipeline_rlds as ip
    rng = np.random.default_rng(0)
    frames = np.random.default_rng(1).integers(0, 256, (3, 40, 60, 3), dtype=np.uint8)
    crop = ip.DecodeAndRandomResizedCrop(0.95, (18, 32))
    # the factor applies to each side, in whole pixels; offsets stay inside; one window for the whole sequence
    for _ in range(50):
        y0, x0, ch, cw = crop.window(40, 60, rng)
        assert (ch, cw) == (38, 57) and 0 <= y0 < 2 and 0 <= x0 < 3
    out = crop(frames, rng)
    assert out.shape == (3, 18, 32, 3) and 0.0 <= float(out.min()) and float(out.max()) <= 1.0
    assert crop.window(40, 60, rng, train=False) == (1, 1, 38, 57)
    whole = ip.DecodeAndRandomResizedCrop(None, (40, 60))(frames, rng)
    assert torch.allclose(whole, torch.from_numpy(frames).float() / 255.0, atol=1e-6)
    # encoded frames (the RLDS storage format) are decoded first
    def png(a):
        buf = io.BytesIO()
        Image.fromarray(a).save(buf, format="PNG")
        return buf.getvalue()
    assert torch.equal(ip.DecodeAndRandomResizedCrop(None, (40, 60))([png(f) for f in frames], rng), whole)

    x = torch.rand(2, 6, 5, 3)
    hsv = ip.rgb_to_hsv(x)
    ref = torch.tensor([[[colorsys.rgb_to_hsv(*map(float, x[t, i, j])) for j in range(5)] for i in range(6)] for t in range(2)])
    assert torch.allclose(hsv, ref, atol=1e-5) and torch.allclose(ip.hsv_to_rgb(hsv), x, atol=1e-5)
    pd = ip.PhotometricDistortions()
    assert torch.allclose(pd.apply(x), x, atol=1e-5)                                      # neutral levels: identity
    assert torch.allclose(pd.apply(x, brightness=0.05), x + 0.05, atol=1e-5)
    gray = pd.apply(x, saturation=0.0)
    assert torch.allclose(gray, x.max(-1, keepdim=True)[0].expand_as(x), atol=1e-5)       # S = 0 leaves V on every channel
    c = pd.apply(x, contrast=0.0)
    assert torch.allclose(c, x.mean(dim=(1, 2), keepdim=True).expand_as(x), atol=1e-5)    # per-frame, per-channel mean
    third = pd.apply(torch.tensor([[[[1.0, 0.0, 0.0]]]]), hue=1.0 / 3.0)
    assert torch.allclose(third, torch.tensor([[[[0.0, 1.0, 0.0]]]]), atol=1e-5)          # red a third of a turn on: green
    levels = [pd.sample(rng) for _ in range(200)]
    assert all(abs(l["brightness"]) <= 0.1 and 0.8 <= l["contrast"] <= 1.2 and abs(l["hue"]) <= 0.03
               and 0.8 <= l["saturation"] <= 1.2 for l in levels)
    assert ip.PhotometricDistortions(0.0, 1.0, 1.0, 0.0, 1.0, 1.0)(x, rng) is x           # everything switched off


def test_lava_clip_tokens_from_raw_episodes(tmp_path):
    """lava_lang_encoder="clip" end to end on the converter's RAW tree (instruction = the simulator's byte code): the input
    pipeline BPE-tokenises it into ``instruction_tokenized_clip`` [t, 77] (reference input_pipeline_rlds.py:236-262) and the
    frozen text tower turns that into the language embedding.  Tiny merge table and tower; the real ones are path arguments."""
    import gzip
    from language_table.environments.language_table import LanguageTable
    from language_table.train import input_pipeline_rlds as ip
    from language_table.train.networks.clip_text import ClipTextEncoder
    from rt1b200.data.episodes import save_episode, synthetic_episode
    rng = np.random.default_rng(0)
    for i, text in enumerate(["push the red moon to the blue cube", "push the star left"]):
        steps = synthetic_episode(rng, length=7, height=36, width=64)
        for s in steps:
            s["instruction"] = LanguageTable.encode_instruction(text)
        save_episode(str(tmp_path / "raw" / "train" / f"episode_{i}.npy"), steps)
    # a merge file in CLIP's format: a header line, then one merge per line
    merges = [("p", "u"), ("s", "h</w>"), ("pu", "sh</w>"), ("t", "h"), ("th", "e</w>")]
    bpe = tmp_path / "bpe.txt.gz"
    with gzip.open(bpe, "wt", encoding="utf-8") as f:
        f.write("#version: test\n" + "\n".join(" ".join(m) for m in merges) + "\n")

    cfg = _small(language_table_sim_local.get_config())
    cfg.synthetic_data, cfg.dataset_path, cfg.clip_bpe_path = False, str(tmp_path / "raw"), str(bpe)
    cfg.model.lava_lang_encoder = "clip"
    ds, info = ip.create_datasets(cfg, seed=0)
    batch = next(ds.batches(2, shuffle=False))
    tok = batch["observation"]["instruction_tokenized_clip"]
    assert tok.shape == (2, cfg.sequence_length, 77) and tok.dtype == torch.int64 and "clip_embedding" not in batch["observation"]
    vocab = 512 + len(merges) + 2
    # framed by the start / end ids of the 49 408-entry vocabulary the tower expects (the tiny table's own ids sit below)
    assert int(tok[0, 0, 0]) == 49408 - 2 and int(tok[0, 0].max()) == 49408 - 1 and int(tok[0, 0, 1]) == 512 + 2      # "push</w>"
    assert vocab < 49408 and torch.equal(tok[0, 0], tok[0, -1])                     # one instruction per episode
    # without a tokenizer the pipeline says what is wrong instead of feeding byte codes to the model
    cfg2 = _small(language_table_sim_local.get_config())
    cfg2.synthetic_data, cfg2.dataset_path = False, str(tmp_path / "raw")
    with pytest.raises(ValueError, match="clip_tokenizer"):
        next(ip.create_datasets(cfg2, seed=0)[0].batches(2))

    tower = ClipTextEncoder(features=32, num_heads=2, num_layers=1, out_features=512).freeze()
    from language_table.train.networks import lava                                    # (the 63 M-parameter default tower is not needed here)
    model = lava.SequenceLAVMSE(action_size=2, text_encoder=tower, **dict(cfg.model))
    out = model(batch["observation"])
    assert out.shape == (2, 2) and torch.isfinite(out).all()
    agent = train_lib.bc.BCAgent(model, cfg.sequence_length, 1e-3, {}, {}, None, None, freeze_keys=["text_encoder"])
    before = tower.token_embedding.weight.clone()
    loss = agent.train({"observation": batch["observation"], "action": batch["action"]})["loss"]
    assert np.isfinite(loss) and torch.equal(tower.token_embedding.weight, before)    # the tower stays frozen


def test_clip_token_wrapper_and_policy(tmp_path):
    """ClipTokenWrapper (named by the reference's eval/main.py:85, absent from its wrappers.py): BPE ids of the environment's
    instruction, history-stacked, consumed by the LAVA policy adapter of a model with the in-graph text tower."""
    from language_table.environments import blocks, language_table
    from language_table.environments.rewards import block2block
    from language_table.eval import wrappers
    from language_table.train.networks import lava
    from language_table.train.networks.clip_text import ClipTextEncoder
    from language_table.train.policy import BCJaxPyPolicy
    merges = [("p", "u"), ("s", "h</w>"), ("pu", "sh</w>"), ("t", "h"), ("th", "e</w>")]
    tok = clip_tokenizer.ClipTokenizer(*clip_tokenizer.create_vocab(merges=merges))
    base = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, block2block.BlockToBlockReward, seed=0)
    env = wrappers.HistoryWrapper(wrappers.CentralCropImageWrapper(wrappers.ClipTokenWrapper(base, tokenizer=tok), 36, 64, 0.95),
                                  history_length=4, tile_first_step_obs=True)
    obs = env.reset()
    ids = obs["instruction_tokenized_clip"]
    assert ids.shape == (4, 77) and ids.dtype == np.int64 and ids[0, 0] == 49406 and ids[0].max() == 49407
    text = base.instruction_str
    assert np.array_equal(ids[0], clip_tokenizer.tokenize_text(text, tok)[0]) and np.array_equal(ids[0], ids[3])
    obs2, *_ = env.step(np.zeros(2, np.float32))
    assert np.array_equal(obs2["instruction_tokenized_clip"][-1], ids[0])                  # cached until the next reset
    with pytest.raises(ValueError, match="BPE merge file"):
        wrappers.ClipTokenWrapper(base)
    cfg = _small(language_table_sim_local.get_config())
    cfg.model.lava_lang_encoder = "clip"
    tower = ClipTextEncoder(features=32, num_heads=2, num_layers=1, out_features=512).freeze()
    model = lava.SequenceLAVMSE(action_size=2, text_encoder=tower, **dict(cfg.model)).eval()
    norm = {"action_statistics": {"mean": np.zeros(2, np.float32), "std": np.ones(2, np.float32)},
            "action_min": -np.ones(2, np.float32) * 0.03, "action_max": np.ones(2, np.float32) * 0.03}
    act = BCJaxPyPolicy(model, norm, device="cpu").action(obs)
    assert act.shape == (2,) and np.all(np.abs(act) <= 0.03 + 1e-6)
