"""RLDS → npy converter against an RLDS-shaped source (reference rlds_np_convert.py:13-72) and the loader that reads it."""
import os

import numpy as np

import rlds_np_convert
from rt1b200.data import episodes as E


class _Eager:
    """Stands in for a TF eager tensor: the converter must call ``.numpy()``."""

    def __init__(self, v):
        self._v = np.asarray(v)

    def numpy(self):
        return self._v


def _instruction_bytes(text, n=512):
    raw = np.frombuffer(text.encode("utf-8"), np.uint8).astype(np.int32)
    out = np.zeros(n, np.int32)
    out[:len(raw)] = raw
    return out


def _fake_rlds(n_episodes, rng):
    texts = ["push the red moon to the blue cube", "slide the green star next to the yellow pentagon"]
    for e in range(n_episodes):
        T = int(rng.integers(4, 9))
        steps = []
        for t in range(T):
            steps.append({"observation": {"rgb": _Eager(rng.integers(0, 255, (18, 32, 3), dtype=np.uint8)),
                                          "instruction": _Eager(_instruction_bytes(texts[e % 2])),
                                          "effector_translation": _Eager(rng.random(2).astype(np.float32)),
                                          "effector_target_translation": _Eager(rng.random(2).astype(np.float32))},
                          "action": _Eager(rng.normal(0, 0.01, 2).astype(np.float32)), "reward": _Eager(np.float32(0)),
                          "is_first": _Eager(t == 0), "is_last": _Eager(t == T - 1), "is_terminal": _Eager(t == T - 1)})
        yield {"steps": steps}


def test_convert_rlds_keeps_every_key_and_embeds_the_instruction(tmp_path):
    rng = np.random.default_rng(0)
    calls = []

    def embed(text):
        calls.append(text)
        v = np.zeros(512, np.float32)
        v[len(text) % 512] = 1.0
        return v

    out, raw = str(tmp_path / "emb"), str(tmp_path / "raw")
    written = E.convert_rlds(_fake_rlds(7, rng), out, embed, splits=(("train", 4), ("val", 2), ("test", 2)), raw_root=raw)
    assert written == {"train": 4, "val": 2, "test": 1}                      # the source ran dry in the last split
    ep = E.load_episode(os.path.join(out, "train", "episode_0.npy"))
    ep_raw = E.load_episode(os.path.join(raw, "train", "episode_0.npy"))
    keys = {"rgb", "instruction", "effector_translation", "effector_target_translation", "action", "reward", "is_first",
            "is_last", "is_terminal"}
    assert set(ep[0]) == keys and set(ep_raw[0]) == keys
    assert ep[0]["instruction"].shape == (512,) and ep[0]["instruction"].dtype == np.float32
    assert ep_raw[0]["instruction"].dtype == np.int32 and E.decode_instruction(ep_raw[0]["instruction"]).startswith("push the red")
    assert ep[0]["rgb"].dtype == np.uint8 and bool(ep[0]["is_first"]) and not bool(ep[1]["is_first"])
    assert len(calls) == 7                                                    # one embedding per episode, cached over its steps
    # the reference-format loader reads the converted tree
    from rt1b200.data.dataset import EmbodiedIntelligenceDataset
    ds = EmbodiedIntelligenceDataset(os.path.join(out, "train"), ids=list(range(4)), window_length=3)
    sample = ds[0]
    assert sample["train_observation"]["natural_language_embedding"].shape == (3, 512)


def test_cli_converts_an_exported_directory(tmp_path):
    rng = np.random.default_rng(1)
    src = tmp_path / "export"
    src.mkdir()
    for i, ep in enumerate(_fake_rlds(3, rng)):
        flat_steps = [{**{k: v.numpy() for k, v in s.items() if k != "observation"},
                       "observation": {k: v.numpy() for k, v in s["observation"].items()}} for s in ep["steps"]]
        E.save_episode(str(src / f"episode_{i}.npy"), flat_steps)
    written = rlds_np_convert.main(["--out", str(tmp_path / "out"), "--source", str(src), "--train", "2", "--val", "1", "--test", "0"])
    assert written == {"train": 2, "val": 1, "test": 0}
    ep = E.load_episode(str(tmp_path / "out" / "val" / "episode_0.npy"))
    assert abs(float(np.linalg.norm(ep[0]["instruction"])) - 1.0) < 1e-4     # hashed stand-in embeddings are unit vectors


def test_oracle_episodes_through_the_converter(tmp_path):
    """``rlds_np_convert.py --oracle N``: scripted-oracle episodes from the simulator (the offline stand-in for the published
    *_oracle_sim datasets) in RLDS nesting → the reference's per-episode .npy layout → the training dataset class."""
    import rlds_np_convert
    from language_table.examples.collect_oracle_episodes import oracle_episodes
    from load_np_dataset import DecodeAndRandomResizedCrop, EmbodiedIntelligenceDataset
    from rt1b200.data.episodes import load_episode
    eps = list(oracle_episodes(2, seed=3, max_episode_steps=120))
    again = list(oracle_episodes(1, seed=3, max_episode_steps=120))
    assert len(eps) == 2 and len(eps[0]["steps"]) == len(again[0]["steps"])                      # seeded: reproducible
    steps = eps[0]["steps"]
    assert steps[0]["is_first"] and steps[-1]["is_last"] and steps[-1]["is_terminal"] and float(steps[-1]["reward"]) > 0
    assert not any(s["is_last"] for s in steps[:-1]) and steps[0]["observation"]["rgb"].shape == (180, 320, 3)
    assert all(np.abs(s["action"]).max() <= 0.03 + 1e-6 for s in steps) and np.abs(steps[0]["action"]).max() > 0
    assert steps[0]["observation"]["instruction"].shape == (512,) and steps[0]["observation"]["instruction"].dtype == np.int32

    written = rlds_np_convert.main(["--out", str(tmp_path / "ds"), "--raw_out", str(tmp_path / "raw"), "--oracle", "3", "--seed", "1",
                                    "--train", "2", "--val", "1", "--test", "0"])
    assert written == {"train": 2, "val": 1, "test": 0}
    ep = load_episode(str(tmp_path / "ds" / "train" / "episode_0.npy"))
    raw = load_episode(str(tmp_path / "raw" / "train" / "episode_0.npy"))
    assert np.asarray(ep[0]["instruction"]).shape == (512,) and np.asarray(ep[0]["instruction"]).dtype == np.float32
    assert np.asarray(raw[0]["instruction"]).dtype == np.int32 and len(ep) == len(raw)
    text = bytes(int(v) for v in raw[0]["instruction"] if v).decode()
    assert any(w in text for w in ("block", "moon", "star", "cube", "pentagon"))
    ds = EmbodiedIntelligenceDataset(str(tmp_path / "ds" / "train"), [0, 1], 6, transform=DecodeAndRandomResizedCrop(0.95, (112, 64)))
    sample = ds[len(ds) - 1]
    assert sample["train_observation"]["image"].shape == (6, 3, 64, 112) and int(sample["action_label"]["terminate_episode"][-1]) == 1
