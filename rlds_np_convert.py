"""RLDS → per-episode ``.npy`` conversion (reference ``rlds_np_convert.py``).

    python rlds_np_convert.py --out DIR --tfds_dir /data/language_table_blocktoblock_sim      # needs tensorflow_datasets
    python rlds_np_convert.py --out DIR --source EXPORT_DIR                                   # RLDS episodes exported as .npy
    python rlds_np_convert.py --out DIR --oracle 200 --train 180 --val 10 --test 10            # scripted-oracle episodes from the simulator
    python rlds_np_convert.py --out DIR --synthetic                                           # procedural stand-in frames (benchmarks, tests)

Same output as the reference: the first 8 000 episodes split 7 800 / 100 / 100 into ``DIR/{train,val,test}/episode_<i>.npy``
(pickled lists of flat step dicts) with the instruction replaced by a 512-d sentence embedding; ``--raw_out`` also writes the
un-embedded tree.  The reference hard-codes its paths and embeds with the Universal Sentence Encoder from TF-Hub; here paths
are flags and the encoder is ``--encoder use`` (needs tensorflow_hub + a local USE copy via ``--use_path``) or the deterministic
hashed stand-in of ``language_table.common.rt1_tokenizer`` (default — no network on this box).
"""
import argparse

from rt1b200.data.episodes import convert_rlds, iter_exported_episodes, write_synthetic_dataset  # noqa: F401


NUM_TRAIN, NUM_VAL, NUM_TEST = 7800, 100, 100          # the reference's split of the first 8 000 episodes (``:35-37``)


def decode_inst(inst) -> str:
    """Zero-padded utf-8 byte array → instruction string."""
    from rt1b200.data.episodes import decode_instruction
    return decode_instruction(inst)


def create_episode(path, raw_episode, embed=False, embed_model=None, path_embed=None):
    """One RLDS episode → ``path`` (``.npy`` of flat step dicts); with ``embed`` also ``path_embed`` with every instruction
    replaced by ``embed_model([text])[0]`` (reference ``:13-33``; ``embed_model`` is any callable on a list of strings — the
    Universal Sentence Encoder upstream).  Embeddings are cached per distinct instruction."""
    import copy
    import os

    import numpy as np
    from rt1b200.data.episodes import flatten_rlds_episode, save_episode
    steps = flatten_rlds_episode(raw_episode)
    save_episode(path, steps)
    if embed:
        cache, out = {}, []
        for st in steps:
            st = copy.deepcopy(st)
            text = decode_inst(st["instruction"])
            if text not in cache:
                e = embed_model([text])[0]
                cache[text] = np.asarray(e.numpy() if hasattr(e, "numpy") else e, np.float32)
            st["instruction"] = cache[text]
            out.append(st)
        os.makedirs(os.path.dirname(os.path.abspath(path_embed)), exist_ok=True)
        save_episode(path_embed, out)


def _encoder(kind: str, use_path=None):
    if kind == "use":
        import tensorflow_hub as hub           # not installed offline: raises a clear ImportError
        model = hub.load(use_path)
        return lambda text: model([text])[0].numpy()
    from language_table.common.rt1_tokenizer import tokenize_text
    return tokenize_text


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--out", required=True)
    p.add_argument("--raw_out", default=None, help="also write the episodes with the raw byte instruction here")
    p.add_argument("--tfds_dir", default=None, help="TFDS builder directory of a Language-Table RLDS dataset")
    p.add_argument("--source", default=None, help="directory of RLDS episodes exported as .npy (same nesting as RLDS)")
    p.add_argument("--synthetic", action="store_true")
    p.add_argument("--oracle", type=int, default=0, metavar="N",
                   help="collect N episodes of the scripted RRT* oracle in the simulator as the source (the offline stand-in for the "
                        "published *_oracle_sim datasets; language_table/examples/collect_oracle_episodes.py)")
    p.add_argument("--oracle_reward", default="block2block", help="reward module name: block2block, block2absolutelocation, ..., block1_to_corner")
    p.add_argument("--oracle_block_mode", default="BLOCK_8")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--encoder", choices=["hashed", "use"], default="hashed")
    p.add_argument("--use_path", default=None)
    p.add_argument("--train", type=int, default=7800)
    p.add_argument("--val", type=int, default=100)
    p.add_argument("--test", type=int, default=100)
    a = p.parse_args(argv)
    splits = (("train", a.train), ("val", a.val), ("test", a.test))
    if a.synthetic:
        write_synthetic_dataset(a.out, a.train, a.val, a.test)
        print(f"wrote synthetic dataset to {a.out}")
        return {"train": a.train, "val": a.val, "test": a.test}
    if a.oracle:
        from language_table.examples.collect_oracle_episodes import oracle_episodes
        episodes = oracle_episodes(a.oracle, reward=a.oracle_reward, block_mode=a.oracle_block_mode, seed=a.seed, progress=25)
    elif a.source:
        episodes = iter_exported_episodes(a.source)
    elif a.tfds_dir:
        try:
            import tensorflow_datasets as tfds
        except ImportError as e:
            raise SystemExit(f"--tfds_dir needs tensorflow_datasets ({e}); export the episodes to .npy and use --source, "
                             "or pass --synthetic")
        total = a.train + a.val + a.test
        episodes = tfds.builder_from_directory(builder_dir=a.tfds_dir).as_dataset(split=f"train[0:{total}]", shuffle_files=False)
    else:
        raise SystemExit("give one of --tfds_dir, --source, --oracle N or --synthetic")
    written = convert_rlds(episodes, a.out, _encoder(a.encoder, a.use_path), splits=splits, raw_root=a.raw_out)
    print(f"wrote {written} to {a.out}")
    return written


if __name__ == "__main__":
    main()
