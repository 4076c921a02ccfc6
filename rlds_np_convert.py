"""RLDS → per-episode ``.npy`` conversion (reference ``rlds_np_convert.py``).

    python rlds_np_convert.py --out DIR --tfds_dir /data/language_table_blocktoblock_sim      # needs tensorflow_datasets
    python rlds_np_convert.py --out DIR --source EXPORT_DIR                                   # RLDS episodes exported as .npy
    python rlds_np_convert.py --out DIR --synthetic                                           # synthetic Language-Table episodes

Same output as the reference: the first 8 000 episodes split 7 800 / 100 / 100 into ``DIR/{train,val,test}/episode_<i>.npy``
(pickled lists of flat step dicts) with the instruction replaced by a 512-d sentence embedding; ``--raw_out`` also writes the
un-embedded tree.  The reference hard-codes its paths and embeds with the Universal Sentence Encoder from TF-Hub; here paths
are flags and the encoder is ``--encoder use`` (needs tensorflow_hub + a local USE copy via ``--use_path``) or the deterministic
hashed stand-in of ``language_table.common.rt1_tokenizer`` (default — no network on this box).
"""
import argparse

from rt1b200.data.episodes import convert_rlds, iter_exported_episodes, write_synthetic_dataset  # noqa: F401


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
    if a.source:
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
        raise SystemExit("give one of --tfds_dir, --source or --synthetic")
    written = convert_rlds(episodes, a.out, _encoder(a.encoder, a.use_path), splits=splits, raw_root=a.raw_out)
    print(f"wrote {written} to {a.out}")
    return written


if __name__ == "__main__":
    main()
