"""RLDS → per-episode ``.npy`` conversion (reference ``rlds_np_convert.py``).

The reference script reads ``language_table_blocktoblock_sim`` through TFDS and embeds instructions with the Universal
Sentence Encoder; neither is available offline.  This entry point keeps the output format and split (7 800 / 100 / 100)
and takes its episodes either from an RLDS iterator supplied by the caller (:func:`rt1b200.data.episodes.convert_rlds`)
or, with ``--synthetic``, from the synthetic Language-Table generator.
"""
import argparse

from rt1b200.data.episodes import convert_rlds, write_synthetic_dataset  # noqa: F401


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--out", required=True)
    p.add_argument("--synthetic", action="store_true")
    p.add_argument("--train", type=int, default=7800)
    p.add_argument("--val", type=int, default=100)
    p.add_argument("--test", type=int, default=100)
    a = p.parse_args(argv)
    if not a.synthetic:
        raise SystemExit("TFDS/RLDS is not available offline: pass --synthetic, or call rt1b200.data.episodes.convert_rlds "
                         "with your own episode iterator and sentence encoder")
    write_synthetic_dataset(a.out, a.train, a.val, a.test)
    print(f"wrote synthetic dataset to {a.out}")


if __name__ == "__main__":
    main()
