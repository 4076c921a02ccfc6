"""Walk a Language-Table episode dataset and print what a step contains (reference
``language_table/examples/dataset_example.py``, which streams RLDS from GCS; here the per-episode ``.npy`` files that
``rlds_np_convert.py`` writes, or synthetic episodes when no path is given).

    python -m language_table.examples.dataset_example [--dataset_dir DIR] [--episodes 2]
"""
import argparse
import glob
import os

import numpy as np

from rt1b200.data.episodes import load_episode, synthetic_episode


# where the published RLDS datasets live (reference ``dataset_example.py:31-34``); not reachable offline — convert a
# local TFDS copy with ``rlds_np_convert.py --tfds_dir`` and pass the result as ``--dataset_dir``
dataset_paths = {
    "language_table": "gs://gresearch/robotics/language_table/0.0.1/",
    "language_table_sim": "gs://gresearch/robotics/language_table_sim/0.0.1/",
}


def decode_inst(inst) -> str:
    """Instructions stored as zero-padded utf-8 byte arrays (the simulator's format) → string."""
    return bytes(int(x) for x in np.asarray(inst).tolist() if x != 0).decode("utf-8")


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--dataset_dir", default=None)
    ap.add_argument("--split", default="train")
    ap.add_argument("--episodes", type=int, default=2)
    a = ap.parse_args(argv)
    if a.dataset_dir:
        eps = [load_episode(f) for f in sorted(glob.glob(os.path.join(a.dataset_dir, a.split, "episode_*.npy")))[:a.episodes]]
    else:
        rng = np.random.default_rng(0)
        eps = [synthetic_episode(rng, 6, 90, 160) for _ in range(a.episodes)]
    for i, ep in enumerate(eps):
        print(f"episode {i}: {len(ep)} steps")
        for t, step in enumerate(ep[:3]):
            inst = step["instruction"]
            text = decode_inst(inst) if np.asarray(inst).dtype.kind in "iu" else f"<{np.asarray(inst).shape[0]}-d sentence embedding>"
            print(f"  step {t}: rgb {np.asarray(step['rgb']).shape} action {np.round(np.asarray(step['action']), 4).tolist()} "
                  f"first={bool(step.get('is_first', False))} terminal={bool(step.get('is_terminal', False))} instruction={text}")
    return eps


if __name__ == "__main__":
    main()
