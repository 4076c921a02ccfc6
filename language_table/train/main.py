"""``python -m language_table.train.main --config language_table/train/configs/language_table_sim_local.py --workdir DIR``
(reference ``language_table/train/main.py:31-70``; absl/ml_collections flags replaced by argparse + a config module)."""
import argparse
import importlib.util
import sys


def load_config(path: str):
    spec = importlib.util.spec_from_file_location("lava_config", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.get_config()


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True)
    ap.add_argument("--workdir", required=True)
    ap.add_argument("--num_train_steps", type=int, default=None)
    ap.add_argument("--synthetic_data", action="store_true")
    ap.add_argument("--dataset_path", default=None)
    a = ap.parse_args(argv)
    config = load_config(a.config)
    if a.num_train_steps is not None:
        config.num_train_steps = a.num_train_steps
    if a.synthetic_data:
        config.synthetic_data = True
    if a.dataset_path:
        config.dataset_path = a.dataset_path
    from language_table.train import train as train_lib
    out = train_lib.train(config, a.workdir)
    print(f"[lava] finished {out['steps']} steps, last loss {out['losses'][-1]:.5f}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
