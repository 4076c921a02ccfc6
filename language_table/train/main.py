"""``python -m language_table.train.main --config language_table/train/configs/language_table_sim_local.py --workdir DIR``
(reference ``language_table/train/main.py:31-70``; absl/ml_collections flags replaced by argparse + a config module).

Any field of the config can be overridden on the command line the way ``ml_collections.config_flags`` allows upstream:
``--config.agent.learning_rate=3e-4 --config.model.lava_num_layers 2 --config.data_skip_normalize_keys='["rgb"]'`` (values are
read as Python literals when they parse as one, else as strings; the field must exist).  Under ``torchrun`` every process trains
its own shard order and gradients are averaged (``--tf_data_service_address`` of the reference has no counterpart: there is
no TF data service here)."""
import argparse
import ast
import importlib.util
import sys


def load_config(path: str):
    spec = importlib.util.spec_from_file_location("lava_config", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.get_config()


def apply_config_overrides(config, args):
    """``["--config.a.b=1", "--config.c", "x"]`` → ``config.a.b = 1; config.c = "x"``.  Unknown fields and anything that is not
    a ``--config.`` option are errors (a typo must not train silently with the default)."""
    i = 0
    while i < len(args):
        arg = args[i]
        if not arg.startswith("--config."):
            raise SystemExit(f"unrecognised argument: {arg}")
        if "=" in arg:
            key, raw = arg[2:].split("=", 1)
            i += 1
        else:
            if i + 1 >= len(args):
                raise SystemExit(f"{arg} needs a value")
            key, raw = arg[2:], args[i + 1]
            i += 2
        node, parts = config, key.split(".")[1:]
        for part in parts[:-1]:
            if part not in node:
                raise SystemExit(f"--{key}: the config has no section '{part}'")
            node = node[part]
        if parts[-1] not in node:
            raise SystemExit(f"--{key}: the config has no field '{parts[-1]}'")
        try:
            value = ast.literal_eval(raw)
        except (ValueError, SyntaxError):
            value = raw
        node[parts[-1]] = value
    return config


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True)
    ap.add_argument("--workdir", required=True)
    ap.add_argument("--num_train_steps", type=int, default=None)
    ap.add_argument("--synthetic_data", action="store_true")
    ap.add_argument("--dataset_path", default=None)
    a, extra = ap.parse_known_args(argv)
    config = load_config(a.config)
    apply_config_overrides(config, extra)
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
