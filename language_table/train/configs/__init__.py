"""Training configs of the LAVA baseline (reference ``language_table/train/configs``)."""


class ConfigDict(dict):
    """Attribute-style dict standing in for ``ml_collections.ConfigDict`` (not installed offline)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


class Hyper:
    """The two calls a config's ``get_hyper`` makes on its argument: ``sweep(name, values)`` → a list of one-key
    overrides, ``product(list of sweeps)`` → their cartesian product as merged override dicts."""

    @staticmethod
    def sweep(name, values):
        return [{name: v} for v in values]

    @staticmethod
    def product(sweeps):
        out = [{}]
        for sweep in sweeps:
            out = [{**a, **b} for a in out for b in sweep]
        return out


def apply_overrides(config, overrides):
    """``{"config.a.b": v}`` → ``config.a.b = v`` (one element of ``get_hyper``'s result applied to a config)."""
    for key, value in overrides.items():
        parts = key.split(".")
        node = config
        for part in parts[1 if parts[0] == "config" else 0:-1]:
            node = node[part]
        node[parts[-1]] = value
    return config
