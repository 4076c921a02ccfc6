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
