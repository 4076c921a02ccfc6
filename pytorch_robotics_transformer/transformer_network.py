from rt1b200.models.transformer_network import TransformerNetwork  # noqa: F401
