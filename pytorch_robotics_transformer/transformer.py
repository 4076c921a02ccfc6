from rt1b200.models.transformer import TF_MultiHeadAttention, Transformer, _TransformerLayer, attention  # noqa: F401
