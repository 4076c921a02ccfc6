from rt1b200.models.tokenizers import RT1ActionTokenizer  # noqa: F401
