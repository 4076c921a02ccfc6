from rt1b200.models.tokenizers import RT1ImageTokenizer  # noqa: F401
