from tensorflow import test  # noqa: F401
