"""Import shim: the reference imports skimage.data at module top of preprocessors.py but only uses it under __main__."""
from . import data  # noqa: F401
