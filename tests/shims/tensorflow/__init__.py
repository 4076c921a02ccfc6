"""Test-harness stand-in: see ../README.md."""
from . import test  # noqa: F401
