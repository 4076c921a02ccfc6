"""Test-harness stand-in: see ../README.md."""
from . import io, test  # noqa: F401
