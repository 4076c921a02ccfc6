"""Test-harness stand-in: see ../README.md."""
