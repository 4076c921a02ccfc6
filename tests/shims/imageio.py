"""Test-harness stand-in: imported by the reference's oracles/plot.py, never called when plotting is off."""
