"""Test-harness stand-in: see README.md.  Connection modes and the one visualiser flag the reference's tests name."""
DIRECT, GUI, SHARED_MEMORY = 2, 1, 3
COV_ENABLE_GUI = 1
