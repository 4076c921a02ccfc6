"""Import shim: the reference imports ``gym.spaces`` only for space descriptors (Dict/Box/Discrete/MultiDiscrete).
gym is not installable offline; this stands in for it on the reference arm of the benchmark and in parity tests."""
from . import spaces  # noqa: F401
