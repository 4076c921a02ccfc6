"""rt1b200 — a Blackwell-native (B200, sm_100a) framework for training and serving the RT-1 robotics
transformer.  Same capabilities and public surface as ``tanhuajie/Pytorch-RT1-for-Distributed-Training``;
different engine: PyTorch as the framework layer, hand-written CUDA kernels for the hot ops and a P2P
(NVLink/NVSwitch) fused gradient all-reduce + AdamW in place of NCCL DDP + a separate optimizer.
"""
__version__ = "0.1.0"

from . import spaces  # noqa: F401
