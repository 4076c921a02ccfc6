"""Symmetric (peer-mapped) device memory and the NVLink P2P collectives built on it.

``torch.distributed`` is used for bootstrap only (exchanging 64-byte ``cudaIpc`` handles, barriers at set-up);
every byte of gradient/parameter traffic afterwards moves through the kernels in ``csrc/p2p.cu`` as direct
loads/stores on peer memory over NVLink 5 / NVSwitch — no NCCL call on the training hot path.

``SymmetricComm`` (one process per GPU) and ``EmulatedComm`` (all ranks' buffers on ONE GPU, all ranks played by
one launch — the single-GPU test double) share the kernel entry points.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ..ops import _lib
from ..ops._lib import call, f32, i32, i64, stream

_DT = {torch.bfloat16: ("<i2", 2, torch.int16), torch.float32: ("<f4", 4, torch.float32),
       torch.int32: ("<i4", 4, torch.int32), torch.uint8: ("|u1", 1, torch.uint8)}


class _DevBuf:
    """Exposes a raw device allocation through ``__cuda_array_interface__`` so torch can alias it zero-copy."""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2,
                                         "strides": None}


def tensor_from_ptr(ptr: int, n: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    typestr, _, carrier = _DT[dtype]
    t = torch.as_tensor(_DevBuf(ptr, n, typestr), device=device)
    return t.view(dtype) if carrier != dtype else t


def _ptr_array(ptrs: List[int]):
    return (ctypes.c_void_p * len(ptrs))(*[ctypes.c_void_p(p) for p in ptrs])


class _CommBase:
    world: int
    rank: int
    ctas_per_rank: int = 32

    def __init__(self):
        self.epoch = 0
        self._grad_ptrs: List[int] = []
        self._param_ptrs: List[int] = []
        self._flag_ptrs: List[int] = []

    def _next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch


class SymmetricComm(_CommBase):
    """One process per GPU.  All ranks must make the same sequence of ``alloc_tensor`` calls."""

    def __init__(self, process_group=None, ctas_per_rank: int = 96):
        super().__init__()
        import torch.distributed as dist
        self.dist, self.group = dist, process_group
        self.world, self.rank = dist.get_world_size(process_group), dist.get_rank(process_group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.ctas_per_rank = ctas_per_rank
        self._allocs = []          # (local ptr, peer ptr list)
        self._opened: List[int] = []
        nflag = _lib.lib().rt1_p2p_flag_words()
        self.flags, self._flag_ptrs = self._alloc(nflag * 4, torch.int32, nflag)
        self._small, self._small_ptrs = self._alloc(2 * 1024 * 4, torch.float32, 2 * 1024)   # scalar all-reduce pad

    def _alloc(self, nbytes: int, dtype, n):
        nbytes = (nbytes + 255) // 256 * 256
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        call("rt1_symm_alloc", i64(nbytes), ctypes.byref(ptr), handle)
        mine = (bytes(handle), self.device.index)
        gathered = [None] * self.world
        self.dist.all_gather_object(gathered, mine, group=self.group)
        ptrs = []
        for r, (h, _dev) in enumerate(gathered):
            if r == self.rank:
                ptrs.append(ptr.value)
            else:
                p = ctypes.c_void_p()
                hb = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                call("rt1_symm_open", hb, ctypes.byref(p))
                self._opened.append(p.value)
                ptrs.append(p.value)
        self._allocs.append(ptr.value)
        return tensor_from_ptr(ptr.value, n, dtype, self.device), ptrs

    def alloc_tensor(self, n: int, dtype: torch.dtype, device=None) -> torch.Tensor:
        t, ptrs = self._alloc(n * _DT[dtype][1], dtype, n)
        t._rt1_peer_ptrs = ptrs
        return t

    def register_grad_buffer(self, t: torch.Tensor):
        self._grad_ptrs = t._rt1_peer_ptrs

    def register_param_buffer(self, t: torch.Tensor):
        self._param_ptrs = t._rt1_peer_ptrs

    def allreduce_adamw(self, lo: int, hi: int, flat, h: dict):
        one = lambda t: _ptr_array([t.data_ptr()] * self.world)     # noqa: E731  (only [rank] is read)
        call("rt1_p2p_allreduce_adamw", i32(self.world), i32(self.rank), _ptr_array(self._grad_ptrs),
             _ptr_array(self._param_ptrs), _ptr_array(self._flag_ptrs), one(flat.master), one(flat.exp_avg),
             one(flat.exp_avg_sq), i64(lo), i64(hi - lo), ctypes.c_uint(self._next_epoch()), f32(h["lr"]), f32(h["beta1"]),
             f32(h["beta2"]), f32(h["eps"]), f32(h["weight_decay"]), f32(h["bc1"]), f32(h["bc2"]),
             i32(self._ctas_for((hi - lo) * 2 // self.world)), i32(0), stream())

    def _ctas_for(self, bytes_per_rank: int) -> int:
        """Enough CTAs to keep ~2 MB of peer loads in flight (NVLink latency x bandwidth), never more than asked."""
        return max(1, min(self.ctas_per_rank, bytes_per_rank // 16384))

    def allreduce_(self, t: torch.Tensor, scale: float = 1.0, two_shot: Optional[bool] = None, n: Optional[int] = None):
        """In-place sum over ranks of a tensor allocated by :meth:`alloc_tensor` (bf16 or fp32)."""
        n = t.numel() if n is None else n
        nbytes = n * t.element_size()
        if two_shot is None:
            two_shot = nbytes >= 256 * 1024
        call("rt1_p2p_allreduce", i32(self.world), i32(self.rank), _ptr_array(t._rt1_peer_ptrs),
             _ptr_array(self._flag_ptrs), i64(0), i64(n), i32(t.dtype == torch.bfloat16), f32(scale), i32(two_shot),
             ctypes.c_uint(self._next_epoch()),
             i32(self._ctas_for(nbytes // self.world if two_shot else nbytes)), i32(0), stream())
        return t[:n] if two_shot else t[n:2 * n]

    def allreduce_small(self, x: torch.Tensor) -> torch.Tensor:
        """Sum of a small fp32 tensor (≤ 1024 elements) over ranks via the one-shot kernel (metrics, C4)."""
        n = (x.numel() + 3) // 4 * 4
        assert n <= 1024
        self._small[:x.numel()].copy_(x.reshape(-1))
        if n > x.numel():
            self._small[x.numel():n].zero_()
        call("rt1_p2p_allreduce", i32(self.world), i32(self.rank), _ptr_array(self._small_ptrs),
             _ptr_array(self._flag_ptrs), i64(0), i64(n), i32(0), f32(1.0), i32(0), ctypes.c_uint(self._next_epoch()),
             i32(1), i32(0), stream())
        return self._small[n:n + x.numel()].clone().view(x.shape)

    def close(self):
        torch.cuda.synchronize()
        try:
            self.dist.barrier(group=self.group)
        except Exception:
            pass
        for p in self._opened:
            try:
                call("rt1_symm_close", ctypes.c_void_p(p))
            except Exception:
                pass
        self._opened = []


class EmulatedComm(_CommBase):
    """All ``world`` ranks on one GPU (single launch, rank = blockIdx.y).  For tests of the multi-rank logic."""

    def __init__(self, world: int, device=None, ctas_per_rank: int = 4):
        super().__init__()
        self.world, self.rank = world, 0
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.ctas_per_rank = ctas_per_rank
        nflag = _lib.lib().rt1_p2p_flag_words()
        self.flags = [torch.zeros(nflag, dtype=torch.int32, device=self.device) for _ in range(world)]
        self._flag_ptrs = [f.data_ptr() for f in self.flags]

    def allreduce_adamw(self, grads, params, masters, ms, vs, lo: int, hi: int, h: dict):
        call("rt1_p2p_allreduce_adamw", i32(self.world), i32(0), _ptr_array([g.data_ptr() for g in grads]),
             _ptr_array([p.data_ptr() for p in params]), _ptr_array(self._flag_ptrs),
             _ptr_array([t.data_ptr() for t in masters]), _ptr_array([t.data_ptr() for t in ms]),
             _ptr_array([t.data_ptr() for t in vs]), i64(lo), i64(hi - lo), ctypes.c_uint(self._next_epoch()),
             f32(h["lr"]), f32(h["beta1"]), f32(h["beta2"]), f32(h["eps"]), f32(h["weight_decay"]), f32(h["bc1"]),
             f32(h["bc2"]), i32(self.ctas_per_rank), i32(1), stream())

    def allreduce_(self, bufs, n: int, scale: float = 1.0, two_shot: bool = True):
        call("rt1_p2p_allreduce", i32(self.world), i32(0), _ptr_array([b.data_ptr() for b in bufs]),
             _ptr_array(self._flag_ptrs), i64(0), i64(n), i32(bufs[0].dtype == torch.bfloat16), f32(scale), i32(two_shot),
             ctypes.c_uint(self._next_epoch()), i32(self.ctas_per_rank), i32(1), stream())
