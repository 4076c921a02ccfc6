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
        self._pending = None      # bucket whose clearing is deferred to the next fused call
        self._grad_ptrs: List[int] = []
        self._param_ptrs: List[int] = []
        self._flag_ptrs: List[int] = []

    def _next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch


class SymmetricComm(_CommBase):
    """One process per GPU.  All ranks must make the same sequence of ``alloc_tensor`` calls.

    Allocation back ends (``RT1_SYMM=torch|ipc``, default ``torch`` with fall-back to ``ipc``):
      * ``torch`` — ``torch.distributed._symmetric_memory`` (``cuMemCreate`` + fd exchange).  Used for the
        plumbing only; it also binds the allocation to an NVSwitch **multicast** object when the driver exposes
        one, which is what the NVLS kernels (``multimem.ld_reduce`` / ``multimem.st``) need;
      * ``ipc``   — ``cudaMalloc`` + ``cudaIpc`` handles exchanged through the bootstrap group (no multicast).
    Schedule (``RT1_P2P_ALGO=auto|p2p|nvls``): ``auto`` takes NVLS from 4 ranks up when a multicast mapping
    exists (bytes per direction N(1+1/W) vs 2N(W-1)/W for the P2P two-shot schedule)."""

    def __init__(self, process_group=None, ctas_per_rank: Optional[int] = None):
        super().__init__()
        import os
        import torch.distributed as dist
        self.dist, self.group = dist, process_group
        self.world, self.rank = dist.get_world_size(process_group), dist.get_rank(process_group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.ctas_per_rank = int(os.environ.get("RT1_P2P_CTAS", "148")) if ctas_per_rank is None else ctas_per_rank
        self._allocs = []          # cudaIpc: local ptrs;  torch: (tensor, handle) kept alive
        self._opened: List[int] = []
        self.allocator = os.environ.get("RT1_SYMM", "torch")
        self.algo = os.environ.get("RT1_P2P_ALGO", "auto")
        nflag = _lib.lib().rt1_p2p_flag_words()
        self.flags, self._flag_ptrs, _ = self._alloc(nflag * 4, torch.int32, nflag)
        self._small, self._small_ptrs, _ = self._alloc(2 * 1024 * 4, torch.float32, 2 * 1024)   # scalar all-reduce pad
        self._grad_mc = self._param_mc = 0

    @property
    def has_multicast(self) -> bool:
        return bool(self._grad_mc and self._param_mc)

    def use_nvls(self, mc_ptr: int) -> bool:
        if self.algo == "p2p" or not mc_ptr:
            return False
        return self.algo == "nvls" or self.world >= 4

    def _alloc_torch(self, dtype, n):
        import torch.distributed._symmetric_memory as symm_mem
        carrier = _DT[dtype][2]
        t = symm_mem.empty(n, dtype=carrier, device=self.device)
        t.zero_()
        group = self.group if self.group is not None else self.dist.group.WORLD
        hdl = symm_mem.rendezvous(t, group)
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        self._allocs.append((t, hdl))
        torch.cuda.synchronize()
        self.dist.barrier(group=self.group)
        return (t.view(dtype) if carrier != dtype else t), ptrs, mc

    def _alloc(self, nbytes: int, dtype, n):
        if self.allocator == "torch":
            ok, res = 1, None
            try:
                res = self._alloc_torch(dtype, n)
            except Exception as e:                                   # noqa: BLE001
                ok = 0
                self._alloc_error = f"{type(e).__name__}: {e}"
            flags = [None] * self.world
            self.dist.all_gather_object(flags, ok, group=self.group)
            if all(flags):
                return res
            self.allocator = "ipc"          # every rank falls back together
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
        return tensor_from_ptr(ptr.value, n, dtype, self.device), ptrs, 0

    def alloc_tensor(self, n: int, dtype: torch.dtype, device=None) -> torch.Tensor:
        t, ptrs, mc = self._alloc(n * _DT[dtype][1], dtype, n)
        t._rt1_peer_ptrs = ptrs
        t._rt1_mc_ptr = mc
        return t

    def register_grad_buffer(self, t: torch.Tensor):
        self._grad_ptrs, self._grad_mc = t._rt1_peer_ptrs, t._rt1_mc_ptr

    def register_param_buffer(self, t: torch.Tensor):
        self._param_ptrs, self._param_mc = t._rt1_peer_ptrs, t._rt1_mc_ptr

    def allreduce_adamw(self, lo: int, hi: int, flat, h: dict, final: bool = True):
        """Fused reduce + AdamW of bucket ``[lo, hi)``.  Buckets of one step are chained: pass ``final=False`` for all
        but the last one — the closing barrier and the clearing of a bucket then ride on the next call's opening
        barrier (k + 1 barriers per step instead of 2k)."""
        prev_lo, prev_n = self._pending if self._pending is not None else (0, 0)
        self._pending = None if final else (lo, hi - lo)
        one = lambda t: _ptr_array([t.data_ptr()] * self.world)     # noqa: E731  (only [rank] is read)
        nvls = self.has_multicast and self.use_nvls(self._grad_mc)
        call("rt1_p2p_allreduce_adamw", i32(self.world), i32(self.rank), _ptr_array(self._grad_ptrs),
             _ptr_array(self._param_ptrs), _ptr_array(self._flag_ptrs), one(flat.master), one(flat.exp_avg),
             one(flat.exp_avg_sq), ctypes.c_void_p(self._grad_mc if nvls else 0),
             ctypes.c_void_p(self._param_mc if nvls else 0), i64(lo), i64(hi - lo), i64(prev_lo), i64(prev_n),
             i32(1 if final else 0), ctypes.c_uint(self._next_epoch()), f32(h["lr"]), f32(h["beta1"]),
             f32(h["beta2"]), f32(h["eps"]), f32(h["weight_decay"]), f32(h["bc1"]), f32(h["bc2"]),
             i32(self._ctas_for((hi - lo) * 2 // self.world)), i32(0), stream())

    def _ctas_for(self, bytes_per_rank: int) -> int:
        """Enough CTAs to keep ~2 MB of peer loads in flight (NVLink latency x bandwidth), never more than asked."""
        return max(1, min(self.ctas_per_rank, bytes_per_rank // 8192))

    def allreduce_(self, t: torch.Tensor, scale: float = 1.0, two_shot: Optional[bool] = None, n: Optional[int] = None,
                   algo: Optional[str] = None):
        """In-place sum over ranks of a tensor allocated by :meth:`alloc_tensor` (bf16 or fp32).
        ``algo``: ``one_shot`` (result in ``t[n:2n]``), ``two_shot`` (P2P), ``nvls``; default by size and world."""
        n = t.numel() if n is None else n
        nbytes = n * t.element_size()
        if algo is None:
            if two_shot is None:
                two_shot = nbytes >= 256 * 1024
            algo = "one_shot" if not two_shot else ("nvls" if self.use_nvls(t._rt1_mc_ptr) else "two_shot")
        code = {"one_shot": 0, "two_shot": 1, "nvls": 2}[algo]
        call("rt1_p2p_allreduce", i32(self.world), i32(self.rank), _ptr_array(t._rt1_peer_ptrs),
             _ptr_array(self._flag_ptrs), ctypes.c_void_p(t._rt1_mc_ptr if code == 2 else 0), i64(0), i64(n),
             i32(t.dtype == torch.bfloat16), f32(scale), i32(code), ctypes.c_uint(self._next_epoch()),
             i32(self._ctas_for(nbytes // self.world if code else nbytes)), i32(0), stream())
        return t[:n] if code else t[n:2 * n]

    def allreduce_small(self, x: torch.Tensor) -> torch.Tensor:
        """Sum of a small fp32 tensor (≤ 1024 elements) over ranks via the one-shot kernel (metrics, C4)."""
        n = (x.numel() + 3) // 4 * 4
        assert n <= 1024
        self._small[:x.numel()].copy_(x.reshape(-1))
        if n > x.numel():
            self._small[x.numel():n].zero_()
        call("rt1_p2p_allreduce", i32(self.world), i32(self.rank), _ptr_array(self._small_ptrs),
             _ptr_array(self._flag_ptrs), ctypes.c_void_p(0), i64(0), i64(n), i32(0), f32(1.0), i32(0),
             ctypes.c_uint(self._next_epoch()), i32(1), i32(0), stream())
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
        self._allocs = [a for a in self._allocs if not isinstance(a, tuple)]


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
             _ptr_array([t.data_ptr() for t in vs]), ctypes.c_void_p(0), ctypes.c_void_p(0), i64(lo), i64(hi - lo),
             i64(0), i64(0), i32(1), ctypes.c_uint(self._next_epoch()),
             f32(h["lr"]), f32(h["beta1"]), f32(h["beta2"]), f32(h["eps"]), f32(h["weight_decay"]), f32(h["bc1"]),
             f32(h["bc2"]), i32(self.ctas_per_rank), i32(1), stream())

    def allreduce_(self, bufs, n: int, scale: float = 1.0, two_shot: bool = True):
        call("rt1_p2p_allreduce", i32(self.world), i32(0), _ptr_array([b.data_ptr() for b in bufs]),
             _ptr_array(self._flag_ptrs), ctypes.c_void_p(0), i64(0), i64(n), i32(bufs[0].dtype == torch.bfloat16),
             f32(scale), i32(1 if two_shot else 0), ctypes.c_uint(self._next_epoch()), i32(self.ctas_per_rank), i32(1),
             stream())
