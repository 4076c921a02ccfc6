"""Flat parameter / gradient / optimizer-state storage.

All parameters of a model are re-homed into ONE contiguous buffer (and their gradients into a second one)
so that

  * the gradient all-reduce and the AdamW update run as a handful of large fused launches over contiguous
    "buckets" instead of 572 small per-tensor kernels (the reference pays per-tensor ``foreach`` Adam plus
    DDP's 25 MB bucket copies);
  * the gradient buffer can live in NVLink-visible symmetric memory and be read by peer GPUs directly;
  * groups of tensors that one kernel consumes together (the q/k/v projections of an attention layer, all 54
    FiLM projection matrices of the image encoder) sit back to back and are used as ONE GEMM operand.

``FlatParams`` keeps ``nn.Parameter`` objects (and therefore ``state_dict`` keys, autograd, hooks) intact:
each ``p.data`` becomes a view into the flat parameter buffer and ``p.grad`` a view into the flat gradient
buffer.  With ``param_dtype=torch.bfloat16`` the model computes in bf16 while fp32 master weights and the
Adam moments live here (mixed precision with the master copy owned by the optimizer).

``flat_groups()`` protocol: a module may return ``{attr_name: [params...]}``; the listed parameters are packed
back to back (no alignment padding between members) and the module receives ``attr_name`` — a view over all of
them stacked along dim 0 — and ``attr_name + '_grad'``, the same view of the gradient buffer.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Tuple

import torch
import torch.nn as nn

ALIGN = 64  # elements; every tensor / group starts 128-byte aligned in bf16 (TMA/vector-load friendly)


@dataclass
class Slot:
    name: str
    offset: int
    numel: int
    shape: torch.Size


def _groups_of(mod) -> Dict[str, List[nn.Parameter]]:
    fg = getattr(mod, "flat_groups", None)
    if not callable(fg):
        return {}
    groups = fg()
    return groups if isinstance(groups, dict) else {}


def _ordered_params(model: nn.Module) -> List[Tuple[str, nn.Parameter, Optional[int]]]:
    """(name, param, group id) in registration order, except that members of a group are made adjacent."""
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    by_id = {id(p): n for n, p in named}
    grouped: Dict[int, List[nn.Parameter]] = {}
    for mod in model.modules():
        for group in _groups_of(mod).values():
            if all(id(p) in by_id and p.numel() % 8 == 0 for p in group):
                for p in group:
                    grouped[id(p)] = group
    out, done = [], set()
    for n, p in named:
        if id(p) in done:
            continue
        if id(p) in grouped:
            g = grouped[id(p)]
            for q in g:
                if id(q) not in done:
                    out.append((by_id[id(q)], q, id(g)))
                    done.add(id(q))
        else:
            out.append((n, p, None))
            done.add(id(p))
    return out


class FlatParams:
    def __init__(self, model: nn.Module, param_dtype: Optional[torch.dtype] = None, device=None,
                 grad_alloc: Optional[Callable[[int, torch.dtype, torch.device], torch.Tensor]] = None):
        params = _ordered_params(model)
        if not params:
            raise ValueError("model has no trainable parameters")
        first = params[0][1]
        self.device = torch.device(device) if device is not None else first.device
        self.param_dtype = param_dtype or first.dtype
        self.slots: List[Slot] = []
        self._cuts: List[int] = []        # aligned offsets where a bucket may end (never inside a packed group)
        off, prev_gid = 0, None
        for name, p, gid in params:
            if not (gid is not None and gid == prev_gid):
                off = (off + ALIGN - 1) // ALIGN * ALIGN
                self._cuts.append(off)
            self.slots.append(Slot(name, off, p.numel(), p.shape))
            off += p.numel()
            prev_gid = gid
        self.numel = (off + ALIGN - 1) // ALIGN * ALIGN
        self._cuts.append(self.numel)
        self.params: List[nn.Parameter] = [p for _, p, _ in params]
        self.by_name: Dict[str, Slot] = {s.name: s for s in self.slots}

        self.master = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
        # one multi-tensor copy (a handful of launches) instead of one conversion + one copy kernel per parameter
        dst = [self.master[s.offset:s.offset + s.numel] for s in self.slots]
        src = [p.detach().reshape(-1).to(self.device) for p in self.params]
        torch._foreach_copy_(dst, src)
        del dst, src
        if self.param_dtype == torch.float32:
            self.param = self.master                      # fp32 training: parameters ARE the master copy
        elif grad_alloc is not None:                       # peer-visible parameter buffer (P2P all-gather target)
            self.param = grad_alloc(self.numel, self.param_dtype, self.device)
            self.param.copy_(self.master)
        else:
            self.param = self.master.to(self.param_dtype)
        alloc = grad_alloc or (lambda n, dt, dev: torch.zeros(n, dtype=dt, device=dev))
        self.grad = alloc(self.numel, self.param_dtype, self.device)
        self.grad.zero_()
        self.exp_avg = torch.zeros_like(self.master)
        self.exp_avg_sq = torch.zeros_like(self.master)

        for s, p in zip(self.slots, self.params):
            p.data = self.param[s.offset:s.offset + s.numel].view(s.shape)
            p.grad = self.grad[s.offset:s.offset + s.numel].view(s.shape)
        self._install_fused_views(model)

    # -- fused operand views ---------------------------------------------------------------------------
    def _install_fused_views(self, model: nn.Module):
        ids = {id(p): s for s, p in zip(self.slots, self.params)}
        for mod in model.modules():
            for attr, g in _groups_of(mod).items():
                if not all(id(p) in ids for p in g):
                    continue
                contiguous = all(ids[id(g[i + 1])].offset == ids[id(g[i])].offset + g[i].numel()
                                 for i in range(len(g) - 1))
                if not contiguous:
                    continue
                s0 = ids[id(g[0])]
                total = sum(p.numel() for p in g)
                shape = (sum(p.shape[0] for p in g), *g[0].shape[1:])
                setattr(mod, attr, self.param[s0.offset:s0.offset + total].view(shape))
                setattr(mod, attr + "_grad", self.grad[s0.offset:s0.offset + total].view(shape))

    # -- helpers ---------------------------------------------------------------------------------------
    def rebind_grads(self):
        """Re-attach ``p.grad`` views (after something set them to ``None``)."""
        for s, p in zip(self.slots, self.params):
            if p.grad is None or p.grad.data_ptr() != self.grad.data_ptr() + s.offset * self.grad.element_size():
                p.grad = self.grad[s.offset:s.offset + s.numel].view(s.shape)

    def view(self, buf: torch.Tensor, name: str) -> torch.Tensor:
        s = self.by_name[name]
        return buf[s.offset:s.offset + s.numel].view(s.shape)

    def sync_master_from_params(self):
        if self.param is not self.master:
            self.master.copy_(self.param)

    def sync_params_from_master(self):
        if self.param is not self.master:
            self.param.copy_(self.master)

    def buckets(self, bucket_elems: int) -> List[Tuple[int, int]]:
        """Contiguous [start, end) element ranges cut at aligned tensor/group boundaries, each about
        ``bucket_elems`` long."""
        out, start = [], 0
        for cut in self._cuts[1:]:
            if cut - start >= bucket_elems:
                out.append((start, cut))
                start = cut
        if start < self.numel:
            out.append((start, self.numel))
        return out
