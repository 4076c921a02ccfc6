"""Data-parallel training engine: gradient reduction + AdamW as one fused step per bucket.

What it replaces in the reference: Lightning ``DDPStrategy(find_unused_parameters=True)`` → torch DDP (25 MB
buckets, NCCL all-reduce, unused-parameter bitmap all-reduce, per-forward buffer broadcast) followed by a
separate ``torch.optim.Adam`` step (``distribute_train.py:99-110,231-238``; SURVEY §2.5 C1–C5, §2.6 K14).

Design (B200-first):
  * parameters/gradients/moments are flat (:mod:`rt1b200.parallel.flat`); a *bucket* is a contiguous slice;
  * backend ``p2p``  — gradients live in NVLink-visible symmetric memory; ONE kernel per bucket pulls the peers'
    slices over NVSwitch (two-shot: each rank reduces 1/W of the bucket, applies 1/W scaling + bf16→fp32 +
    AdamW on that shard, then pushes the updated bf16 parameters to every peer).  No NCCL call and no separate
    optimizer kernel (``rt1b200/csrc/p2p.cu``);
  * backend ``local`` — world size 1: the same fused AdamW kernel without the exchange;
  * backend ``torchdist`` — ``torch.distributed.all_reduce`` (gloo on CPU, NCCL on GPU) + the fused/reference
    update.  This is the plumbing oracle for CPU tests and the "NCCL baseline" arm of the all-reduce sweep.
  * buckets are launched from gradient-ready hooks on a side stream so the exchange overlaps the rest of
    backward; parameters that never receive a gradient (``_action_token_emb``) simply contribute zeros —
    the reason the reference needs ``find_unused_parameters=True`` does not exist here.
  * BatchNorm running statistics stay per-rank (the reference's DDP re-broadcasts rank 0's every forward;
    we save rank 0's in checkpoints, and ``sync_buffers()`` is available on demand) — SURVEY §8 A4.

Optimizer semantics: decoupled-weight-decay Adam with ``weight_decay=0`` by default, i.e. exactly
``torch.optim.Adam(lr, betas=(0.9, 0.999), eps=1e-8)`` of the reference recipe.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from ..ops import _lib, reference
from .flat import FlatParams


def _dist():
    import torch.distributed as dist
    return dist


def keep_bn_stats_fp32(model: nn.Module):
    """After ``model.to(bfloat16)``: running statistics go back to float32 (bf16 cannot hold a running mean)."""
    for m in model.modules():
        if isinstance(m, nn.modules.batchnorm._BatchNorm):
            if m.running_mean is not None and m.running_mean.dtype != torch.float32:
                m.running_mean = m.running_mean.float()
                m.running_var = m.running_var.float()


class DataParallelEngine:
    def __init__(self, model: nn.Module, lr: float = 5e-4, betas: Tuple[float, float] = (0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 0.0, param_dtype: Optional[torch.dtype] = None,
                 backend: str = "auto", bucket_mb: Optional[float] = None, process_group=None, overlap: bool = True,
                 broadcast_buffers: bool = False):
        self.model = model
        self.broadcast_buffers = broadcast_buffers   # re-broadcast rank 0's buffers after every step (see sync_buffers)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.step_count = 0
        self.group = process_group
        dist = _dist()
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if self.world > 1 else 0
        dev = next(model.parameters()).device
        self.device = dev
        eff_dtype = param_dtype if param_dtype is not None else next(model.parameters()).dtype
        # the fused P2P kernels exchange bf16 gradients and write bf16 parameters into peer-mapped buffers: they exist
        # for the bf16 model only (fp32 parameters alias the master buffer, which is not symmetric memory)
        p2p_ok = dev.type == "cuda" and _lib.available() and eff_dtype == torch.bfloat16
        if backend == "auto":
            if self.world == 1:
                backend = "local"
            elif p2p_ok:
                backend = "p2p"
            else:
                backend = "torchdist"
        elif backend == "p2p" and not p2p_ok:
            raise ValueError("backend='p2p' needs CUDA, the native library and bfloat16 parameters "
                             f"(device {dev}, param dtype {eff_dtype}); use backend='torchdist' for fp32 parameters")
        self.backend = backend
        self.comm = None
        grad_alloc = None
        if backend == "p2p":
            from .symm import SymmetricComm
            self.comm = SymmetricComm(process_group)
            grad_alloc = self.comm.alloc_tensor
        # No `model.to(param_dtype)` here: FlatParams builds the fp32 master from the parameters as they are (full
        # precision) and re-homes every parameter as a `param_dtype` view of its flat buffer; the only other floating
        # tensors of this model are BatchNorm running statistics, which stay fp32 on purpose.
        self.flat = FlatParams(model, param_dtype=param_dtype, device=dev, grad_alloc=grad_alloc)
        keep_bn_stats_fp32(model)
        elem = self.flat.grad.element_size()
        if bucket_mb is None:
            # Few, large buckets: every fused exchange pays two cross-GPU barriers (~15 us) and the whole gradient
            # message of this model moves in ~0.2 ms, so 2-4 buckets give all the overlap with backward there is.
            total_mb = self.flat.numel * elem / 2 ** 20
            bucket_mb = total_mb / min(4, max(1, math.ceil(total_mb / 48.0))) * 0.98
        self.buckets = self.flat.buckets(max(1, int(bucket_mb * 2 ** 20 / elem)))
        if backend == "p2p":
            self.comm.register_grad_buffer(self.flat.grad)
            self.comm.register_param_buffer(self.flat.param)
        if self.world > 1:
            self.broadcast_parameters()
        self.overlap = overlap and dev.type == "cuda"
        self._side_stream = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self._pending: List[int] = []
        self._install_ready_hooks()

    # ------------------------------------------------------------------------------------------ set-up
    def broadcast_parameters(self, src: int = 0):
        """Rank ``src``'s parameters and buffers become everyone's (DDP's constructor broadcast, C5)."""
        dist = _dist()
        dist.broadcast(self.flat.master, src, group=self.group)
        self.flat.sync_params_from_master()
        self.sync_buffers(src)

    def sync_buffers(self, src: int = 0):
        """Rank ``src``'s module buffers (BatchNorm running statistics and step counters) become everyone's — C3 of
        SURVEY §2.5.  torch DDP does this before EVERY forward (``broadcast_buffers=True``, the Lightning default the
        reference runs with, ``distribute_train.py:231-238``); since no rank ever reads another rank's running statistics
        during training, the observable effect is that evaluation and checkpoints use rank ``src``'s.  Here the ~160 buffers
        travel as ONE coalesced broadcast per dtype, (a) at construction, (b) before every evaluation pass and checkpoint
        (``Trainer`` calls this), and (c) after every optimisation step when ``broadcast_buffers=True`` is asked for."""
        if self.world == 1:
            return
        dist = _dist()
        by_dtype: Dict[torch.dtype, List[torch.Tensor]] = {}
        for b in self.model.buffers():
            if b.is_floating_point() or b.dtype in (torch.int64, torch.int32):
                by_dtype.setdefault(b.dtype, []).append(b)
        for bufs in by_dtype.values():
            flat = torch.cat([b.reshape(-1) for b in bufs])
            dist.broadcast(flat, src, group=self.group)
            if self.rank != src:
                torch._foreach_copy_(bufs, [c.view_as(b) for c, b in zip(flat.split([b.numel() for b in bufs]), bufs)])

    def _install_ready_hooks(self):
        """Count gradients as autograd finishes them; when a bucket is complete, reduce+update it right away."""
        self._bucket_of: Dict[int, int] = {}
        self._need: List[int] = [0] * len(self.buckets)
        bi = 0
        for s, p in zip(self.flat.slots, self.flat.params):
            while s.offset >= self.buckets[bi][1]:
                bi += 1
            self._bucket_of[id(p)] = bi
        self._ready_count = [0] * len(self.buckets)
        self._launched = [False] * len(self.buckets)
        self._hook_handles = []
        self._eager = False       # set by begin_step(); hooks are inert otherwise
        if not hasattr(torch.Tensor, "register_post_accumulate_grad_hook"):
            return
        self._expected = [0] * len(self.buckets)
        for p in self.flat.params:
            self._expected[self._bucket_of[id(p)]] += 1

            def hook(param, _self=self):
                if not _self._eager:
                    return
                b = _self._bucket_of[id(param)]
                _self._ready_count[b] += 1
                _self._maybe_launch_ready()
            self._hook_handles.append(p.register_post_accumulate_grad_hook(hook))
        # gradients written directly by the sm_100a backward kernels never pass through autograd's accumulate
        # step; the kernel wrappers announce them through this callback instead
        from ..ops import kernels as _kernels

        def on_ready(params, _self=self):
            if not _self._eager:
                return
            for prm in params:
                b = _self._bucket_of.get(id(prm))
                if b is not None:
                    _self._ready_count[b] += 1
            _self._maybe_launch_ready()
        if self.device.type == "cuda" and _lib.available():
            _kernels.GRAD_READY = on_ready

    def use_single_bucket(self):
        """One bucket spanning the whole flat buffer: ONE fused launch and two cross-GPU barriers per step instead of k
        launches and k + 1 barriers.  Right whenever the exchange does not overlap the backward pass (a captured step hands
        over all gradients at once); the gradient message of RT-1 moves in ~0.2 ms, barriers are ~17 us each."""
        self.buckets = [(0, self.flat.numel)]
        self._bucket_of = {id(p): 0 for p in self.flat.params}
        self._need = [0]
        self._expected = [len(self.flat.params)]
        self._ready_count = [0]
        self._launched = [False]

    def remove_ready_hooks(self):
        """Drop the per-parameter gradient-ready hooks (the bucket-by-bucket overlap with backward).  A captured training
        step does not use them, and a registered hook keeps the parameter's AccumulateGrad node — and the stream it was
        created on — alive, which would tie a later capture to the default stream."""
        for h in self._hook_handles:
            h.remove()
        self._hook_handles = []

    # --------------------------------------------------------------------------------------- stepping
    def zero_grad(self):
        self.flat.grad.zero_()
        self.flat.rebind_grads()

    def begin_step(self, lr: Optional[float] = None, never_ready: Tuple[str, ...] = ("_action_token_emb",)):
        """Arm the overlap path: call right before the step's ONE ``loss.backward()``.  (The kernel wrappers WRITE parameter
        gradients into the flat buffer rather than accumulating, so a step is exactly one backward pass; accumulate
        micro-batches by enlarging the batch.)  Parameters whose names match ``never_ready`` get no gradient in this
        model; they are pre-counted."""
        if not self.overlap:
            return
        self._step_lr = self.lr if lr is None else lr
        self._ready_count = [0] * len(self.buckets)
        self._launched = [False] * len(self.buckets)
        for s, p in zip(self.flat.slots, self.flat.params):
            if any(k in s.name for k in never_ready):
                self._ready_count[self._bucket_of[id(p)]] += 1
        self._eager = True
        self._next_bucket = len(self.buckets) - 1
        self._hyper = self._hyper_for_next_step(self._step_lr)
        self._side_stream.wait_stream(torch.cuda.current_stream(self.device))

    def _maybe_launch_ready(self):
        # buckets complete roughly from the last (decoder) to the first (stem); launch strictly in that
        # order so every rank issues the same sequence of exchanges
        while self._next_bucket >= 0 and self._ready_count[self._next_bucket] >= self._expected[self._next_bucket]:
            b = self._next_bucket
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self._side_stream.wait_event(ev)
            with torch.cuda.stream(self._side_stream):
                self._reduce_and_update(b, self._hyper)
            self._launched[b] = True
            self._next_bucket -= 1

    def _hyper_for_next_step(self, lr: float):
        t = self.step_count + 1
        b1, b2 = self.betas
        return dict(lr=lr, beta1=b1, beta2=b2, eps=self.eps, weight_decay=self.weight_decay,
                    bc1=1.0 - b1 ** t, bc2=1.0 - b2 ** t, step=t)

    def step(self, lr: Optional[float] = None):
        """Reduce every (remaining) bucket across ranks and apply AdamW; gradients are zeroed for the next step."""
        lr = self.lr if lr is None else lr
        if self._eager:
            hyper = self._hyper
            self._eager = False
            remaining = [b for b in range(len(self.buckets)) if not self._launched[b]]
            if remaining:
                self._side_stream.wait_stream(torch.cuda.current_stream(self.device))
                with torch.cuda.stream(self._side_stream):
                    for b in reversed(remaining):
                        self._reduce_and_update(b, hyper)
            torch.cuda.current_stream(self.device).wait_stream(self._side_stream)
        else:
            hyper = self._hyper_for_next_step(lr)
            for b in reversed(range(len(self.buckets))):
                self._reduce_and_update(b, hyper)
        self.step_count += 1
        if self.broadcast_buffers:
            self.sync_buffers(0)

    def _reduce_and_update(self, b: int, h: dict):
        lo, hi = self.buckets[b]
        fl = self.flat
        if self.backend == "p2p":
            self.comm.allreduce_adamw(lo, hi, fl, h, final=(b == 0))   # buckets always run from the last to bucket 0
            return
        g = fl.grad[lo:hi]
        if self.backend == "torchdist" and self.world > 1:
            dist = _dist()
            if g.dtype == torch.bfloat16 and g.device.type == "cpu":
                g32 = g.float()
                dist.all_reduce(g32, group=self.group)
                g.copy_(g32)
            else:
                dist.all_reduce(g, group=self.group)
        scale = 1.0 / self.world
        if g.is_cuda and _lib.available():
            from ..ops import kernels
            kernels.fused_adamw(fl.param[lo:hi] if fl.param is not fl.master else None, fl.master[lo:hi], g,
                                fl.exp_avg[lo:hi], fl.exp_avg_sq[lo:hi], h, grad_scale=scale, zero_grad=True)
        else:
            reference.adamw_step(fl.master[lo:hi], g, fl.exp_avg[lo:hi], fl.exp_avg_sq[lo:hi], h["lr"],
                                 h["beta1"], h["beta2"], h["eps"], h["weight_decay"], h["step"], grad_scale=scale)
            if fl.param is not fl.master:
                fl.param[lo:hi].copy_(fl.master[lo:hi])
            g.zero_()

    # ------------------------------------------------------------------------------ scalars / metrics
    def all_reduce_mean(self, t: torch.Tensor) -> torch.Tensor:
        """Mean of a small tensor over ranks (the reference's ``self.log(..., sync_dist=True)``, C4)."""
        if self.world == 1:
            return t
        if self.backend == "p2p":
            # the small collectives share the flag pad and epoch counter of the fused buckets, which run on the side stream
            # between begin_step() and step(): a collective issued in that window would interleave epochs across streams
            assert not self._eager, "all_reduce_mean() must not be called between begin_step() and step()"
            return self.comm.allreduce_small(t.float()) / self.world
        dist = _dist()
        t = t.clone()
        dist.all_reduce(t, group=self.group)
        return t / self.world

    # ------------------------------------------------------------------------------------ checkpoints
    def _owned_ranges(self, rank: int) -> List[Tuple[int, int]]:
        """Element ranges of the flat buffers whose fp32 state rank ``rank`` updates under the sharded P2P step
        (mirrors ``shard_bounds`` in csrc/p2p.cu)."""
        out = []
        for lo, hi in self.buckets:
            n = hi - lo
            chunk = ((n + self.world - 1) // self.world + 7) & ~7
            a = min(n, chunk * rank)
            b = min(n, a + chunk)
            out.append((lo + a, lo + b))
        return out

    def gather_sharded_state(self):
        """P2P backend: every rank updates only its 1/W shard of master/exp_avg/exp_avg_sq; before a checkpoint
        (or any read of the full fp32 state) the owners' shards are combined on all ranks.  Rare, so it uses the
        bootstrap process group rather than a dedicated kernel."""
        if self.backend != "p2p" or self.world == 1 or self.step_count == 0:
            return
        dist = _dist()
        for buf in (self.flat.master, self.flat.exp_avg, self.flat.exp_avg_sq):
            tmp = torch.zeros_like(buf)
            for a, b in self._owned_ranges(self.rank):
                tmp[a:b] = buf[a:b]
            dist.all_reduce(tmp, group=self.group)
            buf.copy_(tmp)

    def model_state_dict(self, prefix: str = "") -> Dict[str, torch.Tensor]:
        """fp32 state dict with the model's own keys (master weights substituted for bf16 parameters).
        With the P2P backend call :meth:`gather_sharded_state` on ALL ranks first (it is a collective)."""
        if self.device.type == "cuda" and _lib.available():
            from ..ops.kernels_cnn import flush_bn_counters
            flush_bn_counters(self.model)
        sd = {}
        masters = {s.name: self.flat.view(self.flat.master, s.name) for s in self.flat.slots}
        for k, v in self.model.state_dict().items():
            src = masters.get(k, v)
            sd[prefix + k] = src.detach().to("cpu", torch.float32 if src.is_floating_point() else src.dtype).clone()
        return sd

    def load_model_state_dict(self, sd: Dict[str, torch.Tensor], prefix: str = "", strict: bool = True):
        own = self.model.state_dict()
        stripped = {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}
        missing = [k for k in own if k not in stripped]
        unexpected = [k for k in stripped if k not in own]
        if strict and (missing or unexpected):
            raise KeyError(f"state_dict mismatch: missing {missing[:5]} unexpected {unexpected[:5]}")
        with torch.no_grad():
            for k, v in stripped.items():
                if k not in own:
                    continue
                if k in self.flat.by_name:
                    self.flat.view(self.flat.master, k).copy_(v.to(self.device, torch.float32))
                else:
                    own[k].copy_(v.to(own[k].device, own[k].dtype))
        self.flat.sync_params_from_master()

    def optimizer_state_dict(self) -> dict:
        """``torch.optim.Adam``-layout state (what Lightning stores under ``optimizer_states[0]``)."""
        state = {}
        for i, s in enumerate(self.flat.slots):
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self.flat.view(self.flat.exp_avg, s.name).detach().cpu().clone(),
                        "exp_avg_sq": self.flat.view(self.flat.exp_avg_sq, s.name).detach().cpu().clone()}
        group = {"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay,
                 "amsgrad": False, "maximize": False, "foreach": None, "capturable": False,
                 "differentiable": False, "fused": None, "params": list(range(len(self.flat.slots))),
                 "param_names": [s.name for s in self.flat.slots]}
        return {"state": state, "param_groups": [group]}

    def load_optimizer_state_dict(self, osd: dict):
        group = osd["param_groups"][0]
        names = group.get("param_names")
        self.lr = group.get("lr", self.lr)
        for i, s in enumerate(self.flat.slots):
            key = i
            if names is not None and s.name in names:
                key = group["params"][names.index(s.name)]
            st = osd["state"].get(key)
            if st is None:
                continue
            self.flat.view(self.flat.exp_avg, s.name).copy_(st["exp_avg"].to(self.device))
            self.flat.view(self.flat.exp_avg_sq, s.name).copy_(st["exp_avg_sq"].to(self.device))
            self.step_count = int(st["step"]) if not torch.is_tensor(st["step"]) else int(st["step"].item())

    def close(self):
        for h in self._hook_handles:
            h.remove()
        if self.comm is not None:
            self.comm.close()
