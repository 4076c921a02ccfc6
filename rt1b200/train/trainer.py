"""Training loop, RT-1 training module, checkpointing and logging.

What it covers of the reference (``distribute_train.py``):
  * ``RT1Module``       ≙ ``RT1_Lightning`` (``:19-180``): builds the Language-Table RT-1, ``training_step`` /
                          ``validation_step`` / ``test_step`` (``set_actions`` → forward → mean of the ``[b,t]``
                          actor loss), ``debug()``.
  * ``Trainer``         ≙ Lightning ``Trainer(strategy=DDPStrategy(find_unused_parameters=True))`` +
                          ``ModelCheckpoint`` + ``LearningRateMonitor`` + CSV/TensorBoard loggers (``:192-247``):
                          fit (train epoch → validation → MultiStepLR step → checkpoint) and test.
  * checkpoints          Lightning layout (``state_dict`` with ``model.`` prefix, ``optimizer_states``,
                          ``lr_schedulers``, ``epoch``, ``global_step``), file name
                          ``{epoch}-{eval_loss:.6f}-{train_loss_epoch:.6f}.ckpt`` + ``last.ckpt`` (``:214-220``),
                          and — unlike the reference, where resume is a commented hint (``:240``) — ``resume``.

The step itself is the B200 design: bf16 model, flat fp32 master/optimizer state, forward + backward captured as ONE CUDA
graph after two eager warm-up steps (:class:`GraphedTrainStep`), then one fused gradient exchange + AdamW
(:class:`rt1b200.parallel.engine.DataParallelEngine`; on the eager schedule it runs per bucket, overlapped with backward).
The captured step publishes its loss to pinned host memory right after its forward pass, so ``read_loss()`` returns while
the backward is still running and the host has the next step queued before the GPU runs dry; the loss scalar for logging is
averaged over ranks without leaving the device until a log interval is reached.
"""
from __future__ import annotations

import csv
import os
import time
from argparse import Namespace
from dataclasses import asdict
from bisect import bisect_right
from typing import Dict, Iterable, List, Optional

import torch
import torch.nn as nn

from ..models.factory import RT1Config, build_rt1
from ..models.tokenizers import batched_space_sampler, np_to_tensor
from ..parallel.engine import DataParallelEngine
from ..utils import trace


class RT1Module(nn.Module):
    """The RT-1 policy plus its train/val/test step bodies. ``args`` needs ``height, width, seq_len``."""

    def __init__(self, args, cfg: Optional[RT1Config] = None):
        super().__init__()
        self.args = args
        self.HEIGHT, self.WIDTH, self.TIME_SEQUENCE_LENGTH = args.height, args.width, args.seq_len
        cfg = cfg or RT1Config.language_table(height=args.height, width=args.width, seq_len=args.seq_len)
        self.cfg = cfg
        self.model = build_rt1(cfg)
        self.state_space = self.model._input_tensor_space
        self.action_space = self.model._output_tensor_space

    def configure_optimizers(self):
        """The reference's Lightning hook (``distribute_train.py:99-110``): plain ``torch.optim.Adam`` + ``MultiStepLR`` over this
        module's parameters, in Lightning's dict layout.  :class:`Trainer` does not call it — its optimizer is the fused
        reduce + AdamW over flat buffers — but code that drives the module with its own loop (or with Lightning) gets the
        reference's recipe."""
        optimizer = torch.optim.Adam(self.parameters(), lr=self.args.lr)
        scheduler = torch.optim.lr_scheduler.MultiStepLR(optimizer, list(getattr(self.args, "milestones", [])), gamma=0.1)
        return {"optimizer": optimizer, "interval": "epoch",
                "lr_scheduler": {"scheduler": scheduler, "monitor": "val_loss", "frequency": 1}}

    def loss_fn(self, x: torch.Tensor) -> torch.Tensor:
        return torch.mean(x.float())

    def _step(self, batch) -> torch.Tensor:
        obs, labels = batch["train_observation"], batch["action_label"]
        self.model.set_actions(labels)
        # the reference samples a throw-away random network_state on the CPU and ships it to the GPU every
        # step (distribute_train.py:64-65); training never reads it, so none is built here
        self.model(obs, None)
        return self.loss_fn(self.model.get_actor_loss())

    def training_step(self, batch, batch_idx: int = 0) -> torch.Tensor:
        return self._step(batch)

    @torch.no_grad()
    def validation_step(self, batch, batch_idx: int = 0) -> torch.Tensor:
        return self._step(batch)

    @torch.no_grad()
    def test_step(self, batch, batch_idx: int = 0) -> torch.Tensor:
        return self._step(batch)

    def debug(self, training: bool = True, device=None):
        """Constant-input forward, as the reference's ``RT1_Lightning.debug`` (``:126-180``)."""
        device = device or next(self.parameters()).device
        T, H, W = self.TIME_SEQUENCE_LENGTH, self.HEIGHT, self.WIDTH
        if training:
            B = 2
            obs = {"image": torch.full((B, T, 3, H, W), 0.5, device=device),
                   "natural_language_embedding": torch.full((B, T, 512), 1.0, device=device)}
            labels = {"terminate_episode": torch.ones(B, T, dtype=torch.long, device=device),
                      "action": torch.zeros(B, T, 2, device=device)}
            self.model.set_actions(labels)
            state = np_to_tensor(batched_space_sampler(self.model._state_space, B), device)
            pred, _ = self.model(obs, state)
            loss = self.model.get_actor_loss()
            return pred, loss, self.loss_fn(loss)
        obs = {"image": torch.full((1, 3, H, W), 0.5, device=device),
               "natural_language_embedding": torch.full((1, 512), 1.0, device=device)}
        state = self.model.initial_state(1, device=device)
        with torch.no_grad():
            pred, _ = self.model(obs, state)
        return pred


class MultiStepLR:
    """``lr * gamma ** (#milestones <= epoch)``, stepped once per epoch (reference ``:101``)."""

    def __init__(self, base_lr: float, milestones: Iterable[int], gamma: float = 0.1, last_epoch: int = 0):
        self.base_lr, self.milestones, self.gamma, self.last_epoch = base_lr, sorted(milestones), gamma, last_epoch

    def get_lr(self) -> float:
        return self.base_lr * self.gamma ** bisect_right(self.milestones, self.last_epoch)

    def step(self):
        self.last_epoch += 1

    def state_dict(self):
        return {"base_lrs": [self.base_lr], "milestones": {m: 1 for m in self.milestones}, "gamma": self.gamma,
                "last_epoch": self.last_epoch, "_last_lr": [self.get_lr()]}

    def load_state_dict(self, sd):
        self.last_epoch = sd.get("last_epoch", self.last_epoch)


class CSVLogger:
    def __init__(self, save_dir: str, name: str):
        self.dir = os.path.join(save_dir, name)
        os.makedirs(self.dir, exist_ok=True)
        self.path = os.path.join(self.dir, "metrics.csv")
        self._keys: List[str] = []
        self._rows: List[Dict] = []

    def log(self, metrics: Dict[str, float], step: int):
        row = {"step": step, **metrics}
        self._rows.append(row)
        new = [k for k in row if k not in self._keys]
        self._keys.extend(new)
        with open(self.path, "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=self._keys)
            w.writeheader()
            w.writerows(self._rows)


class TensorBoardLogger:
    def __init__(self, save_dir: str, name: str):
        self.writer = None
        try:
            from torch.utils.tensorboard import SummaryWriter
            self.writer = SummaryWriter(os.path.join(save_dir, name))
        except Exception:   # tensorboard is optional
            pass

    def log(self, metrics: Dict[str, float], step: int):
        if self.writer is not None:
            for k, v in metrics.items():
                self.writer.add_scalar(k, v, step)
            self.writer.flush()


def _tree_leaves(b):
    if isinstance(b, dict):
        for v in b.values():
            yield from _tree_leaves(v)
    else:
        yield b


def _tree_map(fn, b):
    return {k: _tree_map(fn, v) for k, v in b.items()} if isinstance(b, dict) else fn(b)


class GraphedTrainStep:
    """Forward + backward of one training step as ONE CUDA-graph replay.

    An RT-1 step is ~1 100 kernel launches of 5-300 µs; issued one by one from Python the host, not the GPU, sets the
    step time (measured: 28.9 ms per step for ~20 ms of GPU work).  The step is therefore captured once — over static
    input buffers, with the weight-gradient lane as a parallel branch of the graph — and replayed with a single launch.
    What varies from step to step lives in device memory and is advanced by nodes of the graph itself: the pad-shift
    crop offset (device RNG → ``rt1_preprocess_nhwc_dev``), the stochastic-depth masks (torch's graph-safe Philox
    state) and the dropout seed epoch (``rt1_set_seed_epoch``).  Gradients land in the engine's flat buffer; the fused
    gradient exchange + AdamW runs right after the replay (its learning rate, bias corrections and cross-GPU epoch
    flags change every step, so it stays outside the graph)."""

    def __init__(self, trainer: "Trainer", template_batch):
        from ..ops import kernels
        self.tr = trainer
        dev = trainer.device
        self.static = _tree_map(lambda t: torch.empty_like(t), template_batch)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        # the step's loss reaches the host from INSIDE the graph (right after the forward): pinned scalar + pinned step counter
        self.loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        self.flag_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.step_ctr = torch.zeros(1, dtype=torch.int32, device=dev)
        self.replays = 0                                          # value flag_host takes when the latest replay's loss is out
        self.seed_epoch = kernels.seed_epoch_tensor(dev)
        self.stream = torch.cuda.Stream(device=dev)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.native_launches = 0
        self._encoders = None

    def _body(self):
        tr = self.tr
        self.seed_epoch.add_(0x61C88647)                         # fresh dropout masks on every replay
        loss = tr.module.training_step(self.static, 0)
        self.loss.copy_(loss.detach().float())
        from ..ops._lib import call, ptr, stream
        call("rt1_publish_scalar", ptr(self.loss), ptr(self.loss_host), ptr(self.step_ctr), ptr(self.flag_host), stream())
        # torch.autograd.grad, not loss.backward(): AccumulateGrad nodes (kept alive by the engine's hooks) remember the
        # stream they were created on — the default stream — and synchronising with it would invalidate the capture.
        # The kernel wrappers write most parameter gradients straight into the flat buffer and hand autograd `None`;
        # the few that do flow through autograd (e.g. the position embedding) are added here.
        params = tr.engine.flat.params
        grads = torch.autograd.grad(loss, params, allow_unused=True)
        dst = [p.grad for p, g in zip(params, grads) if g is not None]
        if dst:
            torch._foreach_add_(dst, [g for g in grads if g is not None])

    def _fill(self, batch):
        for d, s in zip(_tree_leaves(self.static), _tree_leaves(batch)):
            d.copy_(s, non_blocking=True)

    def capture(self, batch):
        """Warm up once on the capture stream (lazy per-stream state: the weight-gradient lane, tensor-map caches, function
        attributes), then record."""
        import gc
        tr = self.tr
        dev = tr.device
        model = tr.module.model
        model._device_rng = True
        # Nothing may keep an autograd graph of an earlier (default-stream) step alive: its AccumulateGrad nodes would be
        # re-used by the captured step and synchronise the capture stream with the default stream, which is illegal
        # inside a capture.  Hooks, the stored loss and the stored aux tensors are such references.
        tr.engine.remove_ready_hooks()
        tr.engine.use_single_bucket()                             # no overlap with backward any more: one fused exchange
        model._loss = None
        model._aux_info = {}
        if hasattr(model, "_attention_scores"):
            model._attention_scores = None
        gc.collect()
        main = torch.cuda.current_stream(dev)
        self._fill(batch)
        self.stream.wait_stream(main)
        with torch.cuda.stream(self.stream):
            tr.engine.zero_grad()
            self._body()
            tr.engine.zero_grad()                                 # the warm-up's gradients are not a training step
        main.wait_stream(self.stream)
        torch.cuda.synchronize(dev)
        self._undo_counters()
        from ..ops import _lib
        self.graph = torch.cuda.CUDAGraph()
        n0 = _lib.launch_count()
        with torch.cuda.graph(self.graph, stream=self.stream):
            self._body()
        self.native_launches = _lib.launch_count() - n0          # native calls recorded into the graph (not executed yet)
        _lib.add_launches(-self.native_launches)
        self._undo_counters()
        torch.cuda.synchronize(dev)
        self.replays = int(self.step_ctr.item())                  # the warm-up pass counted once; replays continue from here

    def wait_loss(self, timeout_s: float = 60.0) -> float:
        """Block until the latest replay has published its loss (i.e. its forward pass is done) and return it."""
        import time
        want = self.replays & 0x7FFFFFFF
        t0 = None
        while (int(self.flag_host[0]) & 0x7FFFFFFF) != want:
            if t0 is None:
                t0 = time.monotonic()
            elif time.monotonic() - t0 > timeout_s:
                torch.cuda.synchronize(self.tr.device)
                raise RuntimeError("the captured step did not publish its loss")
        return float(self.loss_host[0])

    def _bn_encoders(self):
        if self._encoders is None:
            from ..models.film_efficientnet import EfficientNetEncoder
            self._encoders = [m for m in self.tr.module.modules() if isinstance(m, EfficientNetEncoder)]
        return self._encoders

    def _undo_counters(self):
        for m in self._bn_encoders():                             # warm-up / capture are not optimisation steps
            m._nbt_pending = max(0, getattr(m, "_nbt_pending", 0) - 1)

    def run(self, batch) -> torch.Tensor:
        from ..ops import _lib
        self._fill(batch)
        self.graph.replay()
        self.replays += 1
        _lib.add_launches(self.native_launches)
        for m in self._bn_encoders():                             # the Python-side step counter of the BatchNorm buffers
            m._nbt_pending = getattr(m, "_nbt_pending", 0) + 1
            m._bn_stats_epoch = getattr(m, "_bn_stats_epoch", 0) + 1     # running statistics moved: evaluation caches are stale
        return self.loss


class Trainer:
    def __init__(self, module: RT1Module, args: Namespace, device=None, backend: str = "auto",
                 param_dtype: Optional[torch.dtype] = None, loggers=None, verbose: bool = True,
                 graph: Optional[bool] = None):
        self.module, self.args = module, args
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        module.to(self.device)
        if param_dtype is None:
            param_dtype = torch.bfloat16 if self.device.type == "cuda" else torch.float32
        self.engine = DataParallelEngine(module.model, lr=args.lr, param_dtype=param_dtype, backend=backend,
                                         broadcast_buffers=bool(getattr(args, "broadcast_buffers", False)))
        self.sched = MultiStepLR(args.lr, getattr(args, "milestones", []), 0.1)
        self.rank, self.world = self.engine.rank, self.engine.world
        self.epoch = 0
        self.global_step = 0
        self.verbose = verbose and self.rank == 0
        self.loggers = loggers if loggers is not None else (self._default_loggers() if self.rank == 0 else [])
        self.ckpt_dir = os.path.join(args.ckpt_dir, args.exp_name)
        self.log_every = max(1, getattr(args, "log_every_n_steps", 50))
        self.last_metrics: Dict[str, float] = {}
        # CUDA-graph the step (bf16 kernel path only); RT1B200_TRAIN_GRAPH=0 or graph=False keeps the eager schedule
        if graph is None:
            graph = os.environ.get("RT1B200_TRAIN_GRAPH", "1") != "0"
        from ..ops import _lib as _native
        self._graph_on = bool(graph and self.device.type == "cuda" and param_dtype == torch.bfloat16 and _native.available())
        self._graphs: Dict[tuple, GraphedTrainStep] = {}
        self._graph_seen: Dict[tuple, int] = {}
        self.graph_warmup_steps = 2          # eager steps per batch signature before its graph is captured

    def _default_loggers(self):
        a = self.args
        return [CSVLogger(os.path.join(a.log_dir, "csv"), a.exp_name),
                TensorBoardLogger(os.path.join(a.log_dir, "tb"), a.exp_name)]

    def _log(self, metrics: Dict[str, float], step: int):
        self.last_metrics.update(metrics)
        for lg in self.loggers:
            lg.log(metrics, step)

    def _to_device(self, batch):
        from ..data.dataset import to_device
        pf = getattr(self, "_prefetched", None)
        if pf is not None and pf[0] is batch:           # this host batch was staged by prefetch(): just order the streams
            _, dev_batch, ev = pf
            self._prefetched = None
            torch.cuda.current_stream(self.device).wait_event(ev)
        else:
            dev_batch = to_device(batch, self.device)
        tf = getattr(self, "batch_transform", None)      # device-side input transform (e.g. data.gpu_transforms.GpuCropResize)
        return tf(dev_batch) if tf is not None else dev_batch

    def prefetch(self, batch):
        """Start the host→device copy of a (pinned) host batch on a copy stream while the current step computes; the next
        ``train_step(batch)`` with the SAME object picks the staged copy up instead of copying in line.  This is what a
        prefetching data loader does; ``fit()`` uses it with a one-batch look-ahead."""
        from ..data.dataset import to_device, tree_leaves
        if self.device.type != "cuda":
            return
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        with torch.cuda.stream(self._copy_stream):
            dev_batch = to_device(batch, self.device)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        for t in tree_leaves(dev_batch):
            t.record_stream(torch.cuda.current_stream(self.device))
        self._prefetched = (batch, dev_batch, ev)

    # ------------------------------------------------------------------------------------------ steps
    def train_step(self, batch) -> torch.Tensor:
        """One optimisation step; returns the (device-resident, rank-local) loss."""
        batch = self._to_device(batch)
        self.module.train()
        lr = self.sched.get_lr()
        g = self._graph_for(batch) if self._graph_on else None
        if g is not None:
            with trace.range_("train_step/graph replay (forward + backward)"):
                loss = g.run(batch)                   # forward + backward: one graph replay
            self._loss_graph = g                      # the graph itself publishes the loss to pinned memory after its forward
            with trace.range_("train_step/fused exchange + AdamW"):
                self.engine.step(lr)                  # fused gradient exchange + AdamW (clears the gradients)
            self.global_step += 1
            return loss
        with trace.range_("train_step/forward"):
            loss = self.module.training_step(batch, self.global_step)
        self._loss_graph = None
        self._stage_loss(loss)
        self.engine.begin_step(lr)
        with trace.range_("train_step/backward (+ overlapped exchange)"):
            loss.backward()
        with trace.range_("train_step/fused exchange + AdamW"):
            self.engine.step(lr)
        self.global_step += 1
        return loss.detach()

    def _graph_for(self, batch) -> Optional[GraphedTrainStep]:
        """The captured step for this batch signature, once ``graph_warmup_steps`` eager steps have run with it."""
        key = tuple((tuple(t.shape), t.dtype) for t in _tree_leaves(batch))
        g = self._graphs.get(key)
        if g is not None:
            return g
        n = self._graph_seen.get(key, 0)
        self._graph_seen[key] = n + 1
        if n < self.graph_warmup_steps:
            return None
        g = GraphedTrainStep(self, batch)
        try:
            g.capture(batch)
        except Exception as e:                         # keep training on the eager schedule rather than die in set-up
            self._graph_on = False
            try:
                if g.graph is not None:
                    g.graph.reset()
            except Exception:
                pass
            self.module.model._device_rng = False
            torch.cuda.synchronize(self.device)
            self.engine.zero_grad()
            if self.verbose or os.environ.get("RT1B200_TRAIN_GRAPH_STRICT", "0") == "1":
                print(f"[rt1b200] CUDA-graph capture of the training step failed ({type(e).__name__}: {e}); "
                      "continuing with the eager schedule", flush=True)
            if os.environ.get("RT1B200_TRAIN_GRAPH_STRICT", "0") == "1":
                raise
            return None
        self._graphs[key] = g
        return g

    def _stage_loss(self, loss: torch.Tensor):
        """Copy the step's loss to pinned host memory on a side stream as soon as the FORWARD has produced it, so that
        :meth:`read_loss` does not have to wait for the backward pass and the optimizer (a ``float(loss)`` on the main
        stream does, and leaves the GPU idle while the host queues the next step)."""
        if self.device.type != "cuda":
            self._loss_host = loss.detach().float().reshape(1)
            return
        if getattr(self, "_d2h_stream", None) is None:
            self._d2h_stream = torch.cuda.Stream(device=self.device)
            self._loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
            self._loss_ev = torch.cuda.Event()
        l32 = loss.detach().float().reshape(1)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._d2h_stream.wait_event(ev)
        l32.record_stream(self._d2h_stream)
        with torch.cuda.stream(self._d2h_stream):
            self._loss_host.copy_(l32, non_blocking=True)
            self._loss_ev.record(self._d2h_stream)

    def read_loss(self) -> float:
        """The last step's (rank-local) loss as a Python float: device→host read that waits for that step's forward only
        (captured step: the graph writes it to pinned memory itself and the host polls a step counter; eager step: a D2H
        copy on a side stream + event)."""
        if getattr(self, "_loss_graph", None) is not None:
            return self._loss_graph.wait_loss()
        if self.device.type == "cuda":
            self._loss_ev.synchronize()
        return float(self._loss_host[0])

    def _run_eval(self, loader, step_fn) -> float:
        # every rank evaluates with rank 0's BatchNorm running statistics, as under torch DDP's per-forward buffer broadcast
        self.engine.sync_buffers(0)
        self.module.eval()
        total = torch.zeros((), device=self.device)
        n = 0
        for i, batch in enumerate(loader):
            total += step_fn(self._to_device(batch), i).float()
            n += 1
        mean = total / max(n, 1)
        return float(self.engine.all_reduce_mean(mean.reshape(1)).item())

    def fit(self, train_loader, eval_loader=None, max_epochs: Optional[int] = None,
            max_steps: Optional[int] = None, resume: Optional[str] = None):
        if resume:
            self.load_checkpoint(resume)
        max_epochs = max_epochs if max_epochs is not None else self.args.max_epochs
        every = max(1, getattr(self.args, "ckpt_every_n_epochs", 1))
        while self.epoch < max_epochs:
            if hasattr(train_loader, "sampler") and hasattr(train_loader.sampler, "set_epoch"):
                train_loader.sampler.set_epoch(self.epoch)
            if hasattr(train_loader, "batch_sampler") and hasattr(train_loader.batch_sampler, "set_epoch"):
                train_loader.batch_sampler.set_epoch(self.epoch)
            t0 = time.time()
            acc = torch.zeros((), device=self.device)
            window = torch.zeros((), device=self.device)
            n = 0
            it = iter(train_loader)
            nxt = next(it, None)
            while nxt is not None:
                batch, nxt = nxt, None
                loss = self.train_step(batch).float()
                nxt = next(it, None)                      # the loader works while the GPU runs the step just queued
                if nxt is not None:
                    self.prefetch(nxt)
                acc += loss
                window += loss
                n += 1
                if self.global_step % self.log_every == 0:
                    m = float(self.engine.all_reduce_mean((window / self.log_every).reshape(1)).item())
                    window.zero_()
                    self._log({"train_loss_step": m, "lr": self.sched.get_lr()}, self.global_step)
                    if self.verbose:
                        print(f"[epoch {self.epoch} step {self.global_step}] train_loss={m:.6f} "
                              f"lr={self.sched.get_lr():.2e}", flush=True)
                if max_steps is not None and self.global_step >= max_steps:
                    break
            train_loss_epoch = float(self.engine.all_reduce_mean((acc / max(n, 1)).reshape(1)).item())
            metrics = {"train_loss_epoch": train_loss_epoch, "epoch": self.epoch, "lr-Adam": self.sched.get_lr(),
                       "epoch_time_s": time.time() - t0}
            if eval_loader is not None:
                metrics["eval_loss"] = self._run_eval(eval_loader, self.module.validation_step)
            self._log(metrics, self.global_step)
            if self.verbose:
                print(f"[epoch {self.epoch}] " + " ".join(f"{k}={v:.6g}" for k, v in metrics.items()), flush=True)
            self.sched.step()
            self.epoch += 1
            if self.epoch % every == 0:
                self.save_checkpoint(metrics)
            if max_steps is not None and self.global_step >= max_steps:
                break
        return self.last_metrics

    def test(self, test_loader) -> float:
        loss = self._run_eval(test_loader, self.module.test_step)
        self._log({"test_loss": loss}, self.global_step)
        if self.verbose:
            print(f"test_loss={loss:.6f}", flush=True)
        return loss

    # ------------------------------------------------------------------------------------ checkpoints
    def checkpoint_dict(self) -> dict:
        return {
            "epoch": self.epoch, "global_step": self.global_step,
            "pytorch-lightning_version": "2.3.3", "rt1b200_version": "0.1.0",
            "state_dict": self.engine.model_state_dict(prefix="model."),
            "optimizer_states": [self.engine.optimizer_state_dict()],
            "lr_schedulers": [self.sched.state_dict()],
            "loops": {}, "callbacks": {},
            "hyper_parameters": {**{k: v for k, v in vars(self.args).items() if isinstance(v, (int, float, str, list))},
                                 "rt1_config": asdict(self.module.cfg)},
        }

    def save_checkpoint(self, metrics: Optional[Dict[str, float]] = None) -> Optional[str]:
        self.engine.gather_sharded_state()      # collective: every rank contributes the fp32 shards it owns
        if self.rank != 0:
            return None
        os.makedirs(self.ckpt_dir, exist_ok=True)
        metrics = metrics or self.last_metrics
        name = "epoch={}-eval_loss={:.6f}-train_loss_epoch={:.6f}.ckpt".format(
            self.epoch - 1 if self.epoch > 0 else 0, metrics.get("eval_loss", float("nan")),
            metrics.get("train_loss_epoch", float("nan")))
        ck = self.checkpoint_dict()
        path = os.path.join(self.ckpt_dir, name)
        torch.save(ck, path)
        torch.save(ck, os.path.join(self.ckpt_dir, "last.ckpt"))
        return path

    def load_checkpoint(self, path: str, weights_only: bool = False):
        ck = torch.load(path, map_location="cpu", weights_only=False)
        self.engine.load_model_state_dict(ck["state_dict"], prefix="model.")
        if not weights_only:
            if ck.get("optimizer_states"):
                self.engine.load_optimizer_state_dict(ck["optimizer_states"][0])
            if ck.get("lr_schedulers"):
                self.sched.load_state_dict(ck["lr_schedulers"][0])
            self.epoch = ck.get("epoch", 0)
            self.global_step = ck.get("global_step", 0)
        return ck


def load_policy_from_checkpoint(path: str, args=None, device="cpu", dtype: Optional[torch.dtype] = None) -> RT1Module:
    """``RT1_Lightning.load_from_checkpoint(path, args=args)`` equivalent (reference ``main_rt1.py:75``)."""
    ck = torch.load(path, map_location="cpu", weights_only=False)
    hp = ck.get("hyper_parameters", {})
    if args is None:
        args = Namespace(height=hp.get("height", 256), width=hp.get("width", 456), seq_len=hp.get("seq_len", 6))
    cfg = RT1Config(**hp["rt1_config"]) if "rt1_config" in hp else None   # absent in reference-written files
    module = RT1Module(args, cfg)
    sd = {k[len("model."):]: v for k, v in ck["state_dict"].items() if k.startswith("model.")}
    module.model.load_state_dict(sd, strict=True)
    module.to(device)
    if dtype is not None:
        from ..parallel.engine import keep_bn_stats_fp32
        module.to(dtype)
        keep_bn_stats_fp32(module)
    return module.eval()
