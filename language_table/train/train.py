"""Training loop of the LAVA baseline (reference ``language_table/train/train.py:60-266``): build the dataset and its
normalisation statistics, the model and the BC agent, restore the latest checkpoint from ``workdir`` if there is one, then
``num_train_steps`` steps in iterations of ``num_steps_per_train_iter`` (``multi_train_step``) with periodic loss logging
(``metrics.csv``: loss mean / std and learning rate per logging window) and checkpoints (BatchNorm statistics averaged over
the ranks first, ``merge_batch_stats``).  Rank 0 also writes ``parameters.txt`` (the parameter overview) and
``hparams.json``.  One process per device; with an initialised ``torch.distributed`` group gradients are averaged across
ranks (the reference's pmap / pmean)."""
import glob
import itertools
import json
import os
import time
from typing import Dict, Iterable, Optional

import torch

from . import bc, input_pipeline_rlds
from .networks import lava, pixel


def create_model(config, action_size: int = 2):
    if config.model_name == "sequence_lav_mse":
        return lava.SequenceLAVMSE(action_size=action_size, **dict(config.model))
    if config.model_name == "pixel_lang_mse":
        return pixel.PixelLangMSE(action_size=action_size, dense_resnet_width=config.model.dense_resnet_width,
                                  dense_resnet_num_blocks=config.model.dense_resnet_num_blocks,
                                  sequence_length=config.sequence_length)
    raise NotImplementedError(config.model_name)


def create_agent(config, model, norm_info) -> bc.BCAgent:
    if config.agent_name != "bc":
        raise NotImplementedError(config.agent_name)
    return bc.BCAgent(model=model, sequence_length=config.sequence_length, learning_rate=config.agent.learning_rate,
                      observation_statistics=norm_info["observation_statistics"], action_statistics=norm_info["action_statistics"],
                      action_min=norm_info["action_min"], action_max=norm_info["action_max"],
                      pretrained_checkpoints=config.agent.get("pretrained_checkpoints", []), freeze_keys=config.agent.freeze_keys)


def multi_train_step(agent: bc.BCAgent, batches: Iterable[Dict], initial_metrics: Optional[bc.TrainMetrics] = None,
                     on_step=None) -> bc.TrainMetrics:
    """One training iteration = one optimiser step per batch of ``batches`` (``num_steps_per_train_iter`` of them), the
    metrics of the steps merged (reference ``train.py:40-66``, a ``fori_loop`` over the leading axis of a stacked batch;
    here the agent holds the state, so only the metrics are returned).  ``on_step(step, out)`` runs after each step with the
    agent's step counter and its ``{"loss", "learning_rate"}``."""
    metrics = initial_metrics or bc.TrainMetrics.empty()
    for batch in batches:
        out = agent.train(batch)
        metrics = metrics.merge(bc.TrainMetrics.single_from_model_output(loss=out["loss"], learning_rate=out["learning_rate"]))
        if on_step is not None:
            on_step(agent.step, out)
    return metrics


def merge_batch_stats(agent_or_model, group=None):
    """Average the floating-point BatchNorm buffers over the ranks, so the checkpoint rank 0 writes carries everyone's
    statistics and not only its own (reference ``train.py:257-266``, a ``pmean`` over the replicas).  One flat
    all-reduce; a single process (or a model without such buffers) is left as it is."""
    import torch.distributed as dist
    model = getattr(agent_or_model, "model", agent_or_model)
    bufs = [b for b in model.buffers() if b.is_floating_point()]
    if not bufs or not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return agent_or_model
    flat = torch.cat([b.detach().reshape(-1).float() for b in bufs])
    dist.all_reduce(flat, group=group)
    flat /= dist.get_world_size(group)
    at = 0
    with torch.no_grad():
        for b in bufs:
            b.copy_(flat[at:at + b.numel()].view_as(b))
            at += b.numel()
    return agent_or_model


def parameter_overview(model) -> str:
    """Name / shape / size table of the parameters, with totals (the reference writes clu's overview to
    ``<workdir>/parameters.txt``, ``train.py:118-120``)."""
    rows = [(n, tuple(p.shape), p.numel(), p.requires_grad) for n, p in model.named_parameters()]
    w = max([len(r[0]) for r in rows] + [4])
    lines = [f"{'name':<{w}}  {'shape':<22} {'size':>12}  trainable"]
    lines += [f"{n:<{w}}  {str(s):<22} {k:>12,}  {'yes' if t else 'frozen'}" for n, s, k, t in rows]
    total, train = sum(r[2] for r in rows), sum(r[2] for r in rows if r[3])
    return "\n".join(lines + [f"total: {total:,} parameters, {train:,} trainable"]) + "\n"


def latest_checkpoint(workdir: str):
    files = sorted(glob.glob(os.path.join(workdir, "checkpoints", "ckpt_*.pt")), key=lambda f: int(f.split("_")[-1][:-3]))
    return files[-1] if files else None


def train(config, workdir: str, device=None, verbose: bool = True) -> Dict:
    torch.manual_seed(config.seed)
    device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    import torch.distributed as dist
    rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
    # normalisation statistics: computed (or re-loaded) by process 0, read from <workdir>/norm_info by everyone else
    ds, norm_info = input_pipeline_rlds.create_datasets(config, seed=config.seed, normalization_path=os.path.join(workdir, "norm_info"),
                                                        process_index=rank)
    verbose = verbose and rank == 0
    model = create_model(config).to(device)
    agent = create_agent(config, model, norm_info)
    os.makedirs(os.path.join(workdir, "checkpoints"), exist_ok=True)
    ck = latest_checkpoint(workdir)
    if ck:
        agent.load_state_dict(torch.load(ck, map_location=device, weights_only=False))
    if rank == 0:
        with open(os.path.join(workdir, "parameters.txt"), "w") as f:
            f.write(parameter_overview(model))
        if agent.step == 0:
            with open(os.path.join(workdir, "hparams.json"), "w") as f:
                json.dump(config, f, indent=1, default=str)
    a_mean = torch.as_tensor(norm_info["action_statistics"]["mean"], device=device)
    a_std = torch.as_tensor(norm_info["action_statistics"]["std"], device=device)
    per_iter = max(1, int(config.get("num_steps_per_train_iter", 1)))
    losses, t0, window = [], time.time(), bc.TrainMetrics.empty()
    metrics_file = os.path.join(workdir, "metrics.csv")
    if rank == 0 and not os.path.exists(metrics_file):
        with open(metrics_file, "w") as f:
            f.write("step,loss,loss_std,learning_rate,steps_per_sec\n")

    def stream():
        while True:                                           # epochs back to back; the step limit ends the loop
            for batch in ds.batches(config.per_device_batch_size):
                yield {"observation": {k: v.to(device) for k, v in batch["observation"].items()},
                       "action": (batch["action"].to(device) - a_mean) / a_std}

    def after_step(step, out):
        nonlocal n_window
        n_window += 1
        losses.append(out["loss"])
        if step % config.checkpoint_every_steps == 0 or step >= config.num_train_steps:
            merge_batch_stats(agent)                          # a collective: every rank, not only the one that saves
            if rank == 0:
                torch.save(agent.state_dict(), os.path.join(workdir, "checkpoints", f"ckpt_{step}.pt"))

    batches, t_window, n_window = stream(), time.time(), 0
    while agent.step < config.num_train_steps:
        n = min(per_iter, config.num_train_steps - agent.step)
        update = multi_train_step(agent, itertools.islice(batches, n), on_step=after_step)
        window = window.merge(update)
        # the logging period is checked per iteration, like the reference's `step % log_loss_every_steps`
        if agent.step % config.log_loss_every_steps < n or agent.step >= config.num_train_steps:
            vals, rate = window.compute(), n_window / max(time.time() - t_window, 1e-9)
            if rank == 0:
                with open(metrics_file, "a") as f:
                    f.write(f"{agent.step},{vals['loss']:.8g},{vals['loss_std']:.8g},{vals['learning_rate']:.8g},{rate:.4g}\n")
            if verbose:
                print(f"[lava] step {agent.step} loss {vals['loss']:.5f} (std {vals['loss_std']:.5f}) "
                      f"{rate:.2f} steps/s, {(time.time() - t0):.0f} s", flush=True)
            window, t_window, n_window = bc.TrainMetrics.empty(), time.time(), 0
    return {"steps": agent.step, "losses": losses, "agent": agent}
