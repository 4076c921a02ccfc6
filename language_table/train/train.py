"""Training loop of the LAVA baseline (reference ``language_table/train/train.py:60-266``): build the dataset and its
normalisation statistics, the model and the BC agent, restore the latest checkpoint from ``workdir`` if there is one, then
``num_train_steps`` steps with periodic loss logging and checkpoints.  One process per device; with an initialised
``torch.distributed`` group gradients are averaged across ranks (the reference's pmap/pmean)."""
import glob
import os
import time
from typing import Dict

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
                      action_min=norm_info["action_min"], action_max=norm_info["action_max"], freeze_keys=config.agent.freeze_keys)


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
    a_mean = torch.as_tensor(norm_info["action_statistics"]["mean"], device=device)
    a_std = torch.as_tensor(norm_info["action_statistics"]["std"], device=device)
    losses, t0 = [], time.time()
    while agent.step < config.num_train_steps:
        for batch in ds.batches(config.per_device_batch_size):
            batch = {"observation": {k: v.to(device) for k, v in batch["observation"].items()},
                     "action": (batch["action"].to(device) - a_mean) / a_std}
            losses.append(agent.train(batch)["loss"])
            if verbose and agent.step % config.log_loss_every_steps == 0:
                print(f"[lava] step {agent.step} loss {sum(losses[-config.log_loss_every_steps:]) / min(len(losses), config.log_loss_every_steps):.5f} "
                      f"({(time.time() - t0) / max(1, len(losses)):.3f} s/step)", flush=True)
            if rank == 0 and (agent.step % config.checkpoint_every_steps == 0 or agent.step >= config.num_train_steps):
                torch.save(agent.state_dict(), os.path.join(workdir, "checkpoints", f"ckpt_{agent.step}.pt"))
            if agent.step >= config.num_train_steps:
                break
    return {"steps": agent.step, "losses": losses, "agent": agent}
