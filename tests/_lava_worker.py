"""Two-process LAVA training over gloo: every rank ends with the same parameters (averaged gradients), the BatchNorm
statistics are averaged before the checkpoint, only rank 0 writes files.  Launched by tests/test_lava.py."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from language_table.train import train as train_lib  # noqa: E402
from language_table.train.configs import language_table_resnet_sim_local  # noqa: E402


def main():
    workdir = sys.argv[1]
    dist.init_process_group("gloo")
    rank = dist.get_rank()
    torch.set_num_threads(2)
    cfg = language_table_resnet_sim_local.get_config()      # the ResNet pyramid has BatchNorm buffers to merge
    cfg.data_target_width, cfg.data_target_height, cfg.per_device_batch_size = 64, 36, 2
    cfg.model.dense_resnet_width, cfg.model.lava_d_model, cfg.model.lava_num_layers = 64, 32, 1
    cfg.model.lava_temporal_transformer_num_layers = 1
    cfg.agent.freeze_keys = ["text_encoder"]                # leave the ResNet trainable so its statistics move
    cfg.num_train_steps, cfg.checkpoint_every_steps, cfg.log_loss_every_steps = 3, 3, 1
    cfg.seed = 7
    res = train_lib.train(cfg, workdir, device="cpu", verbose=False)
    model = res["agent"].model
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    both = [torch.empty_like(flat) for _ in range(2)]
    dist.all_gather(both, flat)
    assert torch.allclose(both[0], both[1], atol=1e-6), "ranks diverged"
    stats = [b for b in model.buffers() if b.is_floating_point()]
    assert stats, "expected BatchNorm buffers"
    sflat = torch.cat([b.reshape(-1) for b in stats])
    sboth = [torch.empty_like(sflat) for _ in range(2)]
    dist.all_gather(sboth, sflat)
    assert torch.allclose(sboth[0], sboth[1], atol=1e-6), "batch statistics were not merged"
    dist.barrier()
    if rank == 0:
        assert sorted(os.listdir(os.path.join(workdir, "checkpoints"))) == ["ckpt_3.pt"]
        assert os.path.exists(os.path.join(workdir, "norm_info", "norm_info.npy"))
        print("LAVA_WORKER_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
