"""Worker for the world_size=2 data-parallel plumbing test (launched by torchrun; gloo on CPU, NCCL on GPU)."""
import os
import sys
from argparse import Namespace

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rt1b200.data.dataset import SyntheticBatchSource  # noqa: E402
from rt1b200.models.factory import RT1Config  # noqa: E402
from rt1b200.ops import reference  # noqa: E402
from rt1b200.train.trainer import RT1Module, Trainer  # noqa: E402


def main():
    backend = sys.argv[1] if len(sys.argv) > 1 else "torchdist"
    use_cuda = torch.cuda.is_available() and backend != "cpu"
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
    if use_cuda:
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        device = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=device)
    else:
        device = torch.device("cpu")
        dist.init_process_group("gloo")
    torch.manual_seed(1234 + rank)          # different init per rank: the engine must broadcast rank 0's
    args = Namespace(height=64, width=96, seq_len=3, lr=1e-3, milestones=[], exp_name="dp", log_dir="/tmp/rt1_dp/logs",
                     ckpt_dir="/tmp/rt1_dp/ckpt", log_every_n_steps=1000, ckpt_every_n_epochs=1, max_epochs=1)
    cfg = RT1Config(height=64, width=96, seq_len=3, num_layers=1, layer_size=16, num_heads=2, dropout_rate=0.0)
    module = RT1Module(args, cfg)
    p2p = backend == "p2p"          # the fused P2P path is the bf16 path: bf16 model + gradients, sharded fp32 state
    if p2p:
        cfg = RT1Config.language_table(height=64, width=96, seq_len=3, num_layers=1, dropout_rate=0.0)
        module = RT1Module(args, cfg)
    tr = Trainer(module, args, device=device, backend="torchdist" if backend == "cpu" else backend,
                 param_dtype=torch.bfloat16 if p2p else torch.float32, loggers=[], verbose=False)
    eng = tr.engine
    # 1. parameters were broadcast
    flat0 = eng.flat.master.clone()
    gathered = [torch.empty_like(flat0) for _ in range(world)]
    dist.all_gather(gathered, flat0)
    assert all(torch.equal(gathered[0], g) for g in gathered), "parameter broadcast failed"

    src = SyntheticBatchSource(2, 3, 64, 96, pool=3, seed=10 + rank, image_dtype=torch.float32, pin=False)
    for it in range(3):
        batch = next(src)
        # oracle: local grads → all-gather → mean → reference Adam on a copy
        p_before = eng.flat.master.clone(); m_before = eng.flat.exp_avg.clone(); v_before = eng.flat.exp_avg_sq.clone()
        torch.manual_seed(50 + it)
        from rt1b200.data.dataset import to_device
        module.train()
        loss = module.training_step(to_device(batch, device))
        eng.flat.grad.zero_()
        loss.backward()
        g_local = eng.flat.grad.clone().float()
        gl = [torch.empty_like(g_local) for _ in range(world)]
        dist.all_gather(gl, g_local)
        g_mean = torch.stack(gl).sum(0) / world
        reference.adamw_step(p_before, g_mean, m_before, v_before, 1e-3, 0.9, 0.999, 1e-8, 0.0, eng.step_count + 1)
        eng.step(1e-3)
        if p2p:
            eng.gather_sharded_state()      # every rank updates only the shard it owns
            torch.cuda.synchronize()
        torch.testing.assert_close(eng.flat.master, p_before, rtol=1e-4 if p2p else 1e-5, atol=1e-6 if p2p else 1e-7)
        assert float(eng.flat.grad.abs().max()) == 0.0, "gradients must be zeroed by the step"
        assert eng.flat.params[0].grad.data_ptr() == eng.flat.grad.data_ptr()
    # 2. ranks still agree bit-for-bit
    if p2p:
        pb = [torch.empty_like(eng.flat.param) for _ in range(world)]
        dist.all_gather(pb, eng.flat.param.contiguous())           # NCCL moves bf16; the comparison below is bit-wise
        assert all(torch.equal(pb[0].view(torch.int16), q.view(torch.int16)) for q in pb), "bf16 parameters diverged"
    gathered = [torch.empty_like(flat0) for _ in range(world)]
    dist.all_gather(gathered, eng.flat.master.contiguous())
    assert all(torch.equal(gathered[0], g) for g in gathered), "ranks diverged"
    # 3. full train_step path (hooks/overlap) and metric averaging
    l = tr.train_step(next(src))
    m = eng.all_reduce_mean(l.reshape(1).float())
    ls = [torch.empty_like(m) for _ in range(world)]
    dist.all_gather(ls, l.reshape(1).float())
    torch.testing.assert_close(m, torch.stack(ls).mean(0))
    # 4. BatchNorm running statistics are per-rank while training (different data per rank) and rank 0's after sync_buffers()
    def bn_stats():
        return torch.cat([b.detach().float().reshape(-1) for n_, b in module.model.named_buffers() if "running_" in n_])
    mine = bn_stats()
    both = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(both, mine)
    assert not torch.equal(both[0], both[1]), "ranks saw different batches: their running statistics should differ"
    eng.sync_buffers(0)
    after = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(after, bn_stats())
    assert all(torch.equal(both[0], a) for a in after), "sync_buffers must make rank 0's statistics everyone's"
    if rank == 0:
        print("DP_WORKER_OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
