"""One fused all-reduce + AdamW exchange with all 8 "ranks" on ONE GPU (rank = blockIdx.y of a single launch, peers = other
buffers of the same GPU) — the only form in which `ncu --set full` can look at p2p_allreduce_adamw_kernel<8, false> on this
box (ncu is limited to one-GPU commands).  It shows the SM side of the kernel (registers, issue mix, stalls, achieved local
bandwidth: every byte that would cross NVLink is an HBM/L2 access here); link-side numbers come from the 8-rank sweep
(tests/_p2p_worker.py --sweep, timed with CUDA events)."""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch  # noqa: E402
from rt1b200.parallel.symm import EmulatedComm  # noqa: E402

dev = torch.device("cuda", 0)
world, n = 8, 35_324_352
comm = EmulatedComm(world, ctas_per_rank=18)                 # 8 x 18 = 144 CTAs of 512 threads: all co-resident on 148 SMs
grads = [(torch.randn(n, device=dev) * 0.1).to(torch.bfloat16) for _ in range(world)]
params = [torch.zeros(n, dtype=torch.bfloat16, device=dev) for _ in range(world)]
masters = [torch.zeros(n, device=dev) for _ in range(world)]
ms = [torch.zeros(n, device=dev) for _ in range(world)]
vs = [torch.zeros(n, device=dev) for _ in range(world)]
h = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, bc1=0.1, bc2=0.001)
for _ in range(2):
    comm.allreduce_adamw(grads, params, masters, ms, vs, 0, n, h)
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
t0.record()
comm.allreduce_adamw(grads, params, masters, ms, vs, 0, n, h)
t1.record()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"emulated 8-rank exchange of {n} parameters on one GPU: {t0.elapsed_time(t1):.3f} ms")
