"""Where does the persistent GEMM spend its time?  Times one shape with parts of the kernel switched off
(RT1B200_GEMM_DBG bits: 1 no copy-out, 2 no statistics, 4 no TMEM drain/staging, 8 no A loads, 16 no B loads)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rt1b200.ops import kernels

def run(M, N, K, stats, dbg, b_mn=False, iters=10):
    os.environ["RT1B200_GEMM_DBG"] = str(dbg)
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = (torch.randn(K, N, device="cuda") if b_mn else torch.randn(N, K, device="cuda")).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    cs = torch.zeros(N, device="cuda") if stats else None
    cq = torch.zeros(N, device="cuda") if stats else None
    for _ in range(3):
        kernels.gemm(a, b, out, M, N, K, b_mn=b_mn, col_sum=cs, col_sumsq=cq)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        kernels.gemm(a, b, out, M, N, K, b_mn=b_mn, col_sum=cs, col_sumsq=cq)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3

for (M, N, K, stats, b_mn) in [(1400832, 144, 24, True, False), (1400832, 24, 144, False, True), (350208, 192, 32, True, False), (1400832, 128, 64, False, False)]:
    floor = (M * K + M * N) * 2 / 6.55e12 * 1e6
    print(f"M={M} N={N} K={K} stats={stats} b_mn={b_mn} floor={floor:.0f}us")
    for dbg in (0, 32, 2, 1, 3, 7, 31):
        print(f"   dbg={dbg:2d}: {run(M, N, K, stats, dbg, b_mn):8.1f} us")
