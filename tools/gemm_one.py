"""One GEMM shape, a few launches (for ncu).  usage: gemm_one.py M N K stats b_mn"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rt1b200.ops import kernels
M, N, K, stats, b_mn = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
a = torch.randn(M, K, device="cuda").bfloat16()
b = (torch.randn(K, N, device="cuda") if b_mn else torch.randn(N, K, device="cuda")).bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
cs = torch.zeros(N, device="cuda") if stats else None
cq = torch.zeros(N, device="cuda") if stats else None
for _ in range(4):
    kernels.gemm(a, b, out, M, N, K, b_mn=bool(b_mn), col_sum=cs, col_sumsq=cq)
torch.cuda.synchronize()
print("ok")
