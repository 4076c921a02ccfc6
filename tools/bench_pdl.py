"""Does programmatic dependent launch shorten a chain of small dependent kernels inside a captured CUDA graph?
Run twice: RT1B200_PDL=1 and RT1B200_PDL=0 (the switch is read once per process).  Prints one JSON line."""
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch  # noqa: E402
from rt1b200.ops._lib import call, i32, ptr, stream  # noqa: E402

dev = torch.device("cuda", 0)
N, HW, C, CHAIN = 48, 120, 1392, 400                     # a late-block tensor (16 MB): every kernel is ~5 us
x = torch.randn(N, HW, C, device=dev).to(torch.bfloat16)
y = torch.empty_like(x)
a, b = torch.ones(C, device=dev), torch.zeros(C, device=dev)


def chain():
    src, dst = x, y
    for _ in range(CHAIN):
        call("rt1_bn_act_scale", ptr(src), ptr(a), ptr(b), ptr(None), ptr(dst), i32(N), i32(HW), i32(C), i32(0), stream())
        src, dst = dst, src


s = torch.cuda.Stream()
with torch.cuda.stream(s):
    chain()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        chain()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(10):
        g.replay()
    t1.record()
    torch.cuda.synchronize()
print(json.dumps({"pdl": os.environ.get("RT1B200_PDL", "1"), "chain": CHAIN, "us_per_kernel_in_graph": t0.elapsed_time(t1) * 1000 / (10 * CHAIN)}))
