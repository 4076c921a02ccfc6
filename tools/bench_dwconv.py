#!/usr/bin/env python
"""Depthwise-convolution kernels at the encoder's real shapes (48 frames = batch 8 x 6): tile kernels (cnn_dwconv.cu) vs
cuDNN (torch conv2d / convolution_backward, bf16 channels-last — what the reference reaches through ATen), CUDA-event timing
with an L2 flush between launches, clocks sampled.  (profiles/r2_dwconv_bench_vs_round1.json is the same table against the
round-1 row-strip kernels, which were deleted after that measurement.)

    python tools/bench_dwconv.py [--frames 48] [--out gpurun_out/dwconv_bench.json]

Per layer and per kernel: µs, algorithmic GB/s and the fraction of the measured copy bandwidth (MEASURED_PEAKS.json).
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import ClockSampler  # noqa: E402
from rt1b200.ops._lib import call, i32, ptr, stream  # noqa: E402
from rt1b200.ops.kernels import l2_flush  # noqa: E402

# (block indices, C, H, W, k, s, producer-BN) of the RT-1 encoder at 256x456
LAYERS = [("b0", 40, 128, 228, 3, 1, False), ("b1", 24, 128, 228, 3, 1, False), ("b2", 144, 128, 228, 3, 2, True),
          ("b3-4", 192, 64, 114, 3, 1, True), ("b5", 192, 64, 114, 5, 2, True), ("b6-7", 288, 32, 57, 5, 1, True),
          ("b8", 288, 32, 57, 3, 2, True), ("b9-12", 576, 16, 29, 3, 1, True), ("b13", 576, 16, 29, 5, 1, True),
          ("b14-17", 816, 16, 29, 5, 1, True), ("b18", 816, 16, 29, 5, 2, True), ("b19-23", 1392, 8, 15, 5, 1, True),
          ("b24", 1392, 8, 15, 3, 1, True), ("b25", 2304, 8, 15, 3, 1, True)]
COUNT = {"b3-4": 2, "b6-7": 2, "b9-12": 4, "b14-17": 4, "b19-23": 5}


def timeit(fn, flush, reps=5):
    ts = []
    for _ in range(reps + 2):
        l2_flush(flush)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return sorted(ts[2:])[len(ts[2:]) // 2]


def rel(a, b):
    a, b = a.float().flatten(), b.float().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=48)
    ap.add_argument("--out", default="gpurun_out/dwconv_bench.json")
    ap.add_argument("--layers", default="")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    peaks = {}
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm = peaks.get("hbm_gbs", 6650.0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    N = args.frames
    rows = []
    sampler = ClockSampler(0)
    sampler.start()
    want = set(args.layers.split(",")) if args.layers else None
    for name, C, H, W, k, s, has_bn in LAYERS:
        if want and name not in want:
            continue
        pad = (k - 1) // 2
        Ho, Wo = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
        g = torch.Generator(device="cpu").manual_seed(1)
        x = torch.randn(N, H, W, C, generator=g).to(torch.bfloat16).to(dev)
        w = (torch.randn(C, 1, k, k, generator=g) * 0.3).to(torch.bfloat16).to(dev)
        dy = torch.randn(N, Ho, Wo, C, generator=g).to(torch.bfloat16).to(dev)
        a = (torch.rand(C, generator=g) + 0.5).to(dev)
        b = (torch.randn(C, generator=g) * 0.2).to(dev)
        mean, rstd = (torch.randn(C, generator=g) * 0.1).to(dev), (torch.rand(C, generator=g) + 0.5).to(dev)
        in_b, out_b = x.numel() * 2, dy.numel() * 2
        res = {}
        # ---- ours
        y = torch.empty_like(dy)
        s1, s2 = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
        din = torch.empty_like(x)
        sums = torch.zeros(2, C, device=dev)
        scratch = torch.empty(k * k * C, device=dev)
        dw = torch.empty_like(w)

        def fwd(aa=None, bb=None):
            call("rt1_dwconv_fwd", ptr(x), ptr(aa), ptr(bb), ptr(w), ptr(y), ptr(s1), ptr(s2), i32(N), i32(H), i32(W),
                 i32(C), i32(k), i32(s), stream())

        def bwd():
            call("rt1_dwconv_bwd_data", ptr(dy), ptr(w), ptr(x if has_bn else None), ptr(a if has_bn else None),
                 ptr(b if has_bn else None), ptr(mean if has_bn else None), ptr(rstd if has_bn else None), ptr(din),
                 ptr(sums if has_bn else None), i32(N), i32(H), i32(W), i32(C), i32(k), i32(s), stream())

        def wgrad(aa=None, bb=None):
            call("rt1_dwconv_wgrad", ptr(dy), ptr(x), ptr(aa), ptr(bb), ptr(scratch), ptr(dw), i32(N), i32(H), i32(W),
                 i32(C), i32(k), i32(s), stream())

        r = {"fwd_us": timeit(fwd, flush), "bwd_us": timeit(bwd, flush), "wgrad_us": timeit(wgrad, flush)}
        if has_bn:
            r["fwd_act_us"] = timeit(lambda: fwd(a, b), flush)
            r["wgrad_act_us"] = timeit(lambda: wgrad(a, b), flush)
        fwd()
        wgrad()
        torch.cuda.synchronize()
        r["_y"], r["_dw"] = y.clone(), dw.clone()
        res["tile"] = r
        # ---- cuDNN: the same three operations on channels-last views of the same buffers (plain conv: no BN / SiLU terms,
        # so its data gradient is compared on time only — ours also multiplies by SiLU' and takes the BN sums when has_bn)
        xc, dyc = x.permute(0, 3, 1, 2), dy.permute(0, 3, 1, 2)             # NCHW views of NHWC memory = channels_last
        conv = lambda: torch.nn.functional.conv2d(xc, w, None, s, pad, 1, C)  # noqa: E731
        cb = lambda mask: torch.ops.aten.convolution_backward(dyc, xc, w, None, [s, s], [pad, pad], [1, 1], False, [0, 0], C, mask)  # noqa: E731
        v = {"fwd_us": timeit(conv, flush), "bwd_us": timeit(lambda: cb([True, False, False]), flush),
             "wgrad_us": timeit(lambda: cb([False, True, False]), flush)}
        v["_y"] = conv().permute(0, 2, 3, 1)
        v["_dw"] = cb([False, True, False])[1]
        res["cudnn"] = v
        t, v = res["tile"], res["cudnn"]
        row = {"layer": name, "count": COUNT.get(name, 1), "C": C, "H": H, "W": W, "k": k, "s": s, "frames": N,
               "fwd_bytes": in_b + out_b, "bwd_bytes": in_b * (2 if has_bn else 1) + out_b, "wgrad_bytes": in_b + out_b,
               "rel_y_vs_cudnn": rel(t["_y"], v["_y"]), "rel_dw_vs_cudnn": rel(t["_dw"], v["_dw"])}
        for kname, byt in (("fwd", row["fwd_bytes"]), ("bwd", row["bwd_bytes"]), ("wgrad", row["wgrad_bytes"])):
            row[f"{kname}_us"], row[f"{kname}_cudnn_us"] = round(t[f"{kname}_us"], 1), round(v[f"{kname}_us"], 1)
            row[f"{kname}_gbs"] = round(byt / t[f"{kname}_us"] / 1e3, 1)
            row[f"{kname}_frac_of_measured_hbm"] = round(byt / t[f"{kname}_us"] / 1e3 / hbm, 3)
        for kname in ("fwd_act", "wgrad_act"):
            if f"{kname}_us" in t:
                row[f"{kname}_us"] = round(t[f"{kname}_us"], 1)     # with SiLU(BN(.)) applied to the tile (cuDNN has no such mode)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del x, dy, res
        torch.cuda.empty_cache()
    clocks = sampler.stop()
    tot = {k_: round(sum(r[f"{k_}_us"] * r["count"] for r in rows), 1) for k_ in ("fwd", "bwd", "wgrad")}
    tot_v1 = {k_: round(sum(r[f"{k_}_cudnn_us"] * r["count"] for r in rows), 1) for k_ in ("fwd", "bwd", "wgrad")}
    out = {"what": "depthwise conv kernels at RT-1 encoder shapes, per launch, L2 flushed between launches",
           "hbm_gbs_measured": hbm, "frames": N, "rows": rows, "per_step_us_tile": tot, "per_step_us_cudnn": tot_v1, "clocks": clocks}
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    json.dump(out, open(args.out, "w"), indent=1)
    print("per step (us): tile", tot, " cuDNN", tot_v1, " clocks", clocks)


if __name__ == "__main__":
    main()
