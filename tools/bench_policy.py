#!/usr/bin/env python
"""BASELINE config 4: closed-loop policy-step latency (p50 / p99), 1 GPU.

    python tools/bench_policy.py                      # ours: 64 synthetic boards, ONE batched policy call per control step
    python tools/bench_policy.py --impl reference     # the unmodified reference model: batch 1 (its only supported mode),
                                                      # one call per environment → 64 sequential calls per control step

ours      : render (CUDA rasteriser) → crop/resize → `policy_step` (bf16, sm_100a kernels) → clip → board step; the
            latency is the whole control step for all 64 environments, measured with CUDA events.
reference : `RT1_Lightning.model(obs, network_state)` in eval mode under no_grad, exactly what
            `language_table/train/policy.py:81-82` runs per step, on a device-resident 256x456 frame (the reference's
            simulator/renderer — PyBullet + TinyRenderer — is unavailable on this box, so its per-step cost is NOT
            included: the reference number is a lower bound on its control-step time).
Prints one JSON line.
"""
import argparse
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def ours(a):
    import numpy as np
    import torch
    from rt1b200.eval.closed_loop import run_closed_loop
    from rt1b200.models.factory import RT1Config, build_rt1
    from rt1b200.parallel.engine import keep_bn_stats_fp32
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    model = build_rt1(RT1Config.language_table()).to(dev).to(torch.bfloat16)
    keep_bn_stats_fp32(model)
    model.eval()
    res = run_closed_loop(model, num_envs=a.envs, episodes=1, max_episode_steps=a.steps, device=dev, seed=0)
    res_g = run_closed_loop(model, num_envs=a.envs, episodes=1, max_episode_steps=a.steps, device=dev, seed=0, graphed=True)
    # policy call alone (no board / rasteriser), same batch
    state = model.initial_state(a.envs, device=dev, dtype=torch.bfloat16)
    obs = {"image": torch.rand(a.envs, 3, 256, 456, device=dev), "natural_language_embedding": torch.randn(a.envs, 512, device=dev)}
    lat = []
    with torch.no_grad():
        for i in range(a.steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _, state = model.policy_step(obs, state)
            e1.record(); e1.synchronize()
            lat.append(e0.elapsed_time(e1))
    lat = np.asarray(lat[3:])
    return {"impl": "ours", "num_envs": a.envs, "control_steps": a.steps, "dtype": "bf16",
            "control_step_ms_p50": res["step_latency_ms_p50"], "control_step_ms_p99": res["step_latency_ms_p99"],
            "control_step_graphed_policy_ms_p50": res_g["step_latency_ms_p50"],
            "control_step_graphed_policy_ms_p99": res_g["step_latency_ms_p99"],
            "policy_call_ms_p50": float(np.percentile(lat, 50)), "policy_call_ms_p99": float(np.percentile(lat, 99)),
            "env_steps_per_s": a.envs / (res["step_latency_ms_p50"] / 1e3),
            "env_steps_per_s_graphed_policy": a.envs / (res_g["step_latency_ms_p50"] / 1e3),
            "what": "render + crop/resize + batched policy_step + board step for all envs, per control step"}


def graphed(a):
    """Policy call latency, eager vs one CUDA-graph replay, at batch 1 (one robot) and at --envs."""
    import numpy as np
    import torch
    from rt1b200.eval.graphed import GraphedPolicy
    from rt1b200.models.factory import RT1Config, build_rt1
    from rt1b200.parallel.engine import keep_bn_stats_fp32
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    model = build_rt1(RT1Config.language_table(), eval_center_crop=True).to(dev).to(torch.bfloat16)
    keep_bn_stats_fp32(model)
    model.eval()
    out = {"impl": "ours-graphed", "dtype": "bf16", "control_steps": a.steps}

    def time_it(fn):
        lat = []
        for _ in range(a.steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record(); e1.synchronize()
            lat.append(e0.elapsed_time(e1))
        lat = np.asarray(lat[3:])
        return float(np.percentile(lat, 50)), float(np.percentile(lat, 99))

    for b in (1, a.envs):
        img, emb = torch.rand(b, 3, 256, 456, device=dev), torch.randn(b, 512, device=dev)
        state = {"s": model.initial_state(b, device=dev, dtype=torch.bfloat16)}

        def eager():
            with torch.no_grad():
                _, state["s"] = model.policy_step({"image": img, "natural_language_embedding": emb}, state["s"])
        gp = GraphedPolicy(model, b, 256, 456)
        e50, e99 = time_it(eager)
        g50, g99 = time_it(lambda: gp.step(img, emb))
        out[f"batch{b}"] = {"eager_ms_p50": e50, "eager_ms_p99": e99, "graph_ms_p50": g50, "graph_ms_p99": g99}
    return out


def reference(a):
    sys.path.insert(0, os.path.join(REPO, "baseline"))
    import refenv
    if not refenv.available() and not refenv.try_install():
        return {"impl": "reference", "unavailable": refenv.why_unavailable()}
    refenv.activate()
    import numpy as np
    import torch
    from argparse import Namespace
    import distribute_train as ref_train
    from pytorch_robotics_transformer.tokenizers.utils import batched_space_sampler, np_to_tensor
    torch.manual_seed(0)
    module = ref_train.RT1_Lightning(Namespace(height=256, width=456, seq_len=6, lr=5e-4, milestones=[50, 75, 90])).to("cuda")
    module.eval()
    obs = {"image": torch.full([1, 3, 256, 456], 0.5, device="cuda"), "natural_language_embedding": torch.ones(1, 512, device="cuda")}
    state = np_to_tensor(batched_space_sampler(module.model._state_space, batch_size=1), "cuda")
    for k in state:
        state[k] = torch.zeros_like(state[k])
    state["seq_idx"] = state["seq_idx"].reshape(1)           # the caller re-wraps it to shape [1] (policy.py:83)
    lat = []
    with torch.no_grad():
        for i in range(a.steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.envs):                      # batch 1 only: one call per environment
                _, st = module.model(obs, dict(state))   # the model rebinds keys of the dict it is given
            e1.record(); e1.synchronize()
            state = {k: (v.reshape(1) if k == "seq_idx" else v) for k, v in st.items()}
            lat.append(e0.elapsed_time(e1))
    lat = np.asarray(lat[2:])
    return {"impl": "reference", "num_envs": a.envs, "control_steps": a.steps, "dtype": "fp32",
            "policy_calls_ms_p50": float(np.percentile(lat, 50)), "policy_calls_ms_p99": float(np.percentile(lat, 99)),
            "single_call_ms_p50": float(np.percentile(lat, 50)) / a.envs,
            "env_steps_per_s": a.envs / (float(np.percentile(lat, 50)) / 1e3),
            "what": f"{a.envs} sequential batch-1 model calls per control step (policy only; simulator + renderer unavailable on this box)"}


def raster_bench(a):
    """Frames/s of the CUDA z-buffer rasteriser vs its numpy restatement (the oracle and timed CPU baseline; the
    reference's TinyRenderer / any graphics-API path is unavailable on this box)."""
    import time
    import torch
    from rt1b200.envs import raster
    from rt1b200.envs.board import BatchedLanguageTable
    env = BatchedLanguageTable(a.envs, device="cuda", seed=0)
    tris, ntris = env.scene.triangles(env.effector, env.block_xy, env.block_yaw, env.block_mask)
    for _ in range(3):
        img = raster.render(tris, ntris)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        img = raster.render(tris, ntris)
        out = raster.crop_resize(img, 256, 456, 0.95)
    e1.record(); torch.cuda.synchronize()
    gpu_ms = e0.elapsed_time(e1) / 20
    t_np, n_np = tris[:2].cpu().numpy(), ntris[:2].cpu().numpy()
    t0 = time.perf_counter()
    raster.render_reference(t_np, n_np)
    cpu_ms_per_frame = (time.perf_counter() - t0) * 1e3 / 2
    return {"impl": "raster", "num_envs": a.envs, "gpu_ms_per_batch_render_plus_resize": gpu_ms,
            "gpu_frames_per_s": a.envs / (gpu_ms / 1e3), "cpu_restatement_ms_per_frame": cpu_ms_per_frame,
            "cpu_frames_per_s": 1e3 / cpu_ms_per_frame, "frame": "180x320 -> 256x456"}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", choices=["ours", "reference", "raster", "graphed"], default="ours")
    ap.add_argument("--envs", type=int, default=64)
    ap.add_argument("--steps", type=int, default=30)
    a = ap.parse_args()
    from bench import ClockSampler          # every recorded number carries the clocks seen while it was taken
    sampler = ClockSampler(0)
    sampler.start()
    try:
        out = {"ours": ours, "reference": reference, "raster": raster_bench, "graphed": graphed}[a.impl](a)
    except Exception as e:      # noqa: BLE001
        out = {"impl": a.impl, "error": f"{type(e).__name__}: {str(e)[:400]}"}
    out["clocks"] = sampler.stop()
    out["timing"] = "CUDA events per call, p50/p99 over --steps calls after 3 warm-up calls"
    print(json.dumps(out), flush=True)
