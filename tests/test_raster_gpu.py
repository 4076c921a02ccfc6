"""CUDA rasteriser / crop-resize vs their CPU restatements, and the batched closed loop on the GPU."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_rasteriser_matches_cpu_restatement():
    from rt1b200.envs import raster
    from rt1b200.envs.board import BatchedLanguageTable
    env = BatchedLanguageTable(3, device="cuda", seed=0)
    tris, ntris = env.scene.triangles(env.effector, env.block_xy, env.block_yaw, env.block_mask)
    gpu = raster.render(tris, ntris).cpu().numpy()
    cpu = raster.render_reference(tris.cpu().numpy(), ntris.cpu().numpy())
    mismatch = (gpu != cpu).any(-1).mean()
    assert mismatch < 2e-3, mismatch             # only edge pixels may differ (fp32 op order on edge functions)
    a = raster.crop_resize(torch.from_numpy(cpu).cuda(), 256, 456, 0.95).cpu()
    b = raster.crop_resize_reference(torch.from_numpy(cpu), 256, 456, 0.95)
    torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-4)


def test_batched_closed_loop_policy_and_oracle():
    from rt1b200.eval.closed_loop import run_closed_loop
    from rt1b200.models import RT1Config, build_rt1
    from rt1b200.parallel.engine import keep_bn_stats_fp32
    torch.manual_seed(0)
    model = build_rt1(RT1Config.language_table(height=128, width=224, seq_len=3, num_layers=2), eval_center_crop=True).cuda().to(torch.bfloat16)
    keep_bn_stats_fp32(model)
    res = run_closed_loop(model, num_envs=8, episodes=1, max_episode_steps=6, device="cuda", image_hw=(128, 224))
    assert res["episodes"] == 8 and res["step_latency_ms_p50"] > 0
    res = run_closed_loop(model, num_envs=8, episodes=1, max_episode_steps=80, device="cuda", image_hw=(128, 224), use_oracle=True)
    assert res["success_rate"] >= 0.75


def test_graphed_policy_matches_eager_steps():
    """One CUDA-graph replay per control step == the eager policy_step (same fixed crop), over a window roll."""
    from rt1b200.eval.graphed import GraphedPolicy
    from rt1b200.models import RT1Config, build_rt1
    from rt1b200.parallel.engine import keep_bn_stats_fp32
    torch.manual_seed(0)
    H, W, T, B = 128, 224, 3, 4
    model = build_rt1(RT1Config.language_table(height=H, width=W, seq_len=T, num_layers=2), eval_center_crop=True).cuda().to(torch.bfloat16)
    keep_bn_stats_fp32(model)
    model.eval()
    gp = GraphedPolicy(model, B, H, W)
    state = model.initial_state(B, device="cuda", dtype=torch.bfloat16)
    g = torch.Generator(device="cuda").manual_seed(1)
    emb = torch.randn(B, 512, device="cuda", generator=g)
    for step in range(T + 2):                              # past the point where the window starts to roll
        img = torch.rand(B, 3, H, W, device="cuda", generator=g)
        with torch.no_grad():
            want, state = model.policy_step({"image": img, "natural_language_embedding": emb}, state)
        got = gp.step(img, emb)
        torch.cuda.synchronize()
        assert torch.equal(got["terminate_episode"].cpu(), want["terminate_episode"].cpu()), step
        torch.testing.assert_close(got["action"].float().cpu(), want["action"].float().cpu(), rtol=0, atol=2 * 0.2 / 255)
    assert int(gp.state["seq_idx"][0]) == T
    gp.reset()
    assert int(gp.state["seq_idx"].sum()) == 0


def test_crop_resize_boxes_kernel_matches_its_pytorch_statement():
    """rt1_crop_resize_boxes (per-frame crop box + bilinear resize of raw uint8 frames) vs the PyTorch statement of the same
    sampling rule, uint8 and float outputs."""
    from rt1b200.data.gpu_transforms import crop_resize_boxes, crop_resize_boxes_reference
    g = torch.Generator().manual_seed(0)
    frames = torch.randint(0, 256, (5, 45, 80, 3), dtype=torch.uint8, generator=g)
    boxes = torch.tensor([[0, 0, 45, 80], [2, 3, 42.75, 76.0], [1.0, 4.0, 42.75, 76.0], [0, 0, 20, 30], [10, 20, 35, 60]], dtype=torch.float32)
    want8 = crop_resize_boxes_reference(frames, boxes, (64, 114), as_uint8=True)
    got8 = crop_resize_boxes(frames.cuda(), boxes.cuda(), (64, 114), as_uint8=True).cpu()
    assert got8.dtype == torch.uint8 and int((got8.int() - want8.int()).abs().max()) <= 1      # fp rounding at .5 boundaries
    assert float((got8.int() - want8.int()).abs().float().mean()) < 0.01
    wantf = crop_resize_boxes_reference(frames, boxes, (64, 114), as_uint8=False)
    gotf = crop_resize_boxes(frames.cuda(), boxes.cuda(), (64, 114), as_uint8=False).cpu()
    torch.testing.assert_close(gotf, wantf, rtol=1e-5, atol=1e-5)
