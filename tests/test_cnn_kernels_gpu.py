"""NHWC image-tokenizer kernels vs float32 PyTorch statements of the same math (autograd provides the backward oracle)."""
import pytest
import torch
import torch.nn.functional as F

from rt1b200.ops import _lib
from rt1b200.ops._lib import call, f32, i32, i64, ptr, stream

pytestmark = pytest.mark.gpu
DEV, BF = "cuda", torch.bfloat16


def _bf(*shape, scale=1.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(BF).to(DEV)


def _rel(a, b):
    a, b = a.float().flatten(), b.float().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


def _nchw(x_nhwc):
    return x_nhwc.float().permute(0, 3, 1, 2)


@pytest.mark.parametrize("k,s", [(3, 1), (3, 2), (5, 1), (5, 2)])
@pytest.mark.parametrize("with_bn", [True, False])
@pytest.mark.parametrize("C,H,W", [(40, 17, 29), (144, 16, 30), (816, 8, 15)])
def test_dwconv_fwd_bwd_wgrad(k, s, with_bn, C, H, W):
    N = 3
    x = _bf(N, H, W, C, seed=1)
    w = _bf(C, 1, k, k, scale=0.3, seed=2)
    a = (torch.rand(C, device=DEV) + 0.5) if with_bn else None
    b = (torch.randn(C, device=DEV) * 0.2) if with_bn else None
    mean = torch.randn(C, device=DEV) * 0.1
    rstd = torch.rand(C, device=DEV) + 0.5
    pad = (k - 1) // 2
    Ho, Wo = (H + 2 * pad - k) // s + 1, (W + 2 * pad - k) // s + 1
    y = torch.empty(N, Ho, Wo, C, dtype=BF, device=DEV)
    ssum, ssq = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    call("rt1_dwconv_fwd", ptr(x), ptr(a), ptr(b), ptr(w), ptr(y), ptr(ssum), ptr(ssq), i32(N), i32(H), i32(W), i32(C), i32(k),
         i32(s), stream())
    # oracle
    xr = _nchw(x).requires_grad_()
    wr = w.float().requires_grad_()
    h = xr * a[None, :, None, None] + b[None, :, None, None] if with_bn else xr
    act = F.silu(h) if with_bn else h
    yr = F.conv2d(act, wr, None, s, pad, 1, C)
    assert _rel(_nchw(y), yr) < 1e-2
    assert _rel(ssum, yr.sum((0, 2, 3))) < 2e-2 and _rel(ssq, (yr * yr).sum((0, 2, 3))) < 2e-2
    g = _bf(N, Ho, Wo, C, seed=3)
    yr.backward(_nchw(g))
    # weight gradient
    scratch = torch.empty(k * k * C, device=DEV)
    dw = torch.empty_like(w)
    call("rt1_dwconv_wgrad", ptr(g), ptr(x), ptr(a), ptr(b), ptr(scratch), ptr(dw), i32(N), i32(H), i32(W), i32(C), i32(k), i32(s),
         stream())
    assert _rel(dw, wr.grad) < 2e-2
    # data gradient (+ SiLU' and BN-backward sums when the input was BN+SiLU)
    din = torch.empty(N, H, W, C, dtype=BF, device=DEV)
    sums = torch.zeros(2, C, device=DEV)
    call("rt1_dwconv_bwd_data", ptr(g), ptr(w), ptr(x if with_bn else None), ptr(a), ptr(b), ptr(mean if with_bn else None),
         ptr(rstd if with_bn else None), ptr(din), ptr(sums if with_bn else None), i32(N), i32(H), i32(W), i32(C), i32(k), i32(s),
         stream())
    if with_bn:
        dh = xr.grad / a[None, :, None, None]          # d/dh = d/dx / a
        assert _rel(_nchw(din), dh) < 2e-2
        xhat = (_nchw(x) - mean[None, :, None, None]) * rstd[None, :, None, None]
        assert _rel(sums[0], dh.sum((0, 2, 3))) < 3e-2
        assert _rel(sums[1], (dh * xhat).sum((0, 2, 3))) < 3e-2
    else:
        assert _rel(_nchw(din), xr.grad) < 2e-2


def test_stem_im2col_gemm_fwd_wgrad():
    """The stem runs as im2col (K = 27 -> 32) + tcgen05 GEMM; the weight gradient is the split-K GEMM over the same columns."""
    from rt1b200.ops.kernels import gemm
    N, H, W, CO = 2, 64, 96, 40
    x8 = torch.zeros(N, H, W, 8, dtype=BF, device=DEV)
    x8[..., :3] = _bf(N, H, W, 3, seed=1)
    w = _bf(CO, 3, 3, 3, scale=0.3, seed=2)
    Ho, Wo = H // 2, W // 2
    M = N * Ho * Wo
    cols = torch.empty(M, 32, dtype=BF, device=DEV)
    call("rt1_stem_im2col", ptr(x8), ptr(cols), i32(N), i32(H), i32(W), stream())
    w32 = F.pad(w.view(CO, 27), (0, 5))
    y = torch.empty(N, Ho, Wo, CO, dtype=BF, device=DEV)
    s1, s2 = torch.zeros(CO, device=DEV), torch.zeros(CO, device=DEV)
    gemm(cols, w32, y.view(M, CO), M, CO, 32, col_sum=s1, col_sumsq=s2)
    wr = w.float().requires_grad_()
    yr = F.conv2d(_nchw(x8[..., :3]), wr, None, 2, 1)
    assert _rel(_nchw(y), yr) < 1e-2 and _rel(s1, yr.sum((0, 2, 3))) < 2e-2
    g = _bf(N, Ho, Wo, CO, seed=3)
    yr.backward(_nchw(g))
    acc = torch.zeros(CO, 32, device=DEV)
    gemm(g.view(M, CO), cols, acc, CO, 32, M, a_mn=True, b_mn=True, split_k=8)
    assert _rel(acc[:, :27].reshape(CO, 3, 3, 3), wr.grad) < 2e-2


@pytest.mark.parametrize("skip,film,bn", [(True, True, True), (False, True, True), (False, True, False), (True, False, True)])
def test_block_finalize_fwd_bwd(skip, film, bn):
    N, HW, C, GB, off = 4, 77, 48, 200, 16
    y = _bf(N, HW, C, seed=1)
    x_in = _bf(N, HW, C, seed=2) if skip else None
    gb = _bf(N, GB, scale=0.3, seed=3) if film else None
    sd = torch.tensor([1.25, 0.0, 1.25, 1.25], device=DEV) if skip else None
    gamma3, beta3 = _bf(C, seed=4) * 0.2 + 1, _bf(C, seed=5) * 0.2
    mean, rstd = torch.randn(C, device=DEV) * 0.1, torch.rand(C, device=DEV) + 0.5
    a = gamma3.float() * rstd if bn else None
    b = beta3.float() - mean * gamma3.float() * rstd if bn else None
    out = torch.empty_like(y)
    call("rt1_block_finalize", ptr(y), ptr(a), ptr(b), ptr(x_in), ptr(sd), ptr(gb), i64(GB), i64(off), ptr(out), i32(N), i32(HW), i32(C),
         stream())
    # oracle with autograd; BN statistics are *functions of y* in training, so build them from y to get the true gradient
    yr = y.float().requires_grad_()
    xr = x_in.float().requires_grad_() if skip else None
    gbr = gb.float().requires_grad_() if film else None
    g3, b3 = gamma3.float().requires_grad_(), beta3.float().requires_grad_()
    if bn:
        mu = yr.mean((0, 1)); var = yr.var((0, 1), unbiased=False)
        rs = torch.rsqrt(var + 1e-5)
        u = (yr - mu) * rs * g3 + b3
        mean, rstd = mu.detach(), rs.detach()
        a, b = (g3 * rs).detach(), (b3 - mu * g3 * rs).detach()
        call("rt1_block_finalize", ptr(y), ptr(a), ptr(b), ptr(x_in), ptr(sd), ptr(gb), i64(GB), i64(off), ptr(out), i32(N), i32(HW),
             i32(C), stream())
    else:
        u = yr
    z = u * (sd[:, None, None] if sd is not None else 1.0)
    if skip:
        z = z + xr
    outr = (1 + gbr[:, None, off + C:off + 2 * C]) * z + gbr[:, None, off:off + C] if film else z
    assert _rel(out, outr) < 1e-2
    g = _bf(N, HW, C, seed=6)
    outr.backward(g.float())
    S = torch.empty(3 * N * C, device=DEV)
    c1, c2 = torch.zeros(C, device=DEV), torch.zeros(C, device=DEV)
    dgb = torch.zeros(N, GB, dtype=BF, device=DEV)
    dg3, db3 = torch.zeros(C, dtype=BF, device=DEV), torch.zeros(C, dtype=BF, device=DEV)
    dy, dskip = torch.empty_like(y), (torch.empty_like(y) if skip else None)
    call("rt1_block_finalize_bwd", ptr(g), ptr(y), ptr(x_in), ptr(gb), i64(GB), i64(off), ptr(sd), ptr(a), ptr(mean if bn else None),
         ptr(rstd if bn else None), ptr(gamma3 if bn else None), ptr(beta3 if bn else None), ptr(S), ptr(c1), ptr(c2), ptr(dgb),
         ptr(dg3 if bn else None), ptr(db3 if bn else None), ptr(dy), ptr(dskip), i32(N), i32(HW), i32(C), stream())
    assert _rel(dy, yr.grad) < 3e-2
    if skip:
        assert _rel(dskip, xr.grad) < 2e-2
    if film:
        assert _rel(dgb[:, off:off + 2 * C], gbr.grad[:, off:off + 2 * C]) < 3e-2
    if bn:
        assert _rel(dg3, g3.grad) < 3e-2 and _rel(db3, b3.grad) < 3e-2


@pytest.mark.parametrize("use_se", [True, False])
def test_bn_act_bwd_with_se_terms(use_se):
    N, HW, C = 4, 60, 144
    x = _bf(N, HW, C, seed=1)
    gamma, beta = _bf(C, seed=2) * 0.2 + 1, _bf(C, seed=3) * 0.2
    s = torch.rand(N, C, device=DEV) if use_se else None
    xr = x.float().requires_grad_()
    gr, br = gamma.float().requires_grad_(), beta.float().requires_grad_()
    mu, var = xr.mean((0, 1)), xr.var((0, 1), unbiased=False)
    rs = torch.rsqrt(var + 1e-5)
    t = F.silu((xr - mu) * rs * gr + br)
    w_pool = torch.randn(N, C, device=DEV)                  # stands for the gradient flowing through the SE branch
    outr = t * s[:, None, :] if use_se else t
    g = _bf(N, HW, C, seed=4)
    loss = (outr * g.float()).sum() + ((t.mean(1) * w_pool).sum() if use_se else 0.0)
    loss.backward()
    a, b = (gr * rs).detach(), (br - mu * gr * rs).detach()
    dpool = (w_pool / HW) if use_se else None
    sums = torch.empty(2, C, device=DEV)
    dgam, dbet = torch.empty(C, dtype=BF, device=DEV), torch.empty(C, dtype=BF, device=DEV)
    dx = torch.empty_like(x)
    call("rt1_bn_act_bwd", ptr(g), ptr(x), ptr(a), ptr(b), ptr(mu.detach()), ptr(rs.detach()), ptr(s), ptr(dpool), ptr(sums), i32(0),
         ptr(dgam), ptr(dbet), ptr(dx), i32(N), i32(HW), i32(C), i32(1), stream())
    assert _rel(dx, xr.grad) < 3e-2 and _rel(dgam, gr.grad) < 3e-2 and _rel(dbet, br.grad) < 3e-2
    # forward pieces used by the SE path
    pooled = torch.zeros(N, C, device=DEV)
    call("rt1_bn_act_pool", ptr(x), ptr(a), ptr(b), ptr(pooled), i32(N), i32(HW), i32(C), stream())
    assert _rel(pooled / HW, t.mean(1)) < 1e-2
    y = torch.empty_like(x)
    call("rt1_bn_act_scale", ptr(x), ptr(a), ptr(b), ptr(s), ptr(y), i32(N), i32(HW), i32(C), i32(1), stream())
    assert _rel(y, outr) < 1e-2
    if use_se:
        ds = torch.empty(N, C, device=DEV)
        call("rt1_se_scale_bwd_reduce", ptr(g), ptr(x), ptr(a), ptr(b), ptr(ds), i32(N), i32(HW), i32(C), stream())
        assert _rel(ds, (g.float() * t).sum(1)) < 2e-2


@pytest.mark.parametrize("N,C,Cs,HW", [(6, 144, 6, 50), (48, 2304, 96, 120), (5, 40, 10, 33), (12, 816, 34, 64)])
def test_se_fc_fwd_bwd(N, C, Cs, HW):
    """SE FC pair: forward, and the sliced two-launch backward (data + all four parameter gradients + the BN2 sums)."""
    pooled_sum = torch.randn(N, C, device=DEV) * HW * 0.3
    w1, b1, w2, b2 = _bf(Cs, C, scale=0.2, seed=1), _bf(Cs, seed=2), _bf(C, Cs, scale=0.5, seed=3), _bf(C, seed=4)
    z1, s = torch.empty(N, Cs, device=DEV), torch.empty(N, C, device=DEV)
    call("rt1_se_fc_fwd", ptr(pooled_sum), f32(1.0 / HW), ptr(w1), ptr(b1), ptr(w2), ptr(b2), ptr(z1), ptr(s), i32(N), i32(C), i32(Cs),
         stream())
    pr = (pooled_sum / HW).requires_grad_()
    P = [t.float().requires_grad_() for t in (w1, b1, w2, b2)]
    sr = torch.sigmoid(F.linear(F.silu(F.linear(pr, P[0], P[1])), P[2], P[3]))
    assert _rel(s, sr) < 1e-3
    ds = torch.randn(N, C, device=DEV)
    sr.backward(ds)
    parts = torch.randn(5, N, C, device=DEV)
    dpool, sums = torch.empty(N, C, device=DEV), torch.empty(2, C, device=DEV)
    dhid = torch.zeros(N, (Cs + 3) // 4 * 4, device=DEV)
    G = [torch.empty_like(t) for t in (w1, b1, w2, b2)]
    call("rt1_se_bwd", ptr(ds), ptr(s), ptr(z1), ptr(pooled_sum), f32(1.0 / HW), ptr(w1), ptr(w2), ptr(parts), ptr(dhid), ptr(dpool),
         ptr(sums), ptr(G[0]), ptr(G[1]), ptr(G[2]), ptr(G[3]), i32(N), i32(C), i32(Cs), stream())
    assert _rel(dpool * HW, pr.grad) < 1e-2
    for gk, p in zip(G, P):
        assert _rel(gk, p.grad) < 2e-2
    dp = pr.grad / HW
    want = torch.stack([(s * parts[1] + dp * parts[2]).sum(0), (s * parts[3] + dp * parts[4]).sum(0)])
    assert _rel(sums, want) < 1e-2
    # without the BatchNorm parts the sums are left alone
    call("rt1_se_bwd", ptr(ds), ptr(s), ptr(z1), ptr(pooled_sum), f32(1.0 / HW), ptr(w1), ptr(w2), ptr(None), ptr(dhid.zero_()),
         ptr(dpool), ptr(None), ptr(G[0]), ptr(G[1]), ptr(G[2]), ptr(G[3]), i32(N), i32(C), i32(Cs), stream())
    assert _rel(dpool * HW, pr.grad) < 1e-2


@pytest.mark.parametrize("M,K,C", [(48 * 32 * 57, 48, 288), (4 * 16 * 29, 96, 576), (1000, 24, 144)])
def test_gemm_with_batchnorm_tail(M, K, C):
    """The last CTA of the statistics-taking GEMM finalises the BatchNorm: same a/b/mean/rstd/running stats as rt1_bn_finalize."""
    from rt1b200.ops.kernels import gemm
    from rt1b200.ops.kernels_cnn import BnTail
    import ctypes
    x, w = _bf(M, K, seed=1), _bf(C, K, scale=0.3, seed=2)
    gamma, beta = _bf(C, seed=3) * 0.2 + 1, _bf(C, seed=4) * 0.2
    outs = []
    for use_tail in (False, True):
        y = torch.empty(M, C, dtype=BF, device=DEV)
        st = torch.zeros(6, C, device=DEV)
        rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
        ticket = torch.zeros(4, dtype=torch.int32, device=DEV)
        if use_tail:
            t = BnTail(gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(), st[2].data_ptr(), st[3].data_ptr(),
                       st[4].data_ptr(), st[5].data_ptr(), ticket.data_ptr(), float(M), 0.1, 1e-5)
            call("rt1_gemm_bf16_bn", ptr(x), i64(K), i32(0), ptr(w), i64(K), i32(0), ptr(y), i64(C), i32(M), i32(C), i32(K), ptr(st[0]),
                 ptr(st[1]), ctypes.byref(t), stream())
            assert int(ticket[0]) > 0
        else:
            gemm(x, w, y, M, C, K, col_sum=st[0], col_sumsq=st[1])
            call("rt1_bn_finalize", ptr(st[0]), ptr(st[1]), f32(M), ptr(gamma), ptr(beta), ptr(rm), ptr(rv), f32(0.1), f32(1e-5), i32(1),
                 ptr(st[2]), ptr(st[3]), ptr(st[4]), ptr(st[5]), i32(C), stream())
        outs.append((y, st, rm, rv))
    (y0, s0, rm0, rv0), (y1, s1, rm1, rv1) = outs
    assert torch.equal(y0, y1)
    assert _rel(s1[2:], s0[2:]) < 1e-4 and _rel(rm1, rm0) < 1e-4 and _rel(rv1, rv0) < 1e-4
    ref = x.float() @ w.float().t()
    assert _rel(s1[4], ref.mean(0)) < 2e-2 and _rel(s1[5], torch.rsqrt(ref.var(0, unbiased=False) + 1e-5)) < 2e-2


def test_dwconv_with_batchnorm_tail():
    from rt1b200.ops.kernels_cnn import BnTail
    import ctypes
    N, H, W, C, k, s = 4, 32, 57, 288, 5, 1
    x, w = _bf(N, H, W, C, seed=1), _bf(C, k * k, scale=0.2, seed=2)
    gamma, beta = _bf(C, seed=3) * 0.2 + 1, _bf(C, seed=4) * 0.2
    outs = []
    for use_tail in (False, True):
        y = torch.empty(N, H, W, C, dtype=BF, device=DEV)
        st = torch.zeros(6, C, device=DEV)
        rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
        ticket = torch.zeros(4, dtype=torch.int32, device=DEV)
        if use_tail:
            t = BnTail(gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(), st[2].data_ptr(), st[3].data_ptr(),
                       st[4].data_ptr(), st[5].data_ptr(), ticket.data_ptr(), float(N * H * W), 0.1, 1e-5)
            call("rt1_dwconv_fwd_bn", ptr(x), ptr(None), ptr(None), ptr(w), ptr(y), ptr(st[0]), ptr(st[1]), ctypes.byref(t), i32(N), i32(H),
                 i32(W), i32(C), i32(k), i32(s), stream())
        else:
            call("rt1_dwconv_fwd", ptr(x), ptr(None), ptr(None), ptr(w), ptr(y), ptr(st[0]), ptr(st[1]), i32(N), i32(H), i32(W), i32(C),
                 i32(k), i32(s), stream())
            call("rt1_bn_finalize", ptr(st[0]), ptr(st[1]), f32(N * H * W), ptr(gamma), ptr(beta), ptr(rm), ptr(rv), f32(0.1), f32(1e-5),
                 i32(1), ptr(st[2]), ptr(st[3]), ptr(st[4]), ptr(st[5]), i32(C), stream())
        outs.append((y, st, rm, rv))
    (y0, s0, rm0, rv0), (y1, s1, rm1, rv1) = outs
    assert torch.equal(y0, y1)
    assert _rel(s1[2:], s0[2:]) < 1e-4 and _rel(rm1, rm0) < 1e-4 and _rel(rv1, rv0) < 1e-4


@pytest.mark.parametrize("training", [True, False])
def test_full_encoder_matches_fp32_module(training):
    """Whole FiLM-EfficientNet-B3 + TokenLearner through the NHWC kernels vs the fp32 PyTorch module."""
    from rt1b200 import ops
    from rt1b200.models.tokenizers import RT1ImageTokenizer
    from rt1b200.parallel.engine import keep_bn_stats_fp32
    torch.manual_seed(0)
    ref = RT1ImageTokenizer(512, use_token_learner=True, num_tokens=8).to(DEV)
    for n, p in ref.named_parameters():            # non-trivial FiLM and BN parameters
        if "_projection_" in n:
            torch.nn.init.normal_(p, std=0.02)
        elif n.endswith("1.weight") and p.dim() == 1:
            torch.nn.init.uniform_(p, 0.5, 1.5)
    import copy
    fast = copy.deepcopy(ref).to(BF)
    keep_bn_stats_fp32(fast)
    with torch.no_grad():                          # identical (bf16-representable) weights on both sides
        for pr, pf in zip(ref.parameters(), fast.parameters()):
            pr.copy_(pf.float())
    for m in ref.modules():                         # stochastic depth is random: switch it off for the comparison
        if hasattr(m, "drop_rate"):
            m.drop_rate = 0.0
    for m in fast.modules():
        if hasattr(m, "drop_rate"):
            m.drop_rate = 0.0
    ref.train(training); fast.train(training)
    b, t, H, W = 2, 2, 64, 96
    img = torch.rand(b, t, 3, H, W, device=DEV).to(BF)
    ctx = _bf(b, t, 512, seed=5)
    out_r = ref(img.float(), ctx.float())
    x8 = ops.reference.preprocess_nhwc(img.reshape(b * t, 3, H, W).float(), (int(H * 0.07), int(W * 0.07)))
    x8 = x8.to(DEV)
    # reference path sees the un-shifted frame too: centred window == original image
    out_f = fast.forward_nhwc(x8, ctx, b, t)
    assert out_f.shape == out_r.shape
    assert _rel(out_f, out_r) < 6e-2, _rel(out_f, out_r)
    if not training:
        return
    g = _bf(*out_r.shape, seed=6)
    out_r.backward(g.float())
    out_f.backward(g)
    # A 26-block BatchNorm network in bf16 is noisy against fp32 no matter who runs it, so the yardstick is the
    # SAME bf16 model executed with stock PyTorch channels-last ops: the kernels must not be less accurate.
    import os
    torch_bf16 = copy.deepcopy(fast)
    for p_ in torch_bf16.parameters():
        p_.grad = None
    with ops.cnn.force_torch_encoder():
        out_t = torch_bf16.forward_nhwc(x8, ctx, b, t)
        out_t.backward(g)
    bad, n_cmp = [], 0
    for (n, pr), pf, pt in zip(ref.named_parameters(), fast.parameters(), torch_bf16.parameters()):
        assert pf.grad is not None, n
        if float(pr.grad.norm()) < 1e-6:
            continue
        r_k, r_t = _rel(pf.grad, pr.grad), _rel(pt.grad, pr.grad)
        if r_t > 1.0:            # the true gradient is ~0 (e.g. a FiLM shift cancelled by the next BatchNorm)
            continue
        n_cmp += 1
        if r_k > max(2.0 * r_t, 0.1):
            bad.append((n, round(r_k, 3), round(r_t, 3)))
    # a handful of tiny, noise-dominated gradients (SE weights of the first blocks) may land on either side
    assert n_cmp > 300 and len(bad) <= 12, bad[:20]
    # BN running statistics advanced the same way
    for (n, br), bf_ in zip(ref.named_buffers(), fast.buffers()):
        if "running_mean" in n:
            # bf16 activations 20+ blocks deep: a few per cent of drift against the fp32 module is rounding, not a bug
            # (the statistics are accumulated with fp32 atomics, so the last digits also vary from run to run)
            assert _rel(bf_, br) < 1.2e-1 or float(br.norm()) < 1e-3, n


def test_eval_batchnorm_affine_is_cached_and_invalidated_by_training():
    """Evaluation keeps BatchNorm scale/shift on the module; a training pass (which moves the running statistics through raw
    pointers) must invalidate them."""
    from rt1b200 import ops
    from rt1b200.models.tokenizers import RT1ImageTokenizer
    from rt1b200.parallel.engine import keep_bn_stats_fp32
    torch.manual_seed(0)
    net = RT1ImageTokenizer(512, use_token_learner=True, num_tokens=8).to(DEV).to(BF)
    keep_bn_stats_fp32(net)
    b, t, H, W = 2, 2, 64, 96
    img = torch.rand(b * t, 3, H, W, device=DEV)
    x8 = ops.reference.preprocess_nhwc(img, (4, 6)).to(DEV)
    ctx = _bf(b, t, 512, seed=5)
    bns = [m for m in net.modules() if isinstance(m, torch.nn.BatchNorm2d)]

    def drop_caches():
        for m in bns:
            if hasattr(m, "_rt1_eval_affine"):
                del m._rt1_eval_affine

    net.eval()
    with torch.no_grad():
        e0 = net.forward_nhwc(x8, ctx, b, t).clone()
        n0 = _lib.launch_count()
        e1 = net.forward_nhwc(x8, ctx, b, t).clone()
        cached_calls = _lib.launch_count() - n0
        drop_caches()
        n0 = _lib.launch_count()
        net.forward_nhwc(x8, ctx, b, t)
        uncached_calls = _lib.launch_count() - n0
    assert _rel(e1, e0) < 2e-2                                   # (the SE pooling sums with fp32 atomics: not bitwise repeatable)
    assert uncached_calls - cached_calls == len(bns), (uncached_calls, cached_calls, len(bns))
    net.train()
    net.forward_nhwc(x8 * 0.5, ctx, b, t)                      # moves every running mean / variance
    net.eval()
    with torch.no_grad():
        e2 = net.forward_nhwc(x8, ctx, b, t).clone()
        drop_caches()
        e3 = net.forward_nhwc(x8, ctx, b, t).clone()
    assert _rel(e2, e0) > 10 * max(_rel(e1, e0), 1e-3), "the running statistics changed: the evaluation output must change too"
    assert _rel(e2, e3) < 2e-2, "stale cached scale/shift after a training pass"
