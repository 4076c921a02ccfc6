"""Numerics of every sm_100a kernel against a float32 PyTorch statement of the same op (rt1b200.ops.reference).

Inputs are bf16-rounded once and shared by both sides, so the tolerance only has to absorb bf16 output rounding
and accumulation-order differences.
"""
import math

import pytest
import torch

from rt1b200 import ops
from rt1b200.ops import reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF = torch.bfloat16


def _k():
    from rt1b200.ops import kernels
    return kernels


def _bf(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(BF).to(DEV)


def _close(a, b, rtol=2e-2, atol=2e-2, msg=""):
    torch.testing.assert_close(a.float(), b.float(), rtol=rtol, atol=atol, msg=lambda m: f"{msg}: {m}")


def test_library_loads_and_reports_blackwell():
    assert ops.kernels_available(), "native library must be present on a GPU box"
    assert torch.cuda.get_device_capability()[0] >= 10


@pytest.mark.parametrize("M,N,K", [(528, 512, 512), (528, 3072, 512), (528, 512, 1024), (144, 256, 512),
                                   (1000, 24, 40), (4096, 144, 24), (300, 816, 136), (5760, 1536, 384), (77, 40, 72)])
def test_gemm_kmajor_variants(M, N, K):
    k = _k()
    a, b = _bf(M, K, seed=1), _bf(N, K, scale=1 / math.sqrt(K), seed=2)
    out = torch.empty(M, N, dtype=BF, device=DEV)
    k.gemm(a, b, out, M, N, K)
    _close(out, a.float() @ b.float().t(), msg=f"gemm {M}x{N}x{K}")


@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(528, 512, 1024), (512, 3072, 528), (24, 144, 5000), (136, 816, 1392), (200, 72, 96)])
def test_gemm_mn_major_operands(a_mn, b_mn, M, N, K):
    k = _k()
    a_l, b_l = _bf(M, K, seed=3), _bf(N, K, scale=1 / math.sqrt(K), seed=4)
    a = a_l.t().contiguous() if a_mn else a_l          # stored [K, M] when MN-major
    b = b_l.t().contiguous() if b_mn else b_l
    out = torch.empty(M, N, dtype=BF, device=DEV)
    k.gemm(a, b, out, M, N, K, a_mn=a_mn, b_mn=b_mn)
    _close(out, a_l.float() @ b_l.float().t(), msg=f"gemm a_mn={a_mn} b_mn={b_mn} {M}x{N}x{K}")


@pytest.mark.parametrize("M,N,K,stats", [(1000, 144, 24, True), (70000, 144, 24, True), (70000, 24, 144, False),
                                         (20000, 192, 32, True), (4100, 40, 32, True), (333, 200, 72, False)])
def test_gemm_persistent_many_tiles_stats_and_strided_output(M, N, K, stats):
    """The persistent kernel's regimes: long tile runs per CTA, wide + narrow n-tiles (N = 128 + 16), resident weights,
    a ragged last m-tile, BatchNorm column statistics, and an output that is a column slice of a wider buffer."""
    k = _k()
    a, b = _bf(M, K, seed=11), _bf(N, K, scale=1 / math.sqrt(K), seed=12)
    ld = ((N + 7) // 8) * 8 + 16
    wide = torch.full((M, ld), 7.0, dtype=BF, device=DEV)
    out = wide[:, 8:8 + N]                                   # 16-byte aligned column slice, row stride ld
    s = torch.zeros(N, device=DEV) if stats else None
    q = torch.zeros(N, device=DEV) if stats else None
    k.gemm(a, b, out, M, N, K, col_sum=s, col_sumsq=q)
    y = a.float() @ b.float().t()
    _close(out, y, msg=f"persistent {M}x{N}x{K}")
    assert float((wide[:, :8] - 7.0).abs().max()) == 0.0 and float((wide[:, 8 + N:] - 7.0).abs().max()) == 0.0   # nothing outside
    if stats:
        yb = y.to(BF).float()                                # the statistics describe the stored (bf16) values
        _close(s, yb.sum(0), rtol=1e-2, atol=0.02 * math.sqrt(M), msg="col_sum")
        _close(q, (yb * yb).sum(0), rtol=1e-2, atol=0.02 * math.sqrt(M), msg="col_sumsq")


def test_gemm_epilogue_bias_residual_dropout_stats_splitk():
    k = _k()
    M, N, K = 528, 512, 512
    a, b = _bf(M, K, seed=5), _bf(N, K, scale=1 / math.sqrt(K), seed=6)
    bias, res = _bf(N, seed=7), _bf(M, N, seed=8)
    out = torch.empty(M, N, dtype=BF, device=DEV)
    k.gemm(a, b, out, M, N, K, bias=bias, residual=res, dropout_p=0.1, seed=99)
    want = ref.linear(a.float(), b.float(), bias.float(), res.float(), None, 0.1, 99)
    _close(out, want, msg="epilogue")
    # column statistics of the fp32 result
    s, q = torch.zeros(N, device=DEV), torch.zeros(N, device=DEV)
    k.gemm(a, b, out, M, N, K, col_sum=s, col_sumsq=q)
    y = a.float() @ b.float().t()
    _close(s, y.sum(0), rtol=1e-2, atol=0.5, msg="col_sum")
    _close(q, (y * y).sum(0), rtol=1e-2, atol=0.5, msg="col_sumsq")
    # split-K into a zeroed fp32 buffer (weight-gradient shape: tiny output, long reduction)
    M2, N2, K2 = 144, 24, 60000
    a2, b2 = _bf(K2, M2, seed=9), _bf(K2, N2, scale=0.05, seed=10)
    acc = torch.zeros(M2, N2, device=DEV)
    k.gemm(a2, b2, acc, M2, N2, K2, a_mn=True, b_mn=True, split_k=64)
    _close(acc, a2.float().t() @ b2.float(), rtol=2e-2, atol=5e-2, msg="split-k")


def test_linear_autograd_matches_reference():
    torch.manual_seed(0)
    M, K, N = 528, 512, 1024
    x = _bf(M, K, seed=11).requires_grad_()
    lin = torch.nn.Linear(K, N).to(DEV, BF)
    res = _bf(M, N, seed=12).requires_grad_()
    y = ops.linear(x, lin.weight, lin.bias, residual=res)
    g = _bf(M, N, seed=13)
    y.backward(g)
    xr, wr, br, rr = (t.detach().float().requires_grad_() for t in (x, lin.weight, lin.bias, res))
    yr = ref.linear(xr, wr, br, rr)
    yr.backward(g.float())
    _close(y, yr, msg="y")
    _close(x.grad, xr.grad, msg="dx")
    _close(lin.weight.grad, wr.grad, rtol=3e-2, atol=0.15, msg="dw")
    _close(lin.bias.grad, br.grad, rtol=3e-2, atol=0.3, msg="db")
    _close(res.grad, rr.grad, msg="dres")


@pytest.mark.parametrize("D", [512, 1024])
def test_layernorm_fwd_bwd(D):
    rows = 528
    x = (_bf(rows, D, seed=1) * 2 + 0.5).detach().requires_grad_()
    ln = torch.nn.LayerNorm(D).to(DEV, BF)
    with torch.no_grad():
        ln.weight.copy_(_bf(D, seed=2) * 0.2 + 1)
        ln.bias.copy_(_bf(D, seed=3) * 0.2)
    y = ops.layer_norm(x, ln.weight, ln.bias, ln.eps)
    g = _bf(rows, D, seed=4)
    y.backward(g)
    xr, wr, br = (t.detach().float().requires_grad_() for t in (x, ln.weight, ln.bias))
    yr = ref.layer_norm(xr, wr, br, ln.eps)
    yr.backward(g.float())
    _close(y, yr, msg="y")
    _close(x.grad, xr.grad, msg="dx")
    _close(ln.weight.grad, wr.grad, rtol=3e-2, atol=0.3, msg="dw")
    _close(ln.bias.grad, br.grad, rtol=3e-2, atol=0.3, msg="db")


@pytest.mark.parametrize("p_drop", [0.0, 0.1])
@pytest.mark.parametrize("b,s,h,dk", [(8, 66, 8, 128), (2, 33, 4, 16)])
def test_attention_fwd_bwd(b, s, h, dk, p_drop):
    from rt1b200.models import RT1Config, build_rt1
    qkv = _bf(b, s, 3, h, dk, seed=5).requires_grad_()
    if s == 66:
        mask = build_rt1(RT1Config.tiny(seq_len=6))._default_attention_mask.to(DEV)
    else:
        mask = build_rt1(RT1Config.tiny(seq_len=3))._default_attention_mask.to(DEV)
    assert mask.shape == (s, s)
    out = _k().rt1_attention(qkv, mask, p_drop, seed=77)
    g = _bf(b, s, h * dk, seed=6)
    out.backward(g)
    qr = qkv.detach().float().requires_grad_()
    outr = ref.rt1_attention(qr, mask, p_drop, 77)
    outr.backward(g.float())
    _close(out, outr, msg="out")
    _close(qkv.grad, qr.grad, rtol=3e-2, atol=3e-2, msg="dqkv")


def test_attention_is_causal_and_blocks_action_keys():
    from rt1b200.models import RT1Config, build_rt1
    mask = build_rt1(RT1Config.tiny(seq_len=6))._default_attention_mask.to(DEV)
    qkv = _bf(2, 66, 3, 8, 128, seed=8)
    out = _k().rt1_attention(qkv, mask, 0.0)
    qkv2 = qkv.clone()
    qkv2[:, 40:] = 0                       # zero everything from position 40 on
    out2 = _k().rt1_attention(qkv2, mask, 0.0)
    assert torch.equal(out[:, :40], out2[:, :40])


def test_action_tokenizer_kernel_matches_torch_path():
    """rt1_action_tokenize / rt1_action_detokenize (CUDA) against the PyTorch statement of the same rules on the CPU."""
    import collections
    from rt1b200 import spaces
    from rt1b200.models.tokenizers import RT1ActionTokenizer
    space = spaces.Dict(collections.OrderedDict(terminate_episode=spaces.Discrete(2),
                                                action=spaces.Box(low=-0.1, high=0.1, shape=(2,), dtype="float32"),
                                                gripper=spaces.Box(low=0.0, high=1.0, shape=(1,), dtype="float32")))
    tk = RT1ActionTokenizer(space, vocab_size=256)
    g = torch.Generator().manual_seed(3)
    act = {"terminate_episode": torch.randint(0, 2, (8, 6), generator=g),
           "action": (torch.rand(8, 6, 2, generator=g) - 0.5) * 0.3,          # some values outside the box: clamped
           "gripper": torch.rand(8, 6, 1, generator=g) * 1.2 - 0.1}
    want = tk.tokenize(act)                                                   # CPU tensors: PyTorch path
    got = tk.tokenize({k: v.to(DEV) for k, v in act.items()})
    assert got.dtype == torch.int64 and torch.equal(got.cpu(), want)
    toks = torch.randint(0, 256, (8, 6, 4), generator=g)
    toks[..., 0] = torch.randint(0, 4, (8, 6), generator=g)                   # includes out-of-vocabulary discrete tokens
    back_c, back_g = tk.detokenize(toks), tk.detokenize(toks.to(DEV))
    for k in back_c:
        torch.testing.assert_close(back_g[k].cpu().to(back_c[k].dtype), back_c[k], rtol=1e-6, atol=1e-7)


def test_action_ce_loss_fwd_bwd():
    R, A, V = 48, 3, 256
    logits = (_bf(R, A, V, seed=9) * 3).detach().requires_grad_()
    labels = torch.randint(0, V, (R, A), device=DEV)
    num_items = 48 * 11.0
    loss, pred = ops.action_ce_loss(logits, labels, num_items)
    w = torch.rand(R, device=DEV)
    (loss * w).sum().backward()
    lr = logits.detach().float().requires_grad_()
    lossr, predr = ref.action_ce_loss(lr, labels, num_items)
    (lossr * w).sum().backward()
    _close(loss, lossr, rtol=1e-3, atol=1e-6, msg="loss")
    assert torch.equal(pred, predr)
    _close(logits.grad, lr.grad, rtol=2e-2, atol=1e-6, msg="dlogits")


def test_token_learner_fwd_bwd():
    from rt1b200.models.token_learner import TokenLearnerModule
    torch.manual_seed(1)
    n, P, C = 6, 120, 512
    tl = TokenLearnerModule(C, 8).to(DEV, BF)
    feats = _bf(n, P, C, seed=10).requires_grad_()
    tok = ops.token_learner(feats, tl.layerNorm.weight, tl.layerNorm.bias, tl.conv1.weight, tl.conv1.bias,
                            tl.conv2.weight, tl.conv2.bias, tl.layerNorm.eps)
    g = _bf(n, 8, C, seed=11)
    tok.backward(g)
    params = [tl.layerNorm.weight, tl.layerNorm.bias, tl.conv1.weight, tl.conv1.bias, tl.conv2.weight, tl.conv2.bias]
    fr = feats.detach().float().requires_grad_()
    pr = [p.detach().float().requires_grad_() for p in params]
    tokr = ref.token_learner(fr, *pr, tl.layerNorm.eps)
    tokr.backward(g.float())
    _close(tok, tokr, msg="tokens")
    _close(feats.grad, fr.grad, rtol=3e-2, atol=3e-2, msg="dfeats")
    for p, r, name in zip(params, pr, ["ln_w", "ln_b", "w1", "b1", "w2", "b2"]):
        if name == "b2":      # softmax over positions: d(logit) sums to zero per token, so this gradient is ~0
            assert float(p.grad.float().abs().max()) < 2e-2 and float(r.grad.abs().max()) < 1e-3
            continue
        scale = float(r.grad.abs().max()) + 1e-6
        _close(p.grad.reshape(r.grad.shape) / scale, r.grad / scale, rtol=5e-2, atol=3e-2, msg=name)


@pytest.mark.parametrize("dtype", [torch.uint8, torch.float32])
def test_preprocess_nhwc(dtype):
    N, H, W = 4, 64, 96
    if dtype == torch.uint8:
        img = torch.randint(0, 256, (N, 3, H, W), dtype=torch.uint8, device=DEV)
    else:
        img = torch.rand(N, 3, H, W, device=DEV)
    for shift in [(0, 0), (4, 6), (8, 12), (3, 0)]:
        out = ops.preprocess_nhwc(img, shift)
        want = ref.preprocess_nhwc(img, shift)
        assert out.shape == (N, H, W, 8) and out.dtype == BF
        _close(out, want, rtol=1e-2, atol=4e-3, msg=f"shift {shift}")


@pytest.mark.parametrize("grad_dtype", [BF, torch.float32])
def test_fused_adamw_matches_torch_adam(grad_dtype):
    n = 64 * 1000 + 64
    torch.manual_seed(0)
    p0 = torch.randn(n, device=DEV)
    master, m, v = p0.clone(), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    lp = master.to(BF)
    pt = p0.clone().requires_grad_()
    opt = torch.optim.Adam([pt], lr=5e-4)
    for step in range(1, 4):
        g = (torch.randn(n, device=DEV) * 0.1).to(grad_dtype)
        gk = g.clone()
        h = dict(lr=5e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, bc1=1 - 0.9 ** step, bc2=1 - 0.999 ** step)
        _k().fused_adamw(lp, master, gk, m, v, h, grad_scale=1.0, zero_grad=True)
        pt.grad = g.float()
        opt.step()
        torch.testing.assert_close(master, pt.detach(), rtol=1e-5, atol=1e-6)
        assert float(gk.abs().max()) == 0.0
        torch.testing.assert_close(lp.float(), master.to(BF).float())
