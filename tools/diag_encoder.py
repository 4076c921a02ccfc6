"""Per-parameter gradient error of the NHWC-kernel encoder vs the fp32 module (diagnostic)."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rt1b200 import ops
from rt1b200.models.tokenizers import RT1ImageTokenizer
from rt1b200.parallel.engine import keep_bn_stats_fp32
DEV, BF = "cuda", torch.bfloat16
def rel(a, b):
    a, b = a.float().flatten(), b.float().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))
torch.manual_seed(0)
ref = RT1ImageTokenizer(512, use_token_learner=True, num_tokens=8).to(DEV)
for n, p in ref.named_parameters():
    if "_projection_" in n:
        torch.nn.init.normal_(p, std=0.02)
fast = copy.deepcopy(ref).to(BF); keep_bn_stats_fp32(fast)
with torch.no_grad():
    for pr, pf in zip(ref.parameters(), fast.parameters()):
        pr.copy_(pf.float())
for mdl in (ref, fast):
    for m in mdl.modules():
        if hasattr(m, "drop_rate"):
            m.drop_rate = 0.0
ref.train(); fast.train()
b, t, H, W = 4, 2, int(sys.argv[1]) if len(sys.argv) > 1 else 128, int(sys.argv[2]) if len(sys.argv) > 2 else 192
img = torch.rand(b, t, 3, H, W, device=DEV).to(BF)
g0 = torch.Generator().manual_seed(5)
ctx = torch.randn(b, t, 512, generator=g0).to(BF).to(DEV)
out_r = ref(img.float(), ctx.float())
x8 = ops.reference.preprocess_nhwc(img.reshape(b * t, 3, H, W).float(), (int(H * 0.07), int(W * 0.07))).to(DEV)
out_f = fast.forward_nhwc(x8, ctx, b, t)
print("fwd rel", rel(out_f, out_r))
g = torch.randn(*out_r.shape, generator=g0).to(BF).to(DEV)
out_r.backward(g.float()); out_f.backward(g)
# same comparison with the fp32 module run in bf16 channels-last torch ops (what error does bf16 alone give?)
t16 = copy.deepcopy(fast)
for p in t16.parameters():
    p.grad = None
with ops.cnn.force_torch_encoder():
    out_t = t16.forward_nhwc(x8, ctx, b, t); out_t.backward(g)
rows = []
for (n, pr), pf, pt in zip(ref.named_parameters(), fast.parameters(), t16.parameters()):
    if float(pr.grad.norm()) < 1e-8:
        continue
    rows.append((n, rel(pf.grad, pr.grad), rel(pt.grad, pr.grad) if pt.grad is not None else -1))
for n, r1, r2 in rows[::-1]:
    flag = " <<<" if r1 > 2 * max(r2, 0.02) else ""
    print(f"{r1:8.3f} {r2:8.3f}  {n}{flag}")
