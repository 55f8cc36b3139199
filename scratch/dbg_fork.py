import torch
from cloud_b200 import ops
dev = torch.device("cuda:0")
torch.manual_seed(13)
C = 128
x0 = (torch.randn(6, C, 14, 14, device=dev) * 2).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
w1 = torch.randn(C, C, 1, 1, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last) / 12
G = torch.randn(6, C, 14, 14, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
res = {}
for mode in ("fork", "plain", "manual"):
    x = x0.clone().requires_grad_(True)
    gamma = (torch.ones(C, device=dev)).requires_grad_(True)
    beta = torch.zeros(C, device=dev, requires_grad=True)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    a = ops.batch_norm_nhwc(x, gamma, beta, rm, rv, True, 0.1, 1e-5, relu=True)
    if mode == "fork":
        a1, a2 = ops.bn_fork(a)
    else:
        a1, a2 = a, a
    out = torch.nn.functional.conv2d(a1, w1) * 0.5 + a2 * 2.0
    out.backward(G)
    res[mode] = x.grad.float()
d = (res["fork"] - res["plain"]).abs()
print("max err", d.max().item(), "plain-vs-plain", (res["manual"] - res["plain"]).abs().max().item())
per_img = d.amax((1, 2, 3))
print("per image max err", per_img.tolist())
per_ch = d.amax((0, 2, 3))
print("bad channels", (per_ch > 0.05).nonzero().flatten().tolist()[:40])
ratio = res["fork"] / (res["plain"] + 1e-9)
bad = d > 0.05
print("ratio on bad elems: mean", ratio[bad].mean().item(), "count", int(bad.sum()), "of", d.numel())
