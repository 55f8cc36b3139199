import math, torch
from cloud_b200 import ops
dev = torch.device("cuda:0")
torch.manual_seed(8)
N, H, Cin, Cout = 8, 28, 128, 256
x = torch.randn(N, Cin, H, H, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
w = (torch.randn(Cout, Cin, 1, 1, device=dev) / math.sqrt(Cin)).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
gamma = (torch.rand(Cout, device=dev) + 0.5).requires_grad_(True)
beta = (torch.randn(Cout, device=dev) * 0.1).requires_grad_(True)
y1 = ops.conv1x1(x, w, None, bn_stats=True)
s = y1._cb2_bn_sums.view(8, 2, Cout).sum(0)
yf = y1.float()
print("sum err", (s[0] - yf.sum((0, 2, 3))).abs().max().item(), "sq err", (s[1] - (yf * yf).sum((0, 2, 3))).abs().max().item())
y2 = ops.conv1x1(x, w, None, bn_stats=False)
print("y equal", (y1.float() - y2.float()).abs().max().item())
for sums in (y1._cb2_bn_sums, None):
    rm, rv = torch.zeros(Cout, device=dev), torch.ones(Cout, device=dev)
    z = ops.batch_norm_nhwc(y1, gamma, beta, rm, rv, True, 0.1, 1e-5, relu=True, sums=sums)
    ref = torch.relu(torch.nn.functional.batch_norm(yf, None, None, gamma, beta, True, 0.1, 1e-5))
    print("fused" if sums is not None else "plain", "err vs torch", (z.float() - ref).abs().max().item(), "rm", rm[:4].tolist(), "mean", (yf.mean((0, 2, 3))[:4] * 0.1).tolist())
    if sums is not None:
        print("sums after", (sums.view(8, 2, Cout).sum(0)[0] - yf.sum((0, 2, 3))).abs().max().item())
