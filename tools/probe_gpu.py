"""One-off probes of how the reference's stack (torch CUDA) rounds, run on the GPU box."""
import numpy as np
import torch
import torch.nn.functional as F

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
torch.manual_seed(0)
for dt in (torch.float16, torch.bfloat16):
    for cin, cout, hw in ((1, 16, 28), (16, 32, 14)):
        x = torch.rand(32, cin, hw, hw, device="cuda").to(dt)
        w = ((torch.rand(cout, cin, 5, 5, device="cuda") - 0.5) * 0.2).to(dt)
        b = ((torch.rand(cout, device="cuda") - 0.5) * 0.2).to(dt)
        ref = F.conv2d(x, w, b, 1, 2).float()
        acc = F.conv2d(x.float(), w.float(), None, 1, 2)
        single = (acc + b.float().view(1, -1, 1, 1)).to(dt).float()
        double = (acc.to(dt).float() + b.float().view(1, -1, 1, 1)).to(dt).float()
        print(f"conv {dt} {cin}->{cout}: mismatch single {(ref != single).float().mean().item():.5f} "
              f"double {(ref != double).float().mean().item():.5f}")
# SGD foreach add: fused multiply-add or mul then add?
p = torch.randn(1 << 16, device="cuda")
g = torch.randn(1 << 16, device="cuda")
lr = 1e-4
q = p.clone()
torch._foreach_add_([q], [g], alpha=-lr)
pc, gc = p.cpu().numpy(), g.cpu().numpy()
a32 = np.float32(-lr)
fma = (pc.astype(np.float64) + np.float64(a32) * gc.astype(np.float64)).astype(np.float32)
mul_add = (pc + (a32 * gc).astype(np.float32)).astype(np.float32)
qc = q.cpu().numpy()
print("foreach_add mismatches: fma", int((qc != fma).sum()), "mul+add", int((qc != mul_add).sum()))
# BN train fp16: which formula / stats
x = (torch.randn(32, 16, 28, 28, device="cuda") * 0.3 + 0.1).half()
bn = torch.nn.BatchNorm2d(16).cuda()
bn.weight.data.uniform_(0.5, 1.5); bn.bias.data.uniform_(-0.5, 0.5)
with torch.autocast("cuda", dtype=torch.float16):
    out = bn(x)
xf = x.float()
mean = xf.mean((0, 2, 3)); var = xf.var((0, 2, 3), unbiased=False)
invstd = 1 / torch.sqrt(var + 1e-5)
c = lambda t: t.view(1, -1, 1, 1)
f1 = ((xf - c(mean)) * c(invstd) * c(bn.weight) + c(bn.bias)).half()
f2 = (xf * c(invstd * bn.weight) + c(bn.bias - mean * invstd * bn.weight)).half()
print("bn out dtype", out.dtype, "mismatch (x-m)*is*g+b:", (out != f1).float().mean().item(), " scale/shift:", (out != f2).float().mean().item())
print("running_var rel err vs unbiased:", ((bn.running_var - (0.9 + 0.1 * xf.var((0, 2, 3), unbiased=True))).abs().max()).item())
print(torch.cuda.get_device_name(0), torch.cuda.get_device_properties(0).multi_processor_count)
