"""conv6 wgrad at batch 64: ours vs torch GPU fp32 vs CPU fp64; prints the error pattern."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib, ops
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda:0"); _lib.init()
N, H, W, Cin, Cout, R, S, stride, pad = 64, 20, 38, 128, 64, 5, 5, 1, 2
g = torch.Generator().manual_seed(5)
x = (torch.randn(N, H, W, Cin, generator=g) * 0.5).to(torch.bfloat16)
dy = (torch.randn(N, H, W, Cout, generator=g) * 0.5).to(torch.bfloat16)
ref64 = torch.nn.grad.conv2d_weight(x.double().permute(0, 3, 1, 2), (Cout, Cin, R, S),
                                    dy.double().permute(0, 3, 1, 2), stride=stride, padding=pad)
ref64 = ops.w_oihw_to_internal(ref64)
xd, dyd = x.to(dev), dy.to(dev)
for trial in range(3):
    dw = torch.zeros(Cout, R * S, Cin, device=dev)
    ops.conv_wgrad(dyd, xd, dw, R, S, stride, pad)
    torch.cuda.synchronize()
    e = (dw.cpu().double() - ref64)
    print("ours  trial", trial, "rel", float(e.norm() / ref64.norm()), "max", float(e.abs().max()))
    bad = (e.abs() > 0.05).nonzero()
    print("  n_bad", len(bad), "co range", bad[:, 0].min().item() if len(bad) else None,
          bad[:, 0].max().item() if len(bad) else None, "taps", sorted(set(bad[:, 1].tolist()))[:30],
          "ci", sorted(set(bad[:, 2].tolist()))[:10], "...")
tg = torch.nn.grad.conv2d_weight(xd.float().permute(0, 3, 1, 2), (Cout, Cin, R, S),
                                 dyd.float().permute(0, 3, 1, 2), stride=stride, padding=pad)
e = ops.w_oihw_to_internal(tg).cpu().double() - ref64
print("torch-gpu rel", float(e.norm() / ref64.norm()), "max", float(e.abs().max()))
