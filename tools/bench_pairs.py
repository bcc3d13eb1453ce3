"""conv_6 forward: gabril_conv_fwd vs gabril_conv_fwd_pairs (adjacent output pixels stacked along N) at batch B."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gabril_carla_b200  # noqa
from gabril_carla_b200 import _lib, ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
_lib.init()
dev, bf = "cuda:0", torch.bfloat16
x = (torch.randn(B, 20, 38, 128, device=dev) * 0.5).to(bf)
w = (torch.randn(64, 25, 128, device=dev) * 0.05).to(bf)
wp = torch.empty(128, 30, 128, device=dev, dtype=bf)
bias = torch.randn(64, device=dev)
y = torch.empty(B, 20, 38, 64, device=dev, dtype=bf)


def timeit(fn, reps=9):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


print("conv_fwd        %.4f ms" % timeit(lambda: ops.conv_fwd(x, w, bias, y, 5, 5, 1, 2)))
print("pack_wpairs     %.4f ms" % timeit(lambda: ops.conv_pack_wpairs(w, wp, 5, 5)))
print("conv_fwd_pairs  %.4f ms" % timeit(lambda: ops.conv_fwd_pairs(x, wp, bias, y, 5, 5, 2)))
_lib.poll_error()
