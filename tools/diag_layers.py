"""Per-layer forward check: each of our stored activations vs the oracle layer applied to OUR
previous activation (isolates a layer) — bf16 rounding emulated."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import gabril_carla_b200  # noqa
from tests.test_step_gpu import build
from tests.synth import make_batch
from oracle import gabril_oracle as O
spec, arena, eng = build(True, "None", "None")
cpu_batch = make_batch(4, seed=21)
eng.zero_grad(); eng.forward(*[t.to("cuda:0") for t in cpu_batch]); torch.cuda.synchronize()
sd = {k: v.cpu() for k, v in arena.state_dicts()["encoder"].items()}
q = lambda t: t.to(torch.bfloat16).float()
def nchw(name): return eng.buf[name].float().cpu().permute(0, 3, 1, 2)
def rel(a, b): return ((a - b).norm() / b.norm()).item(), (a - b).abs().max().item()
obs = O.format_obs(O.gather_stack(cpu_batch[0], 1, 2), True)
def conv(name, t, s, p): return F.conv2d(t, q(sd[name + ".weight"]), sd.get(name + ".bias"), stride=s, padding=p)
print("a1", rel(nchw("a1"), q(F.relu(conv("_conv_1", q(obs), 2, 1)))))
print("a2", rel(nchw("a2"), q(F.relu(conv("_conv_2", nchw("a1"), 2, 1)))))
print("a3", rel(nchw("a3"), q(F.relu(conv("_conv_3", nchw("a2"), 2, 1)))))
print("r-1", rel(nchw("r-1"), q(F.relu(conv("_conv_4", nchw("a3"), 1, 0)))))
for i in range(2):
    prev = nchw(f"r{i-1}")
    h = q(F.relu(F.conv2d(prev, q(sd[f"_residual_stack._layers.{i}._block.1.weight"]), None, padding=1)))
    print(f"h{i}", rel(nchw(f"h{i}"), h))
    r = q(F.relu(prev + F.conv2d(nchw(f"h{i}"), q(sd[f"_residual_stack._layers.{i}._block.3.weight"]), None)))
    print(f"r{i}", rel(nchw(f"r{i}"), r))
print("a5", rel(nchw("a5"), q(F.relu(conv("_conv_5", nchw("r1"), 1, 2)))))
print("z", rel(nchw("z"), q(conv("_conv_6", nchw("a5"), 1, 2))))
print("---- chained (oracle consumes its own activations) ----")
t = q(F.relu(conv("_conv_1", q(obs), 2, 1))); print("a1", rel(nchw("a1"), t))
t = q(F.relu(conv("_conv_2", t, 2, 1))); print("a2", rel(nchw("a2"), t))
t = q(F.relu(conv("_conv_3", t, 2, 1))); print("a3", rel(nchw("a3"), t))
t = q(F.relu(conv("_conv_4", t, 1, 0))); print("r-1", rel(nchw("r-1"), t))
for i in range(2):
    h = q(F.relu(F.conv2d(t, q(sd[f"_residual_stack._layers.{i}._block.1.weight"]), None, padding=1)))
    print(f"h{i}", rel(nchw(f"h{i}"), h))
    t = q(F.relu(t + F.conv2d(h, q(sd[f"_residual_stack._layers.{i}._block.3.weight"]), None)))
    print(f"r{i}", rel(nchw(f"r{i}"), t))
t = q(F.relu(conv("_conv_5", t, 1, 2))); print("a5", rel(nchw("a5"), t))
t = q(conv("_conv_6", t, 1, 2)); print("z", rel(nchw("z"), t))
z2 = O.encoder_forward(sd, obs, emulate_bf16=True)
print("z vs O.encoder_forward(emu)", rel(nchw("z"), z2), "chain vs O", rel(t, z2))
