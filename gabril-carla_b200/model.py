"""Parameter arena and layer plan of the BC policy (Encoder + pre_actor + actor).

Reference: vlm_gaze/models/linear_models.py:124-207 (Encoder), :78-121 (Residual / ResidualStack),
:6-16 (weight_init), vlm_gaze/train/train_bc.py:57-82 (build_models).

All parameters live in ONE flat fp32 arena (master weights) with parallel flat arenas for the
gradient, Adam moments and the bf16 shadow the tensor-core kernels read.  Entries are laid out in
REVERSE execution order (actor, pre_actor, conv_6 ... conv_1) so that the gradient buckets that are
ready first during backward sit at the front and can be all-reduced while the encoder backward is
still running.  Weights are kept in the kernels' internal layout ([Cout][R*S][Cin], pre_actor
columns permuted to NHWC flatten order, conv_1 as a [Cout][Kp] patch-GEMM matrix) and converted
from / to the reference's state_dict layout (fp32, OIHW, NCHW flatten) only on load / save, so
checkpoints stay drop-in for vlm_gaze/eval/my_agents/bc_agent.py:96-128.
"""
from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch

ALIGN = 64  # elements; 256 B in fp32, 128 B in the bf16 shadow (TMA needs 16 B)


def conv_out(n, k, s, p):
    return (n + 2 * p - k) // s + 1


@dataclass
class Entry:
    name: str            # internal name, e.g. "enc.conv2.w"
    module: str          # "encoder" | "pre_actor" | "actor"
    key: str             # reference state_dict key inside that module
    kind: str            # "conv" | "conv1" | "linear" | "pre_actor" | "head" | "bias" | "head_bias"
    ref_shape: Tuple[int, ...]
    int_shape: Tuple[int, ...]
    offset: int = 0

    @property
    def numel(self):
        n = 1
        for d in self.int_shape:
            n *= d
        return n


class PolicySpec:
    """Static shapes of the network for one configuration."""

    def __init__(self, in_channels: int, H: int = 180, W: int = 320, embedding_dim: int = 64,
                 num_hiddens: int = 128, num_residual_layers: int = 2,
                 num_residual_hiddens: int = 32, z_dim: int = 256, action_dim: int = 7,
                 agil_channels: int = 0, gril_points: int = 0):
        # agil_channels > 0: gaze.method=AGIL's second encoder `encoder_agil` (train_bc.py:70-77) with that many
        # input channels; gril_points > 0: GRIL's gaze-coordinate head on z_flat (train_bc.py:84-86), 2*points outputs
        self.C0a, self.GP = int(agil_channels), int(gril_points)
        if self.GP * 2 > 32:
            raise NotImplementedError("gaze.max_points > 16")
        self.C0, self.H, self.W = in_channels, H, W
        self.E, self.NH, self.NR, self.RH = embedding_dim, num_hiddens, num_residual_layers, num_residual_hiddens
        self.Z, self.A = z_dim, action_dim
        self.C1, self.C2 = num_hiddens // 4, num_hiddens // 2
        for c, what in ((self.C1, "num_hiddens/4"), (self.C2, "num_hiddens/2"), (self.NH, "num_hiddens"),
                        (self.RH, "num_residual_hiddens"), (self.E, "embedding_dim"), (self.Z, "z_dim")):
            if c % 32:
                raise NotImplementedError(f"{what}={c}: the tcgen05 kernels need channel counts that are multiples of 32")
        if action_dim > 32:
            raise NotImplementedError("action_dim > 32")
        self.K1 = 16 * in_channels
        # zero-padded K of the conv_1 patch GEMM: 32, 64 or a multiple of 128 (wgrad N-side rule)
        self.K1p = 32 if self.K1 <= 32 else 64 if self.K1 <= 64 else (self.K1 + 127) // 128 * 128
        self.H1, self.W1 = conv_out(H, 4, 2, 1), conv_out(W, 4, 2, 1)
        self.H2, self.W2 = conv_out(self.H1, 4, 2, 1), conv_out(self.W1, 4, 2, 1)
        self.H3, self.W3 = conv_out(self.H2, 4, 2, 1), conv_out(self.W2, 4, 2, 1)
        self.h, self.w = conv_out(self.H3, 3, 1, 0), conv_out(self.W3, 3, 1, 0)
        self.n = self.h * self.w
        self.flat = self.n * self.E  # 48640 for 180x320 (hard-coded in the reference, train_bc.py:78)
        self.AP = 32                 # action head rows padded to one UMMA N tile

    def entries(self) -> List[Entry]:
        s = self
        e: List[Entry] = []
        # --- reverse execution order ---
        if s.GP:
            g = "gril_gaze_coord_predictor"
            e.append(Entry("gril.2.w", g, "2.weight", "head", (2 * s.GP, s.Z), (s.AP, 1, s.Z)))
            e.append(Entry("gril.2.b", g, "2.bias", "head_bias", (2 * s.GP,), (s.AP,)))
            e.append(Entry("gril.0.w", g, "0.weight", "linear", (s.Z, s.Z), (s.Z, 1, s.Z)))
            e.append(Entry("gril.0.b", g, "0.bias", "bias", (s.Z,), (s.Z,)))
        e.append(Entry("act.2.w", "actor", "2.weight", "head", (s.A, s.Z), (s.AP, 1, s.Z)))
        e.append(Entry("act.2.b", "actor", "2.bias", "head_bias", (s.A,), (s.AP,)))
        e.append(Entry("act.0.w", "actor", "0.weight", "linear", (s.Z, s.Z), (s.Z, 1, s.Z)))
        e.append(Entry("act.0.b", "actor", "0.bias", "bias", (s.Z,), (s.Z,)))
        e.append(Entry("pre.w", "pre_actor", "1.weight", "pre_actor", (s.Z, s.flat), (s.Z, 1, s.flat)))
        e.append(Entry("pre.b", "pre_actor", "1.bias", "bias", (s.Z,), (s.Z,)))

        def conv(name, key, co, ci, k):
            e.append(Entry(f"enc.{name}.w", "encoder", f"{key}.weight", "conv", (co, ci, k, k), (co, k * k, ci)))

        def bias(name, key, co):
            e.append(Entry(f"enc.{name}.b", "encoder", f"{key}.bias", "bias", (co,), (co,)))
        self._encoder_entries(e, conv, bias)
        if s.C0a:   # second encoder: same layers, its own input width, module "encoder_agil"
            self._encoder_entries(e, conv, bias, pfx="agil", module="encoder_agil", c0=s.C0a)
        off = 0
        for ent in e:
            ent.offset = off
            off += (ent.numel + ALIGN - 1) // ALIGN * ALIGN
        self.total = off
        return e

    def k1p_of(self, c0):
        k1 = 16 * c0
        return 32 if k1 <= 32 else 64 if k1 <= 64 else (k1 + 127) // 128 * 128

    def _encoder_entries(self, e, conv, bias, pfx="enc", module="encoder", c0=None):
        s = self
        if pfx != "enc":
            def conv(name, key, co, ci, k):   # noqa: F811
                e.append(Entry(f"{pfx}.{name}.w", module, f"{key}.weight", "conv", (co, ci, k, k), (co, k * k, ci)))

            def bias(name, key, co):          # noqa: F811
                e.append(Entry(f"{pfx}.{name}.b", module, f"{key}.bias", "bias", (co,), (co,)))
        c0 = s.C0 if c0 is None else c0
        conv("conv6", "_conv_6", s.E, s.NH, 5); bias("conv6", "_conv_6", s.E)
        conv("conv5", "_conv_5", s.NH, s.NH, 5); bias("conv5", "_conv_5", s.NH)
        for i in reversed(range(s.NR)):
            conv(f"res{i}.c1", f"_residual_stack._layers.{i}._block.3", s.NH, s.RH, 1)
            conv(f"res{i}.c3", f"_residual_stack._layers.{i}._block.1", s.RH, s.NH, 3)
        conv("conv4", "_conv_4", s.NH, s.NH, 3); bias("conv4", "_conv_4", s.NH)
        conv("conv3", "_conv_3", s.NH, s.C2, 4); bias("conv3", "_conv_3", s.NH)
        # conv_2 (k4/s2/p1) runs as a 2x2 stride-1 convolution over the space-to-depth conv_1 output:
        # its 16 taps are stored as [Cout][(r',s')][(qh,qw,cin)] with (kh, kw) = (2r'+qh, 2s'+qw)
        e.append(Entry(f"{pfx}.conv2.w", module, "_conv_2.weight", "conv_s2d", (s.C2, s.C1, 4, 4),
                       (s.C2, 4, 4 * s.C1)))
        bias("conv2", "_conv_2", s.C2)
        e.append(Entry(f"{pfx}.conv1.w", module, "_conv_1.weight", "conv1", (s.C1, c0, 4, 4), (s.C1, 1, s.k1p_of(c0))))
        bias("conv1", "_conv_1", s.C1)


class GazePredictorSpec(PolicySpec):
    """AutoEncoder(Encoder, Decoder(out_channels=1)) of train_gaze_predictor.py:50-64
    (Decoder: vlm_gaze/models/linear_models.py:210-282).  Entries in reverse execution order:
    decoder (conv_trans_4 ... conv_1) first, then the encoder."""

    def entries(self) -> List[Entry]:
        s = self
        e: List[Entry] = []

        def dconv(name, key, shape_ref, kind="conv"):
            o, i, k, _ = shape_ref
            int_shape = (o, 1, 32) if kind == "patch16" else (o, k * k, i)
            e.append(Entry(f"dec.{name}.w", "decoder", f"{key}.weight", kind, shape_ref, int_shape))

        def dbias(name, key, co):
            e.append(Entry(f"dec.{name}.b", "decoder", f"{key}.bias", "bias", (co,), (co,)))
        # nn.ConvTranspose2d weights are [in, out, k, k]: the "conv" layout of that tensor is exactly the
        # internal weight of the convolution whose data gradient the transposed convolution is
        dconv("ct4", "_conv_trans_4", (s.C1, 1, 4, 4), "patch16"); dbias("ct4", "_conv_trans_4", 1)
        dconv("ct3", "_conv_trans_3", (s.C2, s.C1, 4, 4)); dbias("ct3", "_conv_trans_3", s.C1)
        dconv("ct2", "_conv_trans_2", (s.NH, s.C2, 4, 4)); dbias("ct2", "_conv_trans_2", s.C2)
        dconv("ct1", "_conv_trans_1", (s.NH, s.NH, 3, 3)); dbias("ct1", "_conv_trans_1", s.NH)
        for i in reversed(range(s.NR)):
            dconv(f"res{i}.c1", f"_residual_stack._layers.{i}._block.3", (s.NH, s.RH, 1, 1))
            dconv(f"res{i}.c3", f"_residual_stack._layers.{i}._block.1", (s.RH, s.NH, 3, 3))
        dconv("conv1", "_conv_1", (s.NH, s.E, 3, 3)); dbias("conv1", "_conv_1", s.NH)

        def conv(name, key, co, ci, k):
            e.append(Entry(f"enc.{name}.w", "encoder", f"{key}.weight", "conv", (co, ci, k, k), (co, k * k, ci)))

        def bias(name, key, co):
            e.append(Entry(f"enc.{name}.b", "encoder", f"{key}.bias", "bias", (co,), (co,)))
        self._encoder_entries(e, conv, bias)
        off = 0
        for ent in e:
            ent.offset = off
            off += (ent.numel + ALIGN - 1) // ALIGN * ALIGN
        self.total = off
        return e


def to_internal(ent: Entry, t: torch.Tensor, spec: PolicySpec) -> torch.Tensor:
    """reference-layout fp32 tensor -> internal-layout fp32 tensor (host-side plumbing)."""
    t = t.detach().to(torch.float32)
    if ent.kind == "conv":
        co, ci, r, s_ = t.shape
        return t.permute(0, 2, 3, 1).reshape(co, r * s_, ci).contiguous()
    if ent.kind == "conv_s2d":
        co, ci, r, s_ = t.shape  # r = s_ = 4
        w = t.permute(0, 2, 3, 1).reshape(co, 2, 2, 2, 2, ci)       # [co, r', qh, s', qw, ci]
        return w.permute(0, 1, 3, 2, 4, 5).reshape(co, 4, 4 * ci).contiguous()
    if ent.kind == "conv1":
        co, ci, r, s_ = t.shape
        flat = t.permute(0, 2, 3, 1).reshape(co, r * s_ * ci)
        out = t.new_zeros(co, 1, ent.int_shape[2])
        out[:, 0, :flat.shape[1]] = flat
        return out
    if ent.kind == "patch16":  # ConvTranspose2d(C, 1, 4, 2, 1) weight [C,1,4,4] -> patch-GEMM matrix [C][1][32]
        co = t.shape[0]
        out = t.new_zeros(co, 1, ent.int_shape[2])
        out[:, 0, :16] = t.reshape(co, 16)
        return out
    if ent.kind == "linear":
        return t.reshape(t.shape[0], 1, t.shape[1]).contiguous()
    if ent.kind == "pre_actor":
        # reference flatten order is NCHW (c*n + j); activations are NHWC here (j*C + c)
        z, _ = t.shape
        return t.reshape(z, spec.E, spec.n).permute(0, 2, 1).reshape(z, 1, spec.flat).contiguous()
    if ent.kind == "head":
        out = t.new_zeros(ent.int_shape[0], 1, t.shape[1])
        out[:t.shape[0], 0] = t
        return out
    if ent.kind == "head_bias":
        out = t.new_zeros(ent.int_shape[0])
        out[:t.shape[0]] = t
        return out
    return t.contiguous()


def to_reference(ent: Entry, t: torch.Tensor, spec: PolicySpec) -> torch.Tensor:
    """internal-layout tensor -> reference-layout fp32 tensor."""
    t = t.detach().to(torch.float32).reshape(ent.int_shape)
    if ent.kind == "conv":
        co, ci, r, s_ = ent.ref_shape
        return t.reshape(co, r, s_, ci).permute(0, 3, 1, 2).contiguous()
    if ent.kind == "conv_s2d":
        co, ci, r, s_ = ent.ref_shape
        w = t.reshape(co, 2, 2, 2, 2, ci).permute(0, 1, 3, 2, 4, 5)  # [co, r', qh, s', qw, ci]
        return w.reshape(co, 4, 4, ci).permute(0, 3, 1, 2).contiguous()
    if ent.kind == "conv1":
        co, ci, r, s_ = ent.ref_shape
        return t[:, 0, :r * s_ * ci].reshape(co, r, s_, ci).permute(0, 3, 1, 2).contiguous()
    if ent.kind == "patch16":
        return t[:, 0, :16].reshape(ent.ref_shape).contiguous()
    if ent.kind == "linear":
        return t.reshape(ent.ref_shape).contiguous()
    if ent.kind == "pre_actor":
        z = ent.ref_shape[0]
        return t.reshape(z, spec.n, spec.E).permute(0, 2, 1).reshape(z, spec.flat).contiguous()
    if ent.kind == "head":
        return t[:ent.ref_shape[0], 0].contiguous()
    if ent.kind == "head_bias":
        return t[:ent.ref_shape[0]].contiguous()
    return t.reshape(ent.ref_shape).contiguous()


class ParamArena:
    """Flat fp32 master / grad / Adam-moment arenas + bf16 shadow on one device."""

    def __init__(self, spec: PolicySpec, device):
        self.spec = spec
        self.entries = spec.entries()
        self.by_name: Dict[str, Entry] = {e.name: e for e in self.entries}
        n = (spec.total + 3) // 4 * 4
        self.n = n
        self.device = torch.device(device)
        self.p = torch.zeros(n, device=device, dtype=torch.float32)
        self.g = torch.zeros(n, device=device, dtype=torch.float32)
        self.m = torch.zeros(n, device=device, dtype=torch.float32)
        self.v = torch.zeros(n, device=device, dtype=torch.float32)
        self.wb = torch.zeros(n, device=device, dtype=torch.bfloat16)

    def _view(self, flat, name):
        e = self.by_name[name]
        return flat[e.offset:e.offset + e.numel].view(e.int_shape)

    def P(self, name): return self._view(self.p, name)      # fp32 master
    def G(self, name): return self._view(self.g, name)      # fp32 gradient
    def WB(self, name): return self._view(self.wb, name)    # bf16 shadow

    def offset_of(self, name):
        return self.by_name[name].offset

    def refresh_shadow(self):
        if self.p.is_cuda:
            from . import ops
            ops.cast_f32_bf16(self.p, self.wb)
        else:
            self.wb.copy_(self.p)

    # ---- reference state_dict interchange -------------------------------------------------
    def load_reference(self, sds: Dict[str, Dict[str, torch.Tensor]]):
        """sds = {"encoder": sd, "pre_actor": sd, "actor": sd} with reference keys; prefixes
        `_orig_mod.` / `module.` are stripped like bc_agent.py:107-128 does."""
        for mod, sd in sds.items():
            clean = {k.replace("_orig_mod.", "").replace("module.", ""): v for k, v in sd.items()}
            wanted = [e for e in self.entries if e.module == mod]
            missing = [e.key for e in wanted if e.key not in clean]
            extra = [k for k in clean if k not in {e.key for e in wanted}]
            if missing or extra:
                raise RuntimeError(f"state_dict mismatch for {mod}: missing {missing}, unexpected {extra}")
            for e in wanted:
                t = clean[e.key]
                if tuple(t.shape) != tuple(e.ref_shape):
                    raise RuntimeError(f"{mod}.{e.key}: shape {tuple(t.shape)} != {e.ref_shape}")
                self.P(e.name).copy_(to_internal(e, t, self.spec).to(self.device))
        self.refresh_shadow()

    def _export(self, flat) -> Dict[str, "OrderedDict[str, torch.Tensor]"]:
        out = OrderedDict((m, OrderedDict()) for m in self.modules())
        # reference key order = forward order
        for e in reversed(self.entries):
            out[e.module][e.key] = to_reference(e, self._view(flat, e.name), self.spec)
        # weight before bias inside each layer, like nn.Module.state_dict()
        for mod in out:
            keys = list(out[mod].keys())
            keys.sort(key=lambda k: (_layer_rank(mod, k), 0 if k.endswith("weight") else 1))
            out[mod] = OrderedDict((k, out[mod][k]) for k in keys)
        return out

    def modules(self):
        """Module names in forward order (entries are stored in reverse execution order)."""
        seen = []
        for e in reversed(self.entries):
            if e.module not in seen:
                seen.append(e.module)
        return seen

    def state_dicts(self):
        return self._export(self.p)

    def grad_dicts(self):
        return self._export(self.g)

    def init_reference_style(self, seed: int = None):
        """weight_init (linear_models.py:6-16; applied to the Encoder in its constructor and to
        pre_actor / actor at train_bc.py:80,82): orthogonal (gain sqrt(2) for convs, 1 for
        linears), zero bias."""
        import torch.nn as nn
        if seed is not None:
            torch.manual_seed(seed)
        sds = {m: OrderedDict() for m in self.modules()}
        gain = nn.init.calculate_gain("relu")
        for e in reversed(self.entries):
            t = torch.zeros(e.ref_shape)
            if e.kind in ("conv", "conv1", "conv_s2d", "patch16"):
                nn.init.orthogonal_(t, gain)  # weight_init also covers nn.ConvTranspose2d (linear_models.py:13)
            elif e.kind in ("linear", "pre_actor", "head"):
                nn.init.orthogonal_(t)
            sds[e.module][e.key] = t
        self.load_reference(sds)


def _layer_rank(mod, key):
    if mod not in ("encoder", "decoder", "encoder_agil"):
        return int(key.split(".")[0])
    if key.startswith("_conv_trans_"):
        return 200 + int(key.split(".")[0].split("_")[-1])
    if key.startswith("_conv_"):
        i = int(key.split(".")[0].split("_")[-1])
        return i if i <= 4 else i + 100
    # _residual_stack._layers.{i}._block.{1,3}
    parts = key.split(".")
    return 10 + int(parts[2]) * 4 + int(parts[4])
