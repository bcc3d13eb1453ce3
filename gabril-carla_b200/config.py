"""Configuration surface of the train_bc.py hot path.

Same tree and key names as the reference's Hydra config (vlm_gaze/configs/train_bc_base.yaml:1-121)
so that `key=value` overrides written for the reference work unchanged, plus the legacy argparse
spellings the fork keeps in vlm_gaze/data_utils/utils.py:405-514 (`--gaze_method`, `--dp_method`,
`--result_save_dir`, ...).  Hydra / OmegaConf are not required: this is a small YAML + override
loader (PyYAML reads `5e-4` as a string, so numeric-looking strings are cast; `${a.b}` references
are resolved against the tree).
"""
import copy
import re
from typing import Any, Dict, Iterable, List, Optional


class Cfg(dict):
    """dict with attribute access (cfg.gaze.method), like an OmegaConf DictConfig."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def _wrap(x):
    if isinstance(x, dict):
        return Cfg({k: _wrap(v) for k, v in x.items()})
    if isinstance(x, list):
        return [_wrap(v) for v in x]
    return x


def default_tree() -> Dict[str, Any]:
    """Defaults of train_bc_base.yaml (values, not text)."""
    return {
        "data": {"task": "Mixed_", "hdf5_path": "synthetic:4096", "num_episodes": 200, "batch_size": 4000,
                 "num_workers": 8, "prefetch_factor": 2, "frame_stack": 2, "cache_mode": "low_dim",
                 "cache_getitem": False, "img_height": 180, "img_width": 320, "action_dim": 7,
                 "gaze_key": "gaze_coords"},
        "model": {"grayscale": True, "frame_stack": "${data.frame_stack}", "embedding_dim": 64,
                  "num_hiddens": 128, "num_residual_layers": 2, "num_residual_hiddens": 32, "z_dim": 256},
        "gaze": {"method": "Reg", "mask_sigma": 30.0, "mask_coeff": 0.8, "max_points": 5, "beta": 50.0,
                 "lambda_weight": 10.0, "contrastive_threshold": 10.0, "prob_dist_type": "MSE", "ratio": 1,
                 "temporal_flag": True, "temporal_alpha": 0.7, "temporal_mode": "alpha_decay",
                 "temporal_sigmas": None, "temporal_coeffs": None, "temporal_offset_start": 0,
                 "temporal_use_future": False},
        "dropout": {"method": "None", "num_embeddings": 512, "oreo_num_mask": 4, "oreo_prob": 0.5,
                    "vqvae_path": ""},
        "optimizer": {"type": "adam", "lr": 5e-4, "weight_decay": 0.0},
        "scheduler": {"type": "cosine_warmup", "step_size": 50, "gamma": 0.5, "eta_min": 1e-6,
                      "warmup_steps": 500, "T_0": 10, "T_mult": 1, "pct_start": 0.3,
                      "anneal_strategy": "cos", "div_factor": 25.0, "final_div_factor": 10000.0},
        "training": {"seed": 42, "epochs": 500, "device": "cuda:0", "use_amp": True, "use_compile": True, "cuda_graph": True,
                     "compile_backend": "inductor", "compile_mode": "default",
                     "gradient_accumulation_steps": 1, "save_interval": 50,
                     "distributed": {"enabled": False, "backend": "nccl", "find_unused_parameters": False,
                                     "broadcast_buffers": True, "gradient_as_bucket_view": True}},
        "logging": {"log_dir": "runs", "checkpoint_dir": "runs", "save_params": True},
        "tag": "",
    }


_NUM = re.compile(r"^[+-]?(\d+\.?\d*|\.\d+)([eE][+-]?\d+)?$")


def parse_scalar(s: Any) -> Any:
    """YAML-ish scalar typing for override strings and for PyYAML's '5e-4' strings."""
    if not isinstance(s, str):
        return s
    t = s.strip()
    low = t.lower()
    if low in ("true", "false"):
        return low == "true"
    if low in ("null", "~"):
        return None
    if low == "none":
        return t                                   # "None" (method) / "none" (scheduler) stay strings
    if re.fullmatch(r"[+-]?\d+", t):
        return int(t)
    if _NUM.match(t):
        return float(t)
    if t.startswith("[") and t.endswith("]"):
        inner = t[1:-1].strip()
        return [parse_scalar(x) for x in inner.split(",")] if inner else []
    if len(t) >= 2 and t[0] == t[-1] and t[0] in "'\"":
        return t[1:-1]
    return t


def _cast_numeric_strings(node):
    if isinstance(node, dict):
        return {k: _cast_numeric_strings(v) for k, v in node.items()}
    if isinstance(node, list):
        return [_cast_numeric_strings(v) for v in node]
    if isinstance(node, str) and _NUM.match(node.strip()) and not re.fullmatch(r"[+-]?\d+", node.strip()):
        return float(node)
    return node


def _merge(dst, src):
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dst.get(k), dict):
            _merge(dst[k], v)
        else:
            dst[k] = v


def _set(tree, dotted, value):
    parts = dotted.lstrip("+").split(".")
    node = tree
    for p in parts[:-1]:
        node = node.setdefault(p, {})
    node[parts[-1]] = value


def _get(tree, dotted):
    node = tree
    for p in dotted.split("."):
        node = node[p]
    return node


def _resolve(tree, node=None):
    node = tree if node is None else node
    for k, v in list(node.items()):
        if isinstance(v, dict):
            _resolve(tree, v)
        elif isinstance(v, str):
            m = re.fullmatch(r"\$\{([\w.]+)\}", v.strip())
            if m:
                node[k] = _get(tree, m.group(1))


# legacy argparse flag -> config key (vlm_gaze/data_utils/utils.py:415-511)
LEGACY_FLAGS = {
    "gaze_method": "gaze.method", "dp_method": "dropout.method", "result_save_dir": "logging.checkpoint_dir",
    "seed": "training.seed", "task": "data.task", "datapath": "data.hdf5_path", "hdf5_path": "data.hdf5_path",
    "num_episodes": "data.num_episodes", "stack": "data.frame_stack", "grayscale": "model.grayscale",
    "batch_size": "data.batch_size", "lr": "optimizer.lr", "epochs": "training.epochs", "n_epochs": "training.epochs",
    "save_interval": "training.save_interval", "embedding_dim": "model.embedding_dim",
    "num_hiddens": "model.num_hiddens", "num_residual_layers": "model.num_residual_layers",
    "num_residual_hiddens": "model.num_residual_hiddens", "z_dim": "model.z_dim",
    "num_embeddings": "dropout.num_embeddings", "gaze_mask_sigma": "gaze.mask_sigma",
    "gaze_mask_coeff": "gaze.mask_coeff", "gaze_beta": "gaze.beta", "gaze_lambda": "gaze.lambda_weight",
    "gaze_contrastive_threshold": "gaze.contrastive_threshold", "prob_dist_type": "gaze.prob_dist_type",
    "gaze_ratio": "gaze.ratio", "oreo_num_mask": "dropout.oreo_num_mask", "oreo_prob": "dropout.oreo_prob",
    "device": "training.device", "use_amp": "training.use_amp", "use_compile": "training.use_compile",
    "weight_decay": "optimizer.weight_decay", "scheduler_type": "scheduler.type", "tag": "tag",
    "gradient_accumulation_steps": "training.gradient_accumulation_steps", "num_workers": "data.num_workers",
    "action_dim": "data.action_dim",
}

GAZE_METHODS = ("None", "Teacher", "Reg", "Mask", "Contrastive", "ViSaRL", "AGIL", "GRIL")
DP_METHODS = ("None", "Oreo", "IGMD", "GMD")


def load_config(argv: Optional[Iterable[str]] = None, yaml_path: Optional[str] = None,
                overrides: Optional[Dict[str, Any]] = None) -> Cfg:
    """Build the config from defaults <- yaml file <- CLI (`a.b=c` Hydra-style overrides,
    `--config-name/--config-path` ignored except a yaml path, legacy `--flag value` / `--flag=value`)."""
    tree = default_tree()
    args: List[str] = list(argv or [])
    kv: List = []
    i = 0
    while i < len(args):
        a = args[i]
        if a.startswith("--"):
            name, eq, val = a[2:].partition("=")
            name = name.replace("-", "_")
            if not eq:
                nxt = args[i + 1] if i + 1 < len(args) else None
                if nxt is not None and not nxt.startswith("--") and "=" not in nxt.split(".")[0]:
                    val = nxt
                    i += 1
                else:
                    val = "true"  # store_true style
            if name in ("config", "config_file"):
                yaml_path = val
            elif name in ("config_name", "config_path", "config_dir"):
                pass  # Hydra plumbing; the defaults here are train_bc_base.yaml's
            elif name in LEGACY_FLAGS:
                kv.append((LEGACY_FLAGS[name], parse_scalar(val)))
            else:
                raise SystemExit(f"unknown flag --{name}")
        elif "=" in a:
            k, _, v = a.partition("=")
            kv.append((k, parse_scalar(v)))
        else:
            raise SystemExit(f"cannot parse argument {a!r} (expected key=value or --flag value)")
        i += 1
    if yaml_path:
        import yaml
        with open(yaml_path) as f:
            user = yaml.safe_load(f) or {}
        user.pop("defaults", None)
        user.pop("hydra", None)
        _merge(tree, _cast_numeric_strings(user))
    for k, v in kv:
        _set(tree, k, v)
    for k, v in (overrides or {}).items():
        _set(tree, k, v)
    _resolve(tree)
    # method names are strings even when YAML / the CLI produced a null
    for path in ("gaze.method", "dropout.method"):
        if _get(tree, path) is None:
            _set(tree, path, "None")
    if _get(tree, "gaze.method") not in GAZE_METHODS:
        raise SystemExit(f"gaze.method must be one of {GAZE_METHODS}")
    if _get(tree, "dropout.method") not in DP_METHODS:
        raise SystemExit(f"dropout.method must be one of {DP_METHODS}")
    return _wrap(copy.deepcopy(tree))


def to_plain(cfg) -> Dict[str, Any]:
    if isinstance(cfg, dict):
        return {k: to_plain(v) for k, v in cfg.items()}
    if isinstance(cfg, list):
        return [to_plain(v) for v in cfg]
    return cfg
