"""TEST INFRASTRUCTURE — imports the *unmodified* reference (Gonglitian/GABRIL-CARLA) from
/root/reference on CPU, with inert stubs for the packages that are absent in this image
(hydra, omegaconf, robomimic, matplotlib ...; none of them touches arithmetic — SURVEY.md §8c).

Used only (a) by tests/golden/make_golden.py to generate the committed golden vectors and
(b) by tests that validate oracle/gabril_oracle.py against the real reference when
/root/reference exists (it does not exist on the GPU box).  Never imported by the product.
"""
import importlib.machinery
import os
import sys
import types
from types import SimpleNamespace

REFERENCE_ROOT = os.environ.get("GABRIL_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "vlm_gaze"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, loader=None)
    m.__path__ = []
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules.setdefault(name, m)
    return sys.modules[name]


_loaded = None


def load():
    """Returns a namespace with the reference symbols on the train_bc.py hot path."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    # --- stubs (import-only dependencies) ---
    for n in ("matplotlib", "matplotlib.pyplot", "h5py", "imageio", "termcolor"):
        try:
            __import__(n)
        except Exception:
            _stub(n)
    try:
        import robomimic  # noqa: F401
    except Exception:
        class SequenceDataset:  # dummy base for GABRILSequenceDataset
            def __init__(self, *a, **k):
                pass
        _stub("robomimic")
        _stub("robomimic.utils")
        _stub("robomimic.utils.dataset", SequenceDataset=SequenceDataset)
        _stub("robomimic.utils.obs_utils",
              initialize_obs_utils_with_obs_specs=lambda *a, **k: None)
    try:
        import hydra  # noqa: F401
    except Exception:
        _stub("hydra", main=lambda **kw: (lambda f: f))
    try:
        import omegaconf  # noqa: F401
    except Exception:
        class _OC:
            @staticmethod
            def to_yaml(cfg):
                return str(cfg)
        _stub("omegaconf", DictConfig=dict, OmegaConf=_OC)
    try:
        import tqdm  # noqa: F401
    except Exception:
        _stub("tqdm", tqdm=lambda *a, **k: None)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)

    from vlm_gaze.data_utils.data_loader_robomimic import GazePreprocessor
    from vlm_gaze.data_utils.gaze_utils import apply_gmd_dropout, get_gaze_mask
    from vlm_gaze.models.linear_models import (AutoEncoder, Decoder, Encoder, Residual,
                                               ResidualStack, weight_init)
    from vlm_gaze.train.common.optim import build_optimizer, build_scheduler
    from vlm_gaze.train.train_bc import BCTrainer

    _loaded = SimpleNamespace(
        GazePreprocessor=GazePreprocessor, get_gaze_mask=get_gaze_mask,
        apply_gmd_dropout=apply_gmd_dropout, Encoder=Encoder, Decoder=Decoder,
        AutoEncoder=AutoEncoder, Residual=Residual, ResidualStack=ResidualStack,
        weight_init=weight_init, BCTrainer=BCTrainer, build_optimizer=build_optimizer,
        build_scheduler=build_scheduler)
    return _loaded


def _ns(d):
    if isinstance(d, dict):
        return SimpleNamespace(**{k: _ns(v) for k, v in d.items()})
    return d


def default_cfg(**overrides):
    """cfg tree of vlm_gaze/configs/train_bc_base.yaml as nested SimpleNamespace.
    overrides use dotted keys: default_cfg(**{'gaze.method': 'Reg', 'dropout.method': 'GMD'})."""
    import yaml
    with open(os.path.join(REFERENCE_ROOT, "vlm_gaze/configs/train_bc_base.yaml")) as f:
        raw = yaml.safe_load(f)
    raw.pop("defaults", None)
    raw.pop("hydra", None)
    raw["model"]["frame_stack"] = raw["data"]["frame_stack"]
    raw["optimizer"]["lr"] = float(raw["optimizer"]["lr"])
    raw["scheduler"]["eta_min"] = float(raw["scheduler"]["eta_min"])
    raw["training"]["use_amp"] = False
    raw["training"]["use_compile"] = False
    raw["training"]["device"] = "cpu"
    for k, v in overrides.items():
        node = raw
        parts = k.split(".")
        for p in parts[:-1]:
            node = node[p]
        node[parts[-1]] = v
    if "model.frame_stack" not in overrides:
        raw["model"]["frame_stack"] = raw["data"]["frame_stack"]
    return _ns(raw)


def make_trainer(cfg, seed=42):
    """A reference BCTrainer without data loading / logging (SURVEY.md §8c recipe)."""
    import torch
    import torch.nn as nn
    ref = load()
    t = ref.BCTrainer.__new__(ref.BCTrainer)
    t.cfg = cfg
    t.device = torch.device("cpu")
    t.rank = 0
    t.is_distributed = False
    t.local_rank = 0
    t.world_size = 1
    torch.manual_seed(seed)
    t.build_models()
    t.gaze_preprocessor = ref.GazePreprocessor(
        img_height=cfg.data.img_height, img_width=cfg.data.img_width,
        gaze_sigma=cfg.gaze.mask_sigma, gaze_coeff=cfg.gaze.mask_coeff,
        maxpoints=cfg.gaze.max_points, device="cpu",
        temporal_alpha=float(cfg.gaze.temporal_alpha),
        temporal_mode=str(cfg.gaze.temporal_mode),
        temporal_sigmas=cfg.gaze.temporal_sigmas, temporal_coeffs=cfg.gaze.temporal_coeffs,
        temporal_offset_start=int(cfg.gaze.temporal_offset_start))
    t.criterion = nn.MSELoss()
    t.gaze_ratio = float(cfg.gaze.ratio)
    return t
