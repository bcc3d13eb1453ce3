"""gabril-carla_b200 — B200-native GABRIL / GABRIL+GMD training step.

Host-side mirror of the reference's Python API for the `train_bc.py` hot path
(vlm_gaze/train/train_bc.py, vlm_gaze/data_utils/{data_loader_robomimic,gaze_utils}.py,
vlm_gaze/models/linear_models.py) over the C ABI of libgabril_b200.so (include/gabril_b200.h).
"""
__version__ = "0.1.0"
