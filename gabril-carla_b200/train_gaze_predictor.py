"""CLI entry point with the reference's name: `python -m gabril_carla_b200.train_gaze_predictor
data.batch_size=256 training.epochs=2 ...` (vlm_gaze/train/train_gaze_predictor.py:124-140)."""
from .gaze_predictor import main  # noqa: F401

if __name__ == "__main__":
    main()
