"""CLI entry point with the reference's name: `python -m gabril_carla_b200.train_bc gaze.method=Reg
dropout.method=GMD ...` or the legacy `--gaze_method Reg --dp_method GMD --result_save_dir DIR`
(vlm_gaze/train/train_bc.py:337-351, vlm_gaze/data_utils/utils.py:405-514); under torchrun the
RANK / LOCAL_RANK / WORLD_SIZE environment switches data-parallel training on."""
from .trainer import BCTrainer, main  # noqa: F401

if __name__ == "__main__":
    main()
