"""Inference path, CPU side: oracle.policy_predict against golden vectors produced by
the unmodified reference modules (tests/golden/make_golden_agent.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import gabril_oracle as O
from tests.golden.make_golden_agent import synthetic_g
from tests.synth import make_batch, make_weights

GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_vectors_agent.npz")


@pytest.mark.parametrize("gray", [True, False])
@pytest.mark.parametrize("dp", ["None", "GMD", "IGMD"])
def test_policy_predict_matches_reference_golden(gray, dp):
    gold = np.load(GOLD)
    tag = "gray" if gray else "rgb"
    img, _, _ = make_batch(2, seed=41)
    w = make_weights(2 if gray else 6, seed=13)
    g = synthetic_g(2, 5) if dp != "None" else None      # unclipped: the oracle clips like :278-279
    with torch.no_grad():
        a = O.policy_predict(w["encoder"], w["pre_actor"], w["actor"], img, gray, dp, g)
    ref = torch.from_numpy(gold[f"{tag}/{dp}/action"])
    assert torch.allclose(a, ref, rtol=1e-5, atol=1e-6), (a - ref).abs().max().item()
