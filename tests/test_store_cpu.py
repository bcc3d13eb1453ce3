"""Device-resident store, CPU side: the oracle's window restatement and the host
mirror DemoStore.window_indices against golden vectors produced by the unmodified reference
(robomimic SequenceDataset.get_sequence_from_demo, tests/golden/make_golden_store.py). Integer work:
bit-exact."""
import os

import numpy as np
import torch

from oracle import gabril_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden", "reference_vectors_store.npz")


def _gold():
    return np.load(GOLD)


def test_oracle_windows_match_reference_golden():
    g = _gold()
    lens = g["demo_lengths"].tolist()
    for S in (1, 2, 3, 4):
        ref = g[f"windows_S{S}"]
        got = np.asarray([O.demo_window_indices(lens, i, S) for i in range(sum(lens))], dtype=np.int64)
        assert got.shape == ref.shape and (got == ref).all(), S


def test_host_store_windows_match_reference_golden():
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200.data import DemoStore
    g = _gold()
    lens = g["demo_lengths"]
    n = int(lens.sum())
    starts = np.concatenate([[0], np.cumsum(lens)[:-1]])
    for S in (1, 2, 3, 4):
        st = DemoStore(torch.zeros(n, 2, 4, 3, dtype=torch.uint8), torch.zeros(n, 10), torch.zeros(n, 7),
                       torch.from_numpy(starts), S)
        w = st.window_indices(torch.arange(n))
        assert (w.numpy() == g[f"windows_S{S}"]).all(), S


def test_fused_loader_yields_batch_dicts_on_cpu():
    """fused=True only changes what an HBM-resident store yields; a host store keeps batch dicts."""
    import gabril_carla_b200  # noqa: F401
    from gabril_carla_b200.data import ShardedBatches, synthetic_store
    st = synthetic_store(16, 2, H=8, W=8, demo_len=5)
    b = next(iter(ShardedBatches(st, 4, fused=True)))
    assert "obs" in b and b["obs"]["image"].shape == (4, 2, 8, 8, 3)
