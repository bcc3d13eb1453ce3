"""-m gpu, needs >= 2 GPUs (skipped on a 1-GPU box): N ranks of the CUDA path against one rank on the
concatenated batch, eager and CUDA-graph replay — reduced gradients, post-Adam weights, mean losses,
weights bit-identical across ranks, both buckets reduced every step; GMD per rank.  The
checks live in tools/ddp_parity.py (also run by hand under torchrun; log in profiles/r2_ddp_parity_n*.txt).
Reference semantics: vlm_gaze/train/common/distributed.py:55-68, train_bc.py:116-123; the reference's
own check is test/test_ddp_train_one_epoch.py:23-63."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
@pytest.mark.timeout(600)
def test_two_rank_parity_eager_and_graph():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tools", "ddp_parity.py"), "--global-batch", "64", "--steps", "3"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=560, cwd=ROOT)
    assert r.returncode == 0 and "RESULT: PASS" in r.stdout, r.stdout[-4000:] + r.stderr[-2000:]
