"""P x Q tiled DGEMM over several GPUs (NCCL row / column tile broadcasts) against the single-rank oracle."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _gpus():
    import torch
    return torch.cuda.device_count()


PLAIN = (1, 1, 1, 1, 0, 0)


@pytest.mark.parametrize("world,N,NB,layout", [
    (2, 1024, 128, PLAIN), (4, 1280, 128, PLAIN), (8, 2048, 256, PLAIN),
    # k-cyclic super-tiles and a shifted grid origin: (kp, kq, a_kq, b_kp, ip, jq)
    (2, 1152, 128, (1, 2, 2, 1, 0, 1)), (4, 1280, 128, (2, 3, 1, 2, 1, 1))])
def test_distributed_tiled_dgemm_matches_oracle(world, N, NB, layout):
    if _gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + world), os.path.join(HERE, "dist_worker.py"),
           str(N), str(NB)] + [str(x) for x in layout]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count(" ok: ") == world
