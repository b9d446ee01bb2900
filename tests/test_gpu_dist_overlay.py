"""The distributed overlay with GPU bodies: P x Q ranks of the reference runtime on this box (overlay/pb2run), each
with the B200 device module (overlay/device_cuda_b200_shim.c + libparsec_b200.so) under it and the comm engine of
overlay/comm_engine_b200.c between them.  A(m,k) and B(k,n) reach the other ranks through the reference's own
remote-dependency protocol as host tiles (the arena copies of remote_dep_copy_allocate, remote_dep_mpi.c:593-607) and are
staged into HBM by the device module like any other tile; every C chain runs on the GPU of the rank that owns the tile.

With one GPU on the box the ranks share it (each capped through the reference's own device_cuda_memory_use); with
several, one GPU per rank.  Checked against the oracle: north-star norm, task counts per rank (owner computes), versions.
The CPU-body flavour of the same run, bit-exact, is tests/test_dist_overlay.py."""
import pytest

from test_dist_overlay import _need, check, run_ranks, APP
import json

pytestmark = pytest.mark.gpu


def gpus_on_the_box():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("P,Q,N,NB", [(1, 2, 2048, 256), (2, 2, 3072, 512)])
def test_jdf_dgemm_on_a_grid_of_gpu_ranks(tmp_path, P, Q, N, NB):
    _need()
    n = P * Q
    per_rank = gpus_on_the_box() >= n
    dump = str(tmp_path / "C.bin")
    env = {"PARSEC_MCA_device_cuda_memory_use": "90" if per_rank else str(80 // n)}
    rc, out, err = run_ranks(n, [APP, "-N", str(N), "-t", str(NB), "-c", "4", "-P", str(P), "-Q", str(Q), "--dump", dump,
                                 "--", "--mca", "device_cuda_enabled", "1"], env, timeout=300, gpu_per_rank=per_rank)
    assert rc == 0, (rc, out[-1500:], err[-3000:])
    lines = sorted((json.loads(l) for l in out.splitlines() if l.startswith("{")), key=lambda d: d["rank"])
    assert [l["rank"] for l in lines] == list(range(n))
    for l in lines:
        assert l["gpus"] == 1, err[-2000:]
        assert l["gemm_cpu"] == 0 and l["gemm_gpu"] == l["device_executed_tasks"] > 0, l      # every chain on the GPU
        assert l["h2d_bytes"] > 0 and l["d2h_bytes"] > 0
    check(lines, dump, N, NB, P, Q, gpu=True)
    print("P x Q = %d x %d, N=%d NB=%d: %s GFLOP/s per rank (first run, includes page-locking)" %
          (P, Q, N, NB, [round(l["gflops_first"], 1) for l in lines]))
