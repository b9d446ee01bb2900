"""Tiled Cholesky on the B200 through the overlay: the second algorithm over the same device module.
POTRF, TRSM, SYRK and GEMM run as CUDA bodies calling parsec_b200_dpotrf / _dtrsm / _dsyrk / _dgemm; with --cpu-potrf the
POTRF body declines and every diagonal tile travels GPU -> CPU -> GPU.  Checked with LAPACK's test ratio ||A - L L^T|| / (||A|| N eps) < 10 and against
numpy's factor, plus the task counts of the DAG and where the tasks ran."""
import pytest

from test_overlay_potrf import check_factor, run_potrf

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,NB", [(2048, 256), (1000, 192), (8192, 1024)])
def test_cholesky_on_the_gpu_over_the_overlay(tmp_path, N, NB):
    dump = str(tmp_path / "L.bin")
    out, err = run_potrf(N, NB, dump, mca=("device_cuda_enabled", "1"))
    assert out["gpus"] == 1, err[-2000:]
    nt = -(-N // NB)
    assert out["trsm_gpu"] == nt * (nt - 1) // 2 and out["trsm_cpu"] == 0       # scratch from parsec_gpu_pop_workspace
    assert out["potrf_gpu"] == nt
    assert out["device_executed_tasks"] == out["potrf_gpu"] + out["trsm_gpu"] + out["syrk_gpu"] + out["gemm_gpu"]
    resid = check_factor(out, dump, N, NB, gpu=True)
    print("dpotrf N=%d NB=%d: %.1f GFLOP/s (first run), residual ratio %.3g, h2d %.0f MiB, d2h %.0f MiB" %
          (N, NB, out["gflops"], resid, out["h2d_bytes"] / 2 ** 20, out["d2h_bytes"] / 2 ** 20))


def test_potrf_on_the_cpu_between_gpu_tasks(tmp_path):
    """--cpu-potrf: SYRK leaves the diagonal tile on the GPU, POTRF factors it on the CPU (the module writes it back when
    the CUDA body declines), TRSM reads the factor on the GPU again."""
    N, NB = 2048, 256
    dump = str(tmp_path / "L.bin")
    out, err = run_potrf(N, NB, dump, mca=("device_cuda_enabled", "1"), extra=["--cpu-potrf"])
    assert out["gpus"] == 1 and out["potrf_gpu"] == 0 and out["potrf"] == N // NB and out["trsm_cpu"] == 0
    check_factor(out, dump, N, NB, gpu=True)


def test_trsm_bodies_that_decline_fall_back_to_the_cpu(tmp_path):
    """--cpu-trsm: the CUDA body of TRSM returns PARSEC_HOOK_RETURN_NEXT, its inputs (the POTRF tile, the panel tile
    updated on the GPU) are handed back and the CPU body runs; SYRK and GEMM stay on the GPU."""
    N, NB = 1536, 256
    dump = str(tmp_path / "L.bin")
    out, err = run_potrf(N, NB, dump, mca=("device_cuda_enabled", "1"), extra=["--cpu-trsm"])
    nt = N // NB
    assert out["gpus"] == 1 and out["trsm_gpu"] == 0 and out["trsm_cpu"] == nt * (nt - 1) // 2
    check_factor(out, dump, N, NB, gpu=True)
