"""Tiled Cholesky (overlay/apps/dpotrf_L.jdf) under the reference runtime with CPU bodies: the DAG, the generator and
the driver are right before any GPU is involved.  No GPU needed.  The GPU flavour is tests/test_gpu_overlay_potrf.py."""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
APP = os.path.join(REF, "bin", "dpotrf_L_app")
EPS = 2.0 ** -53


def run_potrf(N, NB, dump, mca=("device_cuda_enabled", "0"), extra=(), lib="overlay", env_extra=None, timeout=300):
    if not (os.path.exists(APP) and os.path.exists(os.path.join(REF, lib, "libparsec.so.4"))):
        pytest.skip("baseline/_ref is not built (overlay/build_overlay.sh needs the reference tree)")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.pathsep.join([os.path.join(REF, lib), os.path.join(ROOT, "parsec-1_b200"),
                                              env.get("LD_LIBRARY_PATH", "")])
    env.update(env_extra or {})
    cmd = [APP, "-N", str(N), "-t", str(NB), "-c", "4", "--dump", dump] + list(extra) + ["--", "--mca", mca[0], mca[1]]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, (p.returncode, p.stdout[-1500:], p.stderr[-3000:])
    return json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1]), p.stderr


def spd_matrix(N, NB):
    """What dpotrf_main.c builds: lower triangle from the oracle's tile generator (seed 3872), N on the diagonal."""
    d = O.square(N, NB)
    dense = O.tiled_to_dense(d, O.fill(d, 3872))
    A = np.tril(dense) + np.tril(dense, -1).T
    A[np.arange(N), np.arange(N)] += N
    return d, A


def check_factor(out, dump, N, NB, gpu):
    d, A = spd_matrix(N, NB)
    nt = -(-N // NB)
    assert out["potrf"] == nt and out["potrf_failures"] == 0
    assert out["trsm_cpu"] + out["trsm_gpu"] == nt * (nt - 1) // 2 == out["syrk_cpu"] + out["syrk_gpu"]
    assert out["gemm_cpu"] + out["gemm_gpu"] == nt * (nt - 1) * (nt - 2) // 6
    L = np.tril(O.tiled_to_dense(d, np.fromfile(dump, dtype=np.float64)))
    inf = lambda X: np.abs(X).sum(axis=1).max()
    resid = inf(A - L @ L.T) / (inf(A) * N * EPS)                     # LAPACK's test ratio for a Cholesky factor
    assert resid < 10, resid
    Lref = np.linalg.cholesky(A)
    assert inf(L - Lref) / (inf(Lref) * np.linalg.cond(A, np.inf) * N * EPS) < 10
    if gpu:
        assert out["syrk_cpu"] == 0 and out["gemm_cpu"] == 0
    return resid


@pytest.mark.parametrize("N,NB", [(512, 128), (1000, 192), (768, 64)])
def test_cholesky_dag_with_cpu_bodies(tmp_path, N, NB):
    dump = str(tmp_path / "L.bin")
    out, err = run_potrf(N, NB, dump)
    assert out["gpus"] == 0 and out["trsm_gpu"] == out["syrk_gpu"] == out["gemm_gpu"] == 0
    check_factor(out, dump, N, NB, gpu=False)
