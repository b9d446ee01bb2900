"""The comm-engine seam for real (piece 3 / row a22): the reference runtime built DISTRIBUTED
(overlay/build_dist.sh: remote_dep.c and remote_dep_mpi.c as they are, overlay/comm_engine_b200.c in place of
parsec_comm_engine.c + parsec_mpi_funnelled.c, libpb2mpi as the datatype layer) runs the parsec-ptgpp-compiled
dgemm_NN JDF on P x Q ranks of this box, started by overlay/pb2run.  Every A and B tile reaches the ranks that need it
through the reference's own ACTIVATE / GET_DATA / put protocol over pb2_ce_*.

CPU bodies, so this runs without a GPU (here, every round): C is compared BIT FOR BIT with the oracle (the CPU body is the
oracle's arithmetic, see tests/test_oracle.py), task counts per class are summed over the ranks, and every rank must see
version KT on the C tiles it owns.  Also: the reference's own comm-engine test, tests/dsl/dtd/dtd_test_ce.c, unmodified,
over the new engine; and both put paths (one copy through process_vm_writev, staged through the rings).
The GPU flavour of the same run is tests/test_gpu_dist_overlay.py.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
APP = os.path.join(REF, "bin", "dgemm_NN_dist")
PB2RUN = os.path.join(ROOT, "overlay", "pb2run")


def _need():
    if not (os.path.exists(APP) and os.path.exists(os.path.join(REF, "dist", "libparsec.so.4"))):
        pytest.skip("baseline/_ref/dist is not built (overlay/build_dist.sh needs the reference tree)")


def run_ranks(n, prog_args, env_extra=None, timeout=300, gpu_per_rank=False):
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.pathsep.join([os.path.join(REF, "dist"), os.path.join(ROOT, "parsec-1_b200"),
                                              env.get("LD_LIBRARY_PATH", "")])
    env["PARSEC_CUCORES_LIB"] = os.path.join(ROOT, "parsec-1_b200", "libparsec_b200.so")
    env.update(env_extra or {})
    cmd = [PB2RUN, "-n", str(n), "--timeout", str(timeout - 20)] + (["--gpu-per-rank"] if gpu_per_rank else []) + prog_args
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    return p.returncode, p.stdout, p.stderr


def run_gemm(n, P, Q, N, NB, dump, mca=("device_cuda_enabled", "0"), env_extra=None, extra=()):
    rc, out, err = run_ranks(n, [APP, "-N", str(N), "-t", str(NB), "-c", "2", "-P", str(P), "-Q", str(Q), "--dump", dump]
                             + list(extra) + ["--", "--mca", mca[0], mca[1]], env_extra)
    assert rc == 0, (rc, out[-1500:], err[-3000:])
    lines = sorted((json.loads(l) for l in out.splitlines() if l.startswith("{")), key=lambda d: d["rank"])
    assert [l["rank"] for l in lines] == list(range(n)), out[-1500:]
    return lines


def check(lines, dump, N, NB, P, Q, gpu=False):
    d = O.square(N, NB)
    A, B, Cm = O.fill(d, 1789), O.fill(d, 1805), O.fill(d, 1901)
    cnt, ver = O.tiled_dgemm(N, N, N, NB, NB, NB, A, B, Cm, nthreads=os.cpu_count() or 8)
    KT = -(-N // NB)
    assert sum(l["read_a"] for l in lines) == cnt.read_a and sum(l["read_b"] for l in lines) == cnt.read_b
    assert sum(l["gemm_cpu"] + l["gemm_gpu"] for l in lines) == cnt.gemm
    mt = -(-N // NB)
    for l in lines:                                    # owner computes: rank (m % P) * Q + n % Q runs the chain of C(m, n)
        p, q = divmod(l["rank"], Q)
        mine = len(range(p, mt, P)) * len(range(q, mt, Q))
        assert l["gemm_cpu"] + l["gemm_gpu"] == mine * KT, l
        assert (l["P"], l["Q"], l["world"]) == (P, Q, P * Q)
        if mine:
            assert l["c_version_min"] == l["c_version_max"] == KT == int(ver.max()) and l["ab_version_max"] == 0, l
    got = np.fromfile(dump, dtype=np.float64)
    assert got.size == Cm.size
    if gpu:
        dense = lambda m: O.tiled_to_dense(d, m)
        norm = O.parity_norm(dense(got), dense(Cm), dense(A), dense(B), N)
        assert norm < 10.0, norm
    else:
        assert np.array_equal(got, Cm)                 # same operations in the same order as the oracle: same bits


@pytest.mark.parametrize("P,Q,N,NB", [(1, 2, 512, 128), (2, 1, 512, 128), (2, 2, 1024, 128), (2, 2, 1000, 192), (1, 3, 768, 128)])
def test_jdf_dgemm_on_a_grid_of_ranks_matches_the_oracle_bit_for_bit(tmp_path, P, Q, N, NB):
    _need()
    dump = str(tmp_path / "C.bin")
    lines = run_gemm(P * Q, P, Q, N, NB, dump)
    check(lines, dump, N, NB, P, Q)


def test_the_same_run_with_puts_staged_through_the_rings(tmp_path):
    """A box that refuses cross-memory attach: tiles travel through the 1 MiB rings in pieces."""
    _need()
    dump = str(tmp_path / "C.bin")
    lines = run_gemm(4, 2, 2, 1024, 256, dump, env_extra={"PB2_CE_NO_CMA": "1"})
    check(lines, dump, 1024, 256, 2, 2)


def test_puts_handed_to_copier_threads(tmp_path):
    """PB2_CE_COPY_THREADS=3: tiles are copied by helper threads while the communication thread keeps answering; the
    completion of each put still reaches the runtime on the communication thread."""
    _need()
    dump = str(tmp_path / "C.bin")
    lines = run_gemm(4, 2, 2, 2048, 256, dump, env_extra={"PB2_CE_COPY_THREADS": "3"})
    check(lines, dump, 2048, 256, 2, 2)


def test_small_rings_make_the_engine_park_and_resume_messages(tmp_path):
    """64 KiB rings (the smallest that hold the longest activation message the runtime registers, 4 x 4 KiB + header):
    activations and tile pieces pile up faster than the peer drains them; the engine parks messages and keeps order."""
    _need()
    dump = str(tmp_path / "C.bin")
    lines = run_gemm(4, 2, 2, 1536, 64, dump, env_extra={"PB2_CE_RING_BYTES": "65536", "PB2_CE_NO_CMA": "1"})
    check(lines, dump, 1536, 64, 2, 2)


def test_reference_comm_engine_test_runs_over_the_new_engine():
    """tests/dsl/dtd/dtd_test_ce.c, unmodified: active messages both ways, a GET of nine integers, a PUT of nine
    integers, with registrations forwarded inside messages and released through copies of themselves."""
    exe = os.path.join(REF, "reftests", "dtd_test_ce")
    if not os.path.exists(exe):
        pytest.skip("baseline/_ref/reftests/dtd_test_ce is not built")
    for env in ({}, {"PB2_CE_NO_CMA": "1"}):
        rc, out, err = run_ranks(2, [exe], env, timeout=120)
        assert rc == 0, (rc, out[-1500:], err[-1500:])
        assert out.count("[1] In callback for tag 2, message sent from 0 size: 12 message: [10,11,12]") == 2
        assert out.count("[0] In callback for tag 3, message sent from 1 size: 8 message: [9.500000,19.500000]") == 2
        assert "[1] GET is over, message:\n[0,1,2,3,4,5,6,7,8]" in out and "[0] Notification of GET over received" in out
        assert "[1] PUT is over, message:\n[0,2,4,6,8,10,12,14,16]" in out and "[0] PUT is finished" in out


@pytest.mark.parametrize("P,Q,N,NB", [(2, 2, 768, 128), (1, 3, 1000, 192), (2, 1, 512, 64)])
def test_cholesky_on_a_grid_of_ranks(tmp_path, P, Q, N, NB):
    """A DAG with more kinds of remote edges than the GEMM: the factored diagonal tile goes down its panel, every panel
    tile goes to one SYRK and to the GEMMs of a whole row and a whole column of the trailing matrix -- several
    destinations per activation, forwarded along the reference's own propagation trees."""
    from test_overlay_potrf import check_factor
    exe = os.path.join(REF, "bin", "dpotrf_L_dist")
    if not os.path.exists(exe):
        pytest.skip("baseline/_ref/bin/dpotrf_L_dist is not built")
    n = P * Q
    dump = str(tmp_path / "L.bin")
    rc, out, err = run_ranks(n, [exe, "-N", str(N), "-t", str(NB), "-c", "2", "-P", str(P), "-Q", str(Q), "--dump", dump,
                                 "--", "--mca", "device_cuda_enabled", "0"])
    assert rc == 0, (rc, out[-1500:], err[-3000:])
    lines = [json.loads(l) for l in out.splitlines() if l.startswith("{")]
    assert sorted(l["rank"] for l in lines) == list(range(n))
    total = {k: sum(l[k] for l in lines) for k in lines[0] if k not in ("rank", "world", "N", "NB", "cores", "gpus", "seconds", "gflops")}
    check_factor(total, dump, N, NB, gpu=False)
    assert sum(1 for l in lines if l["trsm_cpu"] + l["syrk_cpu"] + l["gemm_cpu"] + l["potrf"] > 0) == n   # every rank worked
