"""The drop-in, for real: the reference runtime (libparsec.so built from /root/reference by overlay/build_overlay.sh)
with overlay/device_cuda_b200_shim.c in place of its CUDA module runs a parsec-ptgpp-compiled JDF
(overlay/apps/dgemm_NN.jdf, BODY [type=CUDA dyld=parsec_b200_dgemm]) on the B200.

What is checked against the oracle (CPU restatement, pinned on the reference runtime run with CPU bodies, see
tests/test_overlay_cpu.py): the values of C under the north-star norm, the per-class task counts, that every GEMM ran on
the GPU (device->executed_tasks, parsec/mca/device/device.h:136), and the version of every host parsec_data_copy_t
(KT for C, 0 for A and B).  The binaries ship with the repository snapshot (baseline/_ref/ is
git-ignored, not gpurun-ignored); nothing here reads /root/reference.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
APP = os.path.join(REF, "bin", "dgemm_NN_app")


def _need_binaries():
    if not (os.path.exists(APP) and os.path.exists(os.path.join(REF, "overlay", "libparsec.so.4"))):
        pytest.skip("baseline/_ref is not built (overlay/build_overlay.sh needs the reference tree)")


def run_app(lib, args, mca=("device_cuda_enabled", "1"), timeout=180):
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.pathsep.join([os.path.join(REF, lib), os.path.join(ROOT, "parsec-1_b200"),
                                              env.get("LD_LIBRARY_PATH", "")])
    env["PARSEC_CUCORES_LIB"] = os.path.join(ROOT, "parsec-1_b200", "libparsec_b200.so")
    cmd = [APP] + [str(a) for a in args] + ["--", "--mca", mca[0], mca[1]]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, (p.returncode, p.stdout[-2000:], p.stderr[-2000:])
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line), p.stderr


def oracle_result(N, NB):
    d = O.square(N, NB)
    A, B, Cm = O.fill(d, 1789), O.fill(d, 1805), O.fill(d, 1901)
    cnt, ver = O.tiled_dgemm(N, N, N, NB, NB, NB, A, B, Cm, nthreads=os.cpu_count() or 8)
    return d, A, B, Cm, cnt, ver


def check_against_oracle(out, dump, N, NB, runs=1):
    d, A, B, Cref, cnt, ver = oracle_result(N, NB)
    KT = N // NB
    assert (out["read_a"], out["read_b"]) == (cnt.read_a * runs, cnt.read_b * runs)
    assert out["gemm_gpu"] == cnt.gemm * runs and out["gemm_cpu"] == 0
    assert out["device_executed_tasks"] == cnt.gemm * runs            # every GEMM went through the device module
    assert out["c_version_min"] == out["c_version_max"] == KT * runs == int(ver.max()) * runs
    assert out["ab_version_max"] == 0
    if runs != 1:
        return None
    got = np.fromfile(dump, dtype=np.float64)
    assert got.size == Cref.size
    dense = lambda m: O.tiled_to_dense(d, m)
    norm = O.parity_norm(dense(got), dense(Cref), dense(A), dense(B), N)
    assert norm < 10.0, norm                                          # north star: < 10; observed < 0.1
    return norm


def test_overlay_exports_the_reference_symbol_names():
    _need_binaries()
    out = subprocess.run(["nm", "-D", "--defined-only", os.path.join(REF, "overlay", "libparsec.so.4")],
                         capture_output=True, text=True, check=True).stdout
    names = {l.split()[-1] for l in out.splitlines() if l.strip()}
    for sym in ("parsec_cuda_kernel_scheduler", "parsec_default_cuda_stage_in", "parsec_default_cuda_stage_out",
                "parsec_cuda_module_init", "parsec_cuda_module_fini", "device_cuda_static_component",
                "parsec_gpu_pop_workspace", "parsec_gpu_push_workspace", "parsec_gpu_free_workspace",
                "parsec_gpu_sort_pending_list"):
        assert sym in names, sym
    needed = subprocess.run(["readelf", "-d", os.path.join(REF, "overlay", "libparsec.so.4")],
                            capture_output=True, text=True, check=True).stdout
    assert "libparsec_b200.so" in needed                              # the module's work is done by the product library


@pytest.mark.parametrize("N,NB", [(2048, 256), (8192, 1024)])
def test_jdf_dgemm_over_the_overlay_matches_the_oracle(tmp_path, N, NB):
    _need_binaries()
    dump = str(tmp_path / "C.bin")
    out, err = run_app("overlay", ["-N", N, "-t", NB, "-c", 4, "--dump", dump])
    assert out["gpus"] == 1, err[-2000:]
    norm = check_against_oracle(out, dump, N, NB)
    print("overlay N=%d NB=%d: %.1f GFLOP/s first run, parity norm %.3g" % (N, NB, out["gflops_first"], norm))


def test_ragged_sizes_over_the_overlay(tmp_path):
    """Edge tiles (N not a multiple of NB) through generated code, the module and the kernel's TMA zero fill."""
    _need_binaries()
    N, NB = 1000, 192
    dump = str(tmp_path / "C.bin")
    out, err = run_app("overlay", ["-N", N, "-t", NB, "-c", 4, "--dump", dump])
    assert out["gpus"] == 1, err[-2000:]
    d, A, B, Cref, cnt, ver = oracle_result(N, NB)
    assert out["gemm_gpu"] == cnt.gemm and out["c_version_min"] == out["c_version_max"] == int(ver.max())
    got = np.fromfile(dump, dtype=np.float64)
    dense = lambda m: O.tiled_to_dense(d, m)
    assert O.parity_norm(dense(got), dense(Cref), dense(A), dense(B), N) < 10.0


def test_repeated_taskpools_keep_versions_and_cache(tmp_path):
    """Three taskpools in a row over the same collections: versions keep counting (3 * KT), tiles stay cached."""
    _need_binaries()
    out, err = run_app("overlay", ["-N", 2048, "-t", 256, "-c", 4, "-r", 3])
    assert out["gpus"] == 1, err[-2000:]
    check_against_oracle(out, None, 2048, 256, runs=3)


def test_per_stream_info_objects_serve_the_cublas_body(tmp_path):
    """The reference's own CUDA bodies fetch a per-stream cuBLAS handle with parsec_info_get(&gpu_stream->infos, ..)
    (tests/dsl/ptg/cuda/nvlink.jdf:140): the same body runs over the overlay's streams and gives the oracle's C."""
    _need_binaries()
    dump = str(tmp_path / "C.bin")
    out, err = run_app("overlay", ["-N", 2048, "-t", 256, "-c", 4, "--cublas", "--dump", dump])
    assert out["gpus"] == 1 and out["cublas"] == 1, err[-2000:]
    check_against_oracle(out, dump, 2048, 256)


def test_mca_parameters_reach_the_module(tmp_path):
    """device_cuda_memory_number_of_blocks caps the HBM heap (device_cuda_component.c:187-191): with 24 MiB of heap
    for a 96 MiB problem the run is out of core -- evictions and write-backs -- and still gives the oracle's C."""
    _need_binaries()
    dump = str(tmp_path / "C.bin")
    env_extra = {"PARSEC_MCA_device_cuda_memory_number_of_blocks": "48"}   # 48 x 512 KiB
    old = {k: os.environ.get(k) for k in env_extra}
    os.environ.update(env_extra)
    try:
        out, err = run_app("overlay", ["-N", 2048, "-t", 256, "-c", 4, "--dump", dump])
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    assert out["gpus"] == 1, err[-2000:]
    check_against_oracle(out, dump, 2048, 256)


def test_stock_reference_module_runs_the_same_binary(tmp_path):
    """The unmodified reference CUDA module resolves the same dyld= kernel (parsec_b200_dgemm_sm100) and the same JDF
    binary gives the oracle's C: the kernel library is a drop-in on its own, and the comparator is sound."""
    _need_binaries()
    if not os.path.exists(os.path.join(REF, "stock", "libparsec.so.4")):
        pytest.skip("stock libparsec not built")
    dump = str(tmp_path / "C.bin")
    out, err = run_app("stock", ["-N", 2048, "-t", 256, "-c", 4, "--dump", dump])
    assert out["gpus"] == 1, err[-2000:]
    check_against_oracle(out, dump, 2048, 256)


def test_declined_bodies_fall_back_to_the_cpu_incarnation(tmp_path):
    """Step k = 3 of every chain is declined by the CUDA body (PARSEC_HOOK_RETURN_NEXT): the module takes the task's
    stage-in back, writes the chain's dirty C tile home, and the runtime runs the CPU incarnation on current host data
    (device_cuda_module.c:2644-2656, :2472-2526 -- the reference leaves the host copy stale there).  The chain then
    goes back to the GPU, which must notice that the host copy is the newer one."""
    _need_binaries()
    N, NB, k = 2048, 256, 3
    dump = str(tmp_path / "C.bin")
    out, err = run_app("overlay", ["-N", N, "-t", NB, "-c", 4, "--decline", k, "--dump", dump])
    assert out["gpus"] == 1, err[-2000:]
    d, A, B, Cref, cnt, ver = oracle_result(N, NB)
    chains = (N // NB) ** 2
    assert out["gemm_cpu"] == chains and out["gemm_gpu"] == cnt.gemm - chains
    assert out["device_executed_tasks"] == cnt.gemm - chains
    assert out["c_version_min"] == out["c_version_max"] == N // NB
    got = np.fromfile(dump, dtype=np.float64)
    dense = lambda m: O.tiled_to_dense(d, m)
    assert O.parity_norm(dense(got), dense(Cref), dense(A), dense(B), N) < 10.0


def test_two_gpus_in_one_process_share_tiles_over_nvlink(tmp_path):
    """--mca device_cuda_enabled 2: the runtime spreads the C chains over two device modules of one process
    (parsec_get_best_device); A and B tiles wanted by both GPUs are staged once from the host and once from the peer
    (device_cuda_module.c:1310-1378), the nvlink.jdf pattern."""
    _need_binaries()
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    N, NB = 4096, 256
    dump = str(tmp_path / "C.bin")
    out, err = run_app("overlay", ["-N", N, "-t", NB, "-c", 6, "--dump", dump], mca=("device_cuda_enabled", "2"))
    assert out["gpus"] == 2, err[-2000:]
    # the reference's load balancer (parsec_get_best_device, device.c:121-171) may hand the first task of a chain to the
    # CPU cores when both GPUs are loaded: every GEMM ran somewhere, and what ran on a GPU went through the module
    d, A, B, Cref, cnt, ver = oracle_result(N, NB)
    assert out["gemm_gpu"] + out["gemm_cpu"] == cnt.gemm and out["gemm_gpu"] > 0.9 * cnt.gemm
    assert out["device_executed_tasks"] == out["gemm_gpu"]
    assert out["c_version_min"] == out["c_version_max"] == N // NB and out["ab_version_max"] == 0
    got = np.fromfile(dump, dtype=np.float64)
    dense = lambda m: O.tiled_to_dense(d, m)
    norm = O.parity_norm(dense(got), dense(Cref), dense(A), dense(B), N)
    assert norm < 10.0, norm
    tile = NB * NB * 8
    tiles = 3 * (N // NB) ** 2
    assert out["d2d_bytes"] > 0                                   # some operand tiles came from the peer GPU
    assert out["h2d_bytes"] + out["d2d_bytes"] >= tiles * tile    # every tile was staged at least once
    assert out["h2d_bytes"] < 2 * tiles * tile                    # and the host was not asked twice for everything
    print("2 GPUs, one process: h2d %.0f MiB, d2d %.0f MiB, norm %.3g" % (out["h2d_bytes"] / 2 ** 20, out["d2d_bytes"] / 2 ** 20, norm))
