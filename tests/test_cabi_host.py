"""CPU-side checks of the C-ABI library: symbols, host-only logic (heap invariants, block-cyclic mapping)."""
import ctypes as C
import json
import os

import pytest

import oracle_lib as O
from conftest import load_hostapi

HERE = os.path.dirname(os.path.abspath(__file__))
H = load_hostapi()


def test_library_exports_every_declared_symbol():
    L = H.lib()
    names = H.declared_symbols()
    assert len(names) > 40
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_heap_invariants_on_the_reference_traces_and_random_ones():
    """The HBM heap has its own placement policy (size-class stacks over a coalescing map of free runs), so what is
    checked is what any allocator owes its caller, on the operation sequences recorded from the reference's
    zone_malloc.c runs and on random ones: extents are unit aligned, inside the heap and disjoint, in_use is the sum of
    the live extents, a freed extent is handed out again to the next request of its size, and once everything is freed
    the whole heap can be taken in one piece (free space merges back)."""
    import random
    L = H.lib()

    def run(nseg, unit, ops):
        z = L.pb2_zone_create(nseg, unit)
        live = {}                      # offset -> bytes (rounded to units)
        for op in ops:
            if op[0] == "malloc":
                size = op[1]
                rounded = -(-size // unit) * unit
                off = L.pb2_zone_malloc(z, size)
                if off == -1:
                    assert rounded > nseg * unit - sum(live.values()) or len(live) > 0   # only when space is short or split
                    continue
                assert off % unit == 0 and 0 <= off and off + rounded <= nseg * unit
                for o, b in live.items():
                    assert off + rounded <= o or o + b <= off, "overlapping extents"
                live[off] = rounded
            else:
                off = op[1]
                if off not in live:
                    continue           # the reference's trace frees its own offsets; replay frees ours instead
                assert L.pb2_zone_free(z, off) == 0
                del live[off]
            assert L.pb2_zone_in_use(z) == sum(live.values())
        # LIFO reuse of an exact size
        if live:
            off, b = next(iter(live.items()))
            assert L.pb2_zone_free(z, off) == 0
            assert L.pb2_zone_malloc(z, b) == off
        for off in list(live):
            assert L.pb2_zone_free(z, off) == 0
        assert L.pb2_zone_in_use(z) == 0
        assert L.pb2_zone_malloc(z, nseg * unit) == 0          # everything merged back into one run
        L.pb2_zone_destroy(z)

    for case in json.load(open(os.path.join(HERE, "golden", "zone_trace.json"))):
        # keep the recorded sizes and the malloc / free rhythm; frees are mapped onto our own live offsets
        ops, mine = [], []
        z = L.pb2_zone_create(case["nseg"], case["unit"])
        for op in case["ops"]:
            if op[0] == "malloc":
                off = L.pb2_zone_malloc(z, op[1])
                ops.append(("malloc", op[1]))
                if off != -1:
                    mine.append(off)
            elif mine:
                off = mine.pop(len(mine) // 2)
                L.pb2_zone_free(z, off)
                ops.append(("free", off))
        L.pb2_zone_destroy(z)
        run(case["nseg"], case["unit"], ops)

    rng = random.Random(7)
    for _ in range(20):
        nseg, unit = rng.choice([64, 333, 1024]), rng.choice([4096, 512 * 1024])
        ops, offs = [], []
        z = L.pb2_zone_create(nseg, unit)
        for _ in range(600):
            if offs and rng.random() < 0.45:
                off = offs.pop(rng.randrange(len(offs)))
                L.pb2_zone_free(z, off)
                ops.append(("free", off))
            else:
                size = rng.choice([1, unit, 2 * unit, 16 * unit, 16 * unit, 3 * unit + 1])
                off = L.pb2_zone_malloc(z, size)
                ops.append(("malloc", size))
                if off != -1:
                    offs.append(off)
        L.pb2_zone_destroy(z)
        run(nseg, unit, ops)


def test_heap_error_paths():
    L = H.lib()
    assert not L.pb2_zone_create(0, 4096)
    z = L.pb2_zone_create(4, 4096)
    assert L.pb2_zone_malloc(z, 5 * 4096) == -1
    off = L.pb2_zone_malloc(z, 1)
    assert off == 0 and L.pb2_zone_free(z, off) == 0
    assert L.pb2_zone_free(z, off) == -3  # PARSEC_ERR_NOT_FOUND: double free
    L.pb2_zone_destroy(z)


@pytest.mark.parametrize("P,Q", [(1, 1), (1, 2), (2, 2), (2, 4), (3, 2)])
def test_block_cyclic_matches_oracle(P, Q):
    N, NB = 1000, 96  # ragged: 11 x 11 tiles
    for rank in range(P * Q):
        m = H.Matrix(H.MATRIX_DOUBLE, N, N, NB, NB, P, Q, rank)
        d = O.square(N, NB, P, Q, rank)
        assert m.nb_local_tiles == d.nb_local_tiles
        L = H.lib()
        assert (L.pb2_matrix_mt(m.h), L.pb2_matrix_nt(m.h)) == (d.mt, d.nt)
        for n in range(d.nt):
            for mm in range(d.mt):
                assert L.pb2_matrix_rank_of(m.h, mm, n) == O.lib().orc_bc_rank_of(C.byref(d), mm, n)
                assert L.pb2_matrix_data_key(m.h, mm, n) == O.lib().orc_bc_key(C.byref(d), mm, n)
                if L.pb2_matrix_rank_of(m.h, mm, n) == rank:
                    assert L.pb2_matrix_position(m.h, mm, n) == O.lib().orc_bc_position(C.byref(d), mm, n)
                else:
                    assert not L.pb2_matrix_data_of(m.h, mm, n)  # not local: no data
        m.close()


def test_descriptor_argument_errors():
    L = H.lib()
    assert not L.pb2_matrix_block_cyclic_new(0, 0, 0, 4, 8, 8, 0, 0, 8, 8, 1, 1, 1, 1, 0, 0)   # mb = 0
    assert not L.pb2_matrix_block_cyclic_new(0, 4, 4, 4, 8, 8, 0, 0, 8, 8, 2, 2, 1, 1, 0, 0)   # rank >= P*Q
    assert not L.pb2_matrix_block_cyclic_new(0, 0, 4, 4, 8, 8, 0, 0, 8, 8, 1, 1, 0, 1, 0, 0)   # kp = 0
    assert not L.pb2_matrix_block_cyclic_new(9, 0, 4, 4, 8, 8, 0, 0, 8, 8, 1, 1, 1, 1, 0, 0)   # unknown mtype


def test_taskpool_conformity_of_k_cyclic_descriptors():
    """Owner-computes with row / column broadcasts needs A and C to deal tile rows alike and B and C tile columns alike,
    on one grid origin; how the k dimension is dealt is free.  Sub-matrix views are refused on a P x Q grid."""
    L = H.lib()

    def mats(P, Q, rank, a, b, c):
        return [H.Matrix(H.MATRIX_DOUBLE, 64, 64, 8, 8, P, Q, rank, kp=x[0], kq=x[1], ip=x[2], jq=x[3]) for x in (a, b, c)]

    def accepted(ms):
        tp = L.pb2_gemm_taskpool_new(ms[0].h, ms[1].h, ms[2].h, C.c_void_p(1))   # dev is only stored by _new
        if tp:
            L.pb2_gemm_taskpool_free(C.c_void_p(tp))
        for m in ms:
            m.close()
        return bool(tp)

    assert accepted(mats(1, 1, 0, (2, 3, 0, 0), (2, 3, 0, 0), (2, 3, 0, 0)))
    assert accepted(mats(2, 2, 1, (2, 1, 0, 0), (1, 3, 0, 0), (2, 3, 0, 0)))      # A: kp of C, B: kq of C, k free
    assert accepted(mats(2, 2, 3, (1, 1, 1, 1), (1, 1, 1, 1), (1, 1, 1, 1)))      # shifted origin, all alike
    assert not accepted(mats(2, 2, 1, (1, 1, 0, 0), (1, 3, 0, 0), (2, 3, 0, 0)))  # A deals rows unlike C
    assert not accepted(mats(2, 2, 1, (2, 1, 0, 0), (1, 1, 0, 0), (2, 3, 0, 0)))  # B deals columns unlike C
    assert not accepted(mats(2, 2, 1, (1, 1, 1, 0), (1, 1, 0, 0), (1, 1, 0, 0)))  # different origins
    sub = [C.c_void_p(L.pb2_matrix_block_cyclic_new(0, 1, 8, 8, 64, 64, 8, 0, 56, 64, 2, 2, 1, 1, 0, 0)) for _ in range(3)]
    assert all(sub) and not L.pb2_gemm_taskpool_new(sub[0], sub[1], sub[2], C.c_void_p(1))
    for m in sub:
        L.pb2_matrix_block_cyclic_free(m)


def test_mca_parameter_names_from_the_environment(monkeypatch):
    """device_cuda_* MCA parameters (device_cuda_component.c:163-206) reach a direct user of the library through the
    reference's own environment names."""
    L = H.lib()
    L.pb2_device_params_from_env.restype = C.c_int
    L.pb2_device_params_from_env.argtypes = [C.POINTER(H.DeviceParams)]
    L.pb2_device_is_enabled.restype = C.c_int
    L.pb2_device_is_enabled.argtypes = [C.c_int]
    for k in list(os.environ):
        if k.startswith("PARSEC_MCA_device_cuda_"):
            monkeypatch.delenv(k)
    p = H.DeviceParams()
    L.pb2_device_params_default(C.byref(p))
    assert (p.memory_use_percent, p.memory_block_size, p.memory_number_of_blocks) == (95, 512 * 1024, -1)
    assert p.max_number_of_ejected_data == 20 and p.sort_pending_tasks == 0
    assert L.pb2_device_params_from_env(C.byref(p)) == 0
    monkeypatch.setenv("PARSEC_MCA_device_cuda_memory_use", "40")
    monkeypatch.setenv("PARSEC_MCA_device_cuda_memory_block_size", "1048576")
    monkeypatch.setenv("PARSEC_MCA_device_cuda_memory_number_of_blocks", "2048")
    monkeypatch.setenv("PARSEC_MCA_device_cuda_max_streams", "8")
    monkeypatch.setenv("PARSEC_MCA_device_cuda_max_number_of_ejected_data", "5")
    monkeypatch.setenv("PARSEC_MCA_device_cuda_sort_pending_tasks", "1")
    assert L.pb2_device_params_from_env(C.byref(p)) == 6
    assert (p.memory_use_percent, p.memory_block_size, p.memory_number_of_blocks) == (40, 1048576, 2048)
    assert (p.max_streams, p.max_number_of_ejected_data, p.sort_pending_tasks) == (8, 5, 1)
    monkeypatch.setenv("PARSEC_MCA_device_cuda_memory_use", "250")      # out of range: ignored
    assert L.pb2_device_params_from_env(C.byref(p)) == 5 and p.memory_use_percent == 40
    # enabled / mask
    assert all(L.pb2_device_is_enabled(i) == 1 for i in range(8))
    monkeypatch.setenv("PARSEC_MCA_device_cuda_enabled", "2")
    assert [L.pb2_device_is_enabled(i) for i in range(4)] == [1, 1, 0, 0]
    monkeypatch.setenv("PARSEC_MCA_device_cuda_enabled", "-1")
    monkeypatch.setenv("PARSEC_MCA_device_cuda_mask", "0x5")
    assert [L.pb2_device_is_enabled(i) for i in range(4)] == [1, 0, 1, 0]
    monkeypatch.setenv("PARSEC_MCA_device_cuda_enabled", "0")
    assert L.pb2_device_is_enabled(0) == 0


def test_blas3_entry_points_validate_their_arguments_before_touching_the_gpu():
    """Argument errors of parsec_b200_dgemm / _dsyrk / _dtrsm / _dpotrf and the scratch sizes are host logic: they answer
    PB2_ERR_BAD_PARAM (-4) or 0 without a CUDA call, so they can be checked here."""
    import ctypes as C
    L = H.lib()
    p = C.c_void_p(0x1000)                      # never dereferenced on these paths
    assert H.dgemm("Q", "N", 8, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8) == -4
    assert H.dgemm("N", "T", 8, 8, 8, 1.0, p, 8, p, 4, 1.0, p, 8) == -4          # op(B) = B^T is stored n x k: ldb >= n
    assert H.dgemm("T", "N", 8, 8, 16, 1.0, p, 8, p, 16, 1.0, p, 8) == -4        # op(A) = A^T is stored k x m: lda >= k
    assert H.dgemm("N", "N", 0, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8) == 0
    assert H.dsyrk("X", "N", 8, 8, 1.0, p, 8, 1.0, p, 8) == -4
    assert H.dsyrk("L", "T", 8, 16, 1.0, p, 8, 1.0, p, 8) == -4                   # A is k x n: lda >= k
    assert H.dsyrk("U", "N", 0, 8, 1.0, p, 8, 1.0, p, 8) == 0
    # scratch: inverted 128 x 128 diagonal blocks of the triangular tile + one m x n image of the solution
    assert H.dtrsm_workspace("R", 1024, 1024) == (8 * 128 * 128 + 1024 * 1024) * 8 + 512
    assert H.dtrsm_workspace("L", 1000, 300) == (8 * 128 * 128 + 1000 * 300) * 8 + 512
    assert H.dtrsm_workspace("R", 1001, 130) == (2 * 128 * 128 + 1002 * 130) * 8 + 512   # odd m: leading dimension made even
    assert H.dtrsm_workspace("Z", 8, 8) == 0
    assert H.dtrsm("R", "L", "T", "N", 8, 8, 1.0, p, 4, p, 8, p, 1 << 20) == -4    # ldt < n
    assert H.dtrsm("R", "L", "T", "N", 8, 8, 1.0, p, 8, p, 8, p, 16) == -4         # scratch too small
    assert H.dtrsm("L", "U", "N", "Q", 8, 8, 1.0, p, 8, p, 8, p, 1 << 20) == -4    # diag
    assert H.dtrsm("L", "U", "N", "U", 0, 8, 1.0, p, 8, p, 8, p, 1 << 20) == 0
    assert L.parsec_b200_dpotrf_workspace(1024) == (128 * 128 + 1024 * 128) * 8 + 512
    assert L.parsec_b200_dpotrf_sm100(b"Q", 8, p, 8, p, 1 << 20, None, None) == -4
    assert L.parsec_b200_dpotrf_sm100(b"L", 8, p, 4, p, 1 << 20, None, None) == -4
    assert L.parsec_b200_dpotrf_sm100(b"L", 8, p, 8, p, 16, None, None) == -4
    assert L.parsec_b200_dpotrf_sm100(b"U", 0, p, 8, p, 16, None, None) == 0
