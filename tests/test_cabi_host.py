"""CPU-side checks of the C-ABI library: symbols, host-only logic (heap placement, block-cyclic mapping)."""
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


def test_heap_placement_matches_reference_golden():
    """The HBM heap makes the same placement decisions as the reference's zone_malloc.c (golden offsets)."""
    L = H.lib()
    for case in json.load(open(os.path.join(HERE, "golden", "zone_trace.json"))):
        z = L.pb2_zone_create(case["nseg"], case["unit"])
        for op in case["ops"]:
            if op[0] == "malloc":
                assert L.pb2_zone_malloc(z, op[1]) == op[2]
                assert L.pb2_zone_in_use(z) == op[3]
            else:
                assert L.pb2_zone_free(z, op[1]) == 0
                assert L.pb2_zone_in_use(z) == op[2]
        L.pb2_zone_destroy(z)


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


def test_k_cyclic_descriptors_map_but_only_plain_cyclic_grids_run_distributed():
    """kp, kq > 1 are mapped exactly as the reference maps them (tests/test_block_cyclic_golden.py); the P x Q taskpool's
    broadcast plan is written for plain cyclic ownership, so it takes k-cyclic descriptors on one rank only."""
    L = H.lib()
    one = [H.Matrix(H.MATRIX_DOUBLE, 64, 64, 8, 8, 1, 1, 0, kp=2, kq=3) for _ in range(3)]
    tp = L.pb2_gemm_taskpool_new(one[0].h, one[1].h, one[2].h, C.c_void_p(1))   # dev is only stored by _new
    assert tp
    L.pb2_gemm_taskpool_free(C.c_void_p(tp))
    grid = [H.Matrix(H.MATRIX_DOUBLE, 64, 64, 8, 8, 2, 2, 1, kp=2, kq=1) for _ in range(3)]
    assert not L.pb2_gemm_taskpool_new(grid[0].h, grid[1].h, grid[2].h, C.c_void_p(1))
    for m in one + grid:
        m.close()


def test_data_master_starts_like_parsec_data_create():
    L = H.lib()
    d = C.c_void_p(L.pb2_data_create(42, 7, None, 1024))
    v = C.c_uint32(99)
    assert L.pb2_data_copy_version(d, 0, C.byref(v)) == 0 and v.value == 0
    assert L.pb2_data_copy_coherency(d, 0) == 1       # OWNED
    assert L.pb2_data_owner_device(d) == 0
    assert L.pb2_data_copy_coherency(d, 2) == -3      # no copy on the first GPU yet
    L.pb2_data_destroy(d)


def test_no_cpu_fallback_without_a_gpu():
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a CUDA device is present")
    except ImportError:
        pass
    with pytest.raises(RuntimeError, match="no CPU path"):
        H.Device(0)
