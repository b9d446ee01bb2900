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
