"""2D block-cyclic mapping against the REFERENCE's own two_dim_rectangle_cyclic.c / matrix.c.

tests/golden/block_cyclic.json holds, for 148 (matrix, tiling, storage, grid, k-cyclicity, grid origin, rank) descriptors, what the reference's
parsec_matrix_block_cyclic_init computed (tile counts, padded sizes, local tile counts) and, for every tile, the owner
rank (rank_of), the data key (data_key) and -- for local tiles -- the byte offset, key and size of the parsec_data_t
its data_of created.  The oracle's restatement and the product's pb2_matrix_* (C ABI) must agree on all of it.  CPU only."""
import ctypes as C
import json
import os

import oracle_lib as O
from conftest import load_hostapi

HERE = os.path.dirname(os.path.abspath(__file__))
H = load_hostapi()


def _cases():
    return json.load(open(os.path.join(HERE, "golden", "block_cyclic.json")))


def test_oracle_block_cyclic_matches_reference():
    L = O.lib()
    n = 0
    for c in _cases():
        if c["lapack"]:
            d = O.bc_lapack(c["rank"], c["mb"], c["nb"], c["M"], c["N"], c["P"], c["Q"], c["kp"], c["kq"], c["ip"], c["jq"],
                            *c["lapack"])
        else:
            d = O.bc(c["rank"], c["mb"], c["nb"], c["M"], c["N"], 0, 0, c["M"], c["N"], c["P"], c["Q"], c["kp"], c["kq"],
                     c["ip"], c["jq"])
        for k, v in c["geo"].items():
            assert getattr(d, k) == v, (k, c["P"], c["Q"], c["rank"])
        for t in c["tiles"]:
            m, nn, owner, key = t[:4]
            assert L.orc_bc_rank_of(C.byref(d), m, nn) == owner
            assert L.orc_bc_key(C.byref(d), m, nn) == key
            if owner == c["rank"]:
                assert L.orc_bc_offset(C.byref(d), m, nn) * 8 == t[4]
                if not c["lapack"]:
                    assert L.orc_bc_position(C.byref(d), m, nn) * d.bsiz * 8 == t[4]
                assert t[5] == L.orc_bc_data_of_key(C.byref(d), m, nn) and t[6] == d.bsiz * 8
                if c["kp"] == 1 and c["kq"] == 1:
                    assert t[5] == key          # k-cyclic data_of builds its key from reduced coordinates (reference quirk)
            n += 1
    assert n > 5000


def test_product_block_cyclic_matches_reference():
    L = H.lib()
    for c in _cases():
        mat = H.Matrix(H.MATRIX_DOUBLE, c["M"], c["N"], c["mb"], c["nb"], c["P"], c["Q"], c["rank"], kp=c["kp"], kq=c["kq"],
                       ip=c["ip"], jq=c["jq"], lapack=c["lapack"])
        assert (L.pb2_matrix_mt(mat.h), L.pb2_matrix_nt(mat.h)) == (c["geo"]["mt"], c["geo"]["nt"])
        assert mat.nb_local_tiles == c["geo"]["nb_local_tiles"]
        base = C.create_string_buffer(8)          # symbolic base address: data_of only does pointer arithmetic on it
        L.pb2_matrix_set_mat(mat.h, C.addressof(base))
        for t in c["tiles"]:
            m, n, owner, key = t[:4]
            assert L.pb2_matrix_rank_of(mat.h, m, n) == owner
            assert L.pb2_matrix_data_key(mat.h, m, n) == key
            if owner == c["rank"]:
                assert L.pb2_matrix_tile_offset(mat.h, m, n) * 8 == t[4]
                if not c["lapack"]:
                    assert L.pb2_matrix_position(mat.h, m, n) * c["mb"] * c["nb"] * 8 == t[4]
                d = C.c_void_p(L.pb2_matrix_data_of(mat.h, m, n))
                assert (L.pb2_data_copy_ptr(d, 0) or 0) - C.addressof(base) == t[4]   # same host address as data_of
            else:
                assert not L.pb2_matrix_data_of(mat.h, m, n)
        mat.close()
