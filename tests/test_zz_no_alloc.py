"""pb2_matrix_fill_lcg_noalloc: the input generator for callers that must not allocate device memory.

Why it exists: once pb2_device_create has taken its 95 % of HBM for the tile heap, a cudaMalloc for a staging buffer
can fail, and while the pipeline is running a cudaMalloc/cudaFree would synchronise the whole device.  This entry point
stages through a fixed 16 MiB buffer that is part of the library image instead.

The file name makes pytest collect it last, so a run that stops at the first failure has already run everything else.
The output is checked bit for bit (the generator is integer arithmetic plus two float roundings) against the CPU oracle,
whose generator is pinned on the reference's own initialize_tile() (tests/golden/lcg_tiles.json), and against the
allocating entry point -- on calls in a row, because every call after the first finds the previous call's bytes in the
shared scratch.  Plain, k-cyclic and origin-shifted grids: the entry point derives tile coordinates from storage
positions, which is where a k-cyclic descriptor differs from a plain one.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()


def _fill(fn_name, mtype, N, NB, seed, dev, P=1, Q=1, rank=0, kp=1, kq=1, ip=0, jq=0):
    import torch
    m = H.Matrix(mtype, N, N, NB, NB, P, Q, rank, kp=kp, kq=kq, ip=ip, jq=jq)
    buf = torch.zeros(max(m.nbytes, 1), dtype=torch.uint8).pin_memory()
    m.set_buffer(buf)
    fn = getattr(H.lib(), fn_name)
    fn.restype = H.C.c_int
    fn.argtypes = [H.C.c_void_p, H.C.c_uint, H.C.c_void_p]
    rc = fn(m.h, seed, dev.h)
    out = buf.numpy()[:m.nbytes].copy()
    m.close()
    return rc, out


def _oracle_fill(N, NB, seed, P, Q, rank, kp, kq, ip, jq):
    d = O.bc(rank, NB, NB, N, N, 0, 0, N, N, P, Q, kp, kq, ip, jq)
    return O.fill(d, seed)


def test_noalloc_generator_matches_the_oracle_twice_in_a_row():
    import torch
    dev = H.Device(0)   # the default: 95 % of HBM goes to the tile heap, which is the situation this entry point is for
    try:
        # first use on this device: resolves the scratch address and loads the kernel image
        assert _fill("pb2_matrix_fill_lcg_noalloc", H.MATRIX_DOUBLE, 128, 64, 1, dev)[0] == 0
        # calls in a row with different shapes, seeds, grids and element types: each one finds a dirty scratch
        cases = [(H.MATRIX_DOUBLE, 1024, 256, 1789, dict(P=1, Q=1, rank=0)),
                 (H.MATRIX_DOUBLE, 1024, 256, 1789, dict(P=1, Q=1, rank=0)),     # the same call again
                 (H.MATRIX_DOUBLE, 1000, 96, 1805, dict(P=2, Q=2, rank=3)),
                 (H.MATRIX_DOUBLE, 1100, 64, 1901, dict(P=2, Q=2, rank=1, kp=2, kq=3, ip=1, jq=0)),   # k-cyclic, shifted origin
                 (H.MATRIX_DOUBLE, 900, 32, 7, dict(P=3, Q=2, rank=4, kp=3, kq=2, ip=2, jq=1)),
                 (H.MATRIX_BF16, 512, 128, 1901, dict(P=1, Q=2, rank=1))]
        for mtype, N, NB, seed, grid in cases:
            free_before = torch.cuda.mem_get_info()[0]
            rc, got = _fill("pb2_matrix_fill_lcg_noalloc", mtype, N, NB, seed, dev, **grid)
            assert rc == 0
            assert torch.cuda.mem_get_info()[0] == free_before      # nothing was allocated or freed on the device
            assert got.any()
            rc, sibling = _fill("pb2_matrix_fill_lcg", mtype, N, NB, seed, dev, **grid)
            assert rc == 0
            assert np.array_equal(got, sibling)                      # same bytes as the allocating entry point
            if mtype == H.MATRIX_DOUBLE:                             # and as the CPU oracle (fp64 generator)
                g = dict(P=1, Q=1, rank=0, kp=1, kq=1, ip=0, jq=0)
                g.update(grid)
                want = _oracle_fill(N, NB, seed, g["P"], g["Q"], g["rank"], g["kp"], g["kq"], g["ip"], g["jq"])
                assert np.array_equal(got.view(np.float64), want)
    finally:
        dev.close()


def test_noalloc_generator_rejects_tiles_beyond_the_scratch():
    dev = H.Device(0, memory_number_of_blocks=64)
    try:
        rc, out = _fill("pb2_matrix_fill_lcg_noalloc", H.MATRIX_DOUBLE, 2048, 2048, 1, dev)   # one 32 MiB tile
        assert rc == -8 and not out.any()          # PARSEC_ERR_VALUE_OUT_OF_BOUNDS, output untouched
        rc, out = _fill("pb2_matrix_fill_lcg_noalloc", H.MATRIX_DOUBLE, 1024, 1024, 1, dev)   # 8 MiB: fits
        assert rc == 0 and out.any()
    finally:
        dev.close()


def test_noalloc_generator_refuses_views_it_cannot_invert():
    """A sub-matrix view stores more tiles than it names; the entry point says so instead of filling the wrong ones."""
    import torch
    dev = H.Device(0, memory_number_of_blocks=64)
    try:
        L = H.lib()
        L.pb2_matrix_block_cyclic_new.restype = C.c_void_p
        h = C.c_void_p(L.pb2_matrix_block_cyclic_new(H.MATRIX_DOUBLE, 0, 64, 64, 512, 512, 0, 0, 256, 256, 1, 1, 1, 1, 0, 0))
        assert h
        buf = torch.zeros(512 * 512 * 8, dtype=torch.uint8).pin_memory()
        L.pb2_matrix_set_mat(h, C.c_void_p(buf.data_ptr()))
        L.pb2_matrix_fill_lcg_noalloc.restype = C.c_int
        L.pb2_matrix_fill_lcg_noalloc.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
        assert L.pb2_matrix_fill_lcg_noalloc(h, 1, dev.h) == -7      # PARSEC_ERR_NOT_SUPPORTED
        assert not buf.numpy().any()
        L.pb2_matrix_block_cyclic_free(h)
    finally:
        dev.close()
