"""PARSEC_MATRIX_LAPACK collections through the device pipeline: parsec_matrix_block_cyclic_lapack_init
descriptors over plain column-major matrices (ld = mloc, not padded to whole tiles).  The reference needs user stage
functions for such tiles (tests/dsl/ptg/cuda/stage_custom.jdf:28-109: cudaMemcpy2DAsync with dc->llm, whole mb x nb
windows, so only N % NB == 0); here the module's own stage-in / stage-out and its write-back of evicted tiles move the
window, short edge tiles included.  Checks: same result, bit for bit, as the same product on tile storage; the
north-star norm against the oracle; nothing written outside the matrix; and the reference's own criterion for custom
versus default staging (stage_main.c:59-93: agreement to 10e-15, here exact)."""
import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()
CANARY = 777.0


def _pinned_f64(n):
    import torch
    t = torch.empty(n, dtype=torch.float64).pin_memory()
    return t, t.numpy()


def _run(N, NB, dev_kw, lapack, **knobs):
    d = O.square(N, NB)
    tiled = [O.fill(d, s) for s in (1789, 1805, 1901)]
    dense = [O.tiled_to_dense(d, t) for t in tiled]
    dev = H.Device(0, **dev_kw)
    keep, mats = [], []
    try:
        for i in range(3):
            if lapack:
                m = H.Matrix(H.MATRIX_DOUBLE, N, N, NB, NB, lapack=(N, N))
                t, v = _pinned_f64(N * N + 64)
                v[:N * N] = dense[i].flatten(order="F")
                v[N * N:] = CANARY                                  # right behind the last column
            else:
                m = H.Matrix(H.MATRIX_DOUBLE, N, N, NB, NB)
                t, v = _pinned_f64(m.nbytes // 8)
                v[:] = tiled[i]
            m.set_buffer(t)
            keep.append((t, v))
            mats.append(m)
        st = H.tiled_gemm(mats[0], mats[1], mats[2], dev, **knobs)
        stats = dev.stats()
        v = keep[2][1]
        if lapack:
            assert np.all(v[N * N:] == CANARY)
            for i in (0, 1):                                         # read-only operands come back untouched
                assert np.array_equal(keep[i][1][:N * N], dense[i].flatten(order="F"))
            got = v[:N * N].reshape((N, N), order="F").copy()
        else:
            got = O.tiled_to_dense(d, v.copy())
        versions = {mats[2].version(m, n) for n in range(d.nt) for m in range(d.mt)}
        return got, dense, st, stats, versions, d
    finally:
        for m in mats:
            m.close()
        dev.close()


@pytest.mark.parametrize("N,NB", [(1024, 128), (1000, 96), (250, 40)])
def test_lapack_storage_equals_tile_storage_and_the_oracle(N, NB):
    kw = dict(memory_block_size=NB * NB * 8, memory_number_of_blocks=512)
    got_l, dense, st, stats, versions, d = _run(N, NB, kw, True)
    got_t, _, st_t, _, versions_t, _ = _run(N, NB, kw, False)
    assert np.array_equal(got_l, got_t)                  # same kernels on the same packed device tiles
    assert st["nb_gemm"] == st_t["nb_gemm"] == d.mt ** 3 and versions == versions_t == {d.mt}
    ref = O.fill(d, 1901)
    O.tiled_dgemm(N, N, N, NB, NB, NB, O.fill(d, 1789), O.fill(d, 1805), ref)
    assert O.parity_norm(got_l, O.tiled_to_dense(d, ref), dense[0], dense[1], N) < 10
    assert stats["data_out_to_host"] == d.mt * d.nt * NB * NB * 8   # counted in device-tile units like the reference


def test_lapack_storage_out_of_core_write_back():
    """HBM smaller than the working set: evicted dirty C tiles go back into their windows of the LAPACK matrix."""
    N, NB = 1024, 128
    kw = dict(memory_block_size=NB * NB * 8, memory_number_of_blocks=40, max_inflight_tasks=8)
    got_l, dense, st, stats, versions, d = _run(N, NB, kw, True, window=0)
    got_t, _, _, _, _, _ = _run(N, NB, dict(memory_block_size=NB * NB * 8, memory_number_of_blocks=512), False)
    assert stats["nb_evictions"] > 0 and stats["nb_w2r_tiles"] > 0
    assert np.array_equal(got_l, got_t) and versions == {d.mt}


def test_host_layout_argument_errors():
    L = H.lib()
    d = L.pb2_data_create(1, 2, None, 1024)
    assert L.pb2_data_set_host_layout(d, 64, 16, 256, 64) == 0
    assert L.pb2_data_set_host_layout(d, 64, 17, 256, 64) != 0       # 17 columns of 64 bytes exceed the 1024-byte tile
    assert L.pb2_data_set_host_layout(d, 64, 16, 32, 64) != 0        # pitch smaller than a column
    assert L.pb2_data_set_host_layout(d, 0, 0, 0, 0) == 0            # back to packed
    L.pb2_data_destroy(d)
    m = H.Matrix(H.MATRIX_DOUBLE, 256, 256, 64, 64, lapack=(256, 256))
    assert L.pb2_matrix_fill_lcg(m.h, 1, None) != 0
    m.close()
