"""BASELINE configs[1] at full size (N = 32768, NB = 1024, fp64, one GPU) through the C ABI, checked with
size-independent properties because the oracle's O(N^3) DGEMM would take hours here:

  * Freivalds' check: for a random +-1 vector x,  C1 x  must equal  C0 x + A (B x)  to within the north-star norm
    ||r||inf / (||A||inf ||B||inf ||x||inf N eps) < 10  (three O(N^2) products on the host, tile by tile);
  * task counts MT*NT*KT / MT*KT / KT*NT as the reference's generated DAG produces them;
  * every C tile's host copy ends at version KT, no eviction, no staging beyond the
    3 * MT * NT cold misses (24 GiB in), each C tile written back once (8 GiB out).
"""
import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()
N, NB = 32768, 1024


def _pinned(nbytes):
    import torch
    return torch.empty(nbytes, dtype=torch.uint8).pin_memory()


def _matvec_tiled(buf, x, nt, mt, nb):
    """y = M x for a tile-stored matrix: tile (m, n) at ((n * mt) + m) * nb * nb, column-major inside the tile."""
    tiles = buf.numpy().view(np.float64).reshape(nt, mt, nb, nb)      # [n][m][col j][row i]
    xb = x.reshape(nt, nb)
    y = np.zeros((mt, nb))
    for n in range(nt):
        y += np.matmul(xb[n][None, None, :], tiles[n])[:, 0, :]       # sum_j x[j] * tile[j][i]
    return y.reshape(-1)


def _inf_norm_tiled(buf, nt, mt, nb):
    tiles = buf.numpy().view(np.float64).reshape(nt, mt, nb, nb)
    rows = np.zeros((mt, nb))
    for n in range(nt):
        rows += np.abs(tiles[n]).sum(axis=1)
    return rows.max()


def test_config2_full_size_freivalds_counts_and_versions():
    import psutil
    import torch
    if psutil.virtual_memory().available < 48 << 30 or torch.cuda.mem_get_info(0)[0] < 40 << 30:
        pytest.skip("needs 24 GiB of pinned host memory and as much HBM")
    T = N // NB
    dev = H.Device(0)
    mats, bufs = [], []
    try:
        for seed in (1789, 1805, 1901):
            m = H.Matrix(H.MATRIX_DOUBLE, N, N, NB, NB)
            b = _pinned(m.nbytes)
            m.set_buffer(b)
            m.fill_lcg(seed, dev)
            mats.append(m)
            bufs.append(b)
        rng = np.random.default_rng(11)
        x = rng.integers(0, 2, N).astype(np.float64) * 2.0 - 1.0
        c0x = _matvec_tiled(bufs[2], x, T, T, NB)
        sampled = [(0, 0), (T - 1, T - 2), (T // 2, T - 1)]     # C tiles recomputed on the host cores after the run
        tile_of = lambda b, m, n: b.numpy().view(np.float64)[(n * T + m) * NB * NB:(n * T + m + 1) * NB * NB]
        c_before = {mn: tile_of(bufs[2], *mn).reshape((NB, NB), order="F").copy(order="F") for mn in sampled}
        dev.stats(reset=True)
        st = H.tiled_gemm(mats[0], mats[1], mats[2], dev)
        s = dev.stats()
        assert (st["nb_gemm"], st["nb_read_a"], st["nb_read_b"]) == (T * T * T, T * T, T * T)
        assert s["executed_tasks"] == T * T * T and s["nb_evictions"] == 0 and s["nb_w2r_tiles"] == 0
        # bytes, as the reference counts them (nb_data_faults += nb_elts): each tile is staged in exactly once
        assert s["nb_data_faults"] == s["data_in_from_host"] == 3 * T * T * NB * NB * 8
        assert s["data_out_to_host"] == T * T * NB * NB * 8           # and each C tile goes back exactly once
        versions = np.array([mats[2].version(m, n) for n in range(T) for m in range(T)])
        assert np.all(versions == T)
        bx = _matvec_tiled(bufs[1], x, T, T, NB)
        abx = _matvec_tiled(bufs[0], bx, T, T, NB)
        c1x = _matvec_tiled(bufs[2], x, T, T, NB)
        eps = np.finfo(np.float64).eps
        denom = _inf_norm_tiled(bufs[0], T, T, NB) * _inf_norm_tiled(bufs[1], T, T, NB) * 1.0 * N * eps
        assert np.max(np.abs(c1x - c0x - abx)) / denom < 10
        # the run really changed C (a no-op pipeline would pass a weaker check)
        assert np.max(np.abs(c1x - c0x)) > 1.0
        # sampled tiles: the chain C(m,n) += A(m,k) B(k,n), k ascending, recomputed tile by tile -- the
        # first with the oracle's tile GEMM, the others with numpy -- against what the pipeline wrote back
        inf = lambda X: np.abs(X).sum(axis=1).max()
        for idx, (m, n) in enumerate(sampled):
            want = c_before[(m, n)]
            a_row = b_col = 0.0
            for k in range(T):
                a = np.asfortranarray(tile_of(bufs[0], m, k).reshape((NB, NB), order="F"))
                b = np.asfortranarray(tile_of(bufs[1], k, n).reshape((NB, NB), order="F"))
                if idx == 0:
                    O.dgemm_tile(1.0, a, b, 1.0, want)
                else:
                    want += a @ b
                a_row, b_col = a_row + np.abs(a).sum(axis=1), b_col + np.abs(b).sum(axis=0)
            got = tile_of(bufs[2], m, n).reshape((NB, NB), order="F")
            # row-wise forward bound of the inner products: |err_ij| <= N eps sum_k |a_ik| |b_kj|
            assert inf(got - want) / (a_row.max() * b_col.max() * N * eps) < 10
    finally:
        for m in mats:
            m.close()
        dev.close()
