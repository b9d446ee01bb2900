"""The device pipeline (stage-in -> kernel -> stage-out, HBM cache, versions) against the oracle."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()


def _pinned(nbytes):
    import torch
    return torch.empty(nbytes, dtype=torch.uint8).pin_memory()


def _as_f64(t):
    return t.numpy().view(np.float64)


def test_generator_matches_oracle_bit_exact():
    dev = H.Device(0, memory_number_of_blocks=64)
    try:
        for (N, NB, seed) in [(512, 128, 1789), (250, 40, 1805)]:
            m = H.Matrix(H.MATRIX_DOUBLE, N, N, NB, NB)
            buf = _pinned(m.nbytes)
            m.set_buffer(buf)
            m.fill_lcg(seed, dev)
            want = O.fill(O.square(N, NB), seed)
            assert np.array_equal(_as_f64(buf), want)
            m.close()
    finally:
        dev.close()


def _run_tiled(M, N, K, nb, dev, seeds=(1789, 1805, 1901), **knobs):
    dA, dB, dC = O.bc(0, nb, nb, M, K, 0, 0, M, K), O.bc(0, nb, nb, K, N, 0, 0, K, N), O.bc(0, nb, nb, M, N, 0, 0, M, N)
    hA, hB, hC = O.fill(dA, seeds[0]), O.fill(dB, seeds[1]), O.fill(dC, seeds[2])
    mats, bufs = [], []
    for (r, c, h) in ((M, K, hA), (K, N, hB), (M, N, hC)):
        m = H.Matrix(H.MATRIX_DOUBLE, r, c, nb, nb)
        b = _pinned(m.nbytes)
        _as_f64(b)[:] = h
        m.set_buffer(b)
        mats.append(m)
        bufs.append(b)
    st = H.tiled_gemm(mats[0], mats[1], mats[2], dev, **knobs)
    ref = hC.copy()
    cnt, ver = O.tiled_dgemm(M, N, K, nb, nb, nb, hA, hB, ref)
    got = _as_f64(bufs[2]).copy()
    Ad, Bd = O.tiled_to_dense(dA, hA), O.tiled_to_dense(dB, hB)
    norm = O.parity_norm(O.tiled_to_dense(dC, got), O.tiled_to_dense(dC, ref), Ad, Bd, K)
    versions = np.array([mats[2].version(m, n) for n in range(dC.nt) for m in range(dC.mt)])
    return st, cnt, ver, versions, norm, mats


def test_config1_parity_counts_versions_and_trace():
    """BASELINE config 1 (N=2048, NB=256): values within the north-star norm, task counts and final versions equal to
    the reference run, and the per-tile data-state sequence bit-identical to the oracle's replay."""
    dev = H.Device(0, memory_number_of_blocks=2048, trace_events=1)
    try:
        st, cnt, ver, versions, norm, mats = _run_tiled(2048, 2048, 2048, 256, dev)
        assert norm < 10
        assert (st["nb_gemm"], st["nb_read_a"], st["nb_read_b"]) == (cnt.gemm, cnt.read_a, cnt.read_b) == (512, 64, 64)
        assert np.array_equal(versions, ver) and np.all(versions == 8)
        s = dev.stats()
        assert s["executed_tasks"] == 512 and s["kernel_launches"] == 512
        assert s["data_in_from_host"] == 3 * 64 * 256 * 256 * 8      # every tile crosses PCIe exactly once
        assert s["data_out_to_host"] == 64 * 256 * 256 * 8
        got, want = dev.events(), O.trace(8, 8, 8, device=dev.index())
        assert len(got) == len(want)
        fields = ["kind", "k", "gpu_version", "cpu_version", "gpu_coh", "cpu_coh", "owner_device", "transfer"]
        for dc in (0, 1, 2):
            for key in range(64):
                g = got[(got["dc"] == dc) & (got["key"] == key)][fields]
                w = want[(want["dc"] == dc) & (want["key"] == key)][fields]
                if dc == 2:   # C chains are totally ordered by the DAG: exact sequence
                    assert np.array_equal(g, w), (dc, key)
                else:         # readers of A/B tiles interleave freely: same multiset of states
                    assert np.array_equal(np.sort(g, order=fields), np.sort(w, order=fields)), (dc, key)
        for m in mats:
            m.close()
    finally:
        dev.close()


def test_ragged_matrix():
    dev = H.Device(0, memory_number_of_blocks=512)
    try:
        st, cnt, ver, versions, norm, mats = _run_tiled(250, 130, 90, 40, dev)
        assert norm < 10 and st["nb_gemm"] == cnt.gemm == 7 * 4 * 3
        assert np.array_equal(versions, ver)
    finally:
        dev.close()


def test_single_tile_and_window():
    dev = H.Device(0, memory_number_of_blocks=512)
    try:
        _, _, ver, versions, norm, _ = _run_tiled(128, 128, 128, 128, dev)
        assert norm < 10 and list(versions) == [1]
        _, cnt, ver, versions, norm, _ = _run_tiled(1024, 1024, 512, 128, dev, window=5)
        assert norm < 10 and cnt.gemm == 8 * 8 * 4 and np.array_equal(versions, ver)
    finally:
        dev.close()


def test_sorted_pending_list_keeps_parity():
    """device_cuda_sort_pending_tasks: reordering the pending list must change neither values, counts nor versions."""
    nb = 128
    dev = H.Device(0, memory_block_size=nb * nb * 8, memory_number_of_blocks=60, max_inflight_tasks=4, sort_pending_tasks=1)
    try:
        st, cnt, ver, versions, norm, mats = _run_tiled(768, 768, 768, nb, dev)
        assert norm < 10 and st["nb_gemm"] == cnt.gemm == 6 ** 3 and np.array_equal(versions, ver)
    finally:
        dev.close()


def test_out_of_core_eviction_and_writeback():
    """HBM smaller than the working set: clean tiles are evicted, dirty C tiles are really written back (the
    reference's W2R copies nothing) and the result still matches the oracle."""
    nb = 128
    tile = nb * nb * 8
    dev = H.Device(0, memory_block_size=tile, memory_number_of_blocks=40, max_inflight_tasks=8)
    try:
        st, cnt, ver, versions, norm, mats = _run_tiled(1024, 1024, 1024, nb, dev)   # 192 tiles of data, 40 slots
        s = dev.stats()
        assert norm < 10
        assert st["nb_gemm"] == cnt.gemm == 512 and np.array_equal(versions, ver)
        assert s["nb_evictions"] > 0 and s["nb_w2r_tiles"] > 0
        assert s["hbm_bytes_in_use"] <= 40 * tile
    finally:
        dev.close()


def test_task_larger_than_hbm_fails_loudly():
    nb = 128
    dev = H.Device(0, memory_block_size=nb * nb * 8, memory_number_of_blocks=2)
    try:
        with pytest.raises(RuntimeError):
            _run_tiled(256, 256, 256, nb, dev)
    finally:
        dev.close()


def test_prefetch_advice_and_memory_release():
    dev = H.Device(0, memory_number_of_blocks=64)
    try:
        m = H.Matrix(H.MATRIX_DOUBLE, 256, 256, 128, 128)
        buf = _pinned(m.nbytes)
        m.set_buffer(buf)
        m.fill_lcg(7, dev)
        d = m.data_of(1, 0)
        assert H.lib().pb2_device_data_advise(dev.h, d, H.ADVICE_PREFETCH) == 0
        assert m.version(1, 0, dev.index()) == 0
        assert H.lib().pb2_data_copy_coherency(d, dev.index()) == 4      # SHARED
        assert H.lib().pb2_data_copy_readers(d, dev.index()) == 0
        import torch
        ptr = H.lib().pb2_data_copy_ptr(d, dev.index())
        back = torch.empty(128 * 128, dtype=torch.float64)
        C.cdll.LoadLibrary("libcudart.so").cudaMemcpy(C.c_void_p(back.data_ptr()), C.c_void_p(ptr), C.c_size_t(128 * 128 * 8), 2)
        pos = H.lib().pb2_matrix_position(m.h, 1, 0)
        assert np.array_equal(back.numpy(), _as_f64(buf)[pos * 128 * 128:(pos + 1) * 128 * 128])
        assert H.lib().pb2_device_data_advise(dev.h, d, H.ADVICE_PREFERRED_DEVICE) == 0
        assert H.lib().pb2_device_memory_release(dev.h) == 0
        assert H.lib().pb2_data_copy_coherency(d, dev.index()) == -3     # copy dropped from the cache
        assert dev.stats()["hbm_bytes_in_use"] == 0
        m.close()
    finally:
        dev.close()
