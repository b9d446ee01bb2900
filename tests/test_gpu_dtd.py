"""DTD caller of the device boundary: the insertion loop of tests/dsl/dtd/dtd_test_simple_gemm.c:280-303 (for i, j, k:
insert GEMM(A(i,k) INPUT, B(k,j) INPUT, C(i,j) INOUT [| PUSHOUT on the last k])) checked against the oracle."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()


def test_dtd_simple_gemm_matches_oracle():
    import torch
    L = H.lib()
    L.pb2_dtd_taskpool_new.restype = C.c_void_p
    L.pb2_dtd_taskpool_new.argtypes = [C.c_void_p]
    L.pb2_dtd_insert_task.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p),
                                      C.POINTER(C.c_uint8), C.c_int32, C.POINTER(C.c_int64), C.c_int]
    L.pb2_dtd_taskpool_wait.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    L.pb2_dtd_taskpool_free.argtypes = [C.c_void_p]
    body = C.cast(L.pb2_dtd_gemm_body, C.c_void_p)

    N, NB = 768, 128
    MT = N // NB
    dev = H.Device(0, memory_number_of_blocks=1024)
    try:
        d = O.square(N, NB)
        host = [O.fill(d, s) for s in (1789, 1805, 1901)]
        mats, bufs = [], []
        for h in host:
            m = H.Matrix(H.MATRIX_DOUBLE, N, N, NB, NB)
            b = torch.empty(m.nbytes, dtype=torch.uint8).pin_memory()
            b.numpy().view(np.float64)[:] = h
            m.set_buffer(b)
            mats.append(m)
            bufs.append(b)
        A, B, Cm = mats
        tp = C.c_void_p(L.pb2_dtd_taskpool_new(dev.h))
        assert tp
        access = (C.c_uint8 * 3)(H.FLOW_ACCESS_READ, H.FLOW_ACCESS_READ, H.FLOW_ACCESS_RW)
        for i in range(MT):
            for j in range(MT):
                for k in range(MT):
                    data = (C.c_void_p * 3)(A.data_of(i, k), B.data_of(k, j), Cm.data_of(i, j))
                    loc = (C.c_int64 * 6)(i, j, k, NB, NB, NB)
                    pushout = (1 << 2) if k == MT - 1 else 0
                    prio = MT * MT * MT - i * MT + j          # the reference's priority expression (:287)
                    assert L.pb2_dtd_insert_task(tp, body, prio, 3, data, access, pushout, loc, 6) == 0
        done = C.c_int64()
        assert L.pb2_dtd_taskpool_wait(tp, C.byref(done)) == 0
        assert done.value == MT ** 3
        L.pb2_dtd_taskpool_free(tp)
        ref = host[2].copy()
        cnt, ver = O.tiled_dgemm(N, N, N, NB, NB, NB, host[0], host[1], ref)
        got = bufs[2].numpy().view(np.float64)
        Ad, Bd = O.tiled_to_dense(d, host[0]), O.tiled_to_dense(d, host[1])
        assert O.parity_norm(O.tiled_to_dense(d, got), O.tiled_to_dense(d, ref), Ad, Bd, N) < 10
        versions = np.array([Cm.version(m, n) for n in range(MT) for m in range(MT)])
        assert np.array_equal(versions, ver)
        s = dev.stats()
        assert s["executed_tasks"] == cnt.gemm == MT ** 3
        assert s["data_out_to_host"] == MT * MT * NB * NB * 8        # PUSHOUT only on the last k of each chain
        for m in mats:
            m.close()
    finally:
        dev.close()


def test_dtd_write_after_read_is_ordered():
    """A reader inserted before a writer of the same tile must run first (WAR), as in the reference's DTD."""
    import torch
    L = H.lib()
    L.pb2_dtd_taskpool_new.restype = C.c_void_p
    L.pb2_dtd_taskpool_new.argtypes = [C.c_void_p]
    L.pb2_dtd_insert_task.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p),
                                      C.POINTER(C.c_uint8), C.c_int32, C.POINTER(C.c_int64), C.c_int]
    L.pb2_dtd_taskpool_wait.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    L.pb2_dtd_taskpool_free.argtypes = [C.c_void_p]
    body = C.cast(L.pb2_dtd_gemm_body, C.c_void_p)
    NB = 128
    dev = H.Device(0, memory_number_of_blocks=256)
    try:
        rng = np.random.default_rng(0)
        tiles = [np.asfortranarray(rng.uniform(-0.5, 0.5, (NB, NB))) for _ in range(3)]   # X, Y, Z
        bufs = [torch.from_numpy(t.flatten(order="F").copy()).pin_memory() for t in tiles]
        datas = [C.c_void_p(L.pb2_data_create(i, 7, b.data_ptr(), NB * NB * 8)) for i, b in enumerate(bufs)]
        X, Y, Z = datas
        tp = C.c_void_p(L.pb2_dtd_taskpool_new(dev.h))
        rr_w = (C.c_uint8 * 3)(H.FLOW_ACCESS_READ, H.FLOW_ACCESS_READ, H.FLOW_ACCESS_RW)
        loc = (C.c_int64 * 6)(0, 0, 0, NB, NB, NB)
        # T1: Z += X * Y   (reads X)        T2: X += Y * Y (writes X: must wait for T1)        T3: Z += X * Y (new X)
        for flows in ((X, Y, Z), (Y, Y, X), (X, Y, Z)):
            data = (C.c_void_p * 3)(*flows)
            assert L.pb2_dtd_insert_task(tp, body, 0, 3, data, rr_w, 1 << 2, loc, 6) == 0
        done = C.c_int64()
        assert L.pb2_dtd_taskpool_wait(tp, C.byref(done)) == 0 and done.value == 3
        L.pb2_dtd_taskpool_free(tp)
        x, y, z = tiles
        z1 = z + x @ y
        x2 = x + y @ y
        z3 = z1 + x2 @ y
        gz = bufs[2].numpy().reshape((NB, NB), order="F")
        gx = bufs[0].numpy().reshape((NB, NB), order="F")
        assert np.allclose(gx, x2, atol=1e-12) and np.allclose(gz, z3, atol=1e-12)
        for d in datas:
            L.pb2_data_destroy(d)
    finally:
        dev.close()
