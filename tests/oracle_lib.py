"""ctypes binding of oracle/liboracle.so -- test infrastructure only (see oracle/pb2_oracle.h)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None


class BC(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "mb nb lm ln i j m n mt nt lmt lnt llm lln bsiz P Q kp kq ip jq rrank crank myrank "
        "nb_elem_r nb_elem_c nb_local_tiles lapack").split()]


class Counts(C.Structure):
    _fields_ = [("gemm", C.c_int64), ("read_a", C.c_int64), ("read_b", C.c_int64)]


class Event(C.Structure):
    _fields_ = [("dc", C.c_uint32), ("key", C.c_uint32), ("kind", C.c_int32), ("k", C.c_int32),
                ("gpu_version", C.c_uint32), ("cpu_version", C.c_uint32), ("gpu_coh", C.c_uint8),
                ("cpu_coh", C.c_uint8), ("owner_device", C.c_int8), ("transfer", C.c_int8)]


EVENT_DTYPE = np.dtype([("dc", "<u4"), ("key", "<u4"), ("kind", "<i4"), ("k", "<i4"), ("gpu_version", "<u4"),
                        ("cpu_version", "<u4"), ("gpu_coh", "u1"), ("cpu_coh", "u1"), ("owner_device", "i1"),
                        ("transfer", "i1")])
assert EVENT_DTYPE.itemsize == C.sizeof(Event) == 28

EV_RESERVE, EV_STAGE_IN, EV_PUSH_DONE, EV_POP, EV_EPILOG, EV_COMPLETE, EV_W2R, EV_EVICT = range(1, 9)
COH_INVALID, COH_OWNED, COH_EXCLUSIVE, COH_SHARED = 0, 1, 2, 4


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
        L.orc_rnd64_jump.restype = C.c_uint64
        L.orc_rnd64_jump.argtypes = [C.c_uint64, C.c_uint64]
        L.orc_init_tile.argtypes = [C.c_void_p] + [C.c_int] * 6 + [C.c_uint]
        L.orc_bc_init.argtypes = [C.POINTER(BC)] + [C.c_int] * 15
        L.orc_bc_init_lapack.argtypes = [C.POINTER(BC)] + [C.c_int] * 17
        L.orc_bc_offset.restype = C.c_int64
        L.orc_bc_offset.argtypes = [C.POINTER(BC), C.c_int, C.c_int]
        L.orc_bc_rank_of.argtypes = [C.POINTER(BC), C.c_int, C.c_int]
        L.orc_bc_position.argtypes = [C.POINTER(BC), C.c_int, C.c_int]
        L.orc_bc_key.restype = C.c_uint64
        L.orc_bc_key.argtypes = [C.POINTER(BC), C.c_int, C.c_int]
        L.orc_bc_data_of_key.restype = C.c_uint64
        L.orc_bc_data_of_key.argtypes = [C.POINTER(BC), C.c_int, C.c_int]
        L.orc_bc_fill.argtypes = [C.POINTER(BC), C.c_void_p, C.c_uint, C.c_int]
        L.orc_dgemm_tile.argtypes = [C.c_int] * 3 + [C.c_double, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                                     C.c_double, C.c_void_p, C.c_int]
        L.orc_dgemm_tile_simple.argtypes = L.orc_dgemm_tile.argtypes
        L.orc_tiled_dgemm.argtypes = [C.c_int] * 6 + [C.c_void_p] * 3 + [C.c_int, C.POINTER(Counts), C.c_void_p]
        L.orc_gemm_tile_bf16.argtypes = [C.c_int] * 3 + [C.c_float, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                                         C.c_float, C.c_void_p, C.c_int]
        L.orc_f32_to_bf16_rn.restype = C.c_uint16
        L.orc_f32_to_bf16_rn.argtypes = [C.c_float]
        L.orc_matrix_inf_norm_tiled.restype = C.c_double
        L.orc_matrix_inf_norm_tiled.argtypes = [C.POINTER(BC), C.c_void_p]
        L.orc_trace_tiled_gemm.restype = C.c_int64
        L.orc_trace_tiled_gemm.argtypes = [C.c_int] * 4 + [C.c_void_p, C.c_int64]
        L.orc_zone_init.restype = C.c_void_p
        L.orc_zone_init.argtypes = [C.c_int, C.c_size_t]
        L.orc_zone_fini.argtypes = [C.c_void_p]
        L.orc_zone_malloc.restype = C.c_int64
        L.orc_zone_malloc.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_zone_free.argtypes = [C.c_void_p, C.c_int64]
        L.orc_zone_in_use.restype = C.c_size_t
        L.orc_zone_in_use.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def bc(myrank, mb, nb, lm, ln, i, j, m, n, P=1, Q=1, kp=1, kq=1, ip=0, jq=0):
    d = BC()
    lib().orc_bc_init(C.byref(d), myrank, mb, nb, lm, ln, i, j, m, n, P, Q, kp, kq, ip, jq)
    return d


def bc_lapack(myrank, mb, nb, lm, ln, P, Q, kp, kq, ip, jq, mloc, nloc):
    d = BC()
    lib().orc_bc_init_lapack(C.byref(d), myrank, mb, nb, lm, ln, 0, 0, lm, ln, P, Q, kp, kq, ip, jq, mloc, nloc)
    return d


def square(N, NB, P=1, Q=1, rank=0):
    return bc(rank, NB, NB, N, N, 0, 0, N, N, P, Q)


def fill(desc, seed, nthreads=8):
    mat = np.empty(desc.nb_local_tiles * desc.bsiz, dtype=np.float64)
    lib().orc_bc_fill(C.byref(desc), mat.ctypes.data, seed, nthreads)
    return mat


def tiled_dgemm(M, N, K, mb, nb, kb, A, B, Cm, nthreads=8):
    cnt = Counts()
    MT, NT = -(-M // mb), -(-N // nb)
    ver = np.zeros(MT * NT, dtype=np.uint32)
    rc = lib().orc_tiled_dgemm(M, N, K, mb, nb, kb, A.ctypes.data, B.ctypes.data, Cm.ctypes.data, nthreads,
                               C.byref(cnt), ver.ctypes.data)
    assert rc == 0
    return cnt, ver


def dgemm_tile(alpha, A, B, beta, Cm):
    """A (mb x kb), B (kb x nb), C (mb x nb): Fortran-ordered float64 arrays; C updated in place."""
    mb, kb = A.shape
    nb = B.shape[1]
    lib().orc_dgemm_tile(mb, nb, kb, alpha, A.ctypes.data, A.strides[1] // 8, B.ctypes.data, B.strides[1] // 8,
                         beta, Cm.ctypes.data, Cm.strides[1] // 8)


def trace(MT, NT, KT, device=2):
    n = lib().orc_trace_tiled_gemm(MT, NT, KT, device, None, 0)
    ev = np.zeros(n, dtype=EVENT_DTYPE)
    lib().orc_trace_tiled_gemm(MT, NT, KT, device, ev.ctypes.data, n)
    return ev


def tiled_to_dense(desc, mat):
    """Tile storage (P = Q = 1) -> dense (m x n) column-major numpy array."""
    out = np.zeros((desc.m, desc.n), order="F")
    for n in range(desc.nt):
        for m in range(desc.mt):
            pos = lib().orc_bc_position(C.byref(desc), m, n)
            tile = mat[pos * desc.bsiz:(pos + 1) * desc.bsiz].reshape((desc.mb, desc.nb), order="F")
            rm = min(desc.mb, desc.m - m * desc.mb)
            rn = min(desc.nb, desc.n - n * desc.nb)
            out[m * desc.mb:m * desc.mb + rm, n * desc.nb:n * desc.nb + rn] = tile[:rm, :rn]
    return out


def parity_norm(Cgot, Cref, A, B, N):
    """north-star acceptance norm: ||C - Cref||inf / (||A||inf ||B||inf N eps), eps = 2^-53."""
    eps = 2.0 ** -53
    inf = lambda X: np.abs(X).sum(axis=1).max()
    return inf(Cgot - Cref) / (inf(A) * inf(B) * N * eps)
