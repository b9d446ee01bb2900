"""ctypes binding of libparsec_b200.so (the C ABI declared in include/parsec_b200.h).

Plumbing only: tests/ and bench.py drive the library through these wrappers.  There is no CPU fallback: if the
shared library is missing the import fails loudly, and every compute entry point needs a CUDA device.
"""
import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_PATH = os.environ.get("PB2_LIB") or os.path.join(HERE, "libparsec_b200.so")   # PB2_LIB: an A/B build (make variant)
HEADER = os.path.join(ROOT, "include", "parsec_b200.h")

SUCCESS = 0
HOOK_RETURN_DONE, HOOK_RETURN_AGAIN, HOOK_RETURN_NEXT, HOOK_RETURN_DISABLE = 0, -1, -2, -3
HOOK_RETURN_ASYNC = -4
HOOK_RETURN_ERROR = -5
FLOW_ACCESS_READ, FLOW_ACCESS_WRITE, FLOW_ACCESS_RW = 0x04, 0x08, 0x0C
MATRIX_DOUBLE, MATRIX_FLOAT, MATRIX_BF16 = 0, 1, 2
MAX_PARAM_COUNT = 20
ADVICE_PREFETCH, ADVICE_PREFERRED_DEVICE, ADVICE_WARMUP = 1, 2, 3


class DeviceParams(C.Structure):
    _fields_ = [("memory_use_percent", C.c_int), ("memory_block_size", C.c_int64),
                ("memory_number_of_blocks", C.c_int64), ("max_streams", C.c_int),
                ("max_number_of_ejected_data", C.c_int), ("max_inflight_tasks", C.c_int), ("trace_events", C.c_int),
                ("profile_launches", C.c_int), ("sort_pending_tasks", C.c_int),
                ("prefetch_deferred", C.c_int), ("use_priorities", C.c_int),
                ("batch_bodies", C.c_int)]


class DeviceStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "executed_tasks data_in_from_host data_in_from_peers data_out_to_host required_data_in required_data_out "
        "nb_data_faults nb_evictions nb_w2r_tasks nb_w2r_tiles nb_stalls kernel_launches hbm_bytes_total "
        "hbm_bytes_in_use host_submit_ns host_retire_ns nb_cleanouts nb_batches").split()] + [("data_in_from_device", C.c_uint64 * 10)]

    def as_dict(self):
        d = {n: int(getattr(self, n)) for n, _ in self._fields_[:-1]}
        d["data_in_from_device"] = [int(x) for x in self.data_in_from_device]
        return d


class GemmStats(C.Structure):
    _fields_ = [("nb_gemm", C.c_int64), ("nb_read_a", C.c_int64), ("nb_read_b", C.c_int64), ("seconds", C.c_double)]


SUBMIT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p)
STAGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint32, C.c_void_p)
COMPLETE_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_void_p)


class GpuTask(C.Structure):
    _fields_ = [("task_type", C.c_int32), ("pushout", C.c_int32), ("submit", SUBMIT_FN), ("stage_in", STAGE_FN),
                ("stage_out", STAGE_FN), ("ec", C.c_void_p), ("priority", C.c_int32), ("nb_flows", C.c_int32),
                ("flow_data", C.c_void_p * MAX_PARAM_COUNT), ("flow_access", C.c_uint8 * MAX_PARAM_COUNT),
                ("flow_nb_elts", C.c_uint32 * MAX_PARAM_COUNT), ("flow_dev_ptr", C.c_void_p * MAX_PARAM_COUNT),
                ("flow_host_ptr", C.c_void_p * MAX_PARAM_COUNT), ("complete", COMPLETE_FN),
                ("complete_cookie", C.c_void_p), ("locals", C.c_int64 * 8), ("trace_dc", C.c_int32 * MAX_PARAM_COUNT),
                ("reserved", C.c_void_p * 4)]


_lib = None


def declared_symbols():
    """Every function name include/parsec_b200.h declares."""
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b((?:pb2|parsec_b200)_[a-z0-9_]+)\s*\(", text)
    return sorted(set(n for n in names if not n.endswith("_fn")))


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("%s is missing: build it with `make -C parsec-1_b200` (no CPU fallback exists)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, u32, u64, dbl, sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_uint64, C.c_double, C.c_size_t
    sig = {
        "parsec_b200_dgemm_sm100": (i32, [C.c_char, C.c_char, i32, i32, i32, dbl, vp, i32, vp, i32, dbl, vp, i32, vp]),
        "parsec_b200_dsyrk_sm100": (i32, [C.c_char, C.c_char, i32, i32, dbl, vp, i32, dbl, vp, i32, vp]),
        "parsec_b200_dtrsm_workspace": (C.c_size_t, [C.c_char, i32, i32]),
        "parsec_b200_dpotrf_workspace": (C.c_size_t, [i32]),
        "parsec_b200_dpotrf_sm100": (i32, [C.c_char, i32, vp, i32, vp, C.c_size_t, vp, vp]),
        "parsec_b200_dtrtri_diag_sm100": (i32, [C.c_char, C.c_char, i32, vp, i32, vp, vp]),
        "parsec_b200_dtrsm_inv_sm100": (i32, [C.c_char, C.c_char, C.c_char, i32, i32, dbl, vp, i32, vp, vp, i32, vp,
                                              C.c_size_t, vp]),
        "parsec_b200_dtrsm_sm100": (i32, [C.c_char, C.c_char, C.c_char, C.c_char, i32, i32, dbl, vp, i32, vp, i32, vp,
                                          C.c_size_t, vp]),
        "parsec_b200_dgemm": (i32, [C.c_char, C.c_char, i32, i32, i32, dbl, vp, i32, vp, i32, dbl, vp, i32, vp]),
        "parsec_b200_gemm_bf16_sm100": (i32, [C.c_char, C.c_char, i32, i32, i32, C.c_float, vp, i32, vp, i32,
                                               C.c_float, vp, i32, vp]),
        "parsec_b200_gemm_tf32_sm100": (i32, [C.c_char, C.c_char, i32, i32, i32, C.c_float, vp, i32, vp, i32,
                                               C.c_float, vp, i32, vp]),
        "pb2_zone_create": (vp, [i32, sz]), "pb2_zone_destroy": (None, [vp]), "pb2_zone_malloc": (i64, [vp, sz]),
        "pb2_zone_free": (i32, [vp, i64]), "pb2_zone_in_use": (sz, [vp]),
        "pb2_data_create": (vp, [u64, u32, vp, u32]), "pb2_data_destroy": (None, [vp]),
        "pb2_data_copy_version": (i32, [vp, i32, C.POINTER(u32)]), "pb2_data_copy_coherency": (i32, [vp, i32]),
        "pb2_data_copy_readers": (i32, [vp, i32]), "pb2_data_owner_device": (i32, [vp]),
        "pb2_data_copy_ptr": (vp, [vp, i32]),
        "pb2_device_params_default": (None, [C.POINTER(DeviceParams)]),
        "pb2_device_create": (i32, [i32, C.POINTER(DeviceParams), C.POINTER(vp)]),
        "pb2_device_destroy": (i32, [vp]), "pb2_device_index": (i32, [vp]), "pb2_device_cuda_index": (i32, [vp]),
        "pb2_device_memory_register": (i32, [vp, vp, sz]), "pb2_device_memory_unregister": (i32, [vp, vp]),
        "pb2_device_memory_release": (i32, [vp]), "pb2_device_data_advise": (i32, [vp, vp, i32]),
        "pb2_device_get_stats": (i32, [vp, C.POINTER(DeviceStats), i32]),
        "pb2_device_get_events": (i64, [vp, vp, i64]),
        "pb2_device_get_launch_ms": (i64, [vp, vp, i64, C.POINTER(i32)]),
        "pb2_kernel_scheduler": (i32, [vp, vp]), "pb2_device_hold": (i32, [vp]), "pb2_device_manage": (i32, [vp]),
        "pb2_default_stage_in": (i32, [vp, u32, vp]), "pb2_default_stage_out": (i32, [vp, u32, vp]),
        "pb2_device_pop_workspace": (vp, [vp, vp, sz]), "pb2_device_push_workspace": (i32, [vp, vp]),
        "pb2_device_free_workspace": (i32, [vp]),
        "pb2_matrix_block_cyclic_new": (vp, [i32] * 16), "pb2_matrix_block_cyclic_free": (None, [vp]),
        "pb2_matrix_block_cyclic_lapack_new": (vp, [i32] * 18), "pb2_matrix_tile_offset": (i64, [vp, i32, i32]),
        "pb2_data_set_host_layout": (i32, [vp, u32, u32, sz, sz]),
        "pb2_matrix_set_mat": (i32, [vp, vp]), "pb2_matrix_nb_local_tiles": (i64, [vp]),
        "pb2_matrix_mt": (i32, [vp]), "pb2_matrix_nt": (i32, [vp]), "pb2_matrix_rank_of": (i32, [vp, i32, i32]),
        "pb2_matrix_position": (i64, [vp, i32, i32]), "pb2_matrix_data_key": (u64, [vp, i32, i32]),
        "pb2_matrix_data_of": (vp, [vp, i32, i32]), "pb2_matrix_register_memory": (i32, [vp, vp]),
        "pb2_matrix_unregister_memory": (i32, [vp, vp]), "pb2_matrix_fill_lcg": (i32, [vp, C.c_uint, vp]),
        "pb2_matrix_fill_lcg_noalloc": (i32, [vp, C.c_uint, vp]),
        "pb2_gemm_taskpool_new": (vp, [vp, vp, vp, vp]), "pb2_gemm_taskpool_set": (i32, [vp, C.c_char_p, i64]),
        "pb2_gemm_taskpool_run": (i32, [vp, C.POINTER(GemmStats)]), "pb2_gemm_taskpool_free": (None, [vp]),
    }
    for name, (res, args) in sig.items():
        if not hasattr(L, name):
            continue  # optional entry points are checked by tests/test_cabi_host.py against the header
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


class Device:
    """pb2_device_t: one GPU's device module (stream manager + HBM tile cache)."""

    def __init__(self, cuda_index=0, **kw):
        p = DeviceParams()
        lib().pb2_device_params_default(C.byref(p))
        for k, v in kw.items():
            setattr(p, k, v)
        h = C.c_void_p()
        rc = lib().pb2_device_create(cuda_index, C.byref(p), C.byref(h))
        if rc != SUCCESS:
            raise RuntimeError("pb2_device_create failed (%d): a CUDA device is required, there is no CPU path" % rc)
        self.h = h

    def stats(self, reset=False):
        s = DeviceStats()
        lib().pb2_device_get_stats(self.h, C.byref(s), int(reset))
        return s.as_dict()

    def events(self):
        import numpy as np
        dt = np.dtype([("dc", "<u4"), ("key", "<u4"), ("kind", "<i4"), ("k", "<i4"), ("gpu_version", "<u4"),
                       ("cpu_version", "<u4"), ("gpu_coh", "u1"), ("cpu_coh", "u1"), ("owner_device", "i1"),
                       ("transfer", "i1")])
        n = lib().pb2_device_get_events(self.h, None, 0)
        ev = np.zeros(n, dtype=dt)
        if n:
            lib().pb2_device_get_events(self.h, ev.ctypes.data, n)
        return ev

    def launch_ms(self, max_samples=1 << 20):
        import numpy as np
        out = np.zeros(max_samples, dtype=np.float32)
        ns = C.c_int()
        n = lib().pb2_device_get_launch_ms(self.h, out.ctypes.data, max_samples, C.byref(ns))
        return out[:max(n, 0)], ns.value

    def index(self):
        return lib().pb2_device_index(self.h)

    def close(self):
        if self.h:
            lib().pb2_device_destroy(self.h)
            self.h = None


class Matrix:
    """parsec_matrix_block_cyclic_t in tile storage, backed by a host buffer (numpy array or raw pointer)."""

    def __init__(self, mtype, M, N, mb, nb, P=1, Q=1, rank=0, buf=None, kp=1, kq=1, ip=0, jq=0, lapack=None):
        """lapack = (mloc, nloc): PARSEC_MATRIX_LAPACK storage, the buffer is a column-major mloc x nloc local matrix."""
        if lapack:
            self.h = C.c_void_p(lib().pb2_matrix_block_cyclic_lapack_new(mtype, rank, mb, nb, M, N, 0, 0, M, N, P, Q, kp, kq,
                                                                         ip, jq, lapack[0], lapack[1]))
        else:
            self.h = C.c_void_p(lib().pb2_matrix_block_cyclic_new(mtype, rank, mb, nb, M, N, 0, 0, M, N, P, Q, kp, kq, ip, jq))
        if not self.h:
            raise ValueError("pb2_matrix_block_cyclic_new rejected the arguments")
        self.mtype, self.M, self.N, self.mb, self.nb = mtype, M, N, mb, nb
        self.elem = {MATRIX_DOUBLE: 8, MATRIX_FLOAT: 4, MATRIX_BF16: 2}[mtype]
        self.nb_local_tiles = lib().pb2_matrix_nb_local_tiles(self.h)
        self.nbytes = lapack[0] * lapack[1] * self.elem if lapack else self.nb_local_tiles * mb * nb * self.elem
        self.buf = None
        if buf is not None:
            self.set_buffer(buf)

    def set_buffer(self, buf):
        self.buf = buf  # keep alive
        ptr = buf.ctypes.data if hasattr(buf, "ctypes") else (buf.data_ptr() if hasattr(buf, "data_ptr") else int(buf))
        lib().pb2_matrix_set_mat(self.h, ptr)

    def fill_lcg(self, seed, dev):
        rc = lib().pb2_matrix_fill_lcg(self.h, seed, dev.h)
        if rc != SUCCESS:
            raise RuntimeError("pb2_matrix_fill_lcg failed (%d)" % rc)

    def data_of(self, m, n):
        return C.c_void_p(lib().pb2_matrix_data_of(self.h, m, n))

    def version(self, m, n, device_index=0):
        v = C.c_uint32()
        rc = lib().pb2_data_copy_version(self.data_of(m, n), device_index, C.byref(v))
        return int(v.value) if rc == SUCCESS else None

    def register(self, dev):
        return lib().pb2_matrix_register_memory(self.h, dev.h)

    def unregister(self, dev):
        return lib().pb2_matrix_unregister_memory(self.h, dev.h)

    def close(self):
        if self.h:
            lib().pb2_matrix_block_cyclic_free(self.h)
            self.h = None


class BcastOp(C.Structure):
    _fields_ = [("which", C.c_int32), ("k", C.c_int32), ("idx", C.c_int32), ("root", C.c_int32),
                ("i_am_root", C.c_int32)]


def comm_plan(P, Q, rank, MT, NT, KT, kp=1, kq=1, a_kq=1, b_kp=1, ip=0, jq=0):
    """pb2_comm_plan[_ex]: this rank's canonical broadcast sequence as a list of (which, k, idx, root, i_am_root)."""
    L = lib()
    L.pb2_comm_plan.restype = C.c_int64
    L.pb2_comm_plan.argtypes = [C.c_int] * 6 + [C.c_void_p, C.c_int64]
    L.pb2_comm_plan_ex.restype = C.c_int64
    L.pb2_comm_plan_ex.argtypes = [C.c_int] * 12 + [C.c_void_p, C.c_int64]
    if (kp, kq, a_kq, b_kp, ip, jq) == (1, 1, 1, 1, 0, 0):
        call = lambda out, cap: L.pb2_comm_plan(P, Q, rank, MT, NT, KT, out, cap)
    else:
        call = lambda out, cap: L.pb2_comm_plan_ex(P, Q, rank, MT, NT, KT, kp, kq, a_kq, b_kp, ip, jq, out, cap)
    n = call(None, 0)
    if n < 0:
        raise ValueError("pb2_comm_plan rejected the arguments (%d)" % n)
    ops = (BcastOp * max(n, 1))()
    call(ops, n)
    return [(o.which, o.k, o.idx, o.root, o.i_am_root) for o in ops[:n]]


def grid_for(world):
    """P x Q with P <= Q, as square as possible: 1 -> 1x1, 2 -> 1x2, 4 -> 2x2, 8 -> 2x4."""
    p = int(world ** 0.5)
    while world % p:
        p -= 1
    return p, world // p


class Comm:
    """pb2_comm_t: row / column NCCL communicators of a P x Q grid; ids are exchanged over torch.distributed."""

    def __init__(self, cuda_index, P, Q, rank):
        import torch
        import torch.distributed as dist
        L = lib()
        L.pb2_comm_unique_id.argtypes = [C.c_void_p]
        L.pb2_comm_create.argtypes = [C.c_int] * 4 + [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
        L.pb2_comm_destroy.argtypes = [C.c_void_p]
        L.pb2_gemm_taskpool_set_comm.argtypes = [C.c_void_p, C.c_void_p]
        # every rank publishes two ids; the row communicator of process row r uses the first id of rank (r, 0), the
        # column communicator of process column c the second id of rank (0, c)
        ids = (C.c_ubyte * 256)()
        for off in (0, 128):
            if L.pb2_comm_unique_id(C.byref(ids, off)) != SUCCESS:
                raise RuntimeError("ncclGetUniqueId failed")
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        t = torch.tensor(list(ids), dtype=torch.uint8, device=dev)
        allids = [torch.empty_like(t) for _ in range(P * Q)]
        dist.all_gather(allids, t)
        myrow, mycol = rank // Q, rank % Q
        self._row = (C.c_ubyte * 128).from_buffer_copy(bytes(allids[myrow * Q].cpu().tolist()[:128]))
        self._col = (C.c_ubyte * 128).from_buffer_copy(bytes(allids[mycol].cpu().tolist()[128:]))
        h = C.c_void_p()
        rc = L.pb2_comm_create(cuda_index, P, Q, rank, self._row, self._col, C.byref(h))
        if rc != SUCCESS:
            raise RuntimeError("pb2_comm_create failed (%d)" % rc)
        self.h, self.P, self.Q, self.rank = h, P, Q, rank

    def close(self):
        if self.h:
            lib().pb2_comm_destroy(self.h)
            self.h = None


def tiled_gemm(A, B, Cm, dev, comm=None, **knobs):
    """C += A * B through the device pipeline (parsec_dgemm_NN taskpool: add_taskpool + start + wait)."""
    tp = C.c_void_p(lib().pb2_gemm_taskpool_new(A.h, B.h, Cm.h, dev.h))
    if not tp:
        raise ValueError("pb2_gemm_taskpool_new rejected the descriptors (non-conforming tilings or dtypes)")
    try:
        if comm is not None:
            lib().pb2_gemm_taskpool_set_comm.argtypes = [C.c_void_p, C.c_void_p]
            if lib().pb2_gemm_taskpool_set_comm(tp, comm.h) != SUCCESS:
                raise ValueError("communicator does not match the descriptors' process grid")
        for k, v in knobs.items():
            if lib().pb2_gemm_taskpool_set(tp, k.encode(), int(v)) != SUCCESS:
                raise KeyError(k)
        st = GemmStats()
        rc = lib().pb2_gemm_taskpool_run(tp, C.byref(st))
        if rc != SUCCESS:
            raise RuntimeError("pb2_gemm_taskpool_run failed (%d)" % rc)
        return {"nb_gemm": st.nb_gemm, "nb_read_a": st.nb_read_a, "nb_read_b": st.nb_read_b, "seconds": st.seconds}
    finally:
        lib().pb2_gemm_taskpool_free(tp)


def dsyrk(uplo, trans, n, k, alpha, A_ptr, lda, beta, C_ptr, ldc, stream=0):
    return lib().parsec_b200_dsyrk_sm100(uplo.encode(), trans.encode(), n, k, alpha, A_ptr, lda, beta, C_ptr, ldc, stream)


def dtrsm_workspace(side, m, n):
    return lib().parsec_b200_dtrsm_workspace(side.encode(), m, n)


def dtrsm(side, uplo, trans, diag, m, n, alpha, T_ptr, ldt, B_ptr, ldb, ws_ptr, ws_bytes, stream=0):
    return lib().parsec_b200_dtrsm_sm100(side.encode(), uplo.encode(), trans.encode(), diag.encode(), m, n, alpha, T_ptr, ldt,
                                         B_ptr, ldb, ws_ptr, ws_bytes, stream)


def dgemm(transa, transb, m, n, k, alpha, A_ptr, lda, B_ptr, ldb, beta, C_ptr, ldc, stream=0):
    return lib().parsec_b200_dgemm_sm100(transa.encode(), transb.encode(), m, n, k, alpha, A_ptr, lda, B_ptr, ldb, beta,
                                         C_ptr, ldc, stream)
