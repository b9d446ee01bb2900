"""Per-stream scratch for task bodies: parsec_gpu_pop_workspace / _push_workspace / _free_workspace
(parsec/mca/device/device_gpu.c:168-246).  The body of stage_custom.jdf (A <- A*A + 2A) is run with the input first
copied into a workspace buffer, W = A; A = W*W + 2A, which also removes the aliasing of the in-place form; results
are checked against numpy and the stack / heap accounting against the reference's rules (two buffers per stream,
taken from the device heap at the first pop, kept until free_workspace)."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()


def _cudart():
    rt = C.CDLL("libcudart.so.12")
    rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    return rt


def test_workspace_body_and_stack_rules():
    import torch
    L = H.lib()
    rt = _cudart()
    NB, MT = 96, 5
    unit = 64 * 1024
    dev = H.Device(0, memory_block_size=unit, memory_number_of_blocks=2048)
    tile_bytes = NB * NB * 8
    units_per_buf = -(-tile_bytes // unit)
    try:
        rng = np.random.default_rng(3)
        host = torch.zeros(MT * MT * NB * NB, dtype=torch.float64).pin_memory()
        hv = host.numpy()
        hv[:] = rng.uniform(-0.5, 0.5, hv.shape)
        before = hv.copy()
        seen_streams, done, faults = set(), [], []

        @H.SUBMIT_FN
        def submit(_dev, task_p, stream):
            t = C.cast(task_p, C.POINTER(H.GpuTask)).contents
            a = t.flow_dev_ptr[0]
            w = L.pb2_device_pop_workspace(dev.h, stream, tile_bytes)
            if not w:
                faults.append("pop returned NULL")
                return -5
            if stream not in seen_streams:            # the stack rules, once per stream
                seen_streams.add(stream)
                w2 = L.pb2_device_pop_workspace(dev.h, stream, tile_bytes)
                if not w2 or w2 == w:
                    faults.append("second pop")
                if L.pb2_device_pop_workspace(dev.h, stream, tile_bytes):
                    faults.append("third pop must fail: PARSEC_GPU_MAX_WORKSPACE is 2")
                if L.pb2_device_push_workspace(dev.h, stream) != 0:
                    faults.append("push of the second buffer")
                if L.pb2_device_pop_workspace(dev.h, stream, tile_bytes) != w2:
                    faults.append("a pop after a push must return the buffer pushed last")
                L.pb2_device_push_workspace(dev.h, stream)
                if L.pb2_device_pop_workspace(dev.h, stream, tile_bytes * 64):
                    faults.append("a larger request than the first one must fail")
            if rt.cudaMemcpyAsync(w, a, tile_bytes, 3, stream) != 0:
                return -5
            rc = L.parsec_b200_dgemm(b"N", b"N", NB, NB, NB, 1.0, w, NB, w, NB, 2.0, a, NB, stream)
            if L.pb2_device_push_workspace(dev.h, stream) != 0:     # stream order protects the buffer (device_gpu.c:207)
                faults.append("push")
            return rc

        @H.COMPLETE_FN
        def complete(task_p, status, cookie):
            done.append(status)

        libc = C.CDLL(None)
        libc.calloc.restype = C.c_void_p
        datas = []
        assert L.pb2_device_pop_workspace(dev.h, None, tile_bytes) is None       # not an exec stream of this device
        assert L.pb2_device_push_workspace(dev.h, None) != 0
        assert L.pb2_device_hold(dev.h) == 1
        for i in range(MT * MT):
            data = C.c_void_p(L.pb2_data_create(i, 98, hv.ctypes.data + i * tile_bytes, tile_bytes))
            datas.append(data)
            raw = libc.calloc(1, C.sizeof(H.GpuTask))
            t = C.cast(raw, C.POINTER(H.GpuTask)).contents
            t.task_type, t.nb_flows, t.pushout = 0, 1, 1
            t.submit = submit
            t.flow_data[0] = data
            t.flow_access[0] = H.FLOW_ACCESS_RW
            t.flow_nb_elts[0] = tile_bytes
            t.complete = complete
            assert L.pb2_kernel_scheduler(dev.h, raw) == H.HOOK_RETURN_ASYNC
        assert L.pb2_device_manage(dev.h) == 0
        assert faults == [] and done == [0] * (MT * MT)
        for i in range(MT * MT):
            a = before[i * NB * NB:(i + 1) * NB * NB].reshape((NB, NB), order="F")
            got = hv[i * NB * NB:(i + 1) * NB * NB].reshape((NB, NB), order="F")
            assert np.max(np.abs(got - (a @ a + 2.0 * a))) <= 1e-13
        for d in datas:
            L.pb2_data_destroy(d)
        # the buffers stay with their streams until free_workspace: 2 per stream that ran a body, nothing else is left
        in_use = dev.stats()["hbm_bytes_in_use"]
        assert in_use == len(seen_streams) * 2 * units_per_buf * unit
        assert L.pb2_device_free_workspace(dev.h) == 0
        assert dev.stats()["hbm_bytes_in_use"] == 0
        assert L.pb2_device_free_workspace(dev.h) == 0            # idempotent
    finally:
        dev.close()
