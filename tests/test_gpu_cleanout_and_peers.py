"""Two parts of the device module that only show when something unusual happens, driven through the C ABI:

  * cleanout (parsec_cuda_kernel_cleanout, device_cuda_module.c:2472-2526): a body that declines
    (PARSEC_HOOK_RETURN_NEXT) after its tiles were staged.  Everything stage-in did is taken back -- readers, the WRITE
    flow's version bump, ownership, LRU membership -- the dirty input is written back so that another incarnation can
    read host memory, the task completes with the body's status and the device stays usable;
  * peer-GPU source (parsec_gpu_data_stage_in, :1310-1378, bookkeeping :1820-1890): two device modules in one process;
    a tile whose valid copy sits on the other GPU is staged device to device, read-only tiles prefer a same-version peer
    copy to the host, a writer takes the tile over.  Values against the oracle's tile GEMM.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()
NB = 128
TILE = NB * NB * 8


def _pinned(nbytes):
    import torch
    return torch.zeros(nbytes, dtype=torch.uint8).pin_memory()


class Tiles:
    """A, B, C tiles in pinned host memory with pb2 masters."""

    def __init__(self, names):
        self.buf = _pinned(TILE * len(names))
        self.host = self.buf.numpy().view(np.float64)
        self.names = list(names)
        rng = np.random.default_rng(5)
        self.host[:] = rng.random(self.host.size) - 0.5
        L = H.lib()
        self.data = {n: C.c_void_p(L.pb2_data_create(i, 7, self.host.ctypes.data + i * TILE, TILE)) for i, n in enumerate(names)}

    def view(self, name):
        i = self.names.index(name)
        return self.host[i * NB * NB:(i + 1) * NB * NB].reshape((NB, NB), order="F")

    def close(self):
        for d in self.data.values():
            H.lib().pb2_data_destroy(d)


def _task(tiles, flows, submit, complete, pushout=0):
    libc = C.CDLL(None)
    libc.calloc.restype = C.c_void_p
    raw = libc.calloc(1, C.sizeof(H.GpuTask))
    t = C.cast(raw, C.POINTER(H.GpuTask)).contents
    t.task_type, t.nb_flows, t.pushout = 0, len(flows), pushout
    t.submit = submit
    for i, (name, access) in enumerate(flows):
        t.flow_data[i] = tiles.data[name]
        t.flow_access[i] = access
        t.flow_nb_elts[i] = TILE
    t.complete = complete
    return raw


def _gemm_body(decline=None):
    L = H.lib()

    @H.SUBMIT_FN
    def submit(_dev, task_p, stream):
        t = C.cast(task_p, C.POINTER(H.GpuTask)).contents
        if decline is not None and decline():
            return H.HOOK_RETURN_NEXT
        rc = L.parsec_b200_dgemm(b"N", b"N", NB, NB, NB, 1.0, t.flow_dev_ptr[0], NB, t.flow_dev_ptr[1], NB, 1.0,
                                 t.flow_dev_ptr[2], NB, stream)
        return 0 if rc == 0 else -5
    return submit


def test_cleanout_takes_stage_in_back_and_writes_dirty_inputs_home():
    L = H.lib()
    dev = H.Device(0, memory_number_of_blocks=64, trace_events=1)
    tiles = Tiles(["A", "B", "C"])
    try:
        idx = dev.index()
        a, b = tiles.view("A").copy(), tiles.view("B").copy()
        want = tiles.view("C").copy(order="F")
        status, calls = [], [0]

        @H.COMPLETE_FN
        def complete(task_p, st, cookie):
            status.append(st)

        def decline_second():
            calls[0] += 1
            return calls[0] == 2

        body = _gemm_body(decline_second)
        flows = [("A", H.FLOW_ACCESS_READ), ("B", H.FLOW_ACCESS_READ), ("C", H.FLOW_ACCESS_RW)]
        # step 0 runs: C is now dirty on the device (no pushout), host copy stale
        assert L.pb2_kernel_scheduler(dev.h, _task(tiles, flows, body, complete)) == H.HOOK_RETURN_ASYNC
        O.dgemm_tile(1.0, np.asfortranarray(a), np.asfortranarray(b), 1.0, want)
        assert status == [0]
        v = C.c_uint32()
        L.pb2_data_copy_version(tiles.data["C"], idx, C.byref(v))
        gpu_v0 = v.value
        L.pb2_data_copy_version(tiles.data["C"], 0, C.byref(v))
        L.pb2_data_test_set(tiles.data["C"], 0, -1, v.value + 1, -1)      # the runtime's complete_hook: data_out->version++
        assert not np.allclose(tiles.view("C"), want)                       # the lie of the epilog: host bytes are old
        # step 1: staged, then the body declines
        assert L.pb2_kernel_scheduler(dev.h, _task(tiles, flows, body, complete)) == H.HOOK_RETURN_ASYNC
        assert status == [0, H.HOOK_RETURN_NEXT]
        s = dev.stats()
        assert s["nb_cleanouts"] == 1 and s["executed_tasks"] == 1 and s["kernel_launches"] == 1
        for name in ("A", "B", "C"):
            assert L.pb2_data_copy_readers(tiles.data[name], idx) == 0
        L.pb2_data_copy_version(tiles.data["C"], idx, C.byref(v))
        assert v.value == gpu_v0                                            # the WRITE bump was taken back
        assert np.max(np.abs(tiles.view("C") - want)) < 1e-12               # ...and the dirty tile went home
        # the device is still healthy: the next incarnation here is the same task again, now accepted
        assert L.pb2_kernel_scheduler(dev.h, _task(tiles, flows, body, complete, pushout=1 << 2)) == H.HOOK_RETURN_ASYNC
        assert status == [0, H.HOOK_RETURN_NEXT, 0]
        O.dgemm_tile(1.0, np.asfortranarray(a), np.asfortranarray(b), 1.0, want)
        assert np.max(np.abs(tiles.view("C") - want)) < 1e-12
        assert L.pb2_device_memory_release(dev.h) == 0
        assert dev.stats()["hbm_bytes_in_use"] == 0                         # nothing leaked out of the lists
    finally:
        tiles.close()
        dev.close()


def test_tiles_move_between_two_gpus_of_one_process():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    L = H.lib()
    d0 = H.Device(0, memory_number_of_blocks=64)
    d1 = H.Device(1, memory_number_of_blocks=64)
    tiles = Tiles(["A", "B", "C"])
    try:
        i0, i1 = d0.index(), d1.index()
        a, b = np.asfortranarray(tiles.view("A").copy()), np.asfortranarray(tiles.view("B").copy())
        want = tiles.view("C").copy(order="F")
        status = []

        @H.COMPLETE_FN
        def complete(task_p, st, cookie):
            status.append(st)

        body = _gemm_body()
        flows = [("A", H.FLOW_ACCESS_READ), ("B", H.FLOW_ACCESS_READ), ("C", H.FLOW_ACCESS_RW)]
        v = C.c_uint32()

        def bump_host_version():
            L.pb2_data_copy_version(tiles.data["C"], 0, C.byref(v))
            L.pb2_data_test_set(tiles.data["C"], 0, -1, v.value + 1, -1)

        # GPU 0 runs step 0: everything comes from the host
        assert L.pb2_kernel_scheduler(d0.h, _task(tiles, flows, body, complete)) == H.HOOK_RETURN_ASYNC
        bump_host_version()
        O.dgemm_tile(1.0, a, b, 1.0, want)
        s0 = d0.stats()
        assert s0["data_in_from_host"] == 3 * TILE and s0["data_in_from_peers"] == 0
        assert L.pb2_data_owner_device(tiles.data["C"]) == i0
        # GPU 1 runs step 1: C's only valid copy is the dirty one on GPU 0; A and B sit on GPU 0 at the host's version
        assert L.pb2_kernel_scheduler(d1.h, _task(tiles, flows, body, complete)) == H.HOOK_RETURN_ASYNC
        bump_host_version()
        O.dgemm_tile(1.0, a, b, 1.0, want)
        s1 = d1.stats()
        assert s1["data_in_from_peers"] == 3 * TILE and s1["data_in_from_host"] == 0
        assert s1["data_in_from_device"][i0] == 3 * TILE
        assert L.pb2_data_owner_device(tiles.data["C"]) == i1
        # GPU 0 takes the tile back and sends it home; its own old copy of C had been demoted, A and B are still there
        assert L.pb2_kernel_scheduler(d0.h, _task(tiles, flows, body, complete, pushout=1 << 2)) == H.HOOK_RETURN_ASYNC
        bump_host_version()
        O.dgemm_tile(1.0, a, b, 1.0, want)
        s0 = d0.stats()
        assert s0["data_in_from_peers"] == TILE and s0["data_in_from_host"] == 3 * TILE
        assert status == [0, 0, 0]
        assert np.max(np.abs(tiles.view("C") - want)) < 1e-12
        for dev, idx in ((d0, i0), (d1, i1)):
            assert L.pb2_device_memory_release(dev.h) == 0
            for name in ("A", "B", "C"):
                assert L.pb2_data_copy_readers(tiles.data[name], idx) in (0, -3)   # no reader left, or no copy left
    finally:
        tiles.close()
        d1.close()
        d0.close()
