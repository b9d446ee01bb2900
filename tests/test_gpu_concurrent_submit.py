"""parsec_cuda_kernel_scheduler is called by every worker thread of the runtime at once: the first caller becomes the
device's manager and drives push -> exec -> pop, every other caller only queues its task and returns
PARSEC_HOOK_RETURN_ASYNC, and a caller that arrives after the manager has left takes the seat again
(device_cuda_module.c:2573-2589, :2742-2753).  Four threads submit tasks concurrently, without pb2_device_hold;
every task must complete exactly once with the right numbers."""
import ctypes as C
import threading

import numpy as np
import pytest

from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()


@pytest.mark.parametrize("pause_us", [0, 150])
def test_four_threads_submit_concurrently(pause_us):
    """pause_us = 0: one manager mostly drives everything.  With a pause between a thread's submissions the device keeps
    draining, so the manager seat changes hands over and over: the decrement that empties the module and the next
    caller's increment race for real (the outgoing manager must not touch the module after that decrement)."""
    import time
    import torch
    L = H.lib()
    NB, PER_THREAD, THREADS = 64, 48 if pause_us == 0 else 160, 4
    total = PER_THREAD * THREADS
    tile_bytes = NB * NB * 8
    dev = H.Device(0, memory_block_size=64 * 1024, memory_number_of_blocks=1024)
    try:
        rng = np.random.default_rng(21)
        host = torch.zeros(total * NB * NB, dtype=torch.float64).pin_memory()
        hv = host.numpy()
        hv[:] = rng.uniform(-0.5, 0.5, hv.shape)
        before = hv.copy()
        done, managers = [], set()

        @H.SUBMIT_FN
        def submit(_dev, task_p, stream):
            managers.add(threading.get_ident())
            t = C.cast(task_p, C.POINTER(H.GpuTask)).contents
            a = t.flow_dev_ptr[0]
            return L.parsec_b200_dgemm(b"N", b"N", NB, NB, NB, 1.0, a, NB, a, NB, 2.0, a, NB, stream)

        @H.COMPLETE_FN
        def complete(task_p, status, cookie):
            t = C.cast(task_p, C.POINTER(H.GpuTask)).contents
            done.append((int(t.locals[0]), status))

        libc = C.CDLL(None)
        libc.calloc.restype = C.c_void_p
        datas = [C.c_void_p(L.pb2_data_create(i, 97, hv.ctypes.data + i * tile_bytes, tile_bytes)) for i in range(total)]
        rcs = []
        start = threading.Barrier(THREADS)

        def worker(w):
            start.wait()
            for j in range(PER_THREAD):
                i = w * PER_THREAD + j
                raw = libc.calloc(1, C.sizeof(H.GpuTask))
                t = C.cast(raw, C.POINTER(H.GpuTask)).contents
                t.task_type, t.nb_flows, t.pushout = 0, 1, 1
                t.submit = submit
                t.flow_data[0] = datas[i]
                t.flow_access[0] = H.FLOW_ACCESS_RW
                t.flow_nb_elts[0] = tile_bytes
                t.complete = complete
                t.locals[0] = i
                rcs.append(L.pb2_kernel_scheduler(dev.h, raw))
                if pause_us:
                    time.sleep(pause_us * 1e-6 * (1 + (i * 7919) % 5))

        threads = [threading.Thread(target=worker, args=(w,)) for w in range(THREADS)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        # every call returned ASYNC; whoever held the manager seat only left once the device had drained
        assert rcs == [H.HOOK_RETURN_ASYNC] * total
        assert sorted(done) == [(i, 0) for i in range(total)]
        assert len(managers) >= 1
        if pause_us:
            print("manager seat held by %d different threads" % len(managers))
        assert dev.stats()["executed_tasks"] == total
        for i in range(total):
            a = before[i * NB * NB:(i + 1) * NB * NB].reshape((NB, NB), order="F")
            got = hv[i * NB * NB:(i + 1) * NB * NB].reshape((NB, NB), order="F")
            assert np.max(np.abs(got - (a @ a + 2.0 * a))) <= 1e-13
        for d in datas:
            L.pb2_data_destroy(d)
    finally:
        dev.close()
