"""One rank of the transport tests (tests/test_comm_shm.py): python ce_worker.py <job> <rank> <world> <scenario>.

Plain ctypes over libparsec_b200.so (pb2_ce_*, include/parsec_b200.h); prints one JSON line and exits 0 on success.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L = C.CDLL(os.environ.get("PB2_LIB", os.path.join(ROOT, "parsec-1_b200", "libparsec_b200.so")))


class Iov(C.Structure):
    _fields_ = [("base", C.c_void_p), ("len", C.c_size_t)]


class Stats(C.Structure):
    _fields_ = [("sent_msgs", C.c_uint64), ("recv_msgs", C.c_uint64), ("put_bytes", C.c_uint64),
                ("staged_bytes", C.c_uint64), ("single_copy", C.c_int32)]


RECV = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_size_t)
L.pb2_ce_attach.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]
L.pb2_ce_detach.argtypes = [C.c_void_p]
L.pb2_ce_send.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_size_t]
L.pb2_ce_poll.argtypes = [C.c_void_p, RECV, C.c_void_p, C.c_int]
L.pb2_ce_set_wait_poll.argtypes = [C.c_void_p, RECV, C.c_void_p]
L.pb2_ce_put.argtypes = [C.c_void_p, C.c_int, C.POINTER(Iov), C.c_int, C.POINTER(Iov), C.c_int]
L.pb2_ce_barrier.argtypes = [C.c_void_p]
L.pb2_ce_allreduce_i64.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.c_int, C.c_int]
L.pb2_ce_allgather.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
L.pb2_ce_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
L.pb2_ce_max_message.argtypes = [C.c_void_p]
L.pb2_ce_max_message.restype = C.c_size_t

TAG_DATA, TAG_ADDR, TAG_DONE = 7, 8, 9


def message(src, dst, i):
    """Deterministic payload of message i from src to dst: length cycles through 0 .. 3000 bytes."""
    n = (i * 37 + src * 11 + dst * 5) % 3001
    rng = np.random.default_rng(1000003 * src + 1009 * dst + i)
    return rng.integers(0, 256, n, dtype=np.uint8).tobytes()


def main():
    job, rank, world, scenario = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
    ce = C.c_void_p()
    ring = 1 << 16 if scenario != "bigring" else 0          # a small ring: wraps and fills all the time
    assert L.pb2_ce_attach(job.encode(), rank, world, ring, C.byref(ce)) == 0
    out = {"rank": rank}
    got = {s: [] for s in range(world)}
    addrs, done = {}, set()

    def on_msg(user, src, tag, msg, n):
        data = C.string_at(msg, n)
        if tag == TAG_DATA:
            got[src].append(data)
        elif tag == TAG_ADDR:
            addrs[src] = np.frombuffer(data, dtype=np.uint64).copy()
        elif tag == TAG_DONE:
            done.add(src)

    cb = RECV(on_msg)
    L.pb2_ce_set_wait_poll(ce, cb, None)

    if scenario == "threads":
        # several threads of a rank send at once (what worker threads do when the runtime runs with
        # MPI_THREAD_MULTIPLE); one thread polls.  Per (source, thread) the order must hold, nothing may be lost or torn.
        import threading
        NT, PER = 4, 800
        seen = {}

        def on_t(user, src, tag, msg, n):
            data = C.string_at(msg, n)
            t, i = int.from_bytes(data[:4], "little"), int.from_bytes(data[4:8], "little")
            assert data[8:] == bytes([(t * 31 + i) % 251]) * (n - 8), "torn message"
            assert seen.get((src, t), -1) == i - 1, "order from rank %d thread %d" % (src, t)
            seen[(src, t)] = i

        cbt = RECV(on_t)
        stop = threading.Event()

        def sender(t):
            for i in range(PER):
                n = 8 + (i * 13 + t * 7) % 900
                data = t.to_bytes(4, "little") + i.to_bytes(4, "little") + bytes([(t * 31 + i) % 251]) * (n - 8)
                buf = C.create_string_buffer(data, n)
                for dst in range(world):
                    while L.pb2_ce_send(ce, dst, 77, buf, n) != 0:
                        pass                                   # full: the polling thread drains

        threads = [threading.Thread(target=sender, args=(t,)) for t in range(NT)]
        for th in threads:
            th.start()
        while len(seen) < world * NT or any(v < PER - 1 for v in seen.values()):
            L.pb2_ce_poll(ce, cbt, None, 0)
        for th in threads:
            th.join()
        assert L.pb2_ce_barrier(ce) == 0
        L.pb2_ce_set_wait_poll(ce, RECV(0), None)
        assert L.pb2_ce_detach(ce) == 0
        print(json.dumps({"rank": rank, "received": world * NT * PER}))
        return

    def send(dst, tag, data):
        buf = C.create_string_buffer(data, len(data)) if data else None
        while True:
            rc = L.pb2_ce_send(ce, dst, tag, buf, len(data))
            if rc == 0:
                return
            assert rc == -1, rc                               # full: let the others drain, and drain ourselves
            L.pb2_ce_poll(ce, cb, None, 0)

    # ---- active messages: every rank sends NMSG messages to every rank (itself included), interleaved
    NMSG = 1500
    for i in range(NMSG):
        for dst in range(world):
            send(dst, TAG_DATA, message(rank, dst, i))
        if i % 7 == 0:
            L.pb2_ce_poll(ce, cb, None, 3)
    while any(len(got[s]) < NMSG for s in range(world)):
        L.pb2_ce_poll(ce, cb, None, 0)
    for s in range(world):
        assert got[s] == [message(s, rank, i) for i in range(NMSG)], "order or content from %d" % s
    assert L.pb2_ce_send(ce, 0, TAG_DATA, C.create_string_buffer(1 << 20), 1 << 20) == -8   # larger than a quarter ring
    out["am"] = NMSG * world

    # ---- put: rank r writes a strided block into a window of rank (r+1) % world, then says so
    rows, cols, ld = 257, 96, 160
    window = np.zeros((rows, ld), dtype=np.float64)
    info = np.array([window.ctypes.data], dtype=np.uint64)
    for dst in range(world):
        send(dst, TAG_ADDR, info.tobytes())
    while len(addrs) < world:
        L.pb2_ce_poll(ce, cb, None, 0)
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    src_block = np.arange(rows * cols, dtype=np.float64).reshape(rows, cols) + 1000.0 * rank
    liov = (Iov * 1)(Iov(src_block.ctypes.data, src_block.nbytes))          # contiguous here ...
    riov = (Iov * rows)(*[Iov(int(addrs[nxt][0]) + r * ld * 8, cols * 8) for r in range(rows)])   # ... strided there
    assert L.pb2_ce_put(ce, nxt, liov, 1, riov, rows) == 0
    send(nxt, TAG_DONE, b"")
    while prv not in done:
        L.pb2_ce_poll(ce, cb, None, 0)
    want = np.arange(rows * cols, dtype=np.float64).reshape(rows, cols) + 1000.0 * prv
    assert np.array_equal(window[:, :cols], want) and not window[:, cols:].any()
    assert L.pb2_ce_put(ce, nxt, liov, 1, riov, rows - 1) == -4             # lengths differ: refused
    # a large contiguous put: 64 MiB in one call
    big = np.zeros(8 << 20, dtype=np.float64)
    done.clear()
    addrs.clear()
    assert L.pb2_ce_barrier(ce) == 0
    for dst in range(world):
        send(dst, TAG_ADDR, np.array([big.ctypes.data], dtype=np.uint64).tobytes())
    while len(addrs) < world:
        L.pb2_ce_poll(ce, cb, None, 0)
    payload = np.full(8 << 20, float(rank + 1))
    assert L.pb2_ce_put(ce, nxt, (Iov * 1)(Iov(payload.ctypes.data, payload.nbytes)), 1,
                        (Iov * 1)(Iov(int(addrs[nxt][0]), payload.nbytes)), 1) == 0
    send(nxt, TAG_DONE, b"")
    while prv not in done:
        L.pb2_ce_poll(ce, cb, None, 0)
    assert big[0] == prv + 1 and big[-1] == prv + 1 and (big == prv + 1).all()

    # ---- collectives
    v = (C.c_int64 * 3)(rank, rank, rank)
    assert L.pb2_ce_allreduce_i64(ce, v, 3, 0) == 0 and list(v) == [world - 1] * 3
    v = (C.c_int64 * 2)(rank + 1, 1)
    assert L.pb2_ce_allreduce_i64(ce, v, 2, 1) == 0 and list(v) == [world * (world + 1) // 2, world]
    mine = np.array([rank, rank * rank], dtype=np.int32)
    everybody = np.zeros(2 * world, dtype=np.int32)
    assert L.pb2_ce_allgather(ce, mine.ctypes.data, mine.nbytes, everybody.ctypes.data) == 0
    assert everybody.tolist() == [x for r in range(world) for x in (r, r * r)]
    for _ in range(200):
        assert L.pb2_ce_barrier(ce) == 0

    st = Stats()
    L.pb2_ce_stats(ce, C.byref(st))
    out.update(put_bytes=st.put_bytes, staged_bytes=st.staged_bytes, single_copy=st.single_copy,
               sent=st.sent_msgs, received=st.recv_msgs)
    assert L.pb2_ce_barrier(ce) == 0
    L.pb2_ce_set_wait_poll(ce, RECV(0), None)
    assert L.pb2_ce_detach(ce) == 0
    print(json.dumps(out))


if __name__ == "__main__":
    main()
