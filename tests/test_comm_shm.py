"""pb2_ce_*: the shared-memory transport under the comm-engine seam (include/parsec_b200.h, csrc/comm_shm.cu), exercised
by real processes on this box -- no GPU needed.  Each rank (tests/ce_worker.py) sends 1500 messages of 0 .. 3000 bytes to
every rank through 64 KiB rings (so the rings wrap and fill all the time) and checks per-pair FIFO order and content,
puts a strided block and a 64 MiB block into its neighbour's memory, and runs the collectives.  Both put paths: one copy
through process_vm_writev, and staged through the ring (PB2_CE_NO_CMA=1, what a box without cross-memory attach does)."""
import json
import os
import subprocess
import sys
import uuid

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def launch(world, scenario="default", env_extra=None, timeout=300):
    job = "t%s" % uuid.uuid4().hex[:12]
    env = dict(os.environ)
    env.update(env_extra or {})
    procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "ce_worker.py"), job, str(r), str(world), scenario],
                              env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
    outs = []
    try:
        for p in procs:
            o, e = p.communicate(timeout=timeout)
            assert p.returncode == 0, (p.returncode, o[-500:], e[-2000:])
            outs.append(json.loads(o.strip().splitlines()[-1]))
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
        leftover = os.path.exists("/dev/shm/pb2ce.%s" % job)
        if leftover:
            os.unlink("/dev/shm/pb2ce.%s" % job)
    assert not leftover                                       # the last rank out removed the segment
    return outs


@pytest.mark.parametrize("world", [1, 2, 4])
def test_messages_puts_and_collectives_between_processes(world):
    outs = launch(world)
    for o in outs:
        assert o["am"] == 1500 * world and o["received"] >= 1500 * world
        if world > 1:
            assert o["single_copy"] == 1 and o["staged_bytes"] == 0 and o["put_bytes"] >= (64 << 20)


def test_several_threads_of_a_rank_send_at_once():
    outs = launch(3, scenario="threads")
    assert all(o["received"] == 3 * 4 * 800 for o in outs)


def test_puts_staged_through_the_rings_when_cross_memory_attach_is_refused():
    outs = launch(3, env_extra={"PB2_CE_NO_CMA": "1"})
    for o in outs:
        assert o["single_copy"] == 0 and o["staged_bytes"] >= (64 << 20)


def test_a_segment_left_by_a_dead_job_is_refused():
    """Same job name twice without cleaning up in between (what a crashed launch leaves under /dev/shm): the second
    job must not mistake the first one's ranks for its own."""
    import ctypes as C
    L = C.CDLL(os.path.join(os.path.dirname(HERE), "parsec-1_b200", "libparsec_b200.so"))
    L.pb2_ce_attach.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]
    job = ("stale%s" % uuid.uuid4().hex[:10]).encode()
    ce = C.c_void_p()
    assert L.pb2_ce_attach(job, 0, 1, 4096, C.byref(ce)) == 0          # a world of one attaches at once ...
    try:
        again = C.c_void_p()                                            # ... and never detaches: the segment stays
        assert L.pb2_ce_attach(job, 0, 1, 4096, C.byref(again)) == -4
    finally:
        os.unlink("/dev/shm/pb2ce.%s" % job.decode())
