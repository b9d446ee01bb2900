"""The coherency / version state machine against the REFERENCE's own parsec/data.c.

tests/golden/ownership_trace.json was produced by driving the reference's parsec_data_copy_attach/_detach and
parsec_data_start/end_transfer_ownership_to_copy (compiled in place into oracle/_ref by `make -C oracle ref`) with
seeded random operation sequences; every step records the full data state.  The same sequences are replayed through
the oracle's restatement and through the product's state machine (C ABI hooks), and must give the same state and the
same "transfer from" answer after every step.  CPU only."""
import ctypes as C
import json
import os

import oracle_lib as O
from conftest import load_hostapi

HERE = os.path.dirname(os.path.abspath(__file__))
H = load_hostapi()
DEVS = (0, 2, 3)
READ, WRITE = 4, 8


class OrcCopy(C.Structure):
    _fields_ = [("present", C.c_uint8), ("coherency", C.c_uint8), ("transfer_status", C.c_uint8),
                ("readers", C.c_int32), ("version", C.c_uint32)]


class OrcData(C.Structure):
    _fields_ = [("key", C.c_uint64), ("owner_device", C.c_int8), ("preferred_device", C.c_int8),
                ("nb_devices", C.c_int), ("copies", OrcCopy * 10)]


def _cases():
    return json.load(open(os.path.join(HERE, "golden", "ownership_trace.json")))


def _apply_set(kind, cur_version, cur_readers, host_version):
    """-> (coherency, version, readers) with -1 = unchanged, as make_golden.py applied it to the reference"""
    if kind == "version++":
        return -1, cur_version + 1, -1
    if kind == "version=host":
        return -1, host_version, -1
    if kind == "release_reader":
        return -1, -1, max(0, cur_readers - 1)
    return 4, -1, -1


def test_oracle_state_machine_matches_reference_data_c():
    L = O.lib()
    L.orc_data_create.argtypes = [C.POINTER(OrcData), C.c_uint64, C.c_int]
    L.orc_data_start_transfer_ownership.argtypes = [C.POINTER(OrcData), C.c_int, C.c_int]
    L.orc_data_end_transfer_ownership.argtypes = [C.POINTER(OrcData), C.c_int, C.c_int]
    oracle_access = lambda a: (1 if a & READ else 0) | (2 if a & WRITE else 0)
    nsteps = 0
    for case in _cases():
        d = OrcData()
        L.orc_data_create(C.byref(d), case["seed"], 4)
        for step in case["steps"]:
            op = step["op"]
            if op[0] == "attach":
                c = d.copies[op[1]]
                c.present, c.coherency, c.version, c.readers, c.transfer_status = 1, 0, 0, 0, 0
            elif op[0] == "detach":
                d.copies[op[1]] = OrcCopy()
            elif op[0] == "start":
                assert L.orc_data_start_transfer_ownership(C.byref(d), op[1], oracle_access(op[2])) == op[3]
            elif op[0] == "end":
                L.orc_data_end_transfer_ownership(C.byref(d), op[1], oracle_access(op[2]))
            else:
                c = d.copies[op[1]]
                coh, ver, rd = _apply_set(op[2], c.version, c.readers, d.copies[0].version)
                if coh >= 0: c.coherency = coh
                if ver >= 0: c.version = ver
                if rd >= 0: c.readers = rd
            got = [d.owner_device] + [[d.copies[v].present, d.copies[v].coherency if d.copies[v].present else -1,
                                       d.copies[v].readers, d.copies[v].version] for v in DEVS]
            assert got == step["state"], (case["seed"], op)
            nsteps += 1
    assert nsteps > 3000


def test_product_state_machine_matches_reference_data_c():
    L = H.lib()
    L.pb2_data_test_attach.argtypes = [C.c_void_p, C.c_int]
    L.pb2_data_test_detach.argtypes = [C.c_void_p, C.c_int]
    L.pb2_data_test_start_transfer.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.pb2_data_test_end_transfer.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.pb2_data_test_end_transfer.restype = None
    L.pb2_data_test_set.argtypes = [C.c_void_p] + [C.c_int] * 4
    L.pb2_data_test_set.restype = None

    def state(d):
        out = [L.pb2_data_owner_device(d)]
        v = C.c_uint32()
        for dev in DEVS:
            coh = L.pb2_data_copy_coherency(d, dev)
            if coh == -3:                                    # PARSEC_ERR_NOT_FOUND: no copy on that device
                out.append([0, -1, 0, 0])
            else:
                L.pb2_data_copy_version(d, dev, C.byref(v))
                out.append([1, coh, L.pb2_data_copy_readers(d, dev), v.value])
        return out

    for case in _cases():
        d = C.c_void_p(L.pb2_data_create(case["seed"], 0, None, 1024))
        for step in case["steps"]:
            op = step["op"]
            if op[0] == "attach":
                assert L.pb2_data_test_attach(d, op[1]) == 0
            elif op[0] == "detach":
                assert L.pb2_data_test_detach(d, op[1]) == 0
            elif op[0] == "start":
                assert L.pb2_data_test_start_transfer(d, op[1], op[2]) == op[3]     # same flag values as the reference
            elif op[0] == "end":
                L.pb2_data_test_end_transfer(d, op[1], op[2])
            else:
                cur = state(d)
                me = cur[1 + DEVS.index(op[1])]
                coh, ver, rd = _apply_set(op[2], me[3], me[2], cur[1][3])
                L.pb2_data_test_set(d, op[1], coh, ver, rd)
            assert state(d) == step["state"], (case["seed"], op)
        for dev in (2, 3):
            L.pb2_data_test_detach(d, dev)
        L.pb2_data_destroy(d)
