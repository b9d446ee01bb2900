"""The GPU input generator (pb2_matrix_fill_lcg, C ABI) against tiles written by the REFERENCE's own initialize_tile()
(tests/golden/lcg_tiles.json, produced from tests/dsl/dtd/dtd_test_simple_gemm.c compiled in place): bit-exact."""
import json
import os

import numpy as np
import pytest

from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()
HERE = os.path.dirname(os.path.abspath(__file__))


def test_gpu_generator_matches_reference_tiles():
    import torch
    dev = H.Device(0, memory_number_of_blocks=64)
    try:
        for c in json.load(open(os.path.join(HERE, "golden", "lcg_tiles.json"))):
            M, N = c["M"], (c["n"] + 1) * c["nb"]                 # just enough columns to contain tile (m, n)
            mat = H.Matrix(H.MATRIX_DOUBLE, M, N, c["mb"], c["nb"])
            buf = torch.zeros(mat.nbytes, dtype=torch.uint8).pin_memory()
            mat.set_buffer(buf)
            mat.fill_lcg(c["seed"], dev)
            pos = H.lib().pb2_matrix_position(mat.h, c["m"], c["n"])
            sz = c["mb"] * c["nb"]
            got = buf.numpy().view(np.float64)[pos * sz:(pos + 1) * sz]
            want = np.array([float.fromhex(v) for v in c["values"]])
            assert np.array_equal(got, want), (c["m"], c["n"], c["M"])
            mat.close()
    finally:
        dev.close()
