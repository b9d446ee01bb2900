"""BASELINE config 5 shape at test size: bf16 A, B and fp32 C tiles through the device pipeline (tcgen05 body)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()


def _bf16_bits(x):
    u = x.astype(np.float32).view(np.uint32).astype(np.uint64)
    u = u + 0x7FFF + ((u >> 16) & 1)
    return (u >> 16).astype(np.uint16)


@pytest.mark.parametrize("N,NB", [(1024, 256), (768, 128)])
def test_tiled_bf16_gemm_matches_oracle(N, NB):
    import torch
    dev = H.Device(0, memory_number_of_blocks=2048)
    try:
        mats, bufs = [], []
        for mtype, seed in ((H.MATRIX_BF16, 1789), (H.MATRIX_BF16, 1805), (H.MATRIX_FLOAT, 1901)):
            m = H.Matrix(mtype, N, N, NB, NB)
            b = torch.zeros(m.nbytes, dtype=torch.uint8).pin_memory()
            m.set_buffer(b)
            m.fill_lcg(seed, dev)
            mats.append(m)
            bufs.append(b)
        d = O.square(N, NB)
        # the generator's values are single precision; bf16 operands are their round-to-nearest-even
        hA = _bf16_bits(O.fill(d, 1789))
        hB = _bf16_bits(O.fill(d, 1805))
        hC = O.fill(d, 1901).astype(np.float32)
        assert np.array_equal(bufs[0].numpy().view(np.uint16), hA)       # GPU generator == oracle generator, in bf16
        assert np.array_equal(bufs[2].numpy().view(np.float32), hC)
        st = H.tiled_gemm(mats[0], mats[1], mats[2], dev)
        MT = N // NB
        assert st["nb_gemm"] == MT ** 3
        got = bufs[2].numpy().view(np.float32)
        worst = 0.0
        back = lambda X: (X.astype(np.uint32) << 16).view(np.float32).astype(np.float64)
        scale = np.abs(back(hA)).max() * np.abs(back(hB)).max() * NB * N * 2.0 ** -24   # |a||b| * row length * K * eps32
        for n in range(MT):
            for m in range(MT):
                pc = O.lib().orc_bc_position(C.byref(d), m, n)
                ref = hC[pc * NB * NB:(pc + 1) * NB * NB].copy()
                for k in range(MT):
                    pa, pb = O.lib().orc_bc_position(C.byref(d), m, k), O.lib().orc_bc_position(C.byref(d), k, n)
                    a = np.ascontiguousarray(hA[pa * NB * NB:(pa + 1) * NB * NB])
                    b = np.ascontiguousarray(hB[pb * NB * NB:(pb + 1) * NB * NB])
                    O.lib().orc_gemm_tile_bf16(NB, NB, NB, 1.0, a.ctypes.data, NB, b.ctypes.data, NB, 1.0, ref.ctypes.data, NB)
                tile = got[pc * NB * NB:(pc + 1) * NB * NB]
                worst = max(worst, np.abs(tile.astype(np.float64) - ref.astype(np.float64)).max() / scale)
                assert mats[2].version(m, n) == MT
        assert worst < 10, worst      # same form as the fp64 acceptance norm, with eps = 2^-24
        for m in mats:
            m.close()
    finally:
        dev.close()


def test_mismatched_types_are_rejected():
    a = H.Matrix(H.MATRIX_BF16, 256, 256, 128, 128)
    c = H.Matrix(H.MATRIX_DOUBLE, 256, 256, 128, 128)
    assert not H.lib().pb2_gemm_taskpool_new(a.h, a.h, c.h, None)
    a.close()
    c.close()


def test_tiled_fp32_gemm_runs_on_tf32_tensor_cores():
    """fp32 A, B, C tiles (PARSEC_MATRIX_FLOAT) go through the kind::tf32 body: the operands lose their low 13 mantissa
    bits in the tensor core, so the bound is 2 * 2^-10 per product, i.e. ||C - Cref||inf <= 2^-9 ||A||inf ||B||inf."""
    import torch
    N, NB = 512, 128
    dev = H.Device(0, memory_number_of_blocks=1024)
    try:
        mats, bufs = [], []
        for seed in (1789, 1805, 1901):
            m = H.Matrix(H.MATRIX_FLOAT, N, N, NB, NB)
            b = torch.zeros(m.nbytes, dtype=torch.uint8).pin_memory()
            m.set_buffer(b)
            m.fill_lcg(seed, dev)
            mats.append(m)
            bufs.append(b)
        d = O.square(N, NB)
        hA, hB, hC = (O.tiled_to_dense(d, O.fill(d, s)) for s in (1789, 1805, 1901))
        st = H.tiled_gemm(mats[0], mats[1], mats[2], dev)
        assert st["nb_gemm"] == (N // NB) ** 3
        got = O.tiled_to_dense(d, bufs[2].numpy().view(np.float32).astype(np.float64))
        want = hC + hA @ hB
        inf = lambda X: np.abs(X).sum(axis=1).max()
        r = inf(got - want) / (inf(hA) * inf(hB))
        assert r < 2.0 ** -9, r
        assert r > 2.0 ** -24      # it really was TF32, not an fp32 FMA path
        for m in mats:
            m.close()
    finally:
        dev.close()
