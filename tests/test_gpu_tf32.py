"""tcgen05 tile GEMM with fp32 operands consumed as TF32 (kind::tf32), through the C ABI."""
import numpy as np
import pytest

from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()


def _tf32(x):
    """keep the 10 mantissa bits a TF32 operand has (drop the low 13): representable inputs make the check exact"""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32) & np.uint32(0xFFFFE000)
    return u.view(np.float32)


def _run(m, n, k, alpha, beta, representable=True, seed=0):
    import torch
    rng = np.random.default_rng(seed)
    A = rng.uniform(-0.5, 0.5, (m, k)).astype(np.float32)
    B = rng.uniform(-0.5, 0.5, (k, n)).astype(np.float32)
    if representable:
        A, B = _tf32(A), _tf32(B)
    Cm = rng.uniform(-0.5, 0.5, (m, n)).astype(np.float32)
    dA = torch.from_numpy(np.ascontiguousarray(A.T)).cuda()      # column-major m x k == row-major k x m
    dB = torch.from_numpy(np.ascontiguousarray(B.T)).cuda()
    dC = torch.from_numpy(np.ascontiguousarray(Cm.T)).cuda()
    rc = H.lib().parsec_b200_gemm_tf32_sm100(b"N", b"N", m, n, k, alpha, dA.data_ptr(), m, dB.data_ptr(), k, beta,
                                              dC.data_ptr(), m, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    got = dC.cpu().numpy().T.astype(np.float64)
    want = alpha * (A.astype(np.float64) @ B.astype(np.float64)) + beta * Cm.astype(np.float64)
    inf = lambda X: np.abs(X).sum(axis=1).max()
    return inf(got - want) / max(abs(alpha) * inf(A.astype(np.float64)) * inf(B.astype(np.float64)) * max(k, 1), 1e-30)


@pytest.mark.parametrize("m,n,k", [(128, 256, 32), (256, 512, 256), (1024, 1024, 1024), (200, 136, 72), (64, 64, 8)])
def test_tf32_representable_inputs(m, n, k):
    # products of TF32-representable inputs are exact in fp32; only the fp32 accumulation rounds: eps = 2^-24
    assert _run(m, n, k, 1.0, 1.0, seed=m + k) < 10 * 2.0 ** -24


def test_tf32_arbitrary_inputs_lose_13_mantissa_bits():
    # full fp32 inputs are cut to 10 mantissa bits by the tensor core: relative error of a product <= 2 * 2^-10
    r = _run(512, 512, 512, 1.0, 0.0, representable=False, seed=3)
    assert 1e-9 < r < 4 * 2.0 ** -10


def test_tf32_alpha_beta_and_arguments():
    assert _run(256, 256, 128, -1.5, 0.5, seed=5) < 10 * 2.0 ** -24
    import torch
    x = torch.zeros(4096, dtype=torch.float32, device="cuda")
    p = x.data_ptr()
    f = H.lib().parsec_b200_gemm_tf32_sm100
    assert f(b"T", b"N", 8, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8, 0) == -7
    assert f(b"N", b"N", 6, 8, 8, 1.0, p, 6, p, 8, 1.0, p, 6, 0) == -7   # lda not a multiple of 4 (16-byte TMA pitch)
    assert f(b"N", b"N", 8, 8, 8, 1.0, p, 4, p, 8, 1.0, p, 8, 0) == -4
