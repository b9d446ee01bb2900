"""tcgen05 / TMEM tile GEMM (bf16 in, fp32 accumulate) through the C ABI against the oracle's bf16 tile GEMM."""
import os

import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()
EPS32 = 2.0 ** -24


def _bf16_bits(x):
    """float32 array -> bf16 bit patterns (round to nearest even), same rule as orc_f32_to_bf16_rn."""
    u = x.astype(np.float32).view(np.uint32).astype(np.uint64)
    u = u + 0x7FFF + ((u >> 16) & 1)
    return (u >> 16).astype(np.uint16)


def _run(m, n, k, alpha, beta, seed=0, ld_pad=0, ldc_extra=0):
    import torch
    rng = np.random.default_rng(seed)
    A = np.asfortranarray(_bf16_bits(rng.uniform(-0.5, 0.5, (m, k))))
    B = np.asfortranarray(_bf16_bits(rng.uniform(-0.5, 0.5, (k, n))))
    Cm = np.asfortranarray(rng.uniform(-0.5, 0.5, (m, n)).astype(np.float32))
    lda, ldb, ldc = m + ld_pad, (k if k else 8) + ld_pad, m + ld_pad + ldc_extra

    def to_dev(X, ld, dtype):
        buf = np.zeros((X.shape[1], ld), dtype=dtype)
        buf[:, :X.shape[0]] = X.T
        return torch.from_numpy(buf.view(np.int16) if dtype == np.uint16 else buf).cuda()

    # C sits between guard columns (4 on each side keep its base 16-byte aligned whenever ldc % 4 == 0, so the TMA
    # epilogue is exercised) and its leading-dimension padding holds a canary: only the m x n elements may change
    CANARY, G = np.float32(12345.0), 4
    hostC = np.full((n + 2 * G, ldc), CANARY, dtype=np.float32)
    hostC[G:n + G, :m] = Cm.T
    dA, dB, dCg = to_dev(A, lda, np.uint16), to_dev(B, ldb, np.uint16), torch.from_numpy(hostC).cuda()
    rc = H.lib().parsec_b200_gemm_bf16_sm100(b"N", b"N", m, n, k, alpha, dA.data_ptr(), lda, dB.data_ptr(), ldb, beta,
                                              dCg.data_ptr() + 4 * G * ldc, ldc, torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    back = dCg.cpu().numpy()
    assert np.all(back[:G] == CANARY) and np.all(back[n + G:] == CANARY) and np.all(back[G:n + G, m:] == CANARY)
    got = back[G:n + G, :m].T
    want = Cm.copy(order="F")
    O.lib().orc_gemm_tile_bf16(m, n, k, alpha, A.ctypes.data, m, B.ctypes.data, k, beta, want.ctypes.data, m)
    back = lambda X: (X.astype(np.uint32) << 16).view(np.float32).astype(np.float64)
    inf = lambda X: np.abs(X).sum(axis=1).max() if X.size else 0.0
    denom = max(abs(alpha) * inf(back(A)) * inf(back(B)) * max(k, 1) * EPS32, 1e-30)
    return inf(got.astype(np.float64) - want.astype(np.float64)) / denom


@pytest.mark.parametrize("m,n,k", [(128, 256, 64), (128, 256, 256), (256, 512, 128), (1024, 1024, 1024), (2048, 2048, 2048)])
def test_aligned_tiles(m, n, k):
    # tolerance: ||C - Cref||inf / (||A||inf ||B||inf K eps32) < 10 with eps32 = 2^-24 (fp32 accumulation)
    assert _run(m, n, k, 1.0, 1.0, seed=m + k) < 10


@pytest.mark.parametrize("m,n,k", [(64, 64, 64), (200, 136, 72), (136, 300, 16), (8, 8, 8), (1000, 520, 264)])
def test_ragged_tiles(m, n, k):
    assert _run(m, n, k, 1.0, 1.0, seed=n) < 10


@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.0, 1.0), (-1.5, 0.5)])
def test_alpha_beta(alpha, beta):
    assert _run(256, 256, 256, alpha, beta, seed=9) < 10


def test_padded_leading_dimension():
    assert _run(256, 256, 128, 1.0, 1.0, seed=4, ld_pad=8) < 10


@pytest.mark.parametrize("m,n,k,beta", [(256, 256, 128, 1.0), (200, 136, 72, 0.5), (136, 300, 16, 0.0)])
def test_c_pitch_not_a_multiple_of_16_bytes_takes_the_register_epilogue(m, n, k, beta):
    # C goes through TMA boxes only when its base is 16-byte aligned and ldc % 4 == 0; any other C is still accepted
    assert _run(m, n, k, 1.0, beta, seed=m, ldc_extra=1) < 10


@pytest.mark.parametrize("beta", [0.0, 1.0, -0.25])
def test_k_zero_scales_c(beta):
    assert _run(200, 300, 0, 1.0, beta, seed=5) < 10


@pytest.mark.parametrize("m,n,k,beta", [(2048, 2048, 512, 1.0), (1000, 520, 264, 0.0), (40, 1000, 64, 2.0)])
def test_more_column_chunks_than_staging_buffers(m, n, k, beta):
    # 256 columns per CTA = 8 boxes of 32 against 6 buffers per warp: the last two recycle a buffer
    assert _run(m, n, k, 0.5, beta, seed=k) < 10


def test_unsupported_arguments():
    import torch
    x = torch.zeros(4096, dtype=torch.float32, device="cuda")
    p = x.data_ptr()
    f = H.lib().parsec_b200_gemm_bf16_sm100
    assert f(b"T", b"N", 8, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8, 0) == -7     # transposes: not supported
    assert f(b"N", b"N", 8, 8, 8, 1.0, p, 4, p, 8, 1.0, p, 8, 0) == -4     # lda < m
    assert f(b"N", b"N", 12, 8, 8, 1.0, p, 12, p, 8, 1.0, p, 12, 0) == -7  # lda not a multiple of 8 (TMA pitch)
    assert f(b"N", b"N", 0, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8, 0) == 0


@pytest.mark.parametrize("variant", ["grouped", "pair"])
def test_large_gemm_both_kernels_agree_with_cublas(variant, monkeypatch):
    """4096^3 and a ragged 3000 x 2100 x 1100: many tiles per CTA pair (the persistent walk, both accumulator stages,
    the raster) against torch.matmul in fp32; PB2_GEMM_BF16_CTAS is read once per process, so the per-tile kernel runs
    in a child."""
    import subprocess
    import sys
    code = r'''
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.join(%r, "tests"))
from conftest import load_hostapi
H = load_hostapi()
for (m, n, k, beta) in [(4096, 4096, 4096, 1.0), (3000, 2100, 1104, 0.0), (512, 8192, 256, 1.0)]:
    g = torch.Generator(device="cuda"); g.manual_seed(m)
    A = (torch.rand(k, m, device="cuda", generator=g) - 0.5).to(torch.bfloat16)    # column-major m x k
    B = (torch.rand(n, k, device="cuda", generator=g) - 0.5).to(torch.bfloat16)    # column-major k x n
    C = (torch.rand(n, m, device="cuda", generator=g) - 0.5)
    want = beta * C + (B.float() @ A.float())
    rc = H.lib().parsec_b200_gemm_bf16_sm100(b"N", b"N", m, n, k, 1.0, A.data_ptr(), m, B.data_ptr(), k, beta, C.data_ptr(), m,
                                             torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    err = (C - want).abs().max().item()
    assert err < 2e-3 * max(1.0, k / 1024.0), (m, n, k, err)
print("ok")
''' % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, PB2_GEMM_BF16_CTAS=variant)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def test_batched_calls_give_the_results_of_single_calls():
    """parsec_b200_batch_begin / _end: 20 GEMMs of mixed shapes, alpha and beta recorded on one stream and launched as
    grouped kernels (16 + 4), bit-identical to the same calls made one by one."""
    import torch
    L = H.lib()
    L.parsec_b200_batch_begin.argtypes = [H.C.c_void_p]
    L.parsec_b200_batch_end.argtypes = [H.C.c_void_p]
    shapes = [(2048, 2048, 2048), (256, 256, 64), (1000, 520, 264), (136, 300, 16), (512, 768, 1024)] * 4
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    ops = []
    for i, (m, n, k) in enumerate(shapes):
        A = (torch.rand(k, m, device="cuda", generator=g) - 0.5).to(torch.bfloat16)
        B = (torch.rand(n, k, device="cuda", generator=g) - 0.5).to(torch.bfloat16)
        C = torch.rand(n, m, device="cuda", generator=g) - 0.5
        ops.append((m, n, k, 1.0 + 0.25 * (i % 3), float(i % 2), A, B, C))
    s = torch.cuda.current_stream().cuda_stream

    def run(batched):
        outs = [op[7].clone() for op in ops]
        if batched:
            assert L.parsec_b200_batch_begin(s) == 0
            assert L.parsec_b200_batch_begin(s) != 0            # one batch per thread at a time
        for (m, n, k, alpha, beta, A, B, _), C in zip(ops, outs):
            assert L.parsec_b200_gemm_bf16_sm100(b"N", b"N", m, n, k, alpha, A.data_ptr(), m, B.data_ptr(), k, beta,
                                                 C.data_ptr(), m, s) == 0
        if batched:
            assert L.parsec_b200_batch_end(s) == 0
            assert L.parsec_b200_batch_end(s) != 0              # nothing open any more
        torch.cuda.synchronize()
        return outs

    single, batched = run(False), run(True)
    for a, b, op in zip(single, batched, ops):
        assert torch.equal(a, b), op[:3]
        assert not torch.equal(a, op[7])                         # and something was computed
