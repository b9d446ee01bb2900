"""Tile kernel parity: parsec_b200_dgemm_sm100 (through the C ABI) against the oracle's tile GEMM."""
import numpy as np
import pytest

import oracle_lib as O
from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()
EPS = 2.0 ** -53


def _run(transa, transb, m, n, k, alpha, beta, seed=0, lda_pad=0):
    import torch
    rng = np.random.default_rng(seed)
    opA = rng.uniform(-0.5, 0.5, (m, k))
    opB = rng.uniform(-0.5, 0.5, (k, n))
    Cm = np.asfortranarray(rng.uniform(-0.5, 0.5, (m, n)))
    A = np.asfortranarray(opA if transa == "N" else opA.T)
    B = np.asfortranarray(opB if transb == "N" else opB.T)
    lda, ldb, ldc = A.shape[0] + lda_pad, B.shape[0] + lda_pad, m + lda_pad

    def to_dev(X, ld):  # column-major with leading dimension ld
        buf = np.zeros((X.shape[1], ld))
        buf[:, :X.shape[0]] = X.T
        return torch.from_numpy(buf).cuda()

    # C sits between two guard columns and its leading-dimension padding holds a canary: the kernel may write the
    # m x n elements of C and nothing else (compute-sanitizer is not available on the GPU pool)
    CANARY = 12345.0
    hostC = np.full((n + 2, ldc), CANARY)
    hostC[1:n + 1, :m] = Cm.T
    dA, dB, dCg = to_dev(A, lda), to_dev(B, ldb), torch.from_numpy(hostC).cuda()
    rc = H.dgemm(transa, transb, m, n, k, alpha, dA.data_ptr(), lda, dB.data_ptr(), ldb, beta, dCg.data_ptr() + 8 * ldc, ldc,
                 torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    back = dCg.cpu().numpy()
    assert np.all(back[0] == CANARY) and np.all(back[n + 1] == CANARY) and np.all(back[1:n + 1, m:] == CANARY)
    got = back[1:n + 1, :m].T
    want = np.asfortranarray(Cm.copy())
    O.dgemm_tile(alpha, np.asfortranarray(opA), np.asfortranarray(opB), beta, want)
    # north-star norm with N := k (the contraction length of this tile product)
    inf = lambda X: np.abs(X).sum(axis=1).max()
    denom = max(inf(opA) * inf(opB) * max(k, 1) * EPS, 1e-300)
    return inf(got - want) / denom, got, want


@pytest.mark.parametrize("n", [128, 256, 512, 1024])
def test_square_tiles_nn(n):
    r, _, _ = _run("N", "N", n, n, n, 1.0, 1.0, seed=n)
    assert r < 10


@pytest.mark.parametrize("m,n,k", [(1, 1, 1), (40, 40, 40), (240, 200, 40), (129, 127, 17), (128, 256, 48),
                                   (1000, 24, 8), (130, 130, 2048), (256, 128, 16)])
def test_ragged_shapes(m, n, k):
    r, _, _ = _run("N", "N", m, n, k, 1.0, 1.0, seed=m + n + k)
    assert r < 10


@pytest.mark.parametrize("alpha,beta", [(0.0, 1.0), (1.0, 0.0), (-2.5, 0.5), (1.0, 2.0)])
def test_alpha_beta(alpha, beta):
    """alpha = 0, beta = 1 is what nvlink.jdf / stress.jdf use (C must come back unchanged)."""
    r, got, want = _run("N", "N", 256, 256, 256, alpha, beta, seed=3)
    assert r < 10
    if alpha == 0.0 and beta == 1.0:
        assert np.array_equal(got, want)


@pytest.mark.parametrize("ta,tb", [("T", "N"), ("N", "T"), ("T", "T")])
def test_transposes(ta, tb):
    r, _, _ = _run(ta, tb, 96, 160, 72, 1.0, 1.0, seed=5)
    assert r < 10


@pytest.mark.parametrize("ta,tb", [("T", "N"), ("N", "T"), ("T", "T")])
@pytest.mark.parametrize("m,n,k", [(1024, 1024, 1024), (1000, 936, 520), (129, 127, 18), (130, 70, 34), (64, 200, 2048),
                                   (2, 2, 2), (128, 128, 16)])
def test_transposed_operands_on_the_tma_path(ta, tb, m, n, k):
    """Even leading dimensions: N,T / T,N / T,T run the same TMA + DMMA tile code as N,N (the operand only arrives in
    the other shared-memory layout).  Every element of C is accumulated over k in the same order whatever the
    layouts, so the result must be BIT-identical to the N,N call on explicitly transposed copies."""
    r, got, want = _run(ta, tb, m, n, k, -1.0, 1.0, seed=m + 3 * n + 7 * k)      # C -= op(A) op(B): the update of a Cholesky / LU step
    assert r < 10
    r_nn, got_nn, _ = _run("N", "N", m, n, k, -1.0, 1.0, seed=m + 3 * n + 7 * k)  # same seed: same op(A), op(B), C
    assert r_nn < 10 and np.array_equal(got, got_nn)
    assert _run(ta, tb, m, n, k, 0.75, 0.0, seed=1, lda_pad=4)[0] < 10


def test_transposed_operands_run_at_the_speed_of_nn():
    import torch
    n = 4096
    g = torch.Generator(device="cuda").manual_seed(2)
    A = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    B = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    Cm = torch.zeros(n, n, dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    rate = {}
    for ta, tb in (("N", "N"), ("N", "T"), ("T", "N"), ("T", "T")):
        for _ in range(2):
            assert H.dgemm(ta, tb, n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0, Cm.data_ptr(), n, s) == 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            H.dgemm(ta, tb, n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0, Cm.data_ptr(), n, s)
        e1.record()
        torch.cuda.synchronize()
        rate[ta + tb] = 5 * 2.0 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) * 1e-12
    print("fp64 4096^3 TFLOP/s:", {k: round(v, 2) for k, v in rate.items()})
    for k, v in rate.items():
        assert v > 0.9 * rate["NN"], rate


def test_padded_leading_dimensions_and_odd_ld():
    assert _run("N", "N", 256, 128, 64, 1.0, 1.0, seed=7, lda_pad=6)[0] < 10   # even ld: TMA path
    assert _run("N", "N", 256, 128, 64, 1.0, 1.0, seed=8, lda_pad=3)[0] < 10   # odd ld: generic path


def test_bad_arguments_are_rejected():
    import torch
    x = torch.zeros(64, dtype=torch.float64, device="cuda")
    p = x.data_ptr()
    assert H.dgemm("X", "N", 8, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8) == -4
    assert H.dgemm("N", "N", -1, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8) == -4
    assert H.dgemm("N", "N", 8, 8, 8, 1.0, p, 4, p, 8, 1.0, p, 8) == -4   # lda < m
    assert H.dgemm("N", "N", 0, 8, 8, 1.0, p, 8, p, 8, 1.0, p, 8) == 0    # empty: nothing to do


def test_linearity_at_full_tile_size():
    """Size-independent property at the BASELINE tile size: (A1 + A2) B == A1 B + A2 B up to rounding."""
    import torch
    n = 1024
    g = torch.Generator(device="cuda").manual_seed(1)
    A1 = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    A2 = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    B = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
    s = torch.cuda.current_stream().cuda_stream
    outs = []
    for A in (A1, A2, A1 + A2):
        Cm = torch.zeros(n, n, dtype=torch.float64, device="cuda")
        assert H.dgemm("N", "N", n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0, Cm.data_ptr(), n, s) == 0
        outs.append(Cm)
    torch.cuda.synchronize()
    err = (outs[0] + outs[1] - outs[2]).abs().max().item()
    assert err < 1e-11


def test_batched_dgemm_calls_give_the_results_of_single_calls():
    """parsec_b200_batch_begin / _end around 70 fp64 tile GEMMs of mixed shapes (one grouped grid of 64 + one of 6):
    bit-identical to the same calls made one by one; a transposed call inside the batch is launched on its own."""
    import torch
    L = H.lib()
    L.parsec_b200_batch_begin.argtypes = [H.C.c_void_p]
    L.parsec_b200_batch_end.argtypes = [H.C.c_void_p]
    shapes = [(256, 256, 256), (128, 384, 64), (200, 136, 72), (512, 256, 1024), (1000, 24, 16), (256, 256, 256), (130, 130, 130)] * 10
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    ops = []
    for i, (m, n, k) in enumerate(shapes):
        A = torch.rand(k, m, device="cuda", generator=g, dtype=torch.float64) - 0.5      # column-major m x k
        B = torch.rand(n, k, device="cuda", generator=g, dtype=torch.float64) - 0.5
        Cm = torch.rand(n, m, device="cuda", generator=g, dtype=torch.float64) - 0.5
        ops.append((m, n, k, 1.0 + 0.5 * (i % 2), float(i % 3) - 1.0, A, B, Cm))
    s = torch.cuda.current_stream().cuda_stream

    def run(batched):
        outs = [op[7].clone() for op in ops]
        if batched:
            assert L.parsec_b200_batch_begin(s) == 0
        for j, ((m, n, k, alpha, beta, A, B, _), Cm) in enumerate(zip(ops, outs)):
            ta = b"T" if (j == 3 and m == k) else b"N"
            assert L.parsec_b200_dgemm(ta, b"N", m, n, k, alpha, A.data_ptr(), m if ta == b"N" else k, B.data_ptr(), k, beta,
                                       Cm.data_ptr(), m, s) == 0
        if batched:
            assert L.parsec_b200_batch_end(s) == 0
        torch.cuda.synchronize()
        return outs

    single, batched = run(False), run(True)
    for a, b, op in zip(single, batched, ops):
        assert torch.equal(a, b), op[:3]
    m, n, k, alpha, beta, A, B, Cm = ops[0]
    want = beta * Cm + alpha * (B @ A)
    assert (single[0] - want).abs().max().item() < 1e-12
