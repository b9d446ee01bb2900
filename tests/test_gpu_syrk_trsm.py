"""parsec_b200_dsyrk / parsec_b200_dtrsm: the other two BLAS-3 tile bodies of a tiled Cholesky, behind the same kernel
library as the GEMM ("TRSM/SYRK/POTRF-style bodies as DPLASMA would supply").  Checked against
numpy / scipy in fp64 under the same kind of norm the north star uses for GEMM: error relative to the norms of the
operands, the inner dimension and the unit roundoff.  The triangle that BLAS does not reference must come back
untouched, bit for bit, and so must everything outside the matrices (canaries in the padding)."""
import numpy as np
import pytest

from conftest import load_hostapi

pytestmark = pytest.mark.gpu
H = load_hostapi()
EPS = 2.0 ** -53
CANARY = 777.25


def dev_colmajor(X, ld, guard=1):
    """X (rows x cols) as a column-major device matrix with leading dimension ld, one guard column either side."""
    import torch
    host = np.full((X.shape[1] + 2 * guard, ld), CANARY)
    host[guard:guard + X.shape[1], :X.shape[0]] = X.T
    t = torch.from_numpy(host).cuda()
    return t, t.data_ptr() + 8 * ld * guard


def read_back(t, rows, cols, ld, guard=1):
    back = t.cpu().numpy()
    assert np.all(back[:guard] == CANARY) and np.all(back[guard + cols:] == CANARY) and np.all(back[guard:guard + cols, rows:] == CANARY)
    return back[guard:guard + cols, :rows].T.copy()


inf = lambda X: np.abs(X).sum(axis=1).max() if X.size else 0.0


@pytest.mark.parametrize("uplo", ["L", "U"])
@pytest.mark.parametrize("trans", ["N", "T"])
@pytest.mark.parametrize("n,k,pad", [(1024, 1024, 0), (1000, 520, 2), (129, 18, 0), (128, 16, 4), (130, 2048, 0), (7, 5, 1)])
def test_dsyrk_updates_one_triangle_only(uplo, trans, n, k, pad):
    import torch
    rng = np.random.default_rng(n + 7 * k + ord(uplo) + ord(trans))
    A = rng.uniform(-0.5, 0.5, (n, k) if trans == "N" else (k, n))
    C0 = rng.uniform(-0.5, 0.5, (n, n))
    lda, ldc = A.shape[0] + pad, n + pad
    dA, pA = dev_colmajor(A, lda)
    dC, pC = dev_colmajor(C0, ldc)
    alpha, beta = -1.0, 1.0                       # the update of a Cholesky step
    assert H.dsyrk(uplo, trans, n, k, alpha, pA, lda, beta, pC, ldc, torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    got = read_back(dC, n, n, ldc)
    opA = A if trans == "N" else A.T
    full = alpha * (opA @ opA.T) + beta * C0
    tri = np.tril if uplo == "L" else np.triu
    other = (lambda X: np.triu(X, 1)) if uplo == "L" else (lambda X: np.tril(X, -1))
    assert np.array_equal(other(got), other(C0))                                   # not referenced, not touched
    err = inf(tri(got) - tri(full)) / max(inf(opA) * inf(opA.T) * max(k, 1) * EPS, 1e-300)
    assert err < 10, err
    assert read_back(dA, A.shape[0], A.shape[1], lda).tobytes() == A.tobytes()      # the input is an input


def test_dsyrk_is_the_triangle_of_the_gemm_bit_for_bit():
    """Same tile code, same order of accumulation: SYRK must give exactly the bits DGEMM gives with B = A^T."""
    import torch
    n, k = 640, 384
    rng = np.random.default_rng(5)
    A = rng.uniform(-0.5, 0.5, (n, k))
    C0 = rng.uniform(-0.5, 0.5, (n, n))
    s = torch.cuda.current_stream().cuda_stream
    dA, pA = dev_colmajor(A, n)
    d1, p1 = dev_colmajor(C0, n)
    d2, p2 = dev_colmajor(C0, n)
    assert H.dsyrk("L", "N", n, k, 1.5, pA, n, 0.5, p1, n, s) == 0
    assert H.dgemm("N", "T", n, n, k, 1.5, pA, n, pA, n, 0.5, p2, n, s) == 0
    torch.cuda.synchronize()
    assert np.array_equal(np.tril(read_back(d1, n, n, n)), np.tril(read_back(d2, n, n, n)))


def well_conditioned_triangle(rng, n, lower):
    T = rng.uniform(-0.5, 0.5, (n, n)) / max(n, 1) ** 0.5
    T = np.tril(T) if lower else np.triu(T)
    T[np.arange(n), np.arange(n)] = rng.uniform(1.0, 2.0, n) * rng.choice([-1.0, 1.0], n)
    return T


@pytest.mark.parametrize("side,uplo,trans,diag", [("R", "L", "T", "N"),                      # Cholesky, lower
                                                  ("L", "L", "N", "N"), ("R", "U", "N", "N"),  # LU: L \\ and / U
                                                  ("L", "U", "T", "N"), ("R", "L", "N", "U"), ("L", "L", "T", "U"),
                                                  ("R", "U", "T", "N"), ("L", "U", "N", "U")])
@pytest.mark.parametrize("m,n", [(1024, 1024), (1000, 520), (130, 129), (64, 200), (5, 3), (256, 128)])
def test_dtrsm_against_scipy(side, uplo, trans, diag, m, n):
    import torch
    from scipy.linalg import solve_triangular
    rng = np.random.default_rng(m + 13 * n + sum(map(ord, side + uplo + trans + diag)))
    tri = n if side == "R" else m
    T = well_conditioned_triangle(rng, tri, uplo == "L")
    garbage = rng.uniform(5.0, 9.0, (tri, tri))                                     # the triangle BLAS must not read
    stored = T + (np.triu(garbage, 1) if uplo == "L" else np.tril(garbage, -1))
    if diag == "U":
        stored[np.arange(tri), np.arange(tri)] = 55.0                                # nor the diagonal when it is unit
        T[np.arange(tri), np.arange(tri)] = 1.0
    B0 = rng.uniform(-0.5, 0.5, (m, n))
    alpha = 1.0 if (m + n) % 2 else -0.75
    ldt, ldb = tri + 2, m + 4
    dT, pT = dev_colmajor(stored, ldt)
    dB, pB = dev_colmajor(B0, ldb)
    ws_bytes = H.dtrsm_workspace(side, m, n)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    assert H.dtrsm(side, uplo, trans, diag, m, n, alpha, pT, ldt, pB, ldb, ws.data_ptr(), ws_bytes,
                   torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    X = read_back(dB, m, n, ldb)
    opT = T if trans == "N" else T.T
    lower_eff = (uplo == "L") == (trans == "N")
    if side == "L":
        want = solve_triangular(opT, alpha * B0, lower=lower_eff)
        resid = opT @ X - alpha * B0
    else:
        want = solve_triangular(opT.T, (alpha * B0).T, lower=not lower_eff).T
        resid = X @ opT - alpha * B0
    # backward error of the solve, and agreement with LAPACK's substitution up to the conditioning of T
    assert inf(resid) / (inf(opT) * inf(X) * tri * EPS + 1e-300) < 10
    cond = np.linalg.cond(T, np.inf)
    assert inf(X - want) / (inf(want) * cond * tri * EPS + 1e-300) < 10
    assert read_back(dT, tri, tri, ldt).tobytes() == stored.tobytes()
    assert H.dtrsm(side, uplo, trans, diag, m, n, alpha, pT, ldt, pB, ldb, ws.data_ptr(), ws_bytes - 1024, 0) == -4   # scratch too small
    # the two halves on their own (what a DAG that inverts the diagonal tile once per panel calls): same bits
    L = H.lib()
    dB2, pB2 = dev_colmajor(B0, ldb)
    inv = torch.empty(((tri + 127) // 128) * 128 * 128, dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    assert L.parsec_b200_dtrtri_diag_sm100(uplo.encode(), diag.encode(), tri, pT, ldt, inv.data_ptr(), s) == 0
    assert L.parsec_b200_dtrsm_inv_sm100(side.encode(), uplo.encode(), trans.encode(), m, n, alpha, pT, ldt, inv.data_ptr(),
                                         pB2, ldb, ws.data_ptr(), ws_bytes, s) == 0
    torch.cuda.synchronize()
    assert np.array_equal(read_back(dB2, m, n, ldb), X)


@pytest.mark.parametrize("uplo", ["L", "U"])
@pytest.mark.parametrize("n,pad", [(1024, 0), (1000, 2), (130, 0), (128, 4), (64, 0), (5, 1), (2048, 0)])
def test_dpotrf_tile_on_the_gpu(uplo, n, pad):
    import torch
    rng = np.random.default_rng(n + ord(uplo))
    G = rng.uniform(-0.5, 0.5, (n, n))
    A = G + G.T + n * np.eye(n)                                   # symmetric, strictly diagonally dominant
    other = rng.uniform(5.0, 9.0, (n, n))                         # what sits in the triangle that is not referenced
    stored = (np.tril(A) + np.triu(other, 1)) if uplo == "L" else (np.triu(A) + np.tril(other, -1))
    lda = n + pad
    dA, pA = dev_colmajor(stored, lda)
    L = H.lib()
    ws_bytes = L.parsec_b200_dpotrf_workspace(n)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert L.parsec_b200_dpotrf_sm100(uplo.encode(), n, pA, lda, ws.data_ptr(), ws_bytes, info.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    assert info.item() == 0
    got = read_back(dA, n, n, lda)
    F = np.tril(got) if uplo == "L" else np.triu(got).T           # the lower factor either way
    untouched = (lambda X: np.triu(X, 1)) if uplo == "L" else (lambda X: np.tril(X, -1))
    assert np.array_equal(untouched(got), untouched(stored))
    assert inf(A - F @ F.T) / (inf(A) * n * EPS) < 10             # LAPACK's test ratio
    ref = np.linalg.cholesky(A)
    assert inf(F - ref) / (inf(ref) * np.linalg.cond(A, np.inf) * n * EPS) < 10


def test_dpotrf_reports_the_first_bad_pivot():
    import torch
    n = 300
    A = np.eye(n) * 4.0
    A[200, 200] = -1.0                                            # not positive definite from pivot 201 on
    dA, pA = dev_colmajor(A, n)
    L = H.lib()
    ws_bytes = L.parsec_b200_dpotrf_workspace(n)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert L.parsec_b200_dpotrf_sm100(b"L", n, pA, n, ws.data_ptr(), ws_bytes, info.data_ptr(), 0) == 0
    torch.cuda.synchronize()
    assert info.item() == 201
    assert L.parsec_b200_dpotrf_sm100(b"L", n, pA, n, ws.data_ptr(), 100, info.data_ptr(), 0) == -4


def test_bad_arguments():
    import torch
    x = torch.zeros(1 << 16, dtype=torch.float64, device="cuda")
    p = x.data_ptr()
    assert H.dsyrk("X", "N", 8, 8, 1.0, p, 8, 1.0, p, 8) == -4
    assert H.dsyrk("L", "N", 8, 8, 1.0, p, 4, 1.0, p, 8) == -4
    assert H.dsyrk("L", "N", 0, 8, 1.0, p, 8, 1.0, p, 8) == 0
    assert H.dtrsm("Q", "L", "N", "N", 8, 8, 1.0, p, 8, p, 8, p, 1 << 19) == -4
    assert H.dtrsm("L", "L", "N", "N", 8, 8, 1.0, p, 4, p, 8, p, 1 << 19) == -4
    assert H.dtrsm_workspace("R", 1024, 1024) == (8 * 128 * 128 + 1024 * 1024) * 8 + 512


def test_speed_of_the_cholesky_bodies():
    """Reported, loosely asserted.  A single NB = 1024 call is 64 CTAs on 148 SMs, so SYRK's skipped blocks only free
    SMs for the tasks running beside it; at 4096 (many waves) one triangle must cost clearly less than the square."""
    import torch
    s = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device="cuda").manual_seed(3)
    us = {}
    for n in (1024, 4096):
        A = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
        T = torch.tril(torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) / 32) + 2 * torch.eye(n, dtype=torch.float64, device="cuda")
        T = T.T.contiguous()                                     # column-major lower triangle
        Cm = torch.zeros(n, n, dtype=torch.float64, device="cuda")
        ws_bytes = H.dtrsm_workspace("R", n, n)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
        pws_bytes = H.lib().parsec_b200_dpotrf_workspace(n)
        pws = torch.empty(pws_bytes, dtype=torch.uint8, device="cuda")
        P = (A @ A.T + n * torch.eye(n, dtype=torch.float64, device="cuda")).contiguous()
        P0 = P.clone()

        def potrf():
            P.copy_(P0)
            return H.lib().parsec_b200_dpotrf_sm100(b"L", n, P.data_ptr(), n, pws.data_ptr(), pws_bytes, None, s)

        calls = {"potrf L (+ copy)": potrf, "gemm NT": lambda: H.dgemm("N", "T", n, n, n, -1.0, A.data_ptr(), n, A.data_ptr(), n, 1.0, Cm.data_ptr(), n, s),
                 "syrk L": lambda: H.dsyrk("L", "N", n, n, -1.0, A.data_ptr(), n, 1.0, Cm.data_ptr(), n, s),
                 "trsm RLTN": lambda: H.dtrsm("R", "L", "T", "N", n, n, 1.0, T.data_ptr(), n, Cm.data_ptr(), n, ws.data_ptr(), ws_bytes, s)}
        for name, fn in calls.items():
            for _ in range(3):
                assert fn() == 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                fn()
            e1.record()
            torch.cuda.synchronize()
            us["%s %d" % (name, n)] = e0.elapsed_time(e1) * 1e3 / 10
    print("us per call:", {k: round(v, 1) for k, v in us.items()})
    assert us["syrk L 4096"] < 0.7 * us["gemm NT 4096"], us
    assert us["trsm RLTN 4096"] < 1.5 * us["gemm NT 4096"], us      # half the flops of the GEMM, in skinny pieces
    assert us["trsm RLTN 1024"] < 10 * us["gemm NT 1024"], us
