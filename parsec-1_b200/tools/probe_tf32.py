"""Layout probes for the tf32 variant (debug aid)."""
import importlib.util, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
spec = importlib.util.spec_from_file_location("hostapi", os.path.join(ROOT, "parsec-1_b200", "hostapi.py"))
H = importlib.util.module_from_spec(spec); spec.loader.exec_module(H)
f = H.lib().parsec_b200_gemm_tf32_sm100
m, n, k = 128, 256, 32
def run(A, B):
    dA = torch.from_numpy(np.ascontiguousarray(A.T.astype(np.float32))).cuda()
    dB = torch.from_numpy(np.ascontiguousarray(B.T.astype(np.float32))).cuda()
    dC = torch.zeros(n, m, dtype=torch.float32, device="cuda")
    assert f(b"N", b"N", m, n, k, 1.0, dA.data_ptr(), m, dB.data_ptr(), k, 0.0, dC.data_ptr(), m, 0) == 0
    torch.cuda.synchronize()
    return dC.cpu().numpy().T
i = np.arange(m)[:, None]; j = np.arange(n)[None, :]; kk = np.arange(k)
C = run(np.tile(i, (1, k)).astype(float), np.ones((k, n)))
print("A[i,k]=i, B=1: expect k*i. rows 0..5 col0:", C[:6, 0], " row 33,64,97:", C[33,0], C[64,0], C[97,0], "ok" if np.allclose(C, k * i) else "MISMATCH")
C = run(np.ones((m, k)), np.tile(j, (k, 1)).astype(float))
print("A=1, B[k,j]=j: expect k*j. cols 0..5:", C[0, :6], "ok" if np.allclose(C, k * j) else "MISMATCH")
for k0 in (0, 1, 7, 8, 9, 31):
    A = np.zeros((m, k)); A[:, k0] = 1.0
    B = np.zeros((k, n)); B[k0, :] = 1.0
    C = run(A, B)
    print("one-hot k0=%d: expect all ones; got sum=%g min=%g max=%g" % (k0, C.sum(), C.min(), C.max()))
# which k of A pairs with which k of B?
for ka in (0, 1, 8):
    A = np.zeros((m, k)); A[:, ka] = 1.0
    hits = []
    for kb in range(k):
        B = np.zeros((k, n)); B[kb, :] = 1.0
        if run(A, B).max() > 0.5: hits.append(kb)
    print("A column ka=%d multiplies B rows" % ka, hits)
