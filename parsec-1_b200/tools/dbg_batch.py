import os, sys, torch
sys.path.insert(0, "tests")
from conftest import load_hostapi
H = load_hostapi(); L = H.lib()
L.parsec_b200_batch_begin.argtypes = [H.C.c_void_p]; L.parsec_b200_batch_end.argtypes = [H.C.c_void_p]
s = torch.cuda.current_stream().cuda_stream
shapes = [(256, 256, 256), (128, 384, 64), (200, 136, 72), (512, 256, 1024), (1000, 24, 16), (256, 256, 256), (130, 130, 130)] * 10
g = torch.Generator(device="cuda"); g.manual_seed(5)
ops = []
for i, (m, n, k) in enumerate(shapes):
    A = torch.rand(k, m, device="cuda", generator=g, dtype=torch.float64) - 0.5
    B = torch.rand(n, k, device="cuda", generator=g, dtype=torch.float64) - 0.5
    Cm = torch.rand(n, m, device="cuda", generator=g, dtype=torch.float64) - 0.5
    ops.append((m, n, k, 1.0 + 0.5 * (i % 2), float(i % 3) - 1.0, A, B, Cm))
def run(batched):
    outs = [op[7].clone() for op in ops]
    if batched: L.parsec_b200_batch_begin(s)
    for (m, n, k, alpha, beta, A, B, _), Cm in zip(ops, outs):
        L.parsec_b200_dgemm(b"N", b"N", m, n, k, alpha, A.data_ptr(), m, B.data_ptr(), k, beta, Cm.data_ptr(), m, s)
    if batched: L.parsec_b200_batch_end(s)
    torch.cuda.synchronize()
    return outs
for rep in range(2):
    single, batched = run(False), run(True)
    for i, (a, b, op) in enumerate(zip(single, batched, ops)):
        m, n, k, alpha, beta, A, B, Cm = op
        want = beta * Cm + alpha * (B @ A)
        ea, eb = (a - want).abs().max().item(), (b - want).abs().max().item()
        if not torch.equal(a, b) or ea > 1e-11 or eb > 1e-11:
            d = (a - b).abs(); idx = torch.nonzero(d > 0)
            print(rep, i, op[:5], "single err %.3g batched err %.3g; differing %d first %s ptr %x" % (ea, eb, idx.shape[0], idx[:3].tolist(), b.data_ptr()))
print("done")
