"""Times parsec_b200_gemm_bf16_sm100 against torch.matmul (cuBLAS) on bf16 inputs, alone and as a stream of tile tasks.
python parsec-1_b200/tools/bench_bf16.py            (PB2_GEMM_BF16_CTAS=1 selects the one-CTA-per-SM variant)"""
import importlib.util
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
spec = importlib.util.spec_from_file_location("hostapi", os.path.join(ROOT, "parsec-1_b200", "hostapi.py"))
H = importlib.util.module_from_spec(spec)
spec.loader.exec_module(H)
L = H.lib()
print("variant: PB2_GEMM_BF16_CTAS=%s" % os.environ.get("PB2_GEMM_BF16_CTAS", "2 (default)"))


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best


for n in (2048, 4096, 8192):
    A = (torch.rand(n, n, device="cuda") - 0.5).to(torch.bfloat16)
    B = (torch.rand(n, n, device="cuda") - 0.5).to(torch.bfloat16)
    C = torch.zeros(n, n, device="cuda", dtype=torch.float32)
    s = torch.cuda.current_stream().cuda_stream
    reps = 20 if n <= 4096 else 5
    for beta in (0.0, 1.0):
        ours = timed(lambda: L.parsec_b200_gemm_bf16_sm100(b"N", b"N", n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, beta,
                                                           C.data_ptr(), n, s), reps)
        print("n=%5d beta=%.0f ours %.3f ms %.1f TFLOP/s" % (n, beta, ours, 2 * n ** 3 / ours * 1e-9))
    ref = timed(lambda: torch.matmul(B, A), reps)
    print("n=%5d        cuBLAS bf16 %.3f ms %.1f TFLOP/s" % (n, ref, 2 * n ** 3 / ref * 1e-9))

# a stream of independent 2048^3 tile tasks (C += A B), S streams round-robin: the shape the device module produces
nb, ntile = 2048, 24
As = [(torch.rand(nb, nb, device="cuda") - 0.5).to(torch.bfloat16) for _ in range(ntile)]
Bs = [(torch.rand(nb, nb, device="cuda") - 0.5).to(torch.bfloat16) for _ in range(ntile)]
Cs = [torch.zeros(nb, nb, device="cuda", dtype=torch.float32) for _ in range(ntile)]
for S in (1, 2, 4):
    streams = [torch.cuda.Stream() for _ in range(S)]

    def run():
        for i in range(ntile * 4):
            t = i % ntile
            L.parsec_b200_gemm_bf16_sm100(b"N", b"N", nb, nb, nb, 1.0, As[t].data_ptr(), nb, Bs[t].data_ptr(), nb, 1.0,
                                          Cs[t].data_ptr(), nb, streams[i % S].cuda_stream)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for st in streams:
            st.wait_event(e0)
        run()
        for st in streams:
            torch.cuda.current_stream().wait_stream(st)
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("tile stream S=%d: %d x 2048^3 (beta=1) in %.3f ms -> %.1f TFLOP/s" % (S, ntile * 4, best, 2 * nb ** 3 * ntile * 4 / best * 1e-9))

# the same stream of tile tasks, handed over in batches (parsec_b200_batch_begin / _end): one grouped launch per batch
L.parsec_b200_batch_begin.argtypes = [H.C.c_void_p]
L.parsec_b200_batch_end.argtypes = [H.C.c_void_p]
for batch in (4, 8, 16):
    st = torch.cuda.current_stream().cuda_stream

    def run_batched():
        for i0 in range(0, ntile * 4, batch):
            L.parsec_b200_batch_begin(st)
            for i in range(i0, min(i0 + batch, ntile * 4)):
                t = i % ntile
                L.parsec_b200_gemm_bf16_sm100(b"N", b"N", nb, nb, nb, 1.0, As[t].data_ptr(), nb, Bs[t].data_ptr(), nb, 1.0,
                                              Cs[t].data_ptr(), nb, st)
            L.parsec_b200_batch_end(st)
    best = timed(run_batched, 1)
    print("tile stream, batches of %2d: %d x 2048^3 (beta=1) in %.3f ms -> %.1f TFLOP/s" % (
        batch, ntile * 4, best, 2 * nb ** 3 * ntile * 4 / best * 1e-9))
