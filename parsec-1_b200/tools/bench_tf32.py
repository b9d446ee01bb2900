"""Times parsec_b200_gemm_tf32_sm100 against torch.matmul with TF32 enabled (cuBLAS).  python parsec-1_b200/tools/bench_tf32.py"""
import importlib.util
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
spec = importlib.util.spec_from_file_location("hostapi", os.path.join(ROOT, "parsec-1_b200", "hostapi.py"))
H = importlib.util.module_from_spec(spec)
spec.loader.exec_module(H)
L = H.lib()
torch.backends.cuda.matmul.allow_tf32 = True


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
    A = torch.rand(n, n, device="cuda") - 0.5
    B = torch.rand(n, n, device="cuda") - 0.5
    C = torch.zeros(n, n, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    reps = 20 if n <= 4096 else 5
    ours = timed(lambda: L.parsec_b200_gemm_tf32_sm100(b"N", b"N", n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0,
                                                       C.data_ptr(), n, s), reps)
    ref = timed(lambda: torch.matmul(B, A), reps)
    print("n=%5d tf32 ours %.3f ms %.1f TFLOP/s | cuBLAS tf32 %.3f ms %.1f TFLOP/s" % (
        n, ours, 2 * n ** 3 / ours * 1e-9, ref, 2 * n ** 3 / ref * 1e-9))
