"""Times parsec_b200_gemm_bf16_sm100 against torch.matmul (cuBLAS) on bf16 inputs.  python parsec-1_b200/tools/bench_bf16.py"""
import importlib.util
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
spec = importlib.util.spec_from_file_location("hostapi", os.path.join(ROOT, "parsec-1_b200", "hostapi.py"))
H = importlib.util.module_from_spec(spec)
spec.loader.exec_module(H)
L = H.lib()


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
    ours = timed(lambda: L.parsec_b200_gemm_bf16_sm100(b"N", b"N", n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 0.0,
                                                       C.data_ptr(), n, s), 20 if n <= 4096 else 5)
    # column-major C = A B with column-major A, B is row-major C^T = B^T A^T: same flops for cuBLAS
    ref = timed(lambda: torch.matmul(B, A), 20 if n <= 4096 else 5)
    got = C.clone()
    want = torch.matmul(B.float(), A.float())   # row-major view of (A_cm * B_cm)^T
    err = (got - want).abs().max().item()
    print("n=%5d ours %.3f ms %.1f TFLOP/s | cuBLAS bf16 %.3f ms %.1f TFLOP/s | max|diff vs fp32 matmul| %.3e" % (
        n, ours, 2 * n ** 3 / ours * 1e-9, ref, 2 * n ** 3 / ref * 1e-9, err))
