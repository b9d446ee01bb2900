"""Where a 1024-tile TRSM spends its time: the pieces timed alone (CUDA events, 20 repetitions each)."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import torch
from conftest import load_hostapi
H = load_hostapi()
n = 1024
s = torch.cuda.current_stream().cuda_stream
g = torch.Generator(device="cuda").manual_seed(3)
A = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
T = (torch.tril(torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) / 32) + 2 * torch.eye(n, dtype=torch.float64, device="cuda")).T.contiguous()
Cm = torch.zeros(n, n, dtype=torch.float64, device="cuda")
W = torch.zeros(n, n, dtype=torch.float64, device="cuda")
ws_bytes = H.dtrsm_workspace("R", n, n)
ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
def timed(name, fn, reps=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    print("%-40s %8.1f us" % (name, e0.elapsed_time(e1) * 1e3 / reps))
timed("trsm R L T N 1024 x 1024", lambda: H.dtrsm("R", "L", "T", "N", n, n, 1.0, T.data_ptr(), n, Cm.data_ptr(), n, ws.data_ptr(), ws_bytes, s))
timed("trsm with m = 2 (inversions + launches)", lambda: H.dtrsm("R", "L", "T", "N", 2, n, 1.0, T.data_ptr(), n, Cm.data_ptr(), n, ws.data_ptr(), ws_bytes, s))
timed("trsm with n = 128 (one block)", lambda: H.dtrsm("R", "L", "T", "N", n, 128, 1.0, T.data_ptr(), n, Cm.data_ptr(), n, ws.data_ptr(), ws_bytes, s))
timed("gemm NT 1024 x 128 x 128", lambda: H.dgemm("N", "T", n, 128, 128, 1.0, A.data_ptr(), n, T.data_ptr(), n, 0.0, W.data_ptr(), n, s))
timed("gemm NT 1024 x 896 x 128", lambda: H.dgemm("N", "T", n, 896, 128, -1.0, A.data_ptr(), n, T.data_ptr(), n, 1.0, Cm.data_ptr(), n, s))
timed("gemm NT 1024 x 512 x 512", lambda: H.dgemm("N", "T", n, 512, 512, -1.0, A.data_ptr(), n, T.data_ptr(), n, 1.0, Cm.data_ptr(), n, s))
timed("gemm NT 1024^3", lambda: H.dgemm("N", "T", n, n, n, -1.0, A.data_ptr(), n, T.data_ptr(), n, 1.0, Cm.data_ptr(), n, s))
timed("copy 1024 x 128 (2D, device to device)", lambda: Cm[:128].copy_(W[:128]))
