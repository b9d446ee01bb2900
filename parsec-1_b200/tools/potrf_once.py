"""One parsec_b200_dpotrf and one parsec_b200_dtrsm of a 1024 tile, for a launch list under ncu."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import torch
from conftest import load_hostapi
H = load_hostapi()
L = H.lib()
n = 1024
g = torch.Generator(device="cuda").manual_seed(3)
A = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g) - 0.5
P = (A @ A.T + n * torch.eye(n, dtype=torch.float64, device="cuda")).contiguous()
B = torch.rand(n, n, dtype=torch.float64, device="cuda", generator=g)
ws_bytes = max(L.parsec_b200_dpotrf_workspace(n), H.dtrsm_workspace("R", n, n))
ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
assert L.parsec_b200_dpotrf_sm100(b"L", n, P.data_ptr(), n, ws.data_ptr(), ws_bytes, None, 0) == 0
assert H.dtrsm("R", "L", "T", "N", n, n, 1.0, P.data_ptr(), n, B.data_ptr(), n, ws.data_ptr(), ws_bytes, 0) == 0
torch.cuda.synchronize()
print("done")
