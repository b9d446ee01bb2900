"""Six launches of the grouped persistent tcgen05 kernel for an ncu capture: three 8192^3 bf16 GEMMs (beta = 1) and three
batches of sixteen 2048^3 tile tasks (beta = 1, distinct tiles: 1.1 GiB, larger than L2).

    ncu --set full --clock-control none --import-source on -k regex:gemm_umma_grouped -c 6 -o gpurun_out/grouped_r2 \
        python parsec-1_b200/tools/ncu_grouped.py
"""
import importlib.util
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
spec = importlib.util.spec_from_file_location("hostapi", os.path.join(ROOT, "parsec-1_b200", "hostapi.py"))
H = importlib.util.module_from_spec(spec)
spec.loader.exec_module(H)
L = H.lib()
L.parsec_b200_batch_begin.argtypes = [H.C.c_void_p]
L.parsec_b200_batch_end.argtypes = [H.C.c_void_p]
s = torch.cuda.current_stream().cuda_stream

n = 8192
A = (torch.rand(n, n, device="cuda") - 0.5).to(torch.bfloat16)
B = (torch.rand(n, n, device="cuda") - 0.5).to(torch.bfloat16)
C = torch.zeros(n, n, device="cuda", dtype=torch.float32)
for _ in range(3):
    L.parsec_b200_gemm_bf16_sm100(b"N", b"N", n, n, n, 1.0, A.data_ptr(), n, B.data_ptr(), n, 1.0, C.data_ptr(), n, s)
torch.cuda.synchronize()

nb, nt = 2048, 24
As = [(torch.rand(nb, nb, device="cuda") - 0.5).to(torch.bfloat16) for _ in range(nt)]
Bs = [(torch.rand(nb, nb, device="cuda") - 0.5).to(torch.bfloat16) for _ in range(nt)]
Cs = [torch.zeros(nb, nb, device="cuda", dtype=torch.float32) for _ in range(nt)]
for rep in range(3):
    L.parsec_b200_batch_begin(s)
    for i in range(16):
        t = (rep * 16 + i) % nt
        L.parsec_b200_gemm_bf16_sm100(b"N", b"N", nb, nb, nb, 1.0, As[t].data_ptr(), nb, Bs[t].data_ptr(), nb, 1.0,
                                      Cs[t].data_ptr(), nb, s)
    L.parsec_b200_batch_end(s)
torch.cuda.synchronize()
print("done")
