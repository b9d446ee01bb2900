#!/bin/bash
# On a 2-GPU box: the tests that need two GPUs, then the distributed overlay (reference runtime, 2 ranks x 1 GPU, comm
# engine of overlay/comm_engine_b200.c) next to the two other ways of using two GPUs, same problem.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_dist_overlay.py tests/test_gpu_cleanout_and_peers.py tests/test_gpu_overlay.py tests/test_gpu_dist.py \
    -q -s --timeout 300 2>&1 | grep -a "passed\|failed\|skipped\|GFLOP\|Error" | tee gpurun_out/gputest_g2_r2b.log
R=$PWD/baseline/_ref
export PARSEC_CUCORES_LIB=$PWD/parsec-1_b200/libparsec_b200.so
echo "== reference runtime, 2 ranks x 1 GPU (1 x 2 grid), tiles between ranks through the comm engine (host to host)"
LD_LIBRARY_PATH=$R/dist:$PWD/parsec-1_b200 timeout 300 overlay/pb2run -n 2 --gpu-per-rank --timeout 280 $R/bin/dgemm_NN_dist \
    -N 16384 -t 1024 -c 8 -P 1 -Q 2 -r 2 -- --mca device_cuda_enabled 1 2>&1 | grep '"rank"' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print({k: d[k] for k in ('rank', 'gflops_first', 'gflops_best', 'gemm_gpu', 'h2d_bytes', 'd2h_bytes')})"
echo "== reference runtime, 1 process driving 2 GPUs (tiles between GPUs D2D)"
LD_LIBRARY_PATH=$R/overlay:$PWD/parsec-1_b200 timeout 300 $R/bin/dgemm_NN_app -N 16384 -t 1024 -c 8 -r 2 -- --mca device_cuda_enabled 2 2>&1 | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print({k: d[k] for k in ('gflops_first', 'gflops_best', 'gemm_gpu', 'gemm_cpu', 'h2d_bytes', 'd2d_bytes')})"
echo "== reference runtime, 1 rank x 1 GPU, same N"
LD_LIBRARY_PATH=$R/overlay:$PWD/parsec-1_b200 timeout 300 $R/bin/dgemm_NN_app -N 16384 -t 1024 -c 8 -r 2 -- --mca device_cuda_enabled 1 2>&1 | grep '^{' | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print({k: d[k] for k in ('gflops_first', 'gflops_best', 'gemm_gpu')})"
