#!/bin/bash
# Like-for-like on one box: the same JDF binary at BASELINE configs[1] (N=32768, NB=1024 fp64)
#   overlay   reference runtime + overlay shim + libparsec_b200 (sm_100a DMMA kernel)
#   stock     reference runtime + its own CUDA module + its own body (cublasDgemm_v2)       <- BASELINE.md's GPU comparator
#   mixed     the two cross terms (stock module + our kernel, our module + cuBLAS)
# Logs into gpurun_out/overlay_compare.log.  N and NB can be overridden: run_compare.sh 16384 1024
cd "$(dirname "$0")/.."
N=${1:-32768}; NB=${2:-1024}; CORES=${3:-8}; RUNS=${4:-3}
R=$PWD/baseline/_ref
export PARSEC_CUCORES_LIB=$PWD/parsec-1_b200/libparsec_b200.so
mkdir -p gpurun_out
LOG=gpurun_out/overlay_compare.log
: > $LOG
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv >> $LOG
nproc >> $LOG
run() { tag=$1; lib=$2; shift 2
  echo "== $tag: $*" >> $LOG
  LD_LIBRARY_PATH=$R/$lib:$PWD/parsec-1_b200:$LD_LIBRARY_PATH timeout 900 $R/bin/dgemm_NN_app "$@" -- --mca device_cuda_enabled 1 >> $LOG 2>&1
  echo "rc=$?" >> $LOG
}
run overlay_b200 overlay -N $N -t $NB -c $CORES -r $RUNS --release
run stock_cublas stock -N $N -t $NB -c $CORES -r $RUNS --release --cublas
run stock_b200kernel stock -N $N -t $NB -c $CORES -r $RUNS --release
run overlay_cublas overlay -N $N -t $NB -c $CORES -r $RUNS --release --cublas
run overlay_b200_resident overlay -N $N -t $NB -c $CORES -r $RUNS
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv >> $LOG
grep -E "^==|gflops|rc=" $LOG | cut -c1-400
