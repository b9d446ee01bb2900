#!/bin/bash
# Two ranks of the distributed overlay sharing one GPU, with the comm engine's statistics (PB2_CE_STATS=1).
#   usage: dist_stats.sh [N] ; each word of $CONFIGS is a comma-separated list of VAR=value settings for one run
cd "$(dirname "$0")/../.."
R=$PWD/baseline/_ref
export PARSEC_CUCORES_LIB=$PWD/parsec-1_b200/libparsec_b200.so LD_LIBRARY_PATH=$R/dist:$PWD/parsec-1_b200
for cfg in ${CONFIGS:-PB2_CE_COPY_THREADS=0}; do
  echo "== 2 ranks, $cfg"
  env $(echo "$cfg" | tr ',' ' ') PB2_CE_STATS=1 PARSEC_MCA_device_cuda_memory_use=40 timeout 70 overlay/pb2run -n 2 --timeout 60 $R/bin/dgemm_NN_dist \
      -N ${1:-12288} -t 1024 -P 1 -Q 2 -r 2 -- --mca device_cuda_enabled 1 2>&1 | grep '"rank"\|comm engine' | \
      sed 's/.*"rank": \([0-9]\).*"cores": \([0-9]*\).*"seconds_best": \([0-9.]*\).*"gflops_best": \([0-9.]*\).*/rank \1: cores \2, best \3 s, \4 GFLOP\/s/' | cut -c1-200
done
