#!/bin/bash
# Runs the dgemm_NN application on a GPU box over the overlay and the stock libparsec; logs into gpurun_out/.
cd "$(dirname "$0")/.."
R=$PWD/baseline/_ref
export PARSEC_CUCORES_LIB=$PWD/parsec-1_b200/libparsec_b200.so
mkdir -p gpurun_out
run() { # $1 = tag, $2 = lib dir, rest = args
  tag=$1; lib=$2; shift 2
  echo "== $tag: $*" | tee -a gpurun_out/overlay_check.log
  LD_LIBRARY_PATH=$R/$lib:$PWD/parsec-1_b200:$LD_LIBRARY_PATH timeout 600 $R/bin/dgemm_NN_app "$@" >> gpurun_out/overlay_check.log 2>&1
  echo "rc=$?" | tee -a gpurun_out/overlay_check.log
}
: > gpurun_out/overlay_check.log
nvidia-smi -L >> gpurun_out/overlay_check.log
run overlay_cfg1 overlay -N 2048 -t 256 -c 4 --dump gpurun_out/C_overlay_2048.bin
run overlay_8192 overlay -N 8192 -t 1024 -c 4 -r 3 --dump gpurun_out/C_overlay_8192.bin
run stock_b200kernel_cfg1 stock -N 2048 -t 256 -c 4 --dump gpurun_out/C_stock_2048.bin
run stock_cublas_8192 stock -N 8192 -t 1024 -c 4 -r 3 --cublas
run overlay_cublas_8192 overlay -N 8192 -t 1024 -c 4 -r 3 --cublas
tail -40 gpurun_out/overlay_check.log
