#!/bin/bash
# Runs the reference's own CUDA tests (built unmodified by overlay/build_overlay.sh into baseline/_ref/reftests/) on a
# GPU box, over the overlay libparsec (overlay/device_cuda_b200_shim.c + libparsec_b200.so) and, as the control, over
# the stock one.  One log per run under gpurun_out/reftests/, one summary line per run on stdout.
#   usage: overlay/run_reftests.sh [overlay|stock|both] [with-stress|with-stress-both]   (with-stress: over the overlay only)
cd "$(dirname "$0")/.."
R=$PWD/baseline/_ref
OUT=$PWD/gpurun_out/reftests
mkdir -p "$OUT"
export PARSEC_MCA_device_cuda_enabled=${PARSEC_MCA_device_cuda_enabled:-1}   # the reference's default (-1) enables no GPU
PRE=; [ -f "$R/reftests/libsegv_trace.so" ] && PRE=$R/reftests/libsegv_trace.so
export PARSEC_MCA_device_show_statistics=1   # the runtime's own per-device table at parsec_fini (device.c:350-400)
which=${1:-both}
libs="overlay stock"; [ "$which" != both ] && libs=$which

run() { # $1 = lib dir, $2 = limit in seconds, $3 = test, rest = its arguments
  lib=$1; limit=$2; t=$3; shift 3
  log="$OUT/${t}_${lib}.log"
  start=$(date +%s%N)
  LD_PRELOAD=$PRE LD_LIBRARY_PATH=$R/$lib:$PWD/parsec-1_b200:$LD_LIBRARY_PATH timeout "$limit" "$R/reftests/$t" "$@" > "$log" 2>&1
  rc=$?
  ms=$(( ($(date +%s%N) - start) / 1000000 ))
  printf '%-28s %-8s rc=%-4d %7d ms  gpu-kernels=%-6s cpu-fallbacks=%d  %s\n' "$t" "$lib" "$rc" "$ms" \
         "$(grep -a '^|  Dev  2 |' "$log" | tail -n 1 | cut -d'|' -f3 | tr -d ' ')" "$(grep -ac 'running on a CPU' "$log")" "$(grep -a -m1 -o 'TEST PASSED\|TEST FAILED\|Result is Wrong\|-- done' "$log" | head -1)"
}

for lib in $libs; do
  run $lib 120 stage
  run $lib 120 testing_get_best_device -N 400 -t 20 -g 1
  run $lib 120 dtd_test_cuda_task_insert
  run $lib 120 dtd_test_new_tile -- --mca device_cuda_enabled 1
  run $lib 180 dtd_test_simple_gemm -M 4096 -N 4096 -K 4096 -m 512 -n 512 -k 512 -t 1 -- --mca device_cuda_enabled 1
  run $lib 120 nvlink
  if [ "$2" = with-stress-both ] || { [ "$2" = with-stress ] && [ $lib = overlay ]; }; then
    # 4000 tiles of 8 MiB through a heap capped at 10 % of HBM by the reference's own MCA parameter
    PARSEC_MCA_device_cuda_memory_use=10 run $lib 600 stress
  fi
done
