#!/bin/bash
# Runs the reference's own multi-rank tests (unmodified sources, built by overlay/build_dist.sh against the distributed
# tree) over overlay/comm_engine_b200.c, with the rank counts and arguments of the reference's ctest files
# (tests/*/Testings.cmake).  CPU only.  One log per test under gpurun_out/dist_reftests/, one summary line each.
cd "$(dirname "$0")/.."
R=$PWD/baseline/_ref
D=$R/reftests/dist
OUT=$PWD/gpurun_out/dist_reftests
mkdir -p "$OUT"
export LD_LIBRARY_PATH=$R/dist:$PWD/parsec-1_b200:$LD_LIBRARY_PATH
export PARSEC_MCA_device_cuda_enabled=0
PRE=; [ -f "$R/reftests/libsegv_trace.so" ] && PRE=$R/reftests/libsegv_trace.so

run() { # $1 = ranks, $2 = time limit, $3 = test, rest = arguments
  n=$1; limit=$2; t=$3; shift 3
  [ -x "$D/$t" ] || { printf '%-42s not built\n' "$t"; return; }
  log="$OUT/${t}_np$n.log"
  start=$(date +%s%N)
  LD_PRELOAD=$PRE timeout $((limit + 10)) overlay/pb2run -n "$n" --timeout "$limit" "$D/$t" "$@" > "$log" 2>&1
  rc=$?
  printf '%-42s np=%d rc=%-4d %6d ms  %s\n' "$t" "$n" "$rc" $(( ($(date +%s%N) - start) / 1000000 )) \
         "$(grep -a -m1 -io 'error[^"]*\|wrong[^"]*\|fail[a-z]*[^"]*\|signal [0-9]*' "$log" | head -c 70)"
}

run 2 60 dtd_test_pingpong
run 4 60 dtd_test_task_inserting_task
run 4 60 dtd_test_task_insertion
run 4 60 dtd_test_war
run 4 60 dtd_test_interleave_actions
run 4 60 dtd_test_allreduce
run 2 60 dtd_test_new_tile
run 2 60 jdf_forward_RW_NULL
run 2 60 jdf_forward_READ_NULL
run 4 60 write_check
run 4 60 compose
run 4 120 project -x
run 4 60 merge_sort
run 4 120 reshape -N 120 -t 9 -c 10
run 4 60 input_dep_reshape_single_copy -N 12 -t 2 -c 2
run 4 60 remote_multiple_outs_same_pred_flow -N 12 -t 2 -c 2
run 8 180 testing_redistribute -M 2400 -N 2400 -a 2400 -A 2400 -t 300 -T 300 -b 200 -B 200 -m 2000 -n 2000 -I 30 -J 40 -i 100 -j 121 -v -z -x -P 2 -Q 4 -p 4 -q 2
run 8 180 testing_redistribute_random -M 2400 -N 2400 -a 2400 -A 2400 -t 300 -T 300 -b 200 -B 200 -m 2000 -n 2000 -I 30 -J 40 -i 100 -j 121 -v -z -x -P 2 -Q 4 -p 4 -q 2
run 8 180 testing_band -N 3200 -T 160 -P 4 -s 5 -S 10 -p 2 -f 2 -F 10 -b 2
run 8 180 testing_stencil_1D -t 100 -T 100 -N 1000 -M 1000 -I 10 -R 2 -m 1
run 2 120 bw_test -n 10 -f 10 -l 2097152
run 2 120 schedmicro -t 10 -l 8 -n 512
run 2 120 rtt
