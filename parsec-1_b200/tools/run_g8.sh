#!/bin/bash
# One 8-GPU box: multi-GPU parity tests (log kept), then the scaling bench lines with their parity objects.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dist.py "tests/test_gpu_overlay.py::test_two_gpus_in_one_process_share_tiles_over_nvlink" tests/test_gpu_cleanout_and_peers.py -v -s 2>&1 | grep -E "PASS|FAIL|SKIP|passed|failed|rank . ok|2 GPUs|Error|assert" > gpurun_out/gputest_g8_r2.log
tail -12 gpurun_out/gputest_g8_r2.log
run() { # $1 = tag, $2 = ranks, rest = bench args
  tag=$1; n=$2; shift 2
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus $n "$@" > gpurun_out/bench_r2_$tag.json 2> gpurun_out/bench_r2_$tag.err || tail -5 gpurun_out/bench_r2_$tag.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_r2_$tag.json'));print('$tag', round(d['value']), round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'parity', d['parity']['norm'], d['parity']['versions_ok'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
}
run g8 8
run g8_bf16 8 --dtype bf16 --nb 2048 --n 32768
run g4 4
