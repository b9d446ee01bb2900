#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
run() { tag=$1; n=$2; shift 2
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n "$@" > gpurun_out/bench_r2_$tag.json 2> gpurun_out/bench_r2_$tag.err || tail -5 gpurun_out/bench_r2_$tag.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_r2_$tag.json'));print('$tag', round(d['value']), round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'parity', d['parity'] and d['parity']['norm'], d['clocks']['sm_mhz'], d['clocks']['reasons'])"
}
run g8_bf16 8 --dtype bf16 --nb 2048 --order 32768
run g8_w256 8 --window 256 --steps 2 --no-parity
run g8_w0 8 --window 512 --steps 2 --no-parity
