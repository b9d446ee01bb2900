#!/bin/bash
# Tile-size sweep of the FP64 pipeline with and without batched submissions (profiles/tile_size_sweep_r2.md).
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
for cfg in "2048 256 0" "2048 256 32" "2048 256 64" "8192 256 64" "8192 512 0" "8192 512 32" "16384 512 32" "16384 1024 0" "16384 1024 16"; do
  set -- $cfg
  timeout 200 python bench.py --n $1 --nb $2 --batch $3 --no-cpu-baseline > gpurun_out/sweep_$1_$2_b$3.json 2> gpurun_out/tmp.err || tail -3 gpurun_out/tmp.err
  python -c "
import json;d=json.load(open('gpurun_out/sweep_$1_$2_b$3.json'));print('N=$1 NB=$2 batch=$3', round(d['value']), round(d['e2e']['value']), d['parity']['norm'], {k: round(v,2) for k,v in d['host_us_per_task'].items()})"
done
