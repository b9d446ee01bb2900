#!/bin/bash
# launch list of the default bench command (ncu, cold caches, serialised: shares only) and an out-of-core run at scale
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-parity > gpurun_out/bench_r2_short.json 2> gpurun_out/bench_r2_short.err && \
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-parity > gpurun_out/ncu_bench.log 2>&1
tail -3 gpurun_out/launches_r2.csv | cut -c1-200
# BASELINE configs[3] mechanism at bench scale: N=65536 NB=2048 fp64 (96 GiB of matrices) with a 40 GiB heap and
# every chain active (window = all), so that dirty C tiles have to be written back to make room
timeout 600 python bench.py --order 65536 --nb 2048 --hbm-gib 40 --window 1024 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2_ooc_n65536.json 2> gpurun_out/bench_r2_ooc_n65536.err || tail -5 gpurun_out/bench_r2_ooc_n65536.err
python -c "
import json;d=json.load(open('gpurun_out/bench_r2_ooc_n65536.json'));print(round(d['value']), round(d['e2e']['value']), d['hbm_cache'], d['parity'])"
