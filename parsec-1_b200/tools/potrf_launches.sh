#!/bin/bash
# Launch list of one POTRF and one TRSM of a 1024 tile (ncu, serialised launches): durations per kernel.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/potrf_launches.csv \
    python parsec-1_b200/tools/potrf_once.py > gpurun_out/potrf_ncu.log 2>&1
tail -n 2 gpurun_out/potrf_ncu.log
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/potrf_launches.csv")) if len(r) > 5]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    name = r[ki][:60]
    v = float(r[vi].replace(",", ""))
    v = v / 1000 if r[ui] in ("ns", "nsecond") else v
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
for k, (c, t) in agg.items():
    print("%-62s n=%3d total %8.1f us avg %7.1f us" % (k, c, t, t / c))
PY
