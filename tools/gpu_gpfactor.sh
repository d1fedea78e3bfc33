#!/bin/bash
tag=${1:-gpf}; out=gpurun_out/$tag; mkdir -p $out
timeout 300 python tools/gp_factor_probe.py 10000 3 > $out/plain.log 2>&1 || { tail -20 $out/plain.log; exit 1; }
cat $out/plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $out/launches.csv python tools/gp_factor_probe.py 10000 1 > $out/ncu.log 2>&1
python - <<PY
import csv, collections
t=collections.Counter(); n=collections.Counter()
for r in csv.reader(open("$out/launches.csv")):
    if len(r) > 14 and r[12] == "gpu__time_duration.sum":
        k = r[4].split("(")[0]; v = float(r[14].replace(",", "")); u = r[13]
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(u, 1.0)
        t[k] += v; n[k] += 1
for k, v in t.most_common(): print("%-50s %6d launches %10.1f us total %8.2f us avg" % (k[:50], n[k], v, v / n[k]))
PY
