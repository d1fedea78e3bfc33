"""profiles/traffic.json: DRAM bytes per launch of the FV pass kernels, per workload, from an `ncu --page raw --csv`
dump of a `--set full` (or dram__bytes_*) capture -- what bench.py reports as roofline.traffic / dram_frac.
    ncu -i prof.ncu-rep --page raw --csv > raw.csv; python tools/ncu_traffic.py raw.csv journal8192 "source text"
"""
import csv
import json
import os
import sys

raw, key, src = sys.argv[1], sys.argv[2], " ".join(sys.argv[3:])
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
ki, ri, wi = hdr.index('Kernel Name'), hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum')
scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
out = {}
for r in rows[2:]:
    name = r[ki]
    if 'fv_pass_kernel' not in name:
        continue
    targs = name[name.index('<') + 1:name.index('>')].replace(' ', '').split(',')
    corr, gen = targs[0] == '1', targs[4] == '1'
    if gen:
        continue                                   # the general launch returns at once
    b = float(r[ri]) * scale[units[ri]] + float(r[wi]) * scale[units[wi]]
    out.setdefault('corrector' if corr else 'predictor', []).append(b)
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'profiles', 'traffic.json')
table = json.load(open(path)) if os.path.exists(path) else {}
table[key] = {k: sum(v) / len(v) for k, v in out.items()}
table[key]['source'] = src
json.dump(table, open(path, 'w'), indent=1)
print(table[key])
