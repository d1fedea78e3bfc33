"""SASS opcode histogram of libgapflow_b200.so per kernel family (static instruction counts from
`cuobjdump -sass`): the evidence that the library's kernels use TMA (UTMALDG), mbarriers (SYNCS), the FP64
tensor path (DMMA) and cp.async (LDGSTS).  tcgen05 / TMEM (UTC*MMA, LDTM) do not apply: FP64 has no tcgen05 kind.

    python tools/sass_histogram.py [path/to/lib.so] > profiles/<round>_sass_histogram.txt
"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "gapflow_b200/libgapflow_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
KEY = ["UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "DMMA", "HMMA", "LDGSTS", "DFMA", "DMUL", "DADD", "MUFU", "REDUX", "SHFL",
       "LDS", "STS", "LDG", "STG", "BAR", "ATOM", "RED", "MEMBAR", "CCTL", "UTCHMMA", "LDTM"]
fam = collections.OrderedDict()
cur = None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name)
        name = re.sub(r"^void\s+", "", name)
        name = re.sub(r"gpf::(exact|fastm)::", r"\1::", name).replace("gpf::", "")
        base = name.split("<")[0]
        cur = fam.setdefault(base, {"variants": 0, "ops": collections.Counter(), "total": 0})
        cur["variants"] += 1
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur is not None:
        cur["ops"][m.group(1)] += 1
        cur["total"] += 1
print(f"# static SASS instruction counts per kernel family of {lib} (cuobjdump -sass, sm_100a); variants = template instances")
print("kernel,variants,total," + ",".join(KEY))
tot = collections.Counter()
for k, v in fam.items():
    print(f"{k},{v['variants']},{v['total']}," + ",".join(str(v['ops'].get(o, 0)) for o in KEY))
    tot.update(v["ops"])
print("ALL,,%d," % sum(v["total"] for v in fam.values()) + ",".join(str(tot.get(o, 0)) for o in KEY))
