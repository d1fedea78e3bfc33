"""Per-kernel SASS histogram from `ncu --page source --csv --print-source sass`: executed warp
instructions by opcode class, plus the top stall-sample instructions."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
kern = None; hdr = None; data = collections.OrderedDict()
for r in rows:
    if r and r[0] == "Kernel Name": kern = r[1]; data[kern] = []; hdr = None; continue
    if r and r[0] == "Address": hdr = r; continue
    if hdr and kern and len(r) >= len(hdr) - 2: data[kern].append(dict(zip(hdr, r)))
for k, ins in data.items():
    print("==", k, len(ins), "SASS instructions")
    tot = sum(int(i["Instructions Executed"]) for i in ins)
    h = collections.Counter(); samp = collections.Counter()
    for i in ins:
        op = i["Source"].split()
        op = [o for o in op if not o.startswith("@")][0].split(".")[0]
        h[op] += int(i["Instructions Executed"]); samp[op] += int(i["# Samples"])
    st = sum(samp.values())
    print(f"total warp inst {tot}, samples {st}")
    for op, n in h.most_common(28):
        print(f"  {op:10s} {n:12d} {100*n/tot:5.1f}%   samples {100*samp[op]/max(st,1):5.1f}%")
    if len(sys.argv) > 2:
        top = sorted(ins, key=lambda i: -int(i["# Samples"]))[:int(sys.argv[2])]
        for i in top:
            print(f"   {i['# Samples']:>6s} {i['Instructions Executed']:>9s}  {i['Source'].strip()[:90]}  long_sb={i.get('stall_long_sb')} wait={i.get('stall_wait')} math={i.get('stall_math')} bar={i.get('stall_barrier')}")
