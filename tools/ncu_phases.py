#!/usr/bin/env python
"""Per-phase (barrier-to-barrier) instruction and stall breakdown from `ncu --page source --csv` output.
usage: ncu -i rep.ncu-rep --page source --csv > src.csv ; python tools/ncu_phases.py src.csv <kernel-substring>"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else ""
secs, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = dict(name=r[1], rows=[]); secs.append(cur)
    elif cur is not None:
        cur["rows"].append(r)
for sec in secs:
    if want not in sec["name"]:
        continue
    hdr = sec["rows"][0]
    data = [r for r in sec["rows"][1:] if len(r) == len(hdr)]
    isrc, iex, ist = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot_ex = sum(int(r[iex]) for r in data); tot_st = sum(int(r[ist]) for r in data)
    print(sec["name"], "instructions", tot_ex, "stall samples", tot_st)
    start = 0
    for i, r in enumerate(data):
        if "BAR.SYNC" in r[isrc] or i == len(data) - 1:
            seg = data[start:i + 1]
            ex = sum(int(x[iex]) for x in seg); st = sum(int(x[ist]) for x in seg)
            tot = {}
            for x in seg:
                for c in stall_cols:
                    tot[hdr[c]] = tot.get(hdr[c], 0) + int(x[c])
            top = sorted(tot.items(), key=lambda kv: -kv[1])[:4]
            if ex / max(tot_ex, 1) > 0.005 or st / max(tot_st, 1) > 0.01:
                print(f"  instr {start:5d}-{i:5d}: exec {ex / tot_ex * 100:5.1f}% ({ex / 1e6:6.1f}M) stalls {st / tot_st * 100:5.1f}%  {top}")
            start = i + 1
    break
