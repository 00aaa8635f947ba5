#!/usr/bin/env python
"""Hot regions of an `ncu --page source --csv` export (SASS view): executed warp-instructions and stall samples per
block of consecutive SASS lines, with the dominant opcodes.  Usage: sass_hot.py src.csv [block=60] [section=0]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
blk = int(sys.argv[2]) if len(sys.argv) > 2 else 60
want = int(sys.argv[3]) if len(sys.argv) > 3 else 0
secs, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "rows": []}; secs.append(cur); continue
    if r and r[0] == "Address": cur["hdr"] = r; continue
    if cur is not None and len(r) > 10: cur["rows"].append(r)
s = secs[want]; hdr, data = s["hdr"], s["rows"]
ie, ss = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
tot = sum(int(r[ie]) for r in data); tots = sum(int(r[ss]) for r in data)
print(s["name"][:60], "| warp-instructions", tot, "| samples", tots, "| SASS lines", len(data), "| sections", len(secs))
for b in range(0, len(data), blk):
    c = sum(int(r[ie]) for r in data[b:b + blk]); sm = sum(int(r[ss]) for r in data[b:b + blk])
    if c > tot * 0.01 or sm > tots * 0.01:
        ops = {}
        for r in data[b:b + blk]:
            t = r[1].split(); op = t[1] if t[0].startswith('@') else t[0]
            ops[op] = ops.get(op, 0) + int(r[ie])
        top = sorted(ops.items(), key=lambda kv: -kv[1])[:7]
        print("%5d-%5d inst %5.1f%% samples %5.1f%%  %s" % (b, b + blk, 100 * c / tot, 100 * sm / tots, ' '.join("%s:%.1f" % (k, 100 * v / tot) for k, v in top)))
