#!/bin/bash
out=gpurun_out
MASG_PROBE_ONE=1 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 4000 --csv --log-file $out/launches_c2b.csv python tools/latency_probe.py > $out/ncu_c2b.log 2>&1
python - <<'PY'
import csv, re, collections
rows = list(csv.DictReader([l for l in open('gpurun_out/launches_c2b.csv') if not l.startswith('==')]))
agg = collections.OrderedDict(); tot = 0
for i, r in enumerate(rows):
    n = re.sub(r'\(.*', '', r['Kernel Name']).replace('void <unnamed>::', '').replace('void ', ''); v = float(r['Metric Value'].replace(',', '')) / 1e3
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += v; tot += v
    print(i, n[:34], r['Grid Size'], round(v, 1))
print("C2 one call: %d launches, kernel sum %.1f us" % (len(rows), tot))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]: print("  %-40s %3d %8.1f us  (%.1f us each)" % (k[:40], v[0], v[1], v[1] / v[0]))
PY
