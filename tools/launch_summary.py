#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: one step of bench.py (the launches between the
last two k_conv_post), grouped by kernel and, for k_conv_tc, by position in the graph."""
import csv, re, sys, collections, signal
signal.signal(signal.SIGPIPE, signal.SIG_DFL)
path = sys.argv[1]
with open(path) as f:
    rows = list(csv.DictReader([l for l in f if not l.startswith('==')]))
def ms(x):
    v = float(x['Metric Value'].replace(',', '')); u = x['Metric Unit']
    return v / 1e6 if u == 'ns' else v / 1e3 if u == 'us' else v
names = [re.sub(r'\(.*', '', x['Kernel Name']).replace('void <unnamed>::', '').replace('<unnamed>::', '').replace('void ', '') for x in rows]
posts = [i for i, n in enumerate(names) if 'k_conv_post' in n]
start, end = (posts[-2] + 1, posts[-1]) if len(posts) > 1 else (0, len(rows) - 1)   # a --profile-range capture holds exactly one step
agg = collections.OrderedDict(); tot = 0.0
for i in range(start, end + 1):
    a = agg.setdefault(names[i], [0, 0.0]); a[0] += 1; a[1] += ms(rows[i]); tot += ms(rows[i])
print("one step: %d launches, %.3f ms (ncu-serialised)" % (end - start + 1, tot))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]): print("  %-45s %4d  %8.3f ms  %5.1f%%" % (k[:45], v[0], v[1], 100 * v[1] / tot))
if len(sys.argv) > 2:
    for i in range(start, end + 1): print(i - start, names[i][:40], rows[i]['Grid Size'], "%.3f" % ms(rows[i]))
