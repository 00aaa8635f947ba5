#!/usr/bin/env python
"""Per-kernel table for the memory-bound kernels of one C3 step from tools/gpu_ncu_mem.sh's raw csv: launches, total ms,
DRAM bytes, achieved DRAM GB/s, and the ALGORITHMIC bytes of SURVEY.md §8(d) (C3: B = 64, L = 200, T = 600, fp32
activations, e = 4) over the duration = achieved algorithmic GB/s against the measured HBM peak (MEASURED_PEAKS.json).
Usage: ncu_mem_summary.py raw.csv [out.csv]"""
import collections, csv, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = list(csv.reader([l for l in open(sys.argv[1]) if not l.startswith("==")]))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
def get(r, name):
    i = col[name]; v = r[i].replace(',', ''); v = float(v) if v not in ('', 'n/a') else 0.0
    return v * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'us': 1e-3, 'ns': 1e-6, 'ms': 1.0, 'second': 1e3}.get(units[i], 1.0)
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6452.2
B, L, T, e = 64, 200, 600, 4
# algorithmic bytes per LAUNCH (SURVEY.md §8(d) formulas); N = rows the launch processes, read from the grid where it varies
def algo(name, grid):
    g = [int(x) for x in re.findall(r"\d+", grid)]
    if "layernorm" in name or "k_ln_" in name:
        n = g[0] * 8                                   # 8 rows per block
        return 2 * n * 256 * e
    if "bucket_add" in name: return B * L * (4 + 3 * 256 * e)
    if "length_regulate" in name: return B * L * (4 + 4 + 256 * e) + B * T * 256 * e + B * T
    if "k_embed" in name: return B * L * (16 + 256 * e) + (429 + 5 + L) * 256 * 4
    if "conv_post" in name: return B * 256 * T * (32 * 2 + 2)      # bf16 planar in (e = 2), int16 out
    if "gmm_sample" in name: return 32 * (16 + 1024) + 7695 * 4
    if "add_pos" in name: return 2 * B * T * 256 * e
    if "add_rowvec" in name: return 2 * B * L * 256 * e
    if "rowdot" in name: return B * L * (256 * e + 4)
    return None
agg = collections.OrderedDict()
for r in data:
    name = re.sub(r'\(.*', '', r[col['Kernel Name']]).replace('void <unnamed>::', '').replace('void ', '')
    a = agg.setdefault(name, dict(n=0, ms=0.0, dram=0.0, l2=0.0, algo=0.0, pct=0.0, known=True))
    ms = get(r, 'gpu__time_duration.sum')
    a['n'] += 1; a['ms'] += ms; a['dram'] += get(r, 'dram__bytes_read.sum') + get(r, 'dram__bytes_write.sum')
    a['l2'] += get(r, 'lts__t_bytes.sum'); a['pct'] += get(r, 'dram__throughput.avg.pct_of_peak_sustained_elapsed') * ms
    ab = algo(name, r[col['Grid Size']])
    if ab is None: a['known'] = False
    else: a['algo'] += ab
out = [("kernel", "launches", "ms_total", "dram_MB", "dram_GBps", "dram_pct_time_weighted", "l2_MB", "algorithmic_MB", "algorithmic_GBps", "frac_of_hbm_peak_%.0f" % peak)]
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]['ms']):
    ag = a['algo'] / a['ms'] / 1e6 if a['known'] and a['ms'] else 0.0
    out.append((k, a['n'], "%.4f" % a['ms'], "%.2f" % (a['dram'] / 1e6), "%.0f" % (a['dram'] / a['ms'] / 1e6), "%.1f" % (a['pct'] / a['ms']),
                "%.2f" % (a['l2'] / 1e6), "%.2f" % (a['algo'] / 1e6) if a['known'] else "", "%.0f" % ag if a['known'] else "", "%.3f" % (ag / peak) if a['known'] else ""))
w = csv.writer(open(sys.argv[2], 'w') if len(sys.argv) > 2 else sys.stdout)
w.writerows(out)
for o in out: print("  ".join(str(x) for x in o))
