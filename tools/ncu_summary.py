#!/usr/bin/env python
"""Per-launch table from an `ncu --set full ... --page raw --csv` export: duration, DRAM bytes, DRAM and tensor-pipe
utilisation, shared-memory wavefronts.  Usage: ncu_summary.py prof.csv [out.csv]"""
import csv, re, sys
rows = list(csv.reader([l for l in open(sys.argv[1]) if not l.startswith('==')]))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
def scale(v, u):
    v = float(v.replace(',', '')) if v not in ('', 'n/a') else 0.0
    m = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1, 'us': 1e-3, 'ns': 1e-6, 'ms': 1.0, 'second': 1e3}
    return v * m.get(u, 1.0)
def get(r, name):
    i = col[name]; return scale(r[i], units[i])
out = [("idx", "kernel", "grid", "ms", "dram_rd_MB", "dram_wr_MB", "dram_GBps", "dram_pct", "tensor_pct_elapsed", "smem_KB", "regs")]
tot = [0.0, 0.0, 0.0]
for k, r in enumerate(data):
    name = re.sub(r'\(.*', '', r[col['Kernel Name']]).replace('void <unnamed>::', '')
    ms = get(r, 'gpu__time_duration.sum'); rd = get(r, 'dram__bytes_read.sum'); wr = get(r, 'dram__bytes_write.sum')
    out.append((k, name, r[col['Grid Size']].replace(' ', ''), "%.4f" % ms, "%.1f" % (rd / 1e6), "%.1f" % (wr / 1e6), "%.0f" % ((rd + wr) / ms / 1e6),
                "%.1f" % get(r, 'dram__throughput.avg.pct_of_peak_sustained_elapsed') if 'dram__throughput.avg.pct_of_peak_sustained_elapsed' in col else "",
                "%.1f" % get(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed'),
                "%.1f" % (get(r, 'launch__shared_mem_per_block_dynamic') / 1e3), r[col['launch__registers_per_thread']]))
    tot[0] += ms; tot[1] += rd; tot[2] += wr
w = csv.writer(open(sys.argv[2], 'w') if len(sys.argv) > 2 else sys.stdout)
w.writerows(out)
print("# %d launches, %.3f ms, DRAM read %.2f GB, written %.2f GB" % (len(data), tot[0], tot[1] / 1e9, tot[2] / 1e9), file=sys.stderr)
