"""Per-kernel table from `ncu --csv --page raw` output of tools/ncu_step.py: launches, total time, DRAM bytes, achieved
GB/s, tensor-pipe activity.  usage: python tools/ncu_table.py raw.csv [hbm_peak_GBs]"""
import csv
import re
import sys
from collections import OrderedDict

import gzip
opener = gzip.open if sys.argv[1].endswith('.gz') else open
rows = [r for r in csv.reader(l for l in opener(sys.argv[1], 'rt') if not l.startswith('==') and l.strip() != 'ok')]
peak = float(sys.argv[2]) if len(sys.argv) > 2 else 6548.8
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
units = dict(zip(hdr, rows[1]))
def val(r, k):
    try:
        return float(r[col[k]].replace(',', ''))
    except Exception:
        return 0.0
def to_bytes(v, u):
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1)
def to_us(v, u):
    return v * {'ns': 1e-3, 'us': 1, 'ms': 1e3, 's': 1e6}.get(u, 1)
agg = OrderedDict()
for r in rows[2:]:
    name = r[col['Kernel Name']]
    name = re.sub(r'\(.*', '', name)
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0, 0.0])
    t = to_us(val(r, 'gpu__time_duration.sum'), units['gpu__time_duration.sum'])
    a[0] += 1
    a[1] += t
    a[2] += to_bytes(val(r, 'dram__bytes_read.sum'), units['dram__bytes_read.sum'])
    a[3] += to_bytes(val(r, 'dram__bytes_write.sum'), units['dram__bytes_write.sum'])
    a[4] += val(r, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed') * t
tot = sum(a[1] for a in agg.values())
print('%-62s %6s %10s %6s %9s %9s %8s %7s %8s' % ('kernel', 'calls', 'time us', 'share', 'rd MB', 'wr MB', 'GB/s', '%HBM', 'tensor%'))
for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    gbs = (a[2] + a[3]) / (a[1] * 1e-6) / 1e9 if a[1] > 0 else 0
    print('%-62s %6d %10.1f %5.1f%% %9.1f %9.1f %8.0f %6.1f%% %7.1f%%' % (name[:62], a[0], a[1], 100 * a[1] / tot, a[2] / 1e6, a[3] / 1e6, gbs,
                                                                 100 * gbs / peak, a[4] / a[1] if a[1] > 0 else 0))
print('total %.1f us over %d launches (serialised, cold-cache ncu replays: compare shares and per-kernel rates, not the sum)' % (tot, sum(a[0] for a in agg.values())))
if len(sys.argv) > 3:                       # tensor-family DRAM traffic for bench.py's roofline.traffic
    import json
    fam = [(n, a) for n, a in agg.items() if re.search(r'gemm_kernel|attn_kernel|wgrad_slab_kernel|attn_bwd', n)]
    launches = sum(a[0] for _, a in fam)
    total = sum(a[2] + a[3] for _, a in fam)
    json.dump({'source': sys.argv[4] if len(sys.argv) > 4 else sys.argv[1], 'launches': launches, 'dram_bytes_per_step': total,
               'dram_bytes_per_launch_mean': total / max(launches, 1), 'time_us': sum(a[1] for _, a in fam),
               'kernels': {n: {'launches': a[0], 'dram_bytes': a[2] + a[3], 'time_us': a[1]} for n, a in fam}},
              open(sys.argv[3], 'w'), indent=1)
