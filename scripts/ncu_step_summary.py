"""One training step out of an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py: the launches between
two consecutive kNN kernels (one per step), per-kernel count / device time / share.
Usage: python scripts/ncu_step_summary.py launches.csv [out.md] [step_index] [title note]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
hdr = rows[hi]
kn, mv = hdr.index('Kernel Name'), hdr.index('Metric Value')
L = []
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    name = r[kn].replace('(anonymous namespace)::', '').replace('<unnamed>::', '').replace('void ', '')
    L.append((re.sub(r'\(.*', '', name), float(r[mv].replace(',', ''))))
ks = [i for i, (n, _) in enumerate(L) if n.startswith('knn_kernel')]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 1
step = L[ks[which]:ks[which + 1]]
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for n, t in step:
    agg[n][0] += 1
    agg[n][1] += t
    tot += t
note = sys.argv[4] if len(sys.argv) > 4 else 'serialised under ncu: compare shares'
out = [f'One eager training step (configs[1], 32x1024 pts, k=20) between two knn launches: {len(step)} launches, '
       f'{tot / 1e6:.3f} ms device time ({note})', '', '| kernel | launches | ms | share |', '|---|---:|---:|---:|']
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f'| `{n[:110]}` | {c} | {t / 1e6:.3f} | {100 * t / tot:.1f}% |')
text = '\n'.join(out)
print(text)
if len(sys.argv) > 2 and sys.argv[2] != '-':
    open(sys.argv[2], 'w').write(text + '\n')
