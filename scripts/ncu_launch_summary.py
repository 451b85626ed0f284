"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total
device time and share.  Usage: python scripts/ncu_launch_summary.py launches.csv [out.md]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
hdr = rows[hi]
kn, mv = hdr.index('Kernel Name'), hdr.index('Metric Value')
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    name = r[kn].replace('(anonymous namespace)::', '').replace('void ', '')
    name = re.sub(r'\(.*', '', name)
    t = float(r[mv].replace(',', ''))
    agg[name][0] += 1
    agg[name][1] += t
    tot += t
lines = [f'launches {sum(v[0] for v in agg.values())}, total device time {tot / 1e6:.3f} ms (cold-cache, serialised under ncu: compare shares)', '',
         '| kernel | launches | ms | share |', '|---|---:|---:|---:|']
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    lines.append(f'| `{n[:110]}` | {c} | {t / 1e6:.3f} | {100 * t / tot:.1f}% |')
text = '\n'.join(lines)
print(text)
if len(sys.argv) > 2:
    open(sys.argv[2], 'w').write(text + '\n')
