"""Summarise an ncu CSV (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum[,...]) per kernel name:
launches, total and mean duration, DRAM bytes per launch.  Usage: python scripts/ncu_traffic_summary.py file.csv [regex]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1], errors='replace')))
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
hdr = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
per = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) < 15:
        continue
    name = re.sub(r'\(anonymous namespace\)::|<unnamed>::', '', r[4]).split('(')[0].replace('void ', '')
    if (pat and not pat.search(name)) or 'spin_kernel' in name:
        continue
    per.setdefault((r[0], name), {})[r[12]] = float(r[14].replace(',', ''))
agg = collections.OrderedDict()
for (_, name), m in per.items():
    a = agg.setdefault(name, collections.Counter())
    a['launches'] += 1
    for k, v in m.items():
        a[k] += v
total_ns = sum(a.get('gpu__time_duration.sum', 0) for a in agg.values())
print('| kernel | launches | total ms | share | mean us | DRAM read MB/launch | DRAM write MB/launch |')
print('|---|---:|---:|---:|---:|---:|---:|')
for name, a in sorted(agg.items(), key=lambda kv: -kv[1].get('gpu__time_duration.sum', 0)):
    n = a['launches']
    t = a.get('gpu__time_duration.sum', 0)
    print(f"| `{name[:80]}` | {n} | {t / 1e6:.4f} | {100 * t / max(total_ns, 1):.1f}% | {t / n / 1e3:.1f} | "
          f"{a.get('dram__bytes_read.sum', 0) / n / 1e6:.2f} | {a.get('dram__bytes_write.sum', 0) / n / 1e6:.2f} |")
print(f'\ntotal {total_ns / 1e6:.4f} ms over {sum(a["launches"] for a in agg.values())} launches')
