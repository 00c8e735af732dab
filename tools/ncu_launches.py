"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, time and share."""
import collections
import csv
import re
import sys

path = sys.argv[1]
forwards = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
with open(path) as f:
    lines = [ln for ln in f if not ln.startswith('==')]
rows = []
for x in csv.DictReader(lines):
    if x.get('Metric Name') == 'gpu__time_duration.sum':
        rows.append((x['Kernel Name'], float(x['Metric Value'].replace(',', ''))))
agg = collections.defaultdict(lambda: [0, 0.0])
for name, v in rows:
    name = re.sub(r'<.*', '', name)[:64]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print(f'{len(rows)} launches, {tot / 1e3:.1f} us total; per forward ({forwards:g} forwards): '
      f'{len(rows) / forwards:.1f} launches, {tot / forwards / 1e3:.1f} us')
for name, (c, v) in sorted(agg.items(), key=lambda t: -t[1][1])[:30]:
    print(f'{name:66s} {c / forwards:7.1f}/fwd {v / forwards / 1e3:9.1f} us/fwd {100 * v / tot:5.1f}%')
