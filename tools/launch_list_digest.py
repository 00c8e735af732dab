"""ncu launch list (`--metrics gpu__time_duration.sum --csv --log-file x.csv`, tools/gpu_launchlist.sh) -> the per-kernel
digest kept next to it in profiles/.   usage: python tools/launch_list_digest.py x.csv forwards > x.txt"""
import collections
import csv
import re
import sys

path, forwards = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 2
rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"')) if len(r) > 14]
hdr, rows = rows[0], rows[1:]
ix = {h: i for i, h in enumerate(hdr)}
acc = collections.OrderedDict()
for r in rows:
    name = re.sub(r'^void ', '', r[ix['Kernel Name']])
    name = re.sub(r'\(.*$', '', name)
    key = (name, r[ix['Grid Size']], r[ix['Block Size']])
    acc.setdefault(key, []).append(float(r[ix['Metric Value']]) / 1e3)
total = sum(sum(v) for v in acc.values()) / forwards
print(f'# {path}: ncu --metrics gpu__time_duration.sum --clock-control none over {forwards} eager forwards')
print('# (bench.py --launch-list --steps 2: the kernels of the timed step\'s schedule; cold caches, serialised: compare shares).')
print('# kernel, grid, block, launches per forward, mean us, us per forward, share')
for (name, grid, block), v in sorted(acc.items(), key=lambda kv: -sum(kv[1])):
    per = sum(v) / forwards
    print(f'{name:48s} grid {grid:18s} block {block:13s} x {len(v) / forwards:3.0f}  {sum(v) / len(v):8.1f} us  {per:8.1f} us  {100 * per / total:5.1f} %')
print(f'total per forward: {total:.1f} us in {len(rows) / forwards:.0f} launches')
