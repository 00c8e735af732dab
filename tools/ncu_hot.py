"""Top stall lines of an `ncu --page source --csv` dump (SASS view): samples, executed count, main stall reason."""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
data = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    try:
        s = int(r[ix['# Samples']] or 0)
    except ValueError:
        continue
    data.append((s, r))
tot = sum(s for s, _ in data)
print('total samples', tot)
agg = {h: 0 for h in stalls}
for s, r in data:
    for h in stalls:
        try:
            agg[h] += int(r[ix[h]] or 0)
        except ValueError:
            pass
print('stall totals:', ', '.join(f'{h[6:]}={v}' for h, v in sorted(agg.items(), key=lambda t: -t[1])[:8]))
for s, r in sorted(data, key=lambda t: -t[0])[:top]:
    best = max(stalls, key=lambda h: int(r[ix[h]] or 0))
    print(f'{s:6d} {100 * s / tot:5.1f}%  exec={r[ix["Instructions Executed"]]:>9s}  {best[6:]:14s} {r[ix["Source"]][:90]}')
