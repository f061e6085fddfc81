"""Top stalled SASS instructions of an `ncu --page source --csv --print-source sass` export.
Usage: ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv; python scripts/ncu_hot.py src.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
h = rows[hdr]
body = [r for r in rows[hdr + 1:] if len(r) == len(h)]
ci = {k: h.index(k) for k in h}
stalls = [k for k in h if k.startswith('stall_') and 'Not Issued' not in k]
tot = sum(int(r[ci['# Samples']] or 0) for r in body)
print('total samples', tot, 'instructions', len(body))
agg = {k: sum(int(r[ci[k]] or 0) for r in body) for k in stalls}
print('by reason:', {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
order = sorted(range(len(body)), key=lambda i: -int(body[i][ci['# Samples']] or 0))[:n]
for i in sorted(order):
    r = body[i]
    top = sorted(((int(r[ci[k]] or 0), k[6:]) for k in stalls), reverse=True)[:3]
    print(f"{i:5d} {int(r[ci['# Samples']]):6d} {100 * int(r[ci['# Samples']]) / tot:5.1f}%  {r[ci['Source']][:70]:70s} " +
          ' '.join(f'{k}={v}' for v, k in top if v))
