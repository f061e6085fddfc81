#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/bench_conv.py $(cat ${SPECS:-scripts/pair_sweep_specs.txt}) > gpurun_out/sweep.txt 2> gpurun_out/sweep.err; echo "rc=$?"; tail -3 gpurun_out/sweep.err
python - <<'PY'
rows = [l.split() for l in open('gpurun_out/sweep.txt') if ': ' in l]
by = {}
for r in rows:
    spec = r[0].rstrip(':').split(',')
    by.setdefault(','.join(spec[:5]), {})[spec[-1]] = float(r[1])
ts = tp = tb = 0
for k, v in by.items():
    a, b = v.get('-1'), v.get('1')
    if a is None or b is None: continue
    ts += a; tp += b; tb += min(a, b)
    print(f'{k:22s} single {a:.4f}  pair {b:.4f}  ratio {b / a:.2f}')
print(f'sum single {ts:.3f} pair {tp:.3f} best-of {tb:.3f}')
PY
