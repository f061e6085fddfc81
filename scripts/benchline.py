"""One-line digest of a bench.py JSON line read from stdin: python scripts/benchline.py TAG"""
import json
import sys

d = json.loads(sys.stdin.read().strip().splitlines()[-1])
c = d.get('clocks', {})
print(' '.join(sys.argv[1:]), '| value', round(d['value'], 1), 'ms/step', round(d['ms_per_step'], 4), '| e2e',
      round(d['e2e']['value'], 1), '|', c.get('sm_mhz'), 'MHz', c.get('power_w'), 'W', c.get('reasons'))
