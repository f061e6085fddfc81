"""Per-kernel counts of the Blackwell-native SASS mnemonics in libyolo_b200.so (cuobjdump -sass):
UTC*MMA (tcgen05.mma), LDTM/STTM (tcgen05.ld/st), UTMALDG/UTMASTG/UBLKCP (TMA), UTCBAR (tcgen05.commit), HMMA (legacy
mma.sync).  Usage: python scripts/sass_summary.py > profiles/sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, 'yolo_series_b200', 'libyolo_b200.so')
txt = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True).stdout
pat = re.compile(r'\b(UTC[A-Z]*MMA(?:\.2CTA)?|LDTM|STTM|UTMALDG(?:\.\dD)?(?:\.2CTA)?|UTMASTG|UBLKCP|UTCBAR(?:\.2CTA)?(?:\.MULTICAST)?|HMMA|UTCATOMSWS|MUFU\.TANH)')
per = collections.OrderedDict()
cur, arch = None, set()
for line in txt.splitlines():
    m = re.search(r'Function : (\S+)', line)
    if m:
        name = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.replace('(anonymous namespace)::', '').replace('yb::', '')
        name = re.sub(r'\((?!int\)|bool\)).*$', '', name)
        cur = per.setdefault(name, collections.Counter())
        continue
    m = re.search(r'arch = (sm_\w+)', line)
    if m:
        arch.add(m.group(1))
    if cur is not None:
        for k in pat.findall(line):
            cur[k] += 1
print('libyolo_b200.so -- cubin architectures:', ', '.join(sorted(arch)))
cols = ['UTCHMMA', 'UTCHMMA.2CTA', 'LDTM', 'UTMALDG', 'UTMALDG.2CTA', 'UTMASTG', 'UBLKCP', 'UTCBAR', 'UTCBAR.2CTA', 'HMMA', 'MUFU.TANH']
print(f"{'kernel':58s}" + ''.join(f'{c:>13s}' for c in cols))
tot = collections.Counter()
for name, c in per.items():
    agg = collections.Counter()
    for k, v in c.items():
        if k.startswith('UTMALDG'):
            agg['UTMALDG.2CTA' if k.endswith('2CTA') else 'UTMALDG'] += v
        elif k.startswith('UTCBAR'):
            agg['UTCBAR.2CTA' if '2CTA' in k else 'UTCBAR'] += v
        else:
            agg[k] += v
    if not any(agg[k] for k in cols):
        continue
    tot.update(agg)
    print(f'{name[:57]:58s}' + ''.join(f'{agg[k]:13d}' for k in cols))
print(f"{'TOTAL':58s}" + ''.join(f'{tot[k]:13d}' for k in cols))
