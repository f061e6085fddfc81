"""Summarise an ncu launch list (scripts/gpu_launch_list.sh) into per-kernel totals for ONE bench step and the
DRAM traffic of the conv kernels (bench.py's roofline.traffic).  Usage: summarise_launches.py launches.csv out.json"""
import csv
import json
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith('=='))]
hdr, data = rows[0], rows[1:]
ik, iv, iid, im, iu = (hdr.index(k) for k in ('Kernel Name', 'Metric Value', 'ID', 'Metric Name', 'Metric Unit'))
per = defaultdict(dict)
names = {}
for r in data:
    v = float(r[iv].replace(',', ''))
    u = r[iu]
    scale = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1.0)
    per[int(r[iid])][r[im]] = v * scale
    names[int(r[iid])] = r[ik].split('(')[0].replace('void ', '').replace('yb::', '').replace('<unnamed>::', '')
ids = sorted(per)
last = max(i for i in ids if 'sort_nms' in names[i])
prev = max(i for i in ids if 'sort_nms' in names[i] and i < last)
step = [i for i in ids if prev < i <= last]
agg = defaultdict(lambda: [0, 0.0, 0.0])
for i in step:
    a = agg[names[i]]
    a[0] += 1
    a[1] += per[i].get('gpu__time_duration.sum', 0.0)
    a[2] += per[i].get('dram__bytes_read.sum', 0.0) + per[i].get('dram__bytes_write.sum', 0.0)
tot = sum(a[1] for a in agg.values())
out = {'launches_per_step': len(step), 'sum_kernel_us': tot, 'note': 'ncu per-launch times are cold-cache and serialised',
       'kernels': [{'kernel': k, 'launches': a[0], 'us': round(a[1], 1), 'share': round(a[1] / tot, 4),
                    'dram_bytes': int(a[2])} for k, a in sorted(agg.items(), key=lambda x: -x[1][1])]}
conv = [k for k in agg if k.startswith('k_conv') or k.startswith('k_stem')]
out['conv_dram_bytes_per_step'] = int(sum(agg[k][2] for k in conv))
out['conv_share_of_step'] = round(sum(agg[k][1] for k in conv) / tot, 4)
json.dump(out, open(sys.argv[2], 'w'), indent=1)
for k in out['kernels']:
    print(f"{k['kernel']:40s} {k['launches']:4d} {k['us']:9.1f} us {100 * k['share']:5.1f}%  {k['dram_bytes'] / 1e6:9.1f} MB")
print('step kernels', len(step), 'sum', round(tot, 1), 'us; conv DRAM bytes', out['conv_dram_bytes_per_step'] / 1e9, 'GB')
