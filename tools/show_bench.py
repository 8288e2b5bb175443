#!/usr/bin/env python3
"""pretty-print a bench.py JSON line: python tools/show_bench.py gpurun_out/x.json"""
import json, sys
d = json.load(open(sys.argv[1]))
print({k: v for k, v in d.items() if k not in ('stages', 'roofline', 'decode_cfg3', 'config', 'train_cfg1', 'cpu_baseline')})
if 'roofline' in d:
    r = d['roofline']; print('roofline frac', round(r['frac'], 4), r.get('kernels_ms'), 'share', r.get('share_of_step'))
for s in d.get('stages', []):
    print(f"{s['stage'][:58]:58s} ms={s['ms']:8.3f} frac={s['frac']:.3f} {s['bound']:9s} gbs={s['achieved_gbs']:6.0f} mhz={s['clocks']['sm_mhz']} n={s['clocks']['samples']} {','.join(s['clocks']['reasons'])}")
for k in ('train_cfg1', 'cfg5', 'decode_cfg3', 'strong_scaling', 'cpu_baseline'):
    if k in d: print(k, json.dumps(d[k])[:700])
