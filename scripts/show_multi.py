#!/usr/bin/env python
import json, sys, glob
d = sys.argv[1]
print(open(d + '/status.txt').read().strip())
for f in sorted(glob.glob(d + '/*.json')):
    try:
        txt = open(f).read()
        line = [l for l in txt.splitlines() if l.startswith('{')][-1]
        j = json.loads(line)
        print(f"{f.split('/')[-1]:28s} {j['ms_per_step']*1000:7.1f} us flushed {j['steady_state']['ms_per_step']*1000:7.1f} steady  e2e {j['e2e']['ms_per_step']*1000:8.1f}  {j['config'].get('comm')} {j['clocks']['reasons']} value {j['value']:.4g}")
    except Exception as e:
        print(f, 'ERR', e)
