#!/usr/bin/env python
"""One line per bench JSON of a job directory: times, clocks, and -- loudly -- whether the run's own checks passed."""
import json, os, sys, glob
d = sys.argv[1]
if os.path.isfile(d + '/status.txt'):
    print(open(d + '/status.txt').read().strip())
for f in sorted(glob.glob(d + '/*.json')):
    try:
        txt = open(f).read()
        line = [l for l in txt.splitlines() if l.startswith('{')][-1]
        j = json.loads(line)
        print(f"{f.split('/')[-1]:28s} {j['ms_per_step']*1000:7.1f} us flushed {j['steady_state']['ms_per_step']*1000:7.1f} steady  e2e {j['e2e']['ms_per_step']*1000:8.1f}  {j.get('impl_config', {}).get('comm')} {j['clocks']['reasons']} value {j['value']:.4g}"
              + ("" if j.get('checks', {}).get('ok', True) else f"   *** CHECKS FAILED: golden {j['checks']['golden_losses'].get('status')} "
                 f"(max err {j['checks']['golden_losses'].get('max_abs_err')}), drift {j['checks'].get('replica_drift')} ***"))
    except Exception as e:
        print(f, 'ERR', e)
