#!/usr/bin/env python
"""Print the kernels of the last complete training step from an `ncu --csv` per-launch duration log."""
import csv, sys
for path in sys.argv[1:]:
    with open(path) as f:
        lines = [l for l in f if not l.startswith('==')]
    lines = [lines[0]] + [l for l in lines[1:] if not l.startswith('"ID"')]
    rows = []
    for row in csv.DictReader(lines):
        if row.get('Metric Name') == 'gpu__time_duration.sum':
            name = row['Kernel Name'].replace('<unnamed>::', '').replace('void ', '')
            rows.append((name[:64], float(row['Metric Value'].replace(',', '')), row['Grid Size'], row['Block Size']))
    idx = [i for i, x in enumerate(rows) if x[0].startswith('conv12') or x[0].startswith('conv1_fwd') or (x[0].startswith('conv_fwd_tc_kernel') and ', 1>' in x[0])]
    print(path, 'rows', len(rows))
    if len(idx) < 2:
        continue
    a, b = idx[-2], idx[-1]
    tot = 0
    for x in rows[a:b]:
        print(f'  {x[0]:64s} {x[1]/1000:8.2f} us  {x[2]} {x[3]}')
        tot += x[1]
    print(f'  sum {tot/1000:.1f} us')
