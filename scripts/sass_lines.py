#!/usr/bin/env python
"""Attribute ncu's per-SASS-instruction stall samples to CUDA source lines.

ncu's CSV export of the source page has the samples per SASS instruction but no line numbers; nvdisasm -g has the
line of every instruction.  Both list a kernel's instructions in address order, so they are joined by position.

    scripts/sass_lines.py <cubin> <kernel-substring> <ncu --page source --csv (sass)> [top]
"""
import csv, re, subprocess, sys

cubin, kname, src_csv = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dis = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout.splitlines()
# find the function section for the kernel
lines_of = []
cur_line, in_fn = None, False
for l in dis:
    m = re.match(r'\s*\.section\s+\.text\.(\S+)', l)
    if m:
        in_fn = kname in m.group(1)
        continue
    if not in_fn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur_line = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        lines_of.append((int(m.group(1), 16), cur_line, m.group(2).strip()))
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
si, ai, so = hdr.index('# Samples'), hdr.index('Address'), hdr.index('Source')
body = [r for r in rows[2:] if len(r) > si and r[ai].startswith('0x')]
base = int(body[0][ai], 16)
samples = {}
for r in body:
    samples[int(r[ai], 16) - base] = (int(r[si] or 0), r[so].strip())
by_line, tot = {}, 0
for off, line, text in lines_of:
    n = samples.get(off, (0, ''))[0]
    tot += n
    by_line[line] = by_line.get(line, 0) + n
print(f'{kname}: {len(lines_of)} instructions, {tot} samples')
src_cache = {}
def src(line):
    if not line: return ''
    f, n = line
    if f not in src_cache:
        try: src_cache[f] = open('/root/repo/csrc/' + f).read().splitlines()
        except Exception: src_cache[f] = []
    t = src_cache[f]
    return t[n - 1].strip()[:100] if 0 < n <= len(t) else ''
for line, n in sorted(by_line.items(), key=lambda kv: -kv[1])[:top]:
    print(f'{n:6d} {100*n/max(tot,1):5.1f}%  {line}  {src(line)}')
