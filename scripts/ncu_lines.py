#!/usr/bin/env python
"""Executed instructions / stall samples per SOURCE LINE of one kernel:
   python scripts/ncu_lines.py <source-page csv from ncu> <nvdisasm -g output> <kernel substring> [top]"""
import collections, csv, re, sys
src_csv, dis, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
cur = fn = None
off2line = {}
for l in open(dis):
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\.text\.(\S+):', l)
    if m: fn = m.group(1); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/', l)
    if m and fn and kname in fn: off2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[2:]:
    if r and r[0] == 'Kernel Name': break
    if len(r) == len(hdr): data.append(r)
base = int(data[0][0], 16)
agg = collections.Counter(); smp = collections.Counter()
for r in data:
    ln = off2line.get(int(r[0], 16) - base)
    agg[ln] += int(r[ix['Instructions Executed']] or 0); smp[ln] += int(r[ix['# Samples']] or 0)
tot = sum(agg.values()); ts = sum(smp.values())
print('total warp-instructions', tot, 'samples', ts)
for ln, c in agg.most_common(top): print('%-34s %9d %6.3f   samples %6d %6.3f' % (ln, c, c / tot, smp[ln], smp[ln] / max(1, ts)))
