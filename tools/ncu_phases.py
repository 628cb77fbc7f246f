#!/usr/bin/env python3
"""Sums the per-SASS 'Instructions Executed' / stall samples of an `ncu --page source --csv` export between BAR.SYNCs (kernel phases)."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if 'Source' in r)
data = [r for r in rows if r and r[0].startswith('0x')]
iS, iI, iSamp = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
phase, tot, samp, per = 0, {}, {}, []
for r in data:
    n, s = int(r[iI]), int(r[iSamp])
    tot[phase] = tot.get(phase, 0) + n
    samp[phase] = samp.get(phase, 0) + s
    per.append((phase, r[iS].strip(), n, s))
    if 'BAR.SYNC' in r[iS]:
        phase += 1
print("instructions per phase", tot, "total", sum(tot.values()))
print("samples per phase", samp)
print("warps launched", int(data[0][iI]))
if len(sys.argv) > 2:
    for p, s, n, sm in per:
        print(p, n, sm, s)
