#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: samples per barrier wait / per role region."""
import csv, sys, re
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]; data = rows[2:]
isrc = h.index('Source'); ins = h.index('# Samples'); iex = h.index('Instructions Executed')
tot = sum(int(r[ins] or 0) for r in data)
print("total samples", tot, "instructions", len(data))
# role regions are delimited by USETMAXREG
marks = [i for i, r in enumerate(data) if 'USETMAXREG' in r[isrc]] + [len(data)]
prev = 0
for m in marks:
    s = sum(int(r[ins] or 0) for r in data[prev:m])
    print(f"region [{prev:5d},{m:5d}) samples={s:7d} {100*s/tot:5.1f}%  starts: {data[prev][isrc][:50]}")
    prev = m
print("--- waits (SYNCS TRYWAIT + following BRA) ---")
for i, r in enumerate(data):
    if 'TRYWAIT' in r[isrc]:
        s = int(r[ins] or 0) + int(data[i + 1][ins] or 0)
        if s > 0.002 * tot:
            m = re.search(r'\[(.*?)\]', r[isrc])
            print(f"{i:5d} {s:7d} {100*s/tot:5.1f}% ex={r[iex]:>9s} {m.group(1) if m else ''}")
print("--- other hot instructions ---")
for i, r in sorted(enumerate(data), key=lambda t: -int(t[1][ins] or 0))[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print(f"{i:5d} {int(r[ins]):7d} {100*int(r[ins])/tot:5.1f}% {r[isrc][:80]}")
