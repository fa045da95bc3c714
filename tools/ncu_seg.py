#!/usr/bin/env python
"""Per-role / per-segment sample breakdown of an `ncu --page source --csv` dump (segments are
delimited by mbarrier waits / arrives and tensor-core commits)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]; d = rows[2:]
isrc = h.index('Source'); ins = h.index('# Samples'); iex = h.index('Instructions Executed')
tot = sum(int(r[ins] or 0) for r in d)
marks = [i for i, r in enumerate(d) if 'USETMAXREG' in r[isrc]] + [len(d)]
prev = 0
for m in marks:
    rs = sum(int(r[ins] or 0) for r in d[prev:m])
    if rs > 0.02 * tot:
        print(f"== region [{prev},{m}) {100*rs/tot:.1f}% of samples")
        segs = [i for i in range(prev, m) if any(k in d[i][isrc] for k in ('TRYWAIT', 'SYNCS.ARRIVE', 'UTCBAR'))]
        p2 = prev
        for s in segs + [m]:
            t = sum(int(r[ins] or 0) for r in d[p2:s])
            w = 0
            if s < m and 'TRYWAIT' in d[s][isrc]:
                w = int(d[s][ins] or 0) + int(d[s + 1][ins] or 0)
            if t + w > 0.003 * tot:
                print(f"   [{p2:5d},{s:5d}) {s-p2:4d} instrs  busy {100*t/rs:5.1f}%  then wait {100*w/rs:5.1f}%  {d[s][isrc].strip()[:60] if s < m else ''}")
            p2 = s + (2 if w else 0)
    prev = m
