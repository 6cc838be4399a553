#!/usr/bin/env python
"""Prints the event timeline of the large-head forward's MMA warp and epilogue warp 0 (CTA 0, tiles 2-3) from the
`[big trace]` lines a run with ADLM_BIG_EXP=64 writes to stderr.
    ADLM_BIG_EXP=64 python tools/cfg5_bench.py --reps 1 --hw 512x1024 2> trace.log; python tools/big_trace.py trace.log
Events: 10/11 waiting for / got the A tile; 1bk own prototype group k of block b present, 2bk both CTAs' present,
3bk MMA1 group issued; 4b0 similarities of block b ready; 5bw last-layer group present; 6bw MMA2 group issued;
7b0 (epilogue) D1 of block b complete; 8b0 similarities written; 900 logits of the previous tile stored."""
import sys
lines = [l.split() for l in open(sys.argv[1]) if "big trace" in l]
ent = [(l[-3], int(l[-2]), int(l[-1])) for l in lines]
dumps, cur = [], []
for e in ent:
    if e[0] == "mma" and e[1] == 10 and e[2] == 0 and cur:
        dumps.append(cur)
        cur = []
    cur.append(e)
dumps.append(cur)
d = sorted(dumps[-1], key=lambda e: e[2])
limit = int(sys.argv[2]) if len(sys.argv) > 2 else 70000
prev = {"mma": None, "epi": None}
for role, ev, t in d:
    if t > limit:
        break
    dt = t - prev[role] if prev[role] is not None else 0
    prev[role] = t
    print(f"{t:8d} {role} {ev:4d} (+{dt})")
