#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: hot SASS ranges by executed instructions."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
secs = []
for r in rows:
    if r and r[0] == 'Address':
        secs.append((r, []))
    elif secs and len(r) == len(secs[-1][0]):
        secs[-1][1].append(r)
H, R = secs[which]
ie = H.index('Instructions Executed'); ss = H.index('# Samples'); wf = H.index('L1 Wavefronts Shared'); te = H.index('Thread Instructions Executed')
tot = sum(int(r[ie]) for r in R); tots = sum(int(r[ss]) for r in R); totw = sum(int(r[wf]) for r in R)
print(f"sections={len(secs)} instrs={len(R)} warp-inst={tot} samples={tots} smem-wavefronts={totw}")
cur = None
def flush(end):
    if cur is not None and (acc > 0.004 * tot or accs > 0.004 * tots):
        print(f"{start:5d}-{end:5d} n={end-start+1:4d} exec/instr={curv:10d} thr={acct/max(acc,1):5.1f} inst={acc/tot:.3f} samples={accs/tots:.3f} wf={accw/max(totw,1):.3f}")
for i, r in enumerate(R):
    v = int(r[ie])
    if cur is None or abs(v - curv) > 0.2 * max(v, curv, 1):
        flush(i - 1)
        cur = 1; curv = v; acc = 0; accs = 0; start = i; accw = 0; acct = 0
    acc += v; accs += int(r[ss]); accw += int(r[wf]); acct += int(r[te])
flush(len(R) - 1)
if len(sys.argv) > 4:
    a, b = int(sys.argv[3]), int(sys.argv[4])
    for i in range(a, b + 1):
        r = R[i]
        print(i, r[1][:90], r[ie], r[ss], r[wf])
