#!/usr/bin/env python
"""Condense an ncu source-page export (SASS + PC samples) into basic-block rows.
usage: ncu -i X.ncu-rep --page source --csv --print-source sass > src.csv; python scripts/ncu_blocks.py src.csv [min_samples]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 40
first, seen, hdr = [], set(), None
for r in rows:
    if r and r[0] == "Address": hdr = r; continue
    if hdr is None or len(r) < 10 or r[0] == "Kernel Name": continue
    if r[0] in seen: break
    seen.add(r[0]); first.append(r)
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot_s = sum(int(r[ix['# Samples']]) for r in first); tot_i = sum(int(r[ix['Instructions Executed']]) for r in first)
print("instructions", len(first), "samples", tot_s, "warp-inst", tot_i)
blk, cur = [], None
for n, r in enumerate(first):
    e = int(r[ix['Instructions Executed']]); s = int(r[ix['# Samples']])
    if cur and cur['e'] == e: cur['n'] += 1; cur['s'] += s; cur['end'] = n
    else: cur = {'e': e, 'n': 1, 's': s, 'start': n, 'end': n}; blk.append(cur)
for b in blk:
    if b['s'] >= thr or b['e'] * b['n'] > tot_i * 0.01:
        d = {h: sum(int(first[n][ix[h]] or 0) for n in range(b['start'], b['end'] + 1)) for h in stalls}
        top = sorted(d.items(), key=lambda x: -x[1])[:4]
        print("%5d-%5d n=%4d exec=%8d inst=%4.1f%% samples=%5d (%4.1f%%) %-34s %s" % (
            b['start'], b['end'], b['n'], b['e'], 100 * b['e'] * b['n'] / tot_i, b['s'], 100 * b['s'] / tot_s,
            first[b['start']][1].strip()[:34], ' '.join("%s=%d" % (k[6:], v) for k, v in top)))
