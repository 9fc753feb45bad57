#!/usr/bin/env python
"""Summarise the per-instruction stall samples of an `ncu --page source --csv` dump:
top SASS instructions by samples with their dominant stall reason, and totals per reason."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
reasons = [h for h in hdr if h.startswith('stall_') and '(' not in h]
tot = {r: 0 for r in reasons}
items = []
for k, r in enumerate(rows[2:]):
    if len(r) < len(hdr): continue
    n = int(r[ix['# Samples']] or 0)
    st = {q: int(r[ix[q]] or 0) for q in reasons}
    for q in reasons: tot[q] += st[q]
    items.append((n, k, r[ix['Source']].strip(), st))
all_n = sum(i[0] for i in items)
print('total samples', all_n)
for q, v in sorted(tot.items(), key=lambda x: -x[1]):
    if v: print(f'  {q:24s} {v:8d} {100.0*v/all_n:5.1f} %')
print('top instructions (index in the listing, samples, share, dominant reasons, SASS):')
for n, k, src, st in sorted(items, key=lambda x: -x[0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    dom = ', '.join(f'{q[6:]} {v}' for q, v in sorted(st.items(), key=lambda x: -x[1])[:2] if v)
    print(f'  #{k:5d} {n:6d} {100.0*n/all_n:5.1f} %  [{dom}]  {src[:90]}')
