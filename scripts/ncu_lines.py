#!/usr/bin/env python
"""Summarise an `ncu --page source --print-source cuda,sass --csv` export per CUDA source line.
usage: ncu_lines.py export.csv [kernel-regex] [top]"""
import csv, re, sys
path = sys.argv[1]; pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
rows = csv.reader(open(path)); fname = func = None; hdr = None; lines = {}; seen = set(); active = False
for r in rows:
    if not r: continue
    if r[0] == 'File Path': fname = r[1]; continue
    if r[0] == 'Function Name':
        func = r[1]; key = func
        active = (pat is None or pat.search(func)) and (func not in seen or active_func == func)
        if active: active_func = func; seen.add(func)
        continue
    if r[0] == 'Line No': hdr = r; continue
    if not active or hdr is None or len(r) < len(hdr): continue
    if r[0] == '' : continue   # sass row
    try: s = float(r[hdr.index('# Samples')])
    except ValueError: continue
    k = (fname.split('/')[-1], int(r[0]))
    e = lines.setdefault(k, {'src': r[1], 's': 0.0, 'st': {}, 'inst': 0.0})
    e['s'] += s
    try: e['inst'] += float(r[hdr.index('Instructions Executed')])
    except ValueError: pass
    for i, h in enumerate(hdr):
        if h.startswith('stall_') and 'Not' not in h:
            try: e['st'][h] = e['st'].get(h, 0) + float(r[i])
            except ValueError: pass
tot = sum(e['s'] for e in lines.values()) or 1; toti = sum(e['inst'] for e in lines.values()) or 1
print('total samples', tot, 'warp instructions', toti)
agg = {}
for e in lines.values():
    for h, v in e['st'].items(): agg[h] = agg.get(h, 0) + v
print({h: round(100 * v / tot, 1) for h, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
for k, e in sorted(lines.items(), key=lambda kv: -kv[1]['s'])[:top]:
    st = sorted(e['st'].items(), key=lambda kv: -kv[1])[:3]
    print(f"{100*e['s']/tot:5.1f}% inst {100*e['inst']/toti:4.1f}% {k[0]}:{k[1]:<4d} {e['src'].strip()[:90]:90s} {[(h[6:], int(v)) for h, v in st if v]}")
