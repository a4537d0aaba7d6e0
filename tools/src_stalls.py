"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export: warp-stall samples per CUDA source line.
Usage: python tools/src_stalls.py file.csv [ntop]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
fname = None; agg = collections.Counter(); src = {}; hdr = None; cur = None
for r in rows:
    if r and r[0] == 'File Path': fname = r[1].split('/')[-1]; continue
    if r and r[0] == 'Line No': hdr = r; iA = hdr.index('Warp Stall Sampling (All Samples)'); continue
    if r and r[0] == 'Function Name': continue
    if hdr is None or len(r) < iA + 1: continue
    if r[0].strip(): cur = (fname, r[0]); src[cur] = r[1][:110]
    try: n = int(r[iA])
    except ValueError: n = 0
    if r[2].strip(): agg[cur] += n
tot = sum(agg.values()); print('total samples', tot)
for k, v in agg.most_common(ntop): print(f'{v:6d} {100*v/tot:5.1f}%  {k[0]}:{k[1]}  {src[k]}')
