"""Aggregate an `ncu --page source --csv --print-source sass` export: stall samples by opcode and the top instructions.
Usage: python tools/sass_stalls.py file.csv [ntop]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]; data = [r for r in rows[2:] if len(r) == len(hdr)]
iS = hdr.index('Source'); iA = hdr.index('Warp Stall Sampling (All Samples)'); iE = hdr.index('Instructions Executed')
f = lambda s: int(s) if s.strip().isdigit() else 0
tot = sum(f(r[iA]) for r in data)
print('total samples', tot, 'instrs', len(data))
byop = collections.Counter(); exe = collections.Counter()
for r in data:
    t = r[iS].split()
    if not t: continue
    op = t[1] if t[0].startswith('@') and len(t) > 1 else t[0]
    op = '.'.join(op.split('.')[:2])
    byop[op] += f(r[iA]); exe[op] += f(r[iE])
for op, c in byop.most_common(ntop): print(f'{op:22s} samples {c:8d} {100*c/max(tot,1):5.1f}%  executed {exe[op]}')
print('--- top instructions')
for r in sorted(data, key=lambda r: -f(r[iA]))[:ntop]: print(r[iA], r[iE], r[iS][:100])
