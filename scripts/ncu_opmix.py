"""Executed warp-instructions by opcode from an .ncu-rep source page.  usage: python scripts/ncu_opmix.py x.ncu-rep [ntop]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 30
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[hi]; ci = {h: i for i, h in enumerate(hdr)}
ex = ci.get('# Warp Instructions Executed', ci.get('Instructions Executed'))
data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
mix = collections.Counter(); tot = 0
for r in data:
    try: n = int(r[ex])
    except ValueError: continue
    s = r[ci['Source']].split()
    op = s[1] if s and s[0].startswith('@') and len(s) > 1 else (s[0] if s else '?')
    op = op.split('.')[0]
    mix[op] += n; tot += n
print('total warp instructions', tot, ' static', len(data))
for op, n in mix.most_common(ntop):
    print(f"  {op:12s} {n:12d} {100 * n / tot:6.2f}%")
