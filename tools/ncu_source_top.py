"""Top stall-sampled SASS lines of the first kernel in an ncu source-page CSV.
usage: ncu -i rep --page source --csv > x.csv ; python tools/ncu_source_top.py x.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hs = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hi = hs[0]
end = hs[1] - 1 if len(hs) > 1 else len(rows)
hdr = rows[hi]
data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
ia, isamp, iex = hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
def num(x):
    try: return int(x)
    except Exception: return 0
tot = sum(num(r[isamp]) for r in data)
print('total samples', tot, 'lines', len(data))
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
for r in sorted(data, key=lambda r: -num(r[isamp]))[:n]:
    st = sorted(((num(r[i]), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(r[isamp].rjust(7), r[iex].rjust(9), r[0][-5:], r[ia][:90].ljust(90), st)
