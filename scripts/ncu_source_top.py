"""Top stall lines of one kernel from `ncu -i X --page source --csv --kernel-name ... > file`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = next(r for r in rows if r and r[0] == 'Address')
data = [r for r in rows if len(r) == len(hdr) and r[0].startswith('0x')]
i_src, i_s, i_ex = hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
tot = sum(int(r[i_s] or 0) for r in data)
print('total samples', tot, 'sass lines', len(data))
top = sorted(range(len(data)), key=lambda k: -int(data[k][i_s] or 0))[:n]
for k in sorted(top):
    r = data[k]
    print(f"{k:5d} {int(r[i_s]):6d} ({100*int(r[i_s])/max(tot,1):4.1f}%) ex={r[i_ex]:>8s} {r[i_src][:110]}")
