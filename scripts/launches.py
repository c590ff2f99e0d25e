"""Summarise an ncu --metrics gpu__time_duration.sum CSV: per-kernel time of the last train step."""
import csv, sys
path = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/launches.csv'
marker = sys.argv[2] if len(sys.argv) > 2 else 'batch_window'
with open(path) as f:
    lines = [l for l in f if not l.startswith('==')]
rows = list(csv.DictReader(lines))
names = [x['Kernel Name'] for x in rows]
idx = [i for i, n in enumerate(names) if marker in n]
last = rows[idx[-1]:] if idx else rows
tot = 0
for x in last:
    v = float(x['Metric Value'].replace(',', ''))
    us = v / 1000 if x['Metric Unit'] in ('ns', 'nsecond') else v
    tot += us
    print(f"{us:9.1f} us  {x['Kernel Name'][:72]:72s} grid={x['Grid Size']} block={x['Block Size']}")
print(f"{tot:9.1f} us  total ({len(last)} launches)")
