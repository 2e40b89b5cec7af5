"""Summarise one head step out of an ncu launch-list CSV (gpu__time_duration.sum): per-kernel times in launch order."""
import csv, sys
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith('==')]
for row in csv.DictReader(lines):
    if row.get('Metric Name') == 'gpu__time_duration.sum':
        rows.append((int(row['ID']), row['Kernel Name'], float(row['Metric Value'].replace(',', '')), row['Grid Size'], row.get('Stream', '')))
idx = [i for i, r in enumerate(rows) if 'pixel_fused' in r[1]]
step = rows[idx[0] + 1:idx[1] + 1]
tot = sum(r[2] for r in step)
print('launches per step', len(step), 'sum us', round(tot / 1000, 1))
main = sum(r[2] for r in step if r[4] == step[-1][4])
print('main-stream sum us', round(main / 1000, 1))
short = '-v' not in sys.argv
agg = {}
for r in step:
    name = r[1].split('(')[0].split('::')[-1]
    key = (name, r[3], r[4])
    a = agg.setdefault(key, [0, 0.0]); a[0] += 1; a[1] += r[2]
for (name, grid, st), (n, t) in agg.items():
    print(f'{t/1000:8.1f} us  x{n:<3d} avg {t/n/1000:6.1f}  {name:28s} {grid:14s} stream {st}')
