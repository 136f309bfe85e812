"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel."""
import csv, collections, io, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
r = csv.DictReader(io.StringIO(''.join(lines)))
agg = collections.defaultdict(lambda: [0, 0.0]); tot = 0
for row in r:
    if row.get('Metric Name') != 'gpu__time_duration.sum': continue
    name = row['Kernel Name'].split('(')[0].replace('void ', '').replace('lz77sss::', '').replace('<unnamed>::', '')
    v = float(row['Metric Value'].replace(',', '')); u = row['Metric Unit']
    if u == 'ns': v /= 1e3
    elif u == 'ms': v *= 1e3
    agg[name][0] += 1; agg[name][1] += v; tot += v
print('kernel,launches,total_us,share')
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 20]:
    print(f'{k},{c},{t:.1f},{t / tot:.4f}')
print('total_us', tot)
