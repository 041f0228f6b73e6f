"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and shares."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
tot = collections.defaultdict(lambda: [0, 0.0])
rows = [r for r in csv.DictReader(lines) if r.get('Metric Name') == 'gpu__time_duration.sum']
if '--one-step' in sys.argv:
    # one training step = the launches from one make_target_kernel (first kernel of a step) up to the next one
    starts = [i for i, r in enumerate(rows) if 'make_target_kernel' in r['Kernel Name']]
    rows = rows[starts[-2]:starts[-1]]
for row in rows:
    name = re.sub(r'\(.*', '', row['Kernel Name'])
    v = float(row['Metric Value'].replace(',', ''))
    u = row['Metric Unit']
    v = v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
    tot[name][0] += 1; tot[name][1] += v
T = sum(v for _, v in tot.values())
print(f'total {T:.1f} us over {sum(n for n, _ in tot.values())} launches')
for k, (n, v) in sorted(tot.items(), key=lambda x: -x[1][1]):
    print(f'{v:10.1f} us {100 * v / T:5.1f}%  n={n:3d}  avg={v / n:8.1f}  {k[:110]}')
