"""Hot spots of one kernel in an .ncu-rep: SASS lines with the most stall samples / executed instructions.
    python scripts/ncu_hot.py file.ncu-rep [topN]"""
import csv, subprocess, sys, io
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'source', '--csv'], capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO('\n'.join(lines[start:]))))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
def num(x):
    try: return float(x.replace(',', ''))
    except Exception: return 0.0
tot_s = sum(num(r['# Samples']) for r in rows); tot_i = sum(num(r['Instructions Executed']) for r in rows)
print(f'total samples {tot_s:.0f}, warp instructions executed {tot_i:.0f}, SASS lines {len(rows)}')
stall_cols = [c for c in rows[0] if c.startswith('stall_') and 'Not Issued' not in c]
agg = {c: sum(num(r[c]) for r in rows) for c in stall_cols}
print('stall mix:', ', '.join(f'{c[6:]} {100 * v / max(tot_s, 1):.0f}%' for c, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
for i, r in enumerate(rows): r['_i'] = i
for r in sorted(rows, key=lambda r: -num(r['# Samples']))[:n]:
    top = max(stall_cols, key=lambda c: num(r[c]))
    print(f"{r['_i']:5d} samp {100 * num(r['# Samples']) / max(tot_s, 1):5.1f}%  exec {num(r['Instructions Executed']):9.0f}  {top[6:]:12s} {r['Source'][:90]}")
