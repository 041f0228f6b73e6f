"""Turn gpurun_out/<tag>_*.ncu-rep into the committed evidence: profiles/<tag>_ncu_summary.txt (key metrics per captured
kernel) and profiles/ncu_traffic.json (dram bytes per launch of the attention kernels, read by bench.py)."""
import csv, glob, json, os, subprocess, sys
tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
want = ['launch__grid_size', 'launch__block_size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed.sum']
def to_bytes(v, u):
    v = float(v.replace(',', ''))
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1)
out, traffic = [], {}
for rep in sorted(glob.glob(os.path.join(root, 'gpurun_out', tag + '_*.ncu-rep'))):
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    hdr, units = r[0], r[1]
    out.append('=' * 110 + '\n' + os.path.basename(rep))
    for row in r[2:]:
        name = row[hdr.index('Kernel Name')]
        out.append('-' * 110 + '\n  ' + name[:105])
        for w in want:
            if w in hdr:
                out.append(f'    {w:86s} {row[hdr.index(w)]:>14s} {units[hdr.index(w)]}')
        rd = to_bytes(row[hdr.index('dram__bytes_read.sum')], units[hdr.index('dram__bytes_read.sum')])
        wr = to_bytes(row[hdr.index('dram__bytes_write.sum')], units[hdr.index('dram__bytes_write.sum')])
        out.append(f'    dram read+write per launch: {(rd + wr) / 1e6:.1f} MB')
        if 'gemm_ps_kernel' in name and wr > traffic.get('_ps_wr', 0):      # the vocabulary projection writes the most (logits)
            traffic['_ps_wr'] = wr
            traffic['gemm_ps_logits'] = int(rd + wr)
        for key in ('attn_fwd', 'attn_bwd'):
            if key + '_stream' in name:
                traffic[key] = int(rd + wr)
open(os.path.join(root, 'profiles', tag + '_ncu_summary.txt'), 'w').write('\n'.join(out) + '\n')
traffic.pop('_ps_wr', None)
if traffic:
    json.dump(traffic, open(os.path.join(root, 'profiles', 'ncu_traffic.json'), 'w'))
print('\n'.join(out[:60]))
