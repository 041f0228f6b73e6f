"""Phase time stamps of CTA 0 of the streaming attention kernels (GPU box): where one image's ~18 us go.
    python scripts/attn_prof.py [cfg]
Uses the debug hook ALEXCAP_ATT_PROF_PTR (attention_stream.cu): %globaltimer at kernel entry, after the PDL wait / body
entry, header staged, pass 1 done, softmax done, pass 2 done, partials reduced, finalised."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

prof = torch.zeros(1024, dtype=torch.int64, device='cuda')
os.environ['ALEXCAP_ATT_PROF_PTR'] = hex(prof.data_ptr())
prof_tc = torch.zeros(16, dtype=torch.int64, device='cuda')
os.environ['ALEXCAP_TC_PROF_PTR'] = hex(prof_tc.data_ptr())
import bench

c = bench.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else 'cfg2']
run = bench.TrainRun(c, torch.device('cuda', 0), 0, 1, None, graph=True)
ms, _ = run.time_resident(10, 3)
torch.cuda.synchronize()
p = prof.cpu().tolist()
names_f = ['kernel entry', 'body entry (after PDL wait)', 'header staged', 'pass 1 (scores) done', 'softmax + alpha written',
           'pass 2 (context) done', 'finalised']
names_b = ['kernel entry', 'body entry (after PDL wait)', 'header staged (A)', 'pass B (d alpha over feat) done',
           'dot / de written (C)', 'pass D (score backward over Ws) done', 'finalised']
print(f'{c["name"] if "name" in c else sys.argv[1:]}: step {ms:.4f} ms; stamps of the LAST forward (t = S-1) and LAST backward (t = 0) launch, CTA 0, us since kernel entry')
for title, base, names in (('attn_fwd_stream_kernel', 0, names_f), ('attn_bwd_stream_kernel', 8, names_b)):
    print(title)
    t0 = p[base]
    prev = t0
    for i, n in enumerate(names):
        t = p[base + i]
        print(f'   {n:42s} {(t - t0) / 1e3:7.2f}   (+{(t - prev) / 1e3:5.2f})')
        prev = t

# timeline of the last training step: the S forward launches, then the S backward launches
S = c['T'] + 1 if 'T' in c else 17
for title, base, cnt in (('forward loop', 16, p[7]), ('backward loop', 16 + 256, p[15])):
    rows = []
    for j in range(S):
        n = (cnt - S + j) % 64
        e = p[base + 4 * n: base + 4 * n + 3]
        rows.append(e)
    print(title, '(us): launch | waited for predecessor | body | gap end -> next body entry | period')
    for j, e in enumerate(rows):
        gap = (rows[j + 1][1] - e[2]) / 1e3 if j + 1 < S else float('nan')
        per = (rows[j + 1][1] - e[1]) / 1e3 if j + 1 < S else float('nan')
        print(f'   {j:2d}   {(e[1] - e[0]) / 1e3:6.2f}   {(e[2] - e[1]) / 1e3:6.2f}   {gap:6.2f}   {per:6.2f}')
    print(f'   loop total {(rows[-1][2] - rows[0][1]) / 1e3:.1f} us')

q = prof_tc.cpu().tolist()
if q[0]:
    print('fused LSTM cell -> query GEMM kernel (last launch, CTA (0,0,0), us since kernel entry):')
    for i, n in enumerate(['entry', 'after griddepcontrol.wait', 'gate loads issued + weight tile landed', 'cell math + A tile written',
                           'accumulator complete', 'epilogue done']):
        print(f'   {n:42s} {(q[i] - q[0]) / 1e3:7.2f}')
