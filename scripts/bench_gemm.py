"""Micro-benchmark of the GEMM engines on the training step's shapes (CUDA events, L2-warm like inside a step)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagecaptioning_b200 import functional as F_
dev = 'cuda'
shapes = [  # name, ta, tb, M, N, K
    ('QB    NT', False, True, 128, 1024, 512), ('gates NT', False, True, 128, 2048, 1024),
    ('dgctx NN', False, False, 128, 512, 2048), ('dh    NN', False, False, 128, 512, 3072),
    ('Ws    NT', False, True, 25088, 512, 512), ('logits NT', False, True, 2176, 10004, 512),
    ('dHout NN', False, False, 2176, 512, 10004), ('dW_o  TN', True, False, 10004, 512, 2176),
    ('dW_W  TN', True, False, 512, 512, 25088), ('dWcat TN', True, False, 2048, 1024, 2176),
    ('cfg3 Ws NT', False, True, 25088, 1024, 2048), ('cfg3 gates NT', False, True, 512, 4096, 3072),
    ('cfg3 logits NT', False, True, 8704, 10004, 1024), ('8k^3 NT', False, True, 8192, 8192, 8192),
]
def t(fn, it=30):
    """10 launches per CUDA graph (the python call costs more than a small GEMM), graph replayed and event-timed."""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10): fn()
    g.replay(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(max(it // 10, 1)): g.replay()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / (max(it // 10, 1) * 10) * 1e3
for name, ta, tb, M, N, K in ([] if '--small' in sys.argv else shapes):
    A = torch.randn((K, M) if ta else (M, K), device=dev)
    B = torch.randn((N, K) if tb else (K, N), device=dev)
    C = torch.empty(M, N, device=dev)
    row = []
    engines = [100, 200, 300, 400, 500, 600]
    for eng in engines:
        us = t(lambda: F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=eng))
        row.append(f'e{eng}:{us:7.1f}us')
    tf = 2.0 * M * N * K / 1e12
    us = [float(r.split(':')[1][:-2]) for r in row]
    print(f'{name} M={M} N={N} K={K}  ' + ' '.join(row) + '   TF/s: ' + ' '.join(f'e{e}:{tf / (u * 1e-6):6.1f}' for e, u in zip(engines, us)))

print('--- per-step GEMMs as they run inside the step: accumulate into a pre-zeroed C (no memset node), PDL-chained launches')
for name, ta, tb, M, N, K in shapes[:4]:
    A = torch.randn((K, M) if ta else (M, K), device=dev)
    B = torch.randn((N, K) if tb else (K, N), device=dev)
    C = torch.zeros(M, N, device=dev)
    row = []
    for eng in (1, 100, 200, 500):
        if eng == 500:
            F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=300)
        us = t(lambda: F_.gemm(A, B, transa=ta, transb=tb, out=C, accumulate=True, engine=eng))
        row.append(f'e{eng}:{us:6.2f}us')
    print(f'{name} M={M} N={N} K={K}  ' + ' '.join(row))
