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
]
def t(fn, it=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(it): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / it * 1e3
for name, ta, tb, M, N, K in shapes:
    A = torch.randn((K, M) if ta else (M, K), device=dev)
    B = torch.randn((N, K) if tb else (K, N), device=dev)
    C = torch.empty(M, N, device=dev)
    row = []
    engines = [0, 1] + ([2, 4, 8, 16, 32] if M == 128 or 'dW_W' in name else [])
    for eng in engines:
        us = t(lambda: F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=eng))
        row.append(f'e{eng}:{us:7.1f}us')
    tf = 2.0 * M * N * K / 1e12
    best = min(float(r.split(':')[1][:-2]) for r in row[1:])
    print(f'{name} M={M} N={N} K={K}  ' + ' '.join(row) + f'   best tc {tf / (best * 1e-6):6.1f} TF/s')
