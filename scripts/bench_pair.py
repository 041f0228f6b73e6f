"""Pre-split GEMM kernel alone (operand copies kept: engines 500 = bf16x3, 600 = bf16) on the step's big shapes.
Run twice: ALEXCAP_PS_PAIR=0 / =1 (single-CTA 128 x BN tiles vs CTA-pair 256 x 256 cta_group::2 tiles)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagecaptioning_b200 import functional as F_
dev = 'cuda'
shapes = [
    ('cfg2 Ws     NT', False, True, 25088, 512, 512), ('cfg2 logits NT', False, True, 2176, 10004, 512),
    ('cfg2 dHout  NN', False, False, 2176, 512, 10004), ('cfg2 dW_o   TN', True, False, 10004, 512, 2176),
    ('cfg2 dW_W   TN', True, False, 512, 512, 25088), ('cfg2 dWcat  TN', True, False, 2048, 1024, 2176),
    ('cfg3 Ws     NT', False, True, 25088, 1024, 2048), ('cfg3 logits NT', False, True, 8704, 10004, 1024),
    ('cfg5 logits NT', False, True, 8448, 32004, 1024), ('8k^3        NT', False, True, 8192, 8192, 8192),
]
def t(fn, it=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(5): fn()
    g.replay(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(6): g.replay()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / 30 * 1e3
print('ALEXCAP_PS_PAIR =', os.environ.get('ALEXCAP_PS_PAIR', '(default)'))
for name, ta, tb, M, N, K in shapes:
    A = torch.randn((K, M) if ta else (M, K), device=dev)
    B = torch.randn((N, K) if tb else (K, N), device=dev)
    C = torch.empty(M, N, device=dev)
    out = []
    for eng, mma in ((500, 3), (600, 1)):
        F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=eng - 200)
        us = t(lambda: F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=eng))
        tf = 2.0 * M * N * K / (us * 1e-6) / 1e12
        out.append(f'e{eng}: {us:8.1f} us {tf:7.1f} TF/s (issued {mma * tf:7.1f})')
    print(f'{name} M={M} N={N} K={K}   ' + '   '.join(out), flush=True)
