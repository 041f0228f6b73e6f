"""Pre-split GEMM kernel alone on the cfg2 shapes with plain-store and split-K epilogues (GPU box): python scripts/bench_ps_small.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagecaptioning_b200 import functional as F_
dev = 'cuda'
shapes = [('cfg2 Ws     NT', False, True, 25088, 512, 512), ('cfg2 logits NT', False, True, 2176, 10004, 512),
          ('cfg2 dHout  NN', False, False, 2176, 512, 10004), ('cfg2 dW_o   TN', True, False, 10004, 512, 2176),
          ('cfg2 dWcat  TN', True, False, 2048, 1024, 2176), ('cfg3 logits NT', False, True, 8704, 10004, 1024)]
def t(fn):
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
for name, ta, tb, M, N, K in shapes:
    A = torch.randn((K, M) if ta else (M, K), device=dev)
    B = torch.randn((N, K) if tb else (K, N), device=dev)
    C = torch.empty(M, N, device=dev)
    F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=300)
    us = t(lambda: F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=500))
    F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=400)
    us1 = t(lambda: F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=600))
    print(f'{os.environ.get("TAG", "")} {name} M={M} N={N} K={K}  e500 {us:8.1f} us  {2.0 * M * N * K / us / 1e6:7.1f} TF/s   e600 {us1:8.1f} us {2.0 * M * N * K / us1 / 1e6:7.1f} TF/s', flush=True)
