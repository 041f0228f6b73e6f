"""Micro-benchmark of the PER-STEP GEMMs of the bf16 configurations (cfg3/cfg5: M = B = 512/256) through the pre-split
kernel with the operand copies resident (engine 600 + forced split-K), L2-warm (10 launches per graph) and L2-cold
(a 512 MB write between launches, as inside a step where the attention kernels stream 400 MB between two GEMMs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagecaptioning_b200 import functional as F_
dev = 'cuda'
M = int(sys.argv[1]) if len(sys.argv) > 1 else 512
shapes = [('QB    NT', False, True, M, 3072, 1024, True, False), ('gates NT', False, True, M, 4096, 3072, False, True),
          ('dgctx NN', False, False, M, 2048, 4096, False, True), ('dh    NN', False, False, M, 1024, 7168, False, True)]
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

def warm(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10): fn()
    g.replay(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record(); g.replay(); g.replay(); e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / 20 * 1e3

def cold(fn, n=12):
    ts = []
    for _ in range(n):
        flush.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]

for name, ta, tb, m, N, K, has_bias, acc in shapes:
    A = torch.randn((K, m) if ta else (m, K), device=dev)
    B = torch.randn((N, K) if tb else (K, N), device=dev)
    C = torch.zeros(m, N, device=dev)
    bias = torch.randn(N, device=dev) if has_bias else None
    F_.gemm(A, B, transa=ta, transb=tb, out=C, engine=400)          # makes the copies
    row = []
    for s in (0, 1, 2, 3, 4, 6, 8):
        fn = lambda: F_.gemm(A, B, transa=ta, transb=tb, out=C, bias=bias, accumulate=acc, engine=600 + s)
        row.append(f's{s}: {warm(fn):5.1f}/{cold(fn):5.1f}')
    print(f'{name} M={m} N={N} K={K} GF={2e-9 * m * N * K:5.1f}  warm/cold us  ' + '  '.join(row), flush=True)
