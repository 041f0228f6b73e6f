"""Experiment (GPU box): does the fused attention step run faster per image when fewer images are in flight, and do two
half-batch launches on two streams overlap?  Decides whether pipelining two half-batch recurrences can hide the per-step GEMM
latency.    python scripts/bench_attn_dual.py [B L D H]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imagecaptioning_b200 import functional as F_

B, L, D, H = [int(x) for x in sys.argv[1:5]] if len(sys.argv) >= 5 else (128, 196, 512, 512)
S = 17
dev = 'cuda'
g = torch.Generator().manual_seed(1)
Ws = torch.randn(B, L, H, generator=g).to(dev)
feat = torch.randn(B, L, D, generator=g).relu().to(dev)
q = torch.randn(S, B, H, generator=g).to(dev)
bp = torch.randn(S, B, D, generator=g).to(dev)
wv = (torch.randn(H, generator=g) / H ** 0.5).to(dev)
dg = torch.randn(S, B, D, generator=g).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def graph_time(body, reps=6):
    body()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        body()
    gr.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        gr.replay()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e) * 1e3 / S)
    return min(ts), sum(ts) / len(ts)


def run_range(lo, hi, bwd):
    outs = []
    for t in range(S):
        if bwd is None:
            F_.attn_step_fwd(Ws[lo:hi], feat[lo:hi], q[t, lo:hi], bp[t, lo:hi], wv)
        else:
            a, c, b_, _ = bwd[t]
            F_.attn_step_bwd(Ws[lo:hi], feat[lo:hi], q[t, lo:hi], wv, a[lo:hi], c[lo:hi], b_[lo:hi], dg[t, lo:hi])


outs = [F_.attn_step_fwd(Ws, feat, q[t], bp[t], wv) for t in range(S)]
side = torch.cuda.Stream()


def dual(bwd):
    def body():
        cur = torch.cuda.current_stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            run_range(B // 2, B, bwd)
        run_range(0, B // 2, bwd)
        cur.wait_stream(side)
    return body


for name, bwd in (('fwd', None), ('bwd', outs)):
    full = graph_time(lambda: run_range(0, B, bwd))
    half = graph_time(lambda: run_range(0, B // 2, bwd))
    quarter = graph_time(lambda: run_range(0, B // 4, bwd))
    two = graph_time(dual(bwd))
    print(f'{name}: per launch us (min, mean)  B={B}: {full[0]:.2f} {full[1]:.2f} | B/2 alone: {half[0]:.2f} {half[1]:.2f} | '
          f'B/4 alone: {quarter[0]:.2f} {quarter[1]:.2f} | two B/2 launches on two streams, per pair: {two[0]:.2f} {two[1]:.2f}')
