"""Microbench of the fused attention step kernels at a BASELINE config shape (GPU box).
    python scripts/bench_attn.py [B L D H]      env: ALEXCAP_ATT_ENGINE / _CPS / _ROWS / _NST
Each of the S=17 launches per pass uses a different q/bpre slice over the same Ws/feat, as inside a training step."""
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
outs = [F_.attn_step_fwd(Ws, feat, q[t], bp[t], wv) for t in range(S)]


def timeit(fn, iters):
    """S launches captured into one CUDA graph (the python call costs ~28 us per launch, more than the kernel)."""
    for _ in range(S):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(S):
            fn()
    g.replay()
    torch.cuda.synchronize()
    reps = max(iters // S, 1)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        g.replay()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / (reps * S) * 1e3


st = {'t': 0}
def fwd():
    t = st['t'] = (st['t'] + 1) % S
    F_.attn_step_fwd(Ws, feat, q[t], bp[t], wv)
def bwd():
    t = st['t'] = (st['t'] + 1) % S
    a, c, b_, _ = outs[t]
    F_.attn_step_bwd(Ws, feat, q[t], wv, a, c, b_, dg[t])

bytes_ = B * L * (H + D) * 4
uf, ub = timeit(fwd, 6 * S), timeit(bwd, 6 * S)
cfg = {k: os.environ.get(k) for k in ('ALEXCAP_ATT_ENGINE', 'ALEXCAP_ATT_CPS', 'ALEXCAP_ATT_ROWS', 'ALEXCAP_ATT_NST')}
print(f'{cfg} fwd {uf:.1f} us {bytes_ / uf / 1e3:.0f} GB/s | bwd {ub:.1f} us {bytes_ / ub / 1e3:.0f} GB/s')
