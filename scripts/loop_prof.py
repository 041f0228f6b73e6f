"""Phase breakdown of the persistent time-loop kernels (ALEXCAP_LOOP_PROF=1): per-step %globaltimer stamps of CTA 0.
    ALEXCAP_LOOP_PROF=1 python scripts/loop_prof.py [cfg2]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault('ALEXCAP_LOOP_PROF', '1')
import bench  # noqa: E402

c = bench.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else 'cfg2']
dev = torch.device('cuda', 0)
run = bench.TrainRun(c, dev, 0, 1, None, graph=False)
for _ in range(3):
    run.step.run_resident()
torch.cuda.synchronize()
buf = (C.c_ulonglong * (2 * 64 * 8))()
n = run.lib.alexcap_time_loop_profile(buf, 2 * 64 * 8)
a = np.array(buf[:n], dtype=np.int64).reshape(2, 64, 8)
S = c['T'] + 1
names = {0: ['P1 gemm', 'sync', 'P2 attn', 'sync', 'P3 gemm+cell', 'sync'], 1: ['PB0 cell', 'sync', 'PB1 gemm', 'sync', 'PB2 attn', 'sync', 'PB3 gemm']}
for d, label in ((0, 'forward'), (1, 'backward')):
    x = a[d, :S]
    if not x.any():
        continue
    dt = np.diff(x, axis=1) / 1e3
    print(f'{label}: total {(x.max() - x[x > 0].min()) / 1e3:.1f} us over {S} steps; mean per-step phase durations (us):')
    for i in range(dt.shape[1]):
        col = dt[1:-1, i]
        nm = names[d][i] if i < len(names[d]) else f'seg{i}'
        print(f'   {nm:14s} {col.mean():7.2f}   (min {col.min():6.2f}, max {col.max():6.2f})')
    print('   step period   %.2f' % (np.diff(x[:, 0]).mean() / 1e3 * (1 if d == 0 else -1)))
