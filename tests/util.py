import glob
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def golden_names(prefix):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, prefix + '*.npz')))


def load_golden(name, dtype=torch.float64):
    z = np.load(os.path.join(GOLDEN_DIR, name + '.npz'))
    out = {'params': {}, 'grads': {}}
    for k in z.files:
        v = torch.from_numpy(z[k]) if z[k].ndim else z[k].item()
        if isinstance(v, torch.Tensor) and v.dtype.is_floating_point:
            v = v.to(dtype)
        if k.startswith('p:'):
            out['params'][k[2:]] = v
        elif k.startswith('g:'):
            out['grads'][k[2:]] = v
        else:
            out[k] = v
    B, L, D, H, E, V, T = [int(x) for x in out['dims']]
    out['cfg'] = dict(B=B, L=L, D=D, H=H, E=E, V=V, T=T)
    return out


def rel_err(a, b):
    """max |a-b| / max(|b|_inf, tiny): relative to the tensor's scale (what '1e-4 relative' means for a
    gradient tensor whose individual entries can be ~0)."""
    a = torch.as_tensor(a, dtype=torch.float64).cpu()
    b = torch.as_tensor(b, dtype=torch.float64).cpu()
    scale = max(float(b.abs().max()), 1e-30)
    return float((a - b).abs().max()) / scale
