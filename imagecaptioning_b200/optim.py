"""FusedClipAdam -- the step right after the hot path: clip_grad_norm_ + Adam.step() as one native call.

Reference usage being mirrored (train_LSTMwAttention.py:58-76, :87-88, :109-110):
    optim = Adam(model.parameters(), lr, (beta1, beta2), eps, weight_decay=wd);  scheduler = LambdaLR(optim, lr_lambda)
    loss = model.forward_train(data);  model.clip_gradient(1);  optim.step();  scheduler.step()
FusedClipAdam is a torch.optim.Optimizer (so the reference's LambdaLR drives `param_groups[i]['lr']` unchanged) whose
step() does the global-norm clip (when max_norm is given) and the Adam update of every tensor through
alexcap_clip_adam_step: 3 kernel launches instead of ~10 per parameter tensor.  No CPU path.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib
from . import functional as F_


def warmup_cosine_lambda(warmup_steps: int, max_iter: int, min_lr_ratio: float):
    """The reference's lr_lambda (train_LSTMwAttention.py:66-72): linear warm-up, then cosine decay floored at
    min_lr / learning_rate."""
    def lr_lambda(current_step: int):
        if current_step < warmup_steps:
            return float(current_step) / max(1, warmup_steps)
        cosine_decay = 0.5 * (1.0 + math.cos(math.pi * (current_step - warmup_steps) / (max_iter - warmup_steps)))
        return max(min_lr_ratio, cosine_decay)
    return lr_lambda


class FusedClipAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_norm=None):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1 and 0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError('invalid Adam hyper-parameter')
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self.max_norm = max_norm            # None: no clipping (model.clip_gradient not called)
        self.last_grad_norm = None          # device scalar: total norm before clipping (what clip_grad_norm_ returns)
        self._scratch = None

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        """`grads`: optional list of gradient tensors aligned with the parameters of ALL groups in order (e.g.
        DecoderTrainStep's flat-buffer slices) used instead of `p.grad`."""
        loss = closure() if closure is not None else None
        if self.max_norm is not None and len(self.param_groups) > 1:
            # clip_grad_norm_ is over ALL parameters: one norm for every group
            raise _lib.AlexcapError('FusedClipAdam with max_norm supports a single parameter group')
        gi = 0
        for group in self.param_groups:
            ps, gs, ms, vs = [], [], [], []
            for p in group['params']:
                g = grads[gi] if grads is not None else p.grad
                gi += 1
                if g is None:
                    continue
                if not p.is_cuda:
                    raise _lib.AlexcapError('imagecaptioning_b200 has no CPU path: parameters must live on a CUDA device')
                if p.dtype != torch.float32 or g.dtype != torch.float32 or not p.is_contiguous() or not g.is_contiguous():
                    raise _lib.AlexcapError('FusedClipAdam needs contiguous float32 parameters and gradients')
                st = self.state[p]
                if not st:
                    st['step'] = 0
                    st['exp_avg'] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st['step'] += 1
                ps.append(p); gs.append(g); ms.append(st['exp_avg']); vs.append(st['exp_avg_sq'])
            if not ps:
                continue
            step = self.state[ps[0]]['step']
            dev = ps[0].device
            lib = F_.ensure_device(dev)
            if self._scratch is None or self._scratch.device != dev:
                self._scratch = torch.empty(int(lib.alexcap_clip_adam_scratch_floats()), dtype=torch.float32, device=dev)
                self.last_grad_norm = torch.zeros(1, dtype=torch.float32, device=dev)
            n = len(ps)
            cap = _lib.OPT_MAX_TENSORS
            if self.max_norm is not None and n > cap:
                raise _lib.AlexcapError(f'more than {cap} parameter tensors with clipping')
            for lo in range(0, n, cap):
                hi = min(n, lo + cap)
                arr = lambda ts: (C.c_void_p * (hi - lo))(*[t.data_ptr() for t in ts[lo:hi]])
                numel = (C.c_int64 * (hi - lo))(*[t.numel() for t in ps[lo:hi]])
                _lib.check(lib.alexcap_clip_adam_step(hi - lo, arr(ps), arr(gs), arr(ms), arr(vs), numel,
                                                      float(self.max_norm) if self.max_norm is not None else 0.0,
                                                      float(group['lr']), float(group['betas'][0]), float(group['betas'][1]),
                                                      float(group['eps']), float(group['weight_decay']), int(step),
                                                      self.last_grad_norm.data_ptr(), self._scratch.data_ptr(),
                                                      F_._stream()), 'alexcap_clip_adam_step')
        return loss
