"""-m gpu: FusedClipAdam (alexcap_clip_adam_step) against the reference's own optimiser stack run on the same device:
torch.nn.utils.clip_grad_norm_ (AlexCap/LSTMModel.py:50-51) + torch.optim.Adam (train_LSTMwAttention.py:60) +
LambdaLR warm-up/cosine (train_LSTMwAttention.py:66-76).  fp32 elementwise arithmetic: tolerance 2e-6 relative."""
import pytest
import torch

from tests.gpu_util import DEV, build_decoder, random_attn_params
from tests.util import rel_err

pytestmark = pytest.mark.gpu


def _clone_params(shapes, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(*s, generator=g).to(DEV)) for s in shapes]


@pytest.mark.parametrize('max_norm', [None, 1.0, 1e9])
@pytest.mark.parametrize('wd', [0.0, 1e-6, 0.1])
def test_fused_clip_adam_matches_torch(max_norm, wd):
    from imagecaptioning_b200.optim import FusedClipAdam, warmup_cosine_lambda
    shapes = [(33,), (7, 19), (1,), (300, 257), (5, 4, 3)]
    ours, ref = _clone_params(shapes, 1), _clone_params(shapes, 1)
    kw = dict(lr=3e-3, betas=(0.8, 0.95), eps=1e-8, weight_decay=wd)
    oa, ob = FusedClipAdam(ours, max_norm=max_norm, **kw), torch.optim.Adam(ref, **kw)
    lam = warmup_cosine_lambda(3, 12, 0.1)
    sa, sb = torch.optim.lr_scheduler.LambdaLR(oa, lam), torch.optim.lr_scheduler.LambdaLR(ob, lam)
    g = torch.Generator().manual_seed(2)
    for it in range(8):
        grads = [torch.randn(*s, generator=g).to(DEV) * (10.0 if it % 2 else 0.01) for s in shapes]
        for p, q, gr in zip(ours, ref, grads):
            p.grad, q.grad = gr.clone(), gr.clone()
        if max_norm is not None:
            want_norm = torch.nn.utils.clip_grad_norm_(ref, max_norm)
        oa.step(); ob.step(); sa.step(); sb.step()
        if max_norm is not None:
            assert rel_err(oa.last_grad_norm, want_norm) < 2e-6
            for p, q in zip(ours, ref):
                assert rel_err(p.grad, q.grad) < 2e-6          # .grad scaled in place like clip_grad_norm_
        for p, q in zip(ours, ref):
            assert rel_err(p, q) < 2e-6
    for p, q in zip(ours, ref):
        assert rel_err(oa.state[p]['exp_avg'], ob.state[q]['exp_avg']) < 2e-6
        assert rel_err(oa.state[p]['exp_avg_sq'], ob.state[q]['exp_avg_sq']) < 2e-6


def test_training_loop_with_fused_step_tracks_torch_loop():
    """Five iterations of the reference's loop shape (forward_train -> clip -> Adam.step -> scheduler.step) with the
    native step + FusedClipAdam on the flat gradient slices vs the same native step + torch clip/Adam on .grad."""
    from imagecaptioning_b200.model import DecoderTrainStep
    from imagecaptioning_b200.optim import FusedClipAdam
    from oracle.synth import synth_inputs
    cfg = dict(V=60, E=16, D=32, H=32, T=6)
    p0 = random_attn_params(cfg, seed=6, dtype=torch.float32)
    da, db = build_decoder(cfg, p0), build_decoder(cfg, p0)
    sa, sb = DecoderTrainStep(da), DecoderTrainStep(db)
    oa = FusedClipAdam(da.parameters(), lr=1e-2, weight_decay=1e-6, max_norm=1.0)
    ob = torch.optim.Adam(db.parameters(), lr=1e-2, weight_decay=1e-6)
    la, lb = [], []
    for it in range(5):
        f, g = synth_inputs(6, 9, cfg['D'], cfg['V'], cfg['T'], seed=10 + it)
        oa.zero_grad(set_to_none=True); ob.zero_grad(set_to_none=True)
        sa.stage(f.to(DEV), g.to(DEV))
        la.append(float(sa.run_resident()[0]))
        order = {id(p): i for i, p in enumerate(sa._params)}
        oa.step(grads=[sa._grads[order[id(p)]] for p in da.parameters()])     # no .grad copies at all
        lb.append(float(sb.forward_train(f.to(DEV), g.to(DEV))))
        torch.nn.utils.clip_grad_norm_(db.parameters(), 1.0)
        ob.step()
    assert la[-1] < la[0]
    assert max(abs(a - b) / abs(b) for a, b in zip(la, lb)) < 1e-4
    for (k, p), (_, q) in zip(da.named_parameters(), db.named_parameters()):
        assert rel_err(p, q) < 5e-4, k
