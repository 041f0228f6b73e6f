"""Generate golden vectors from the UNMODIFIED reference modules (run in the build container,
where /root/reference exists):   python tests/golden/make_golden.py
Writes small .npz fixtures next to this file.  Inputs come from oracle/synth.py (seed 123 =
the reference's own seed, AlexCap/LSTM_opts.py:50); weights are the reference's native init under
torch.manual_seed(123).  Dropout is disabled (p=0) for determinism.
The decoder half of forward_train is driven exactly as AlexCap/LSTMwAttentionModel.py:67-72 and
AlexCap/LSTMModel.py:70-73 do."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, '/root/reference')
torch.Tensor.cuda = lambda self, *a, **k: self   # AttentionLanguageModule.py:70-71 hard-codes .cuda()

from AlexCap.AttentionLanguageModule import Decoder          # noqa: E402
from AlexCap.LanguageModule import LanguageModel              # noqa: E402
from AlexCap.CustomLoss import CustomCrossEntropyLoss         # noqa: E402
from oracle.synth import synth_inputs                         # noqa: E402

CASES = {
    # name: (B, L, D, H, E, V, T, full_length)
    'attn_small': (4, 9, 32, 24, 16, 50, 5, False),
    'attn_ragged': (3, 7, 20, 12, 8, 37, 6, False),      # nothing a multiple of the tile sizes
    'attn_mid': (8, 49, 64, 48, 32, 200, 8, False),
    'lstm_small': (4, 1, 32, 24, 32, 50, 5, False),
    'lstm_regions': (3, 5, 20, 12, 20, 37, 6, False),
}


from tests.golden_consts import LSTM_BEAM_END_BIAS        # noqa: E402


def dump(name, **arrs):
    np.savez_compressed(os.path.join(HERE, name + '.npz'), **{k: np.asarray(v) for k, v in arrs.items()})


def run_attn(name, B, L, D, H, E, V, T, full, dtype):
    torch.set_default_dtype(dtype)
    torch.manual_seed(123)
    dec = Decoder(V, E, D, H, T, {}, 0.0, 'cpu')
    dec.dropout.p = 0.0
    feat, gt = synth_inputs(B, L, D, V, T, seed=123, full_length=full, dtype=dtype)
    feat.requires_grad_(True)
    preds, alphas = dec(feat, gt)
    tgt = dec.get_target(gt, make_target=True)
    ce = CustomCrossEntropyLoss()(preds, tgt)
    reg = ((1 - alphas.sum(1)) ** 2).mean()
    loss = ce + reg
    loss.backward()
    out = dict(dims=np.array([B, L, D, H, E, V, T]), feat=feat.detach().numpy(), gt=gt.numpy(),
               in_tok=dec.get_target(gt).numpy(), tgt=tgt.numpy(),
               preds=preds.detach().numpy(), alphas=alphas.detach().numpy(),
               loss=loss.item(), loss_ce=ce.item(), loss_reg=reg.item(), dfeat=feat.grad.numpy())
    for k, v in dec.state_dict().items():
        out['p:' + k] = v.numpy()
    for k, v in dec.named_parameters():
        out['g:' + k] = v.grad.numpy()
    # greedy decode, B=1 per image through the UNPATCHED reference greedy branch (works for B=1 only)
    dec.eval(); dec.use_beam = False
    ids = []
    with torch.no_grad():
        for b in range(B):
            i, _ = dec(feat[b:b + 1].detach(), feat.new())
            ids.append(i.numpy()[0])
    out['greedy_ids'] = np.stack(ids)
    # beam decode (bias END upward so beams can complete with random weights)
    with torch.no_grad():
        dec.deep_output.bias[V + 2] += 3.0
        beams, flags = [], []
        for b in range(B):
            try:
                s, _ = dec.caption(feat[b:b + 1].detach(), 3)
                row = s.numpy()[0].tolist(); flags.append(1)
            except ValueError:
                row = []; flags.append(0)
            beams.append(row + [-1] * (T + 2 - len(row)))
        dec.deep_output.bias[V + 2] -= 3.0
    out['beam3_ids'] = np.array(beams); out['beam3_ok'] = np.array(flags)
    dump(name, **out)
    print(name, 'loss', loss.item())


def run_lstm(name, B, L, D, H, E, V, T, full, dtype):
    torch.set_default_dtype(dtype)
    torch.manual_seed(123)
    lm = LanguageModel(V, E, D, H, T, 1, {}, 0.0, 'cpu')
    feat, gt = synth_inputs(B, L, D, V, T, seed=123, full_length=full, dtype=dtype)
    feat.requires_grad_(True)
    preds = lm(feat, gt)
    tgt = lm.get_target(gt, make_target=True)
    loss = CustomCrossEntropyLoss()(preds, tgt)
    loss.backward()
    out = dict(dims=np.array([B, L, D, H, E, V, T]), feat=feat.detach().numpy(), gt=gt.numpy(),
               in_tok=lm.get_target(gt).numpy(), tgt=tgt.numpy(), preds=preds.detach().numpy(),
               loss=loss.item(), dfeat=feat.grad.numpy())
    for k, v in lm.state_dict().items():
        out['p:' + k] = v.numpy()
    for k, v in lm.named_parameters():
        out['g:' + k] = v.grad.numpy()
    lm.eval()
    out['greedy_ids'] = lm.predict_caption(feat.detach()).numpy()
    # beam decode (LanguageModule.py:152-198; END biased upward so beams complete; sentences keep their leading START)
    lmb = LanguageModel(V, E, D, H, T, 1, {}, 0.0, 'cpu').eval()     # a copy: the arrays in `out` alias lm's parameters
    lmb.load_state_dict(lm.state_dict())
    with torch.no_grad():
        lmb.rnn.linear.bias[V + 2] += LSTM_BEAM_END_BIAS
        beams, flags = [], []
        for b in range(B):
            try:
                row = [int(x) for x in lmb.caption(feat[b:b + 1].detach(), 3)]
                flags.append(1)
            except ValueError:
                row = []; flags.append(0)
            beams.append(row + [-1] * (T + 4 - len(row)))
    out['beam3_ids'] = np.array(beams); out['beam3_ok'] = np.array(flags)
    dump(name, **out)
    print(name, 'loss', loss.item())


if __name__ == '__main__':
    only = sys.argv[1] if len(sys.argv) > 1 else ''          # e.g. `make_golden.py lstm` regenerates the lstm_* files only
    for name, cfg in CASES.items():
        if not name.startswith(only):
            continue
        (run_attn if name.startswith('attn') else run_lstm)(name + '_f64', *cfg, torch.float64)
    torch.set_default_dtype(torch.float32)
