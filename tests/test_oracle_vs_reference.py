"""CPU, build container only (skipped where /root/reference is absent): pins oracle/restated.py and
oracle/ref_port.py against the reference's own modules run live, in fp32 and fp64."""
import pytest
import torch

from oracle import ref_port, restated as R
from oracle.synth import synth_inputs
from tests.util import rel_err


@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-10), (torch.float32, 2e-4)])
def test_attn_restated_vs_live_reference(reference_modules, dtype, tol):
    B, L, D, H, E, V, T = 5, 11, 24, 20, 12, 40, 7
    torch.set_default_dtype(dtype)
    try:
        torch.manual_seed(123)
        dec = reference_modules['Decoder'](V, E, D, H, T, {}, 0.0, 'cpu')
        dec.dropout.p = 0.0
        feat, gt = synth_inputs(B, L, D, V, T, seed=7, dtype=dtype)
        preds, alphas = dec(feat, gt)
        loss = reference_modules['CustomCrossEntropyLoss']()(preds, dec.get_target(gt, True)) \
            + ((1 - alphas.sum(1)) ** 2).mean()
        loss.backward()
        p = {k: v.detach().clone() for k, v in dec.state_dict().items()}
        out = R.attn_train_step(p, feat, gt, V)
        assert abs(float(out['loss']) - float(loss)) < tol * 10
        for k, v in dec.named_parameters():
            if k == 'attention.v.bias':
                continue
            assert rel_err(out['grads'][k], v.grad) < tol, k
    finally:
        torch.set_default_dtype(torch.float32)


def test_ports_are_bit_identical_to_reference(reference_modules):
    """ref_port (the timed CPU baseline) == the reference modules, bit for bit, in fp32."""
    B, L, D, H, E, V, T = 4, 9, 32, 24, 16, 50, 5
    feat, gt = synth_inputs(B, L, D, V, T, seed=3)
    torch.manual_seed(123)
    dec = reference_modules['Decoder'](V, E, D, H, T, {}, 0.0, 'cpu'); dec.dropout.p = 0.0
    port = ref_port.AttnDecoderPort(V, E, D, H, T, dropout_p=0.0)
    port.load_state_dict(dec.state_dict())
    preds, alphas = dec(feat, gt)
    loss = reference_modules['CustomCrossEntropyLoss']()(preds, dec.get_target(gt, True)) + ((1 - alphas.sum(1)) ** 2).mean()
    loss.backward()
    lp = port.train_step(feat, gt)
    assert float(lp) == float(loss)
    pg = dict(port.named_parameters())
    for k, a in dec.named_parameters():
        assert torch.equal(a.grad, pg[k].grad), k

    torch.manual_seed(123)
    lm = reference_modules['LanguageModel'](V, E, D, H, T, 1, {}, 0.0, 'cpu')
    lport = ref_port.LstmDecoderPort(V, E, D, H, T, 1, 0.0)
    lport.load_state_dict(lm.state_dict())
    feat2, gt2 = synth_inputs(B, 3, D, V, T, seed=4)
    pr = lm(feat2, gt2)
    l2 = reference_modules['CustomCrossEntropyLoss']()(pr, lm.get_target(gt2, True)); l2.backward()
    l3 = lport.train_step(feat2, gt2)
    assert float(l2) == float(l3)
    pg = dict(lport.named_parameters())
    for k, a in lm.named_parameters():
        assert torch.equal(a.grad, pg[k].grad), k


@pytest.mark.parametrize('num_layers', [1, 2])
def test_lstm_beam_restated_vs_live_reference(reference_modules, num_layers):
    """oracle lstm_beam == LanguageModel.caption run live (LanguageModule.py:152-198): sentence lists (START included)
    and the ValueError when no beam completes, beam sizes 1..5, one and two LSTM layers."""
    torch.set_default_dtype(torch.float64)
    try:
        seen = {True: 0, False: 0}
        for (V, E, D, H, T, L), bias in (((50, 32, 32, 24, 5, 1), 0.35), ((37, 20, 20, 12, 6, 5), 0.55)):
            torch.manual_seed(123)
            lm = reference_modules['LanguageModel'](V, E, D, H, T, num_layers, {}, 0.0, 'cpu').eval()
            with torch.no_grad():
                lm.rnn.linear.bias[V + 2] += bias
            p = {k: v.detach().clone() for k, v in lm.state_dict().items()}
            feat, _ = synth_inputs(5, L, D, V, T, seed=5, dtype=torch.float64)
            for b in range(5):
                for k in (1, 2, 3, 5):
                    try:
                        ref = lm.caption(feat[b:b + 1], k)
                    except ValueError:
                        ref = None
                    try:
                        mine, _ = R.lstm_beam(p, feat[b:b + 1], V, T, k, num_layers)
                    except ValueError:
                        mine = None
                    assert ref == mine, (V, b, k)
                    seen[ref is not None] += 1
        assert seen[True] > 0
    finally:
        torch.set_default_dtype(torch.float32)
