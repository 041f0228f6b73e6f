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


def test_attn_beam_and_greedy_restated_vs_live_reference(reference_modules):
    """oracle attn_beam / attn_greedy == Decoder.caption (AttentionLanguageModule.py:128-195) / the B=1 greedy branch
    (:88-115) run live: beam sizes 1..5, several images, END bias chosen so that some beams complete and some do not."""
    torch.set_default_dtype(torch.float64)
    try:
        V, E, D, H, T, L = 40, 12, 24, 20, 7, 11
        torch.manual_seed(123)
        dec = reference_modules['Decoder'](V, E, D, H, T, {}, 0.0, 'cpu').eval()
        with torch.no_grad():
            dec.deep_output.bias[V + 2] += 0.3
        p = {k: v.detach().clone() for k, v in dec.state_dict().items()}
        feat, _ = synth_inputs(6, L, D, V, T, seed=11, dtype=torch.float64)
        seen = {True: 0, False: 0}
        with torch.no_grad():
            for b in range(6):
                f1 = feat[b:b + 1]
                for k in (1, 2, 3, 4, 5):
                    try:
                        sent, alpha = dec.caption(f1, k)
                        ref = sent[0].tolist()
                    except ValueError:
                        ref = None
                    try:
                        mine, _ = R.attn_beam(p, f1, V, T, k)
                    except ValueError:
                        mine = None
                    assert ref == mine, (b, k)
                    seen[ref is not None] += 1
                dec.use_beam = False
                ids, alphas = dec(f1, f1.new())                       # B = 1: the reference's own greedy code works
                rid, ral, _ = R.attn_greedy(p, f1, V, T)
                assert torch.equal(ids.long().cpu(), rid) and float((alphas - ral).abs().max()) < 1e-12
                dec.use_beam = True
        assert seen[True] > 0 and seen[False] > 0
    finally:
        torch.set_default_dtype(torch.float32)


@pytest.mark.parametrize('num_layers', [1, 2])
def test_lstm_restated_vs_live_reference(reference_modules, num_layers):
    """oracle lstm_train_step / lstm_greedy == LanguageModel (LanguageModule.py:99-107, :130-150) run live, 1 and 2 layers."""
    torch.set_default_dtype(torch.float64)
    try:
        B, L, D, H, E, V, T = 5, 3, 20, 16, 12, 37, 6
        torch.manual_seed(123)
        lm = reference_modules['LanguageModel'](V, E, D, H, T, num_layers, {}, 0.0, 'cpu')
        feat, gt = synth_inputs(B, L, D, V, T, seed=9, dtype=torch.float64)
        preds = lm(feat, gt)
        loss = reference_modules['CustomCrossEntropyLoss']()(preds, lm.get_target(gt, True))
        loss.backward()
        p = {k: v.detach().clone() for k, v in lm.state_dict().items()}
        out = R.lstm_train_step(p, feat, gt, V, num_layers)
        assert abs(float(out['loss']) - float(loss.detach())) < 1e-10
        for k, v in lm.named_parameters():
            assert rel_err(out['grads'][k], v.grad) < 1e-10, k
        lm.eval()
        assert torch.equal(lm.predict_caption(feat), R.lstm_greedy(p, feat, V, T, num_layers))
    finally:
        torch.set_default_dtype(torch.float32)


def test_make_targets_vs_live_get_target(reference_modules):
    """get_target (AttentionLanguageModule.py:197-212 = LanguageModule.py:113-128) on ragged, full-length, empty and
    zero-in-the-middle captions: START prefix, END at the first 0 at index >= 1, the rest left 0."""
    V, T = 9, 5
    dec = reference_modules['Decoder'](V, 4, 4, 4, T, {}, 0.0, 'cpu')
    lm = reference_modules['LanguageModel'](V, 4, 4, 4, T, 1, {}, 0.0, 'cpu')
    gt = torch.tensor([[1, 2, 3, 4, 5], [1, 2, 0, 0, 0], [0, 0, 0, 0, 0], [3, 0, 4, 0, 0], [0, 7, 0, 0, 0], [9, 9, 9, 9, 0]])
    in_tok, tgt = R.make_targets(gt, V)
    for mod in (dec, lm):
        assert torch.equal(mod.get_target(gt).long(), in_tok.long())
        assert torch.equal(mod.get_target(gt, make_target=True).long(), tgt.long())
