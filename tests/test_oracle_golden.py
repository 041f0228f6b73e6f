"""CPU: the restated oracle (explicit forward + hand-written backward) against the committed
golden vectors that were generated from the UNMODIFIED reference modules."""
import pytest
import torch

from oracle import restated as R
from tests.util import golden_names, load_golden, rel_err


@pytest.mark.parametrize('name', golden_names('attn_'))
def test_attn_train_step_matches_reference_golden(name):
    g = load_golden(name)
    V = g['cfg']['V']
    in_tok, tgt = R.make_targets(g['gt'], V)
    assert torch.equal(in_tok, g['in_tok'].long()) and torch.equal(tgt, g['tgt'].long())
    out = R.attn_train_step(g['params'], g['feat'], g['gt'], V, need_dfeat=True)
    assert rel_err(out['preds'], g['preds']) < 1e-12
    assert rel_err(out['alphas'], g['alphas']) < 1e-12
    assert abs(float(out['loss']) - g['loss']) < 1e-12
    assert abs(float(out['loss_ce']) - g['loss_ce']) < 1e-12
    assert abs(float(out['loss_reg']) - g['loss_reg']) < 1e-12
    for k, ref in g['grads'].items():
        if k == 'attention.v.bias':   # identically zero by softmax shift invariance: absolute tolerance
            assert float(out['grads'][k].abs().max()) < 1e-12
            continue
        assert rel_err(out['grads'][k], ref) < 1e-10, k
    assert rel_err(out['dfeat'], g['dfeat']) < 1e-10


@pytest.mark.parametrize('name', golden_names('lstm_'))
def test_lstm_train_step_matches_reference_golden(name):
    g = load_golden(name)
    V = g['cfg']['V']
    out = R.lstm_train_step(g['params'], g['feat'], g['gt'], V, need_dfeat=True)
    assert rel_err(out['preds'], g['preds']) < 1e-12
    assert abs(float(out['loss']) - g['loss']) < 1e-12
    for k, ref in g['grads'].items():
        assert rel_err(out['grads'][k], ref) < 1e-10, k
    assert rel_err(out['dfeat'], g['dfeat']) < 1e-10


@pytest.mark.parametrize('name', golden_names('attn_'))
def test_attn_decode_matches_reference_golden(name):
    g = load_golden(name)
    c = g['cfg']
    ids, alphas, margins = R.attn_greedy(g['params'], g['feat'], c['V'], c['T'])
    # the reference greedy branch was run per image (B=1), so each row stops at ITS OWN first END;
    # the batched oracle keeps going until all rows emitted END -> compare up to the row's END.
    end = c['V'] + 2
    for b in range(c['B']):
        ref = g['greedy_ids'][b].tolist()
        n = ref.index(end) + 1 if end in ref else len(ref)
        assert ids[b, :n].tolist() == ref[:n]
    p = {k: v.clone() for k, v in g['params'].items()}
    p['deep_output.bias'][end] += 3.0
    for b in range(c['B']):
        if int(g['beam3_ok'][b]):
            got, _ = R.attn_beam(p, g['feat'][b:b + 1], c['V'], c['T'], 3)
            ref = [x for x in g['beam3_ids'][b].tolist() if x >= 0]
            assert got == ref
        else:
            with pytest.raises(ValueError):
                R.attn_beam(p, g['feat'][b:b + 1], c['V'], c['T'], 3)


@pytest.mark.parametrize('name', golden_names('lstm_'))
def test_lstm_greedy_matches_reference_golden(name):
    g = load_golden(name)
    c = g['cfg']
    ids = R.lstm_greedy(g['params'], g['feat'], c['V'], c['T'])
    assert torch.equal(ids, g['greedy_ids'].long())


@pytest.mark.parametrize('name', golden_names('lstm_'))
def test_lstm_beam_matches_reference_golden(name):
    """LanguageModel.caption (LanguageModule.py:152-198), beam 3, END biased as in tests/golden/make_golden.py."""
    from tests.golden_consts import LSTM_BEAM_END_BIAS
    g = load_golden(name)
    c = g['cfg']
    p = {k: v.clone() for k, v in g['params'].items()}
    p['rnn.linear.bias'][c['V'] + 2] += LSTM_BEAM_END_BIAS
    for b in range(c['B']):
        if int(g['beam3_ok'][b]):
            ids, _ = R.lstm_beam(p, g['feat'][b:b + 1], c['V'], c['T'], 3)
            assert ids == [x for x in g['beam3_ids'][b].tolist() if x >= 0]
            assert ids[0] == c['V'] + 1 and ids[-1] == c['V'] + 2
        else:
            with pytest.raises(ValueError):
                R.lstm_beam(p, g['feat'][b:b + 1], c['V'], c['T'], 3)


def test_ce_matches_torch_functional():
    torch.manual_seed(0)
    x = torch.randn(37, 53, dtype=torch.float64)
    t = torch.randint(0, 53, (37,))
    t[::5] = 0
    xr = x.clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(xr, t, ignore_index=0, label_smoothing=0.1)
    ref.backward()
    loss, d, n = R.ce_loss_and_grad(x, t)
    assert abs(float(loss) - float(ref.detach())) < 1e-13 and rel_err(d, xr.grad) < 1e-12 and n == int((t != 0).sum())


def test_make_targets_edge_cases():
    V = 9
    gt = torch.tensor([[0, 0, 0], [1, 2, 3], [4, 0, 5], [0, 7, 0]], dtype=torch.int32)
    in_tok, tgt = R.make_targets(gt, V)
    assert in_tok.tolist() == [[10, 0, 0, 0], [10, 1, 2, 3], [10, 4, 0, 5], [10, 0, 7, 0]]
    # first zero at index >= 1 becomes END; index 0 is never replaced
    assert tgt.tolist() == [[0, 11, 0, 0], [1, 2, 3, 11], [4, 11, 5, 0], [0, 7, 11, 0]]
