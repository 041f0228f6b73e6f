"""-m gpu: BASELINE.json's configurations at FULL size against the CPU oracle (configs[1] and [3] live in
test_gpu_train.py::test_attn_config2_full_size_vs_reference_port and test_gpu_decode.py::test_config4_*).

* configs[0]: plain LSTM, B=16, T=16, H=E=D=512, V=10k, fp32 -- vs oracle/ref_port.LstmDecoderPort in fp64, 1e-4.
* configs[2]: attention-LSTM, 7x7x2048 grid, H=E=1024, V=10k, B=512 -- fp32 mode 1e-4 and bf16 mode 2e-2.
* configs[4]: attention-LSTM, V=32k, T=32, H=E=1024, B=256, bf16 mode 2e-2 (the chunked vocabulary/CE path).
The oracle for the two large ones is oracle/restated.attn_train_step run in fp32 on the host (6-9 s each; its own
rounding error, ~1e-6 of each tensor's scale, is far inside both tolerances)."""
import functools

import pytest
import torch

from oracle import ref_port, restated as R
from oracle.synth import synth_inputs
from tests.gpu_util import DEV, build_decoder, build_lstm, random_attn_params
from tests.util import rel_err

pytestmark = pytest.mark.gpu
TOL, BF16_TOL = 1e-4, 2e-2

CFG3 = dict(B=512, L=49, D=2048, H=1024, E=1024, V=10000, T=16)
CFG5 = dict(B=256, L=49, D=2048, H=1024, E=1024, V=32000, T=32)


@functools.lru_cache(maxsize=1)
def _oracle(name):
    cfg = {'cfg3': CFG3, 'cfg5': CFG5}[name]
    p = random_attn_params(cfg, seed=9, dtype=torch.float32)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=4)
    ref = R.attn_train_step(p, feat, gt, cfg['V'])
    return cfg, p, feat, gt, float(ref['loss']), {k: v.clone() for k, v in ref['grads'].items()}


def _run(name, dtype, tol):
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg, p, feat, gt, ref_loss, ref_grads = _oracle(name)
    dec = build_decoder(cfg, p)
    dec.compute_dtype = dtype
    step = DecoderTrainStep(dec, use_cuda_graph=True)
    step._prepare(feat.to(DEV), gt.to(torch.int32).to(DEV))
    step._alphas = torch.empty(cfg['B'], cfg['T'] + 1, cfg['L'], device=DEV)
    loss = step.forward_train(feat.to(DEV), gt.to(DEV))
    assert abs(float(loss) - ref_loss) < tol * ref_loss
    bad = []
    for k, r in ref_grads.items():
        g = dict(dec.named_parameters())[k].grad
        if k == 'attention.v.bias':              # identically zero (softmax shift invariance)
            assert float(g.abs().max()) < 1e-5
            continue
        e = rel_err(g, r)
        if not e < tol:
            bad.append((k, e))
    assert not bad, bad
    a = step._alphas                             # size-independent properties
    assert float((a.sum(2) - 1).abs().max()) < 1e-4 and float(a.min()) >= 0
    assert abs(float(step._loss[1]) - torch.log(torch.tensor(float(cfg['V'] + 3)))) < 0.5
    reg = float(((1 - a.sum(1)) ** 2).mean())
    assert abs(float(step._loss[2]) - reg) < 1e-4 * max(1.0, reg)
    return float(loss)


@pytest.mark.parametrize('dtype,tol', [('f32', TOL), ('bf16', BF16_TOL)])
def test_config3_full_size(dtype, tol):
    _run('cfg3', dtype, tol)


def test_config5_full_size_bf16():
    _run('cfg5', 'bf16', BF16_TOL)


def test_config1_full_size_plain_lstm():
    from imagecaptioning_b200.model import DecoderTrainStep
    B, L, D, H, E, V, T = 16, 1, 512, 512, 512, 10000, 16
    torch.manual_seed(123)
    port = ref_port.LstmDecoderPort(V, E, D, H, T)
    feat, gt = synth_inputs(B, L, D, V, T, seed=123)
    sd = {k: v.clone() for k, v in port.state_dict().items()}
    port = port.double()
    loss = port.train_step(feat.double(), gt)
    lm = build_lstm(dict(V=V, E=E, D=D, H=H, T=T), sd)
    got = DecoderTrainStep(lm).forward_train(feat.to(DEV), gt.to(DEV))
    assert abs(float(got) - float(loss)) < TOL * float(loss)
    for k, q in port.named_parameters():
        assert rel_err(dict(lm.named_parameters())[k].grad, q.grad) < TOL, k
