"""-m gpu: round-2 additions to the boundary, each through the C ABI against the fp64 oracle.

* dropout p = 0.5 (the reference's real train() mode, AttentionLanguageModule.py:54,79) with the mask DRAWN BY THE LIBRARY
  from a device-side Philox counter: mask statistics, a fresh mask at every replay of the same captured CUDA graph, and the
  fused step against the oracle fed the mask read back from the workspace;
* the stand-alone surface members Attention.forward (AttentionLanguageModule.py:15-22) and Decoder.get_init_lstm_state
  (:116-125);
* the step split into phases 16 / 32 / 4 / 8, two steps interleaved on two workspaces;
* N_valid from the captions, a fractional data-parallel CE divisor, out-of-range tokens, the graph cache key.
"""
import math
import os

import pytest
import torch

from oracle import restated as R
from oracle.synth import synth_inputs
from tests.gpu_util import DEV, build_decoder, random_attn_params
from tests.util import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _check_grads(got, ref, tol=TOL):
    bad = []
    for k, r in ref.items():
        if k == 'attention.v.bias':
            continue
        e = rel_err(got[k], r)
        if not e < tol:
            bad.append((k, e))
    assert not bad, bad


@pytest.mark.parametrize('graph', [False, True])
def test_generated_dropout_mask_fresh_per_replay_and_vs_oracle(graph):
    from imagecaptioning_b200 import functional as F_
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=12, L=10, D=32, H=64, E=16, V=60, T=7)
    S = cfg['T'] + 1
    p = random_attn_params(cfg, seed=4)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=11, dtype=torch.float64)
    dec = build_decoder(cfg, p, dropout_p=0.5)
    dec.train()
    step = DecoderTrainStep(dec, use_cuda_graph=graph)
    masks = []
    for it in range(3):
        for q in dec.parameters():
            q.grad = None
        loss = step.forward_train(feat.float().to(DEV), gt.to(DEV))
        mask = F_.attn_workspace_region(step._dims, step._ws, 'MASK', (S, cfg['B'], cfg['H'])).clone()
        masks.append(mask)
        vals = set(mask.unique().tolist())
        assert vals <= {0.0, 2.0}, vals                                   # 1/(1-p) = 2 or 0
        n = mask.numel()
        keep = float((mask > 0).float().mean())
        assert abs(keep - 0.5) < 5 * math.sqrt(0.25 / n), keep            # Bernoulli(0.5) within 5 sigma
        ref = R.attn_train_step(p, feat, gt, cfg['V'], drop_mask=mask.permute(1, 0, 2).double().cpu())
        assert abs(float(loss) - float(ref['loss'])) < TOL * float(ref['loss'])
        _check_grads({k: q.grad for k, q in dec.named_parameters()}, ref['grads'])
    assert not torch.equal(masks[0], masks[1]) and not torch.equal(masks[1], masks[2])
    if graph:
        assert len(step._graphs) == 1                                     # ONE captured graph served all three steps


def test_standalone_attention_forward_and_init_state():
    cfg = dict(B=9, L=21, D=48, H=40, E=16, V=50, T=5)
    p = random_attn_params(cfg, seed=6)
    dec = build_decoder(cfg, p)
    g = torch.Generator().manual_seed(3)
    feat = torch.randn(cfg['B'], cfg['L'], cfg['D'], generator=g, dtype=torch.float64).relu()
    h = torch.randn(cfg['B'], cfg['H'], generator=g, dtype=torch.float64)
    # AttentionLanguageModule.py:15-22
    U_h = (h @ p['attention.U.weight'].T + p['attention.U.bias']).unsqueeze(1)
    W_s = feat @ p['attention.W.weight'].T + p['attention.W.bias']
    e = (torch.tanh(W_s + U_h) @ p['attention.v.weight'].T + p['attention.v.bias']).squeeze(2)
    alpha = torch.softmax(e, 1)
    ctx = (feat * alpha.unsqueeze(2)).sum(1)
    got_ctx, got_alpha = dec.attention(feat.float().to(DEV), h.float().to(DEV))
    assert rel_err(got_alpha, alpha) < TOL and rel_err(got_ctx, ctx) < TOL
    # :116-125
    m = feat.mean(dim=1)
    c0 = torch.tanh(m @ p['init_c.weight'].T + p['init_c.bias'])
    h0 = torch.tanh(m @ p['init_h.weight'].T + p['init_h.bias'])
    got_h, got_c = dec.get_init_lstm_state(feat.float().to(DEV))
    assert rel_err(got_h, h0) < TOL and rel_err(got_c, c0) < TOL


def test_phases_16_32_4_8_interleaved_on_two_workspaces_equal_fused():
    """The bookkeeping of a step (which bf16 operand copies exist) belongs to its workspace: two steps whose phase calls
    interleave on the same host thread must both equal their fused runs."""
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=40, L=49, D=256, H=192, E=96, V=2100, T=9)       # large enough for the pre-split GEMM path
    p = random_attn_params(cfg, seed=8)
    steps, fused = [], []
    for seed in (5, 6):
        feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=seed)
        dec = build_decoder(cfg, p)
        dec.train()
        st = DecoderTrainStep(dec, need_dfeat=True)
        st.stage(feat.to(DEV), gt.to(DEV))
        st.run_resident(3)
        torch.cuda.synchronize()
        fused.append((st._flat_grads.clone(), st._loss.clone(), st.dfeat.clone()))
        steps.append(st)
    a, b = steps
    a.run_resident(16); b.run_resident(16)
    a.run_resident(32); b.run_resident(32)
    b.run_resident(4); a.run_resident(4)
    a.run_resident(8); b.run_resident(8)
    torch.cuda.synchronize()
    for st, (g0, l0, df0) in zip(steps, fused):
        assert rel_err(st._loss, l0) < 1e-6       # (split-K red.add: a step repeats itself to ~1e-7, not bit for bit)
        assert rel_err(st._flat_grads, g0) < 1e-6 and rel_err(st.dfeat, df0) < 1e-6


@pytest.mark.parametrize('two_streams', [False, True])
def test_phase_128_plus_64_equals_phase_32(two_streams):
    """Phase 32 split into 128 (everything but d deep_output.*) + 64 (those two gradients on the idle SMs), 64 either in
    line or on a second stream beside phase 4 as GradAllReducer.step_overlapped runs it: same loss and gradients as the
    fused step."""
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=40, L=49, D=256, H=192, E=96, V=2100, T=9)       # large enough for the fused-CE / pre-split GEMM path
    p = random_attn_params(cfg, seed=8)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=5)
    dec = build_decoder(cfg, p)
    dec.train()
    st = DecoderTrainStep(dec, need_dfeat=True)
    st.stage(feat.to(DEV), gt.to(DEV))
    st.run_resident(3)
    torch.cuda.synchronize()
    g0, l0, df0 = st._flat_grads.clone(), st._loss.clone(), st.dfeat.clone()
    st._flat_grads.zero_()
    st.run_resident(16)
    st.run_resident(128)
    if two_streams:
        side, main = torch.cuda.Stream(), torch.cuda.current_stream()
        side.wait_stream(main)
        with torch.cuda.stream(side):
            st.run_resident(64)
        st.run_resident(4)
        st.run_resident(8)
        main.wait_stream(side)
    else:
        st.run_resident(64)
        st.run_resident(4)
        st.run_resident(8)
    torch.cuda.synchronize()
    assert rel_err(st._loss, l0) < 1e-6
    assert rel_err(st._flat_grads, g0) < 1e-6 and rel_err(st.dfeat, df0) < 1e-6
    # the deep_output gradients really were produced (the buffer was zeroed before the phased run)
    assert float(st.grad_groups()[0].abs().max()) > 0


def test_count_valid_targets_and_fractional_denominator():
    from imagecaptioning_b200 import functional as F_
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=7, L=10, D=32, H=24, E=16, V=60, T=7)
    p = random_attn_params(cfg, seed=2)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=3, dtype=torch.float64)
    gt[2, 1] = 0                                                    # an interior zero: still exactly one END per row
    _, tgt = R.make_targets(gt, cfg['V'])
    n = F_.count_valid_targets(gt.to(torch.int32).to(DEV))
    assert float(n) == float((tgt != 0).sum())
    # a data-parallel divisor below 1 (N_valid_global / world) must be used as is
    dec = build_decoder(cfg, p)
    dec.train()
    st = DecoderTrainStep(dec)
    st.ce_denominator = torch.tensor([0.25], device=DEV)
    st.forward_train(feat.float().to(DEV), gt.to(DEV))
    ref = R.attn_train_step(p, feat, gt, cfg['V'])
    k = float(ref['n_valid']) / 0.25
    assert abs(float(st._loss[1]) - k * float(ref['loss_ce'])) < TOL * k * float(ref['loss_ce'])
    assert rel_err(dec.deep_output.weight.grad, k * ref['grads']['deep_output.weight']) < TOL


def test_out_of_range_token_gives_nan_loss_not_a_fault():
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=6, L=10, D=32, H=24, E=16, V=60, T=7)
    dec = build_decoder(cfg, random_attn_params(cfg, seed=1))
    dec.train()
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=2)
    st = DecoderTrainStep(dec)
    ok = float(st.forward_train(feat.to(DEV), gt.to(DEV)))
    assert ok == ok
    bad = gt.clone()
    bad[3, 2] = cfg['V'] + 3 + 1000                                 # >= Vp: torch would raise a device-side assert
    loss = float(st.forward_train(feat.to(DEV), bad.to(DEV)))
    torch.cuda.synchronize()
    assert loss != loss                                             # NaN, and the process is still healthy:
    again = float(st.forward_train(feat.to(DEV), gt.to(DEV)))
    assert again == ok


def test_graph_cache_key_follows_settings():
    """a captured graph must not be replayed with stale configuration."""
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=6, L=10, D=32, H=24, E=16, V=60, T=7)
    p = random_attn_params(cfg, seed=1)
    dec = build_decoder(cfg, p)
    dec.train()
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=2)
    st = DecoderTrainStep(dec, use_cuda_graph=True)
    l0 = float(st.forward_train(feat.to(DEV), gt.to(DEV)))
    n_valid = float(st._loss[3])
    st.ce_denominator = torch.tensor([2.0 * n_valid], device=DEV)   # set AFTER the first capture
    st.forward_train(feat.to(DEV), gt.to(DEV))
    assert abs(float(st._loss[1]) * 2.0 - (l0 - float(st._loss[2]))) < 1e-5 * l0
    st.ce_denominator = None
    assert float(st.forward_train(feat.to(DEV), gt.to(DEV))) == l0
    # a second batch shape gets its own slot; coming back to the first one re-uses its graphs
    n_graphs = len(st._graphs)
    st.forward_train(feat[:4].to(DEV), gt[:4].to(DEV))
    assert float(st.forward_train(feat.to(DEV), gt.to(DEV))) == l0 and len(st._graphs) == n_graphs


def test_elementwise_relative_error_of_large_entries():
    """tests/util.rel_err is relative to the TENSOR's scale (max |a-b| / max |b|).  This one check is element-wise: every
    gradient entry above 1e-3 of its tensor's scale must agree with the fp64 oracle to 1e-2 relative (measured worst: 5.3e-3 on
    init_h.weight, i.e. an absolute error of ~5e-6 of the scale on an entry 1e-3 of the scale; bf16x3 GEMM engine)."""
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=24, L=49, D=256, H=192, E=96, V=700, T=9)
    p = random_attn_params(cfg, seed=8)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=5, dtype=torch.float64)
    ref = R.attn_train_step(p, feat, gt, cfg['V'])
    dec = build_decoder(cfg, p)
    dec.train()
    DecoderTrainStep(dec).forward_train(feat.float().to(DEV), gt.to(DEV))
    worst = {}
    for k, q in dec.named_parameters():
        if k == 'attention.v.bias':
            continue
        r = ref['grads'][k]
        big = r.abs() > 1e-3 * r.abs().max()
        e = ((q.grad.double().cpu() - r).abs() / r.abs())[big]
        worst[k] = float(e.max())
    assert max(worst.values()) < 1e-2, worst


@pytest.mark.parametrize('dtype', ['f32', 'bf16'])
def test_fused_vocab_projection_ce_never_materialises_logits(dtype):
    """north_star item 4: vocabulary projection + online-softmax CE + gradient in the tcgen05 epilogue (gemm_ps.cu EPI 1/2).
    Shape chosen so that the three vocabulary GEMMs qualify for the pre-split engine, with a row-tile tail (S*B = 960), a
    class tail (V' = 4003) and ragged captions (ignored rows)."""
    from imagecaptioning_b200 import functional as F_
    from imagecaptioning_b200.model import DecoderTrainStep
    cfg = dict(B=60, L=16, D=64, H=256, E=64, V=4000, T=15)
    S, Vp = cfg['T'] + 1, cfg['V'] + 3
    p = random_attn_params(cfg, seed=12)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=13, dtype=torch.float64)
    ref = R.attn_train_step(p, feat, gt, cfg['V'])
    dec = build_decoder(cfg, p)
    dec.compute_dtype = dtype
    dec.train()
    st = DecoderTrainStep(dec)
    loss = st.forward_train(feat.float().to(DEV), gt.to(DEV))
    tol = TOL if dtype == 'f32' else 2e-2
    assert abs(float(loss) - float(ref['loss'])) < tol * float(ref['loss'])
    assert float(st._loss[3]) == float(ref['n_valid'])
    _check_grads({k: q.grad for k, q in dec.named_parameters()}, ref['grads'], tol)
    # the fused path was taken: the workspace has no room for an fp32 logits matrix next to everything else it holds
    logits_bytes = S * cfg['B'] * ((Vp + 3) // 4 * 4) * 4
    from imagecaptioning_b200 import _lib
    with pytest.raises(_lib.AlexcapError):                              # the layout has NO logits region at all
        F_.attn_workspace_region(st._dims, st._ws, 'LOGITS', (1,))
    assert st._ws.numel() < 3 * logits_bytes + (64 << 20)               # bf16 hi+lo copies of d logits only


@pytest.fixture
def time_loop():
    from imagecaptioning_b200 import functional as F_
    prev = F_.set_time_loop(True)
    yield
    F_.set_time_loop(prev)


@pytest.mark.parametrize('cfg', [
    dict(B=4, L=9, D=32, H=24, E=16, V=50, T=5),            # 6 owning CTAs, 16 query columns each, K tails (24, 32 < 64)
    dict(B=24, L=49, D=256, H=192, E=96, V=700, T=9),       # 48 owning CTAs, two 8-column query chunks on some, none on others
    dict(B=128, L=12, D=512, H=512, E=64, V=300, T=4),      # the cfg2 geometry: 128 CTAs, one image and 4 hidden units each
    dict(B=37, L=20, D=128, H=128, E=32, V=90, T=6),        # more owning CTAs... fewer images than CTAs, odd batch
], ids=['tiny', 'mid', 'cfg2geom', 'oddbatch'])
@pytest.mark.parametrize('graph', [False, True])
def test_persistent_time_loop_forward_vs_oracle(time_loop, cfg, graph):
    """north_star item 3 (time_loop.cu): the whole forward recurrence in one persistent cooperative kernel -- weights resident
    in shared memory, LSTM pointwise in the tcgen05 epilogue, next query GEMM behind a grid barrier -- against the fp64 oracle
    (loss, every gradient, alphas), eager and replayed from a CUDA graph."""
    from imagecaptioning_b200 import _lib
    from imagecaptioning_b200.model import DecoderTrainStep
    p = random_attn_params(cfg, seed=21)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=22, dtype=torch.float64)
    ref = R.attn_train_step(p, feat, gt, cfg['V'], need_dfeat=True)
    dec = build_decoder(cfg, p)
    dec.train()
    st = DecoderTrainStep(dec, use_cuda_graph=graph, need_dfeat=True)
    lib = _lib.load()
    for it in range(2):
        for q in dec.parameters():
            q.grad = None
        n0 = lib.alexcap_launch_count()
        loss = st.forward_train(feat.float().to(DEV), gt.to(DEV))
        n1 = lib.alexcap_launch_count()
    assert abs(float(loss) - float(ref['loss'])) < TOL * float(ref['loss'])
    _check_grads({k: q.grad for k, q in dec.named_parameters()}, ref['grads'])
    assert rel_err(st.dfeat, ref['dfeat']) < TOL
    S = cfg['T'] + 1
    from imagecaptioning_b200 import functional as F_
    alphas = F_.attn_workspace_region(st._dims, st._ws, 'ALPHA', (cfg['B'], S, cfg['L']))
    assert rel_err(alphas, ref['alphas']) < TOL
    if not graph:
        # the 4 * S forward launches of the chain are gone: 3 launches (weight pack, h0 split, the loop) replace them
        F_.set_time_loop(False)
        st2 = DecoderTrainStep(build_decoder(cfg, p).train(), need_dfeat=True)
        st2.forward_train(feat.float().to(DEV), gt.to(DEV))
        m0 = lib.alexcap_launch_count()
        st2.forward_train(feat.float().to(DEV), gt.to(DEV))
        chain = lib.alexcap_launch_count() - m0
        F_.set_time_loop(True)
        assert (n1 - n0) == chain - 4 * S + 3 - 1, (n1 - n0, chain)       # (-1: the weight-concatenation copy is not needed)


def test_lstm_cell_fused_into_next_query_gemm_vs_oracle():
    """north_star item 3 on the per-step chain (opt-in, ALEXCAP_LSTM_Q_FUSED=1): the converter warps of step t+1's query / beta
    GEMM compute the LSTM cell of step t (gemm_tc.cu, LA kernels).  Loss, every gradient and d features against the fp64
    oracle, eager and from a CUDA graph, in a subprocess (the switch is read once per process)."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, ALEXCAP_LSTM_Q_FUSED='1')
    r = subprocess.run([sys.executable, os.path.join(root, 'tests', 'fused_cell_worker.py')], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith('RESULT ')][-1]
    res = json.loads(line[len('RESULT '):])
    env0 = dict(os.environ, ALEXCAP_LSTM_Q_FUSED='0')
    r0 = subprocess.run([sys.executable, os.path.join(root, 'tests', 'fused_cell_worker.py')], env=env0, capture_output=True,
                        text=True, timeout=600)
    assert r0.returncode == 0, r0.stdout[-2000:] + r0.stderr[-2000:]
    base = json.loads([l for l in r0.stdout.splitlines() if l.startswith('RESULT ')][-1][len('RESULT '):])
    for a, b in zip(res, base):
        assert a['loss_rel'] < TOL and a['grad_rel'] < TOL and a['dfeat_rel'] < TOL, a
        if not a['graph']:
            assert a['launches'] == b['launches'] - (a['S'] - 1), (a, b)       # one launch fewer per step but the last


def test_producer_written_bf16_operand_mirrors_equal_split_passes(tmp_path):
    """bf16 mode, per-step GEMMs on the pre-split engine: the attention and LSTM pointwise kernels write their outputs as bf16
    into mirrors of XH / DGQ as well, and the GEMMs read sub-views of those instead of running a split pass per operand per
    time step.  Same roundings, so the step equals the split-pass step to bf16-mode run-to-run noise, with 4 launches fewer per
    time step (ALEXCAP_BF16_MIRROR is read once per process: subprocesses)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = {}
    for tag, val in (('on', '1'), ('off', '0')):
        out = str(tmp_path / f'mirror_{tag}.pt')
        r = subprocess.run([sys.executable, os.path.join(root, 'tests', 'mirror_worker.py'), out],
                           env=dict(os.environ, ALEXCAP_BF16_MIRROR=val), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        res[tag] = torch.load(out)
    on, off = res['on'], res['off']
    # (bf16 mode: a 1e-7 accumulation-order difference in an fp32 intermediate can flip the bf16 rounding of the operand made
    #  from it, so two runs of the SAME code agree to ~1e-3, not to fp32 round-off)
    assert abs(on['loss'] - off['loss']) < 1e-4 * abs(off['loss'])
    assert rel_err(on['flat'], off['flat']) < 1e-3 and rel_err(on['dfeat'], off['dfeat']) < 2e-3
    assert on['launches'] <= off['launches'] - 4 * on['S'], (on['launches'], off['launches'])
