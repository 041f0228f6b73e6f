#!/usr/bin/env python
"""bench.py -- captions/sec of the attention-LSTM decoder training step (fwd+bwd, T=16), BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config cfg2]

* ours: DecoderTrainStep.forward_train (alexcap_attn_train_fused through the C ABI).  `value` = whole-job captions/s
  with inputs resident in HBM; `e2e` = same through the public API from PINNED HOST buffers (H2D of features+captions
  and D2H of the loss inside the timed region).  N>1 (torchrun): one rank per GPU, per-rank batch fixed (weak scaling),
  gradients all-reduced with NCCL inside the timed step.
* reference: the reference's CPU execution path (oracle/ref_port.py: the reference modules' per-step structure, stock
  torch CPU ops + autograd; the reference itself is Python under /root/reference and cannot travel to the GPU box)
  with all host threads, on bounded samples of the same workload.  Rank 0 only.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # BASELINE.json configs[1]: the configuration the metric is quoted on at N=1
    'cfg2': dict(model='LSTMwAttentionModel', B=128, L=196, D=512, H=512, E=512, V=10000, T=16, dtype='f32'),
    'cfg1': dict(ref='BASELINE.json configs[0]', model='LSTMModel', B=16, L=1, D=512, H=512, E=512, V=10000, T=16, dtype='f32'),
    'cfg3_f32': dict(ref='BASELINE.json configs[2] shape in fp32', model='LSTMwAttentionModel', B=512, L=49, D=2048, H=1024, E=1024, V=10000, T=16, dtype='f32'),
    # BASELINE.json configs[2] / configs[4]: bf16 tensor-core GEMMs with fp32 master weights / accumulation (E = 1024 assumed)
    'cfg3': dict(ref='BASELINE.json configs[2]', model='LSTMwAttentionModel', B=512, L=49, D=2048, H=1024, E=1024, V=10000, T=16, dtype='bf16'),
    'cfg5': dict(ref='BASELINE.json configs[4]', model='LSTMwAttentionModel', B=256, L=49, D=2048, H=1024, E=1024, V=32000, T=32, dtype='bf16'),
    # BASELINE.json configs[3]: decode (greedy / beam-5), exact token ids in fp32 -- `--config cfg4` runs the decode arm
    'cfg4': dict(ref='BASELINE.json configs[3]', model='LSTMwAttentionModel', B=1024, L=196, D=512, H=512, E=512, V=10000, T=20, dtype='f32', decode=True),
}
METRIC = 'captions/sec (attn-LSTM fwd+bwd, T=16)'


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(hbm=j['hbm_gbs'], bf16=j['bf16_tflops'], bf16_sus=j['bf16_tflops_sustained'], src='measured')
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src='fallback')


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def __enter__(self):
        q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
             'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
             'clocks_event_reasons.sw_power_cap')
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={q}', '--format=csv,noheader,nounits', '-lms', '100',
                                          '-i', str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        sm.sort()
        return dict(sm_mhz=(sm[len(sm) // 2] if sm else None), sm_max_mhz=mx or None, reasons=sorted(reasons),
                    samples=len(sm))


def algorithmic_work(c):
    """GEMM FLOPs and attention bytes of one training step (frozen features)."""
    B, L, D, H, E, S, Vp = c['B'], c['L'], c['D'], c['H'], c['E'], c['T'] + 1, c['V'] + 3
    f_fwd = (2 * B * L * D * H + 4 * B * D * H + 2 * B * S * E * 4 * H
             + S * (2 * B * H * H + 2 * B * H * D + 2 * B * D * 4 * H + 2 * B * H * 4 * H) + 2 * B * S * H * Vp)
    f_step = 3 * f_fwd - 2 * B * L * D * H - 2 * B * D * H
    e = 2 if c['dtype'] == 'bf16' else 4     # bf16 configs stream Ws / feat as bf16 rows (e = 2)
    attn_fwd_step = B * L * (H + D) * e + B * L * 4 + B * (H + D) * 4
    attn_bwd_step = B * L * (H + D) * e
    return dict(gemm_flops_step=f_step, attn_fwd_bytes_per_launch=attn_fwd_step, attn_bwd_bytes_per_launch=attn_bwd_step)


def make_model(c, dev):
    from imagecaptioning_b200.modules import Decoder, LanguageModel
    torch.manual_seed(123)   # the reference's seed (AlexCap/LSTM_opts.py:50)
    if c['model'] == 'LSTMwAttentionModel':
        m = Decoder(c['V'], c['E'], c['D'], c['H'], c['T'], {}, 0.0, dev)
        m.dropout.p = 0.0    # parity setting; see DESIGN.md (dropout path measured separately)
        m.compute_dtype = 'bf16' if c['dtype'] == 'bf16' else 'f32'
    else:
        m = LanguageModel(c['V'], c['E'], c['D'], c['H'], c['T'], 1, {}, 0.0, dev)
    return m.to(dev)


def time_graph_events(fn, launches_per_graph, reps, warm=2):
    """`launches_per_graph` launches captured into one CUDA graph, replayed `reps` times and timed with CUDA events on
    the launching stream: the Python/ctypes call costs ~25 us per launch, more than the kernel itself."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(launches_per_graph):
                fn()
        g.replay()
        st.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(st)
        for _ in range(reps):
            g.replay()
        e.record(st)
        st.synchronize()
    torch.cuda.current_stream().wait_stream(st)
    return s.elapsed_time(e) / (reps * launches_per_graph)


def dominant_kernel_roofline(c, dev, pk):
    """The fused attention-step kernels (34 launches per training step, each streaming Ws and feat once) timed alone at the
    bench shape, launches replayed from a CUDA graph, CUDA events on the launching stream.  The dominant kernel BY STEP SHARE
    is the backward one (profiles/r2*_step_breakdown.txt), so it is the headline; the forward kernel is nested.
    Two figures each: `achieved` is COLD -- consecutive launches alternate between two distinct Ws/feat sets (2 x 103 MB at
    cfg2 > the 126 MB L2), so every byte comes from HBM, which is what `peak` measures; `l2_warm` replays over ONE set (as
    round 1 did), where part of the 103 MB hits L2: algorithmic bytes over time, NOT HBM bandwidth.  Inside a training step
    the truth lies between (the same set is re-read by the next launch, with the GEMMs' traffic in between)."""
    from imagecaptioning_b200 import functional as F_
    B, L, D, H, S = c['B'], c['L'], c['D'], c['H'], c['T'] + 1
    g = torch.Generator(device='cpu').manual_seed(1)
    sets = []
    esz = 2 if c['dtype'] == 'bf16' else 4
    nsets = 2 if B * L * (H + D) * esz * 2 > 140e6 else max(2, int(160e6 // (B * L * (H + D) * esz)) + 1)
    for _ in range(nsets):
        Ws_, ft_ = torch.randn(B, L, H, generator=g).to(dev), torch.randn(B, L, D, generator=g).relu().to(dev)
        Ws_ = F_.ws_exp_form(Ws_)            # the form the training step streams: e^(2 Ws) (one MUFU op per score element)
        if c['dtype'] == 'bf16':             # the storage type the training step streams in this mode
            Ws_, ft_ = Ws_.bfloat16(), ft_.bfloat16()
        sets.append((Ws_, ft_))
    q = torch.randn(S, B, H, generator=g).to(dev)
    bp = torch.randn(S, B, D, generator=g).to(dev)
    dg = torch.randn(S, B, D, generator=g).to(dev)
    wv = (torch.randn(H, generator=g) / H ** 0.5).to(dev)
    outs = [[F_.attn_step_fwd(Ws, feat, q[t], bp[t], wv, ws_exp=True) for t in range(S)] for Ws, feat in sets]
    state = {'t': 0, 'n': 1}

    def fwd():
        t = state['t'] = (state['t'] + 1) % S
        Ws, feat = sets[t % state['n']]
        F_.attn_step_fwd(Ws, feat, q[t], bp[t], wv, ws_exp=True)

    def bwd():
        t = state['t'] = (state['t'] + 1) % S
        Ws, feat = sets[t % state['n']]
        a, cx, be, _ = outs[t % state['n']][t]
        F_.attn_step_bwd(Ws, feat, q[t], wv, a, cx, be, dg[t], ws_exp=True)

    res = {}
    for name, fn in (('fwd', fwd), ('bwd', bwd)):
        state['n'] = 1
        warm = time_graph_events(fn, 2 * S, 8)
        state['n'] = nsets
        cold = time_graph_events(fn, 2 * S, 8)
        res[name] = (cold, warm)
    w = algorithmic_work(c)
    is2 = c is CONFIGS['cfg2']

    def block(kernel, nbytes, cold, warm, traffic):
        ach = nbytes / (cold * 1e-3) / 1e9
        achw = nbytes / (warm * 1e-3) / 1e9
        return dict(kernel=kernel, bound='hbm', achieved=round(ach, 1), peak=pk['hbm'], unit='GB/s', frac=round(ach / pk['hbm'], 4),
                    traffic=traffic, ms_per_launch=round(cold, 5), algorithmic_bytes_per_launch=nbytes, peak_source=pk['src'],
                    l2_warm=dict(achieved=round(achw, 1), frac=round(achw / pk['hbm'], 4), ms_per_launch=round(warm, 5)))
    out = block('attn_bwd_stream_kernel', w['attn_bwd_bytes_per_launch'], *res['bwd'], NCU_TRAFFIC.get('attn_bwd') if is2 else None)
    out['forward'] = block('attn_fwd_stream_kernel', w['attn_fwd_bytes_per_launch'], *res['fwd'], NCU_TRAFFIC.get('attn_fwd') if is2 else None)
    out['share_of_step'] = 'backward 16.8 %, forward 15.5 % of the cfg2 step (profiles/r2m_step_breakdown.txt: ncu launch list, serialised)'
    out['note'] = ('achieved = cold (launches alternate over %d distinct Ws/feat sets, %.0f MB each: no L2 reuse); l2_warm = one set '
                   're-read by every launch (part of it hits the 126 MB L2)' % (nsets, B * L * (H + D) * esz / 1e6))
    return out


def optimizer_step_timing(model, step, dev, iters=20):
    """SURVEY.md section 8(f) rank 1 (the step right after the hot path): clip_grad_norm_(1) + Adam.step() over the
    decoder's parameters -- FusedClipAdam (alexcap_clip_adam_step, 3 launches) next to stock torch on the same device.
    Informational: it is not part of `value`/`e2e` (BASELINE.json's metric is the fwd+bwd step)."""
    from imagecaptioning_b200.optim import FusedClipAdam
    ps = list(model.parameters())
    order = {id(p): i for i, p in enumerate(step._params)}
    grads = [step._grads[order[id(p)]] for p in ps]
    saved = [p.detach().clone() for p in ps]
    fused = FusedClipAdam(ps, lr=1e-4, weight_decay=1e-6, max_norm=1.0)
    ref_ps = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    ref = torch.optim.Adam(ref_ps, lr=1e-4, weight_decay=1e-6)
    for q, g in zip(ref_ps, grads):
        q.grad = g.clone()

    def t(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / iters

    ms_f = t(lambda: fused.step(grads=grads))
    ms_t = t(lambda: (torch.nn.utils.clip_grad_norm_(ref_ps, 1.0), ref.step()))
    with torch.no_grad():
        for p, s0 in zip(ps, saved):
            p.copy_(s0)
    return dict(what='clip_grad_norm_(1) + Adam.step() on %d tensors / %.1f M parameters' % (len(ps), sum(p.numel() for p in ps) / 1e6),
                fused_ms=round(ms_f, 4), torch_ms=round(ms_t, 4), launches_fused=3)


def gemm_roofline(c, dev, pk):
    """The GEMM group's big-GEMM kernel (ps::gemm_ps_kernel, gemm_ps.cu) on the largest GEMM of the step, the vocabulary
    projection [S*B, H] x [V', H]^T, timed alone from a CUDA graph with its bf16 operand copies already made (inside a step
    the copies are shared by the GEMMs that read the same matrix; `with_split_ms` adds the split pass of both operands).
    `achieved` counts the ALGORITHMIC 2MNK; the fp32 mode issues three tensor-core MMAs per product (bf16x3), so frac is
    against the measured bf16 burst peak with that factor stated, not hidden."""
    from imagecaptioning_b200 import functional as F_
    M, N, K = (c['T'] + 1) * c['B'], (c['V'] + 3 + 3) // 4 * 4, c['H']
    g = torch.Generator(device='cpu').manual_seed(2)
    A = torch.randn(M, K, generator=g).to(dev)
    W = torch.randn(N, K, generator=g).to(dev)
    C = torch.empty(M, N, device=dev)
    bf16 = c['dtype'] == 'bf16'
    F_.gemm(A, W, transa=False, transb=True, out=C, engine=400 if bf16 else 300)        # sizes the arena, makes the copies
    torch.cuda.synchronize()
    ms = time_graph_events(lambda: F_.gemm(A, W, transa=False, transb=True, out=C, engine=600 if bf16 else 500), 8, 6)
    ms_split = time_graph_events(lambda: F_.gemm(A, W, transa=False, transb=True, out=C, engine=400 if bf16 else 300), 8, 6)
    ms_conv = time_graph_events(lambda: F_.gemm(A, W, transa=False, transb=True, out=C, engine=100 if bf16 else 200), 8, 6)
    tf = 2.0 * M * N * K / (ms * 1e-3) / 1e12
    pair = M >= 256 and N >= 256 and K >= (1024 if bf16 else 512)           # gemm_ps.cu: pair_wanted()
    return dict(kernel='ps::gemm_ps2_kernel (CTA-pair 256x256 tiles, cta_group::2)' if pair else 'ps::gemm_ps_kernel', bound='tensor', achieved=round(tf, 1), peak=pk['bf16'], unit='TFLOP/s',
                frac=round(tf / pk['bf16'], 4), traffic=NCU_TRAFFIC.get('gemm_ps_logits') if (c is CONFIGS['cfg2']) else None,
                ms_per_launch=round(ms, 5), shape=[M, N, K],
                mma_per_product=1 if bf16 else 3, issued_frac=round((1 if bf16 else 3) * tf / pk['bf16'], 4),
                with_split_ms=round(ms_split, 5), convert_in_kernel_ms=round(ms_conv, 5), peak_source=pk['src'],
                note='operands pre-split into bf16 hi/lo copies in HBM, kernel = TMA -> smem -> tcgen05.mma only, 16-byte '
                     'vector epilogue; inside the training step the same projection runs with the fused cross-entropy '
                     'epilogues (no logits stored): ncu sm__pipe_tensor_cycles_active 55.5 % / 50.7 % '
                     '(profiles/r2m_ncu_summary.txt); traffic = the round-1 capture of the plain projection; '
                     'convert_in_kernel_ms = the per-step small-GEMM engine (tc::gemm_tc_kernel) on the same shape')


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` captures (profiles/)
NCU_TRAFFIC = {}
_t = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
if os.path.exists(_t):
    NCU_TRAFFIC = json.load(open(_t))


def cpu_port(c):
    from oracle import ref_port
    torch.manual_seed(123)
    if c['model'] == 'LSTMwAttentionModel':
        return ref_port.AttnDecoderPort(c['V'], c['E'], c['D'], c['H'], c['T'], dropout_p=0.0)
    return ref_port.LstmDecoderPort(c['V'], c['E'], c['D'], c['H'], c['T'], 1, 0.0)


def time_cpu_port(c, steps, warmup, budget_s=150.0):
    """captions/s of the reference-structured CPU port with all host threads; bounded per-step sample."""
    from oracle.synth import synth_inputs
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    port = cpu_port(c)
    feat, gt = synth_inputs(c['B'], c['L'], c['D'], c['V'], c['T'], seed=123, full_length=True)
    t0 = time.perf_counter()
    port.zero_grad(set_to_none=True)
    port.train_step(feat, gt)                    # first (cold) step sizes the sample
    t1 = time.perf_counter() - t0
    bs = c['B']
    total = steps + max(warmup - 1, 0)
    if t1 * total > budget_s:
        bs = max(1, int(c['B'] * budget_s / (t1 * total)))
    f, g = feat[:bs].contiguous(), gt[:bs].contiguous()
    for _ in range(max(warmup - 1, 0)):
        port.zero_grad(set_to_none=True); port.train_step(f, g)
    ts = []
    for _ in range(steps):
        port.zero_grad(set_to_none=True)
        a = time.perf_counter(); port.train_step(f, g); ts.append(time.perf_counter() - a)
    ts.sort()
    med = ts[len(ts) // 2]
    return dict(value=round(bs / med, 2), unit='captions/s', cores=cores, kind='port',
                sample=f'{steps} timed steps of batch {bs} (of {c["B"]}) of the same synthetic workload, median step '
                       f'{med * 1e3:.0f} ms, fp32, oracle/ref_port.py (reference-structured torch CPU path)'), med, bs


def run_reference(args, c, rank):
    if rank != 0:
        return
    cb, med, bs = time_cpu_port(c, args.steps, args.warmup)
    line = dict(impl='reference', metric=METRIC, value=cb['value'], unit='captions/s', n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=round(med * 1e3, 3), higher_is_better=True, scaling='weak', vs_baseline=None,
                dtype='f32', data='synthetic', config=workload_config(c, args.gpus), cpu_baseline=cb,
                e2e=dict(value=cb['value'], unit='captions/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


def workload_config(c, n, dropout_p=0.0):
    return dict(workload=f"{c['model']} decoder train step fwd+bwd ({c.get('ref', 'BASELINE.json configs[1]')}): B={c['B']}/GPU, "
                         f"L={c['L']}, D={c['D']}, H={c['H']}, E={c['E']}, V={c['V']}, T={c['T']}, {c['dtype']}, full-length captions, "
                         f"dropout p={dropout_p:g}",
                global_batch=c['B'] * n, parallelism=f'dp{n}',
                l2='no explicit flush: per-step working set (Ws+feat+logits+saved state ~ 400 MB) exceeds the 126 MB L2')


class TrainRun:
    """One model + DecoderTrainStep (+ GradAllReducer) on synthetic inputs of config `c`; times the resident step and the
    end-to-end step (pinned host inputs -> H2D -> step (+ all-reduce) -> loss D2H)."""

    def __init__(self, c, dev, rank, world, dist, graph=True, dropout_p=0.0):
        from imagecaptioning_b200 import _lib
        from imagecaptioning_b200.ddp import GradAllReducer
        from imagecaptioning_b200.model import DecoderTrainStep
        from oracle.synth import synth_inputs       # input generator shared with the oracle arm (not a compute path)
        self.c, self.dev, self.dist, self.world, self.graph = c, dev, dist, world, graph
        self.model = make_model(c, dev)
        if dropout_p > 0:
            self.model.dropout.p = dropout_p        # the reference's train() mode: nn.Dropout() p = 0.5 (AttentionLanguageModule.py:54)
        self.model.train()
        self.step = DecoderTrainStep(self.model, use_cuda_graph=graph)
        feat, gt = synth_inputs(c['B'], c['L'], c['D'], c['V'], c['T'], seed=123 + rank, full_length=True)
        self.feat_h, self.gt_h = feat.pin_memory(), gt.pin_memory()
        self.step.stage(self.feat_h.to(dev), self.gt_h.to(dev))
        self.reducer = GradAllReducer(self.step, world) if world > 1 else None
        self.lib = _lib.load()

    def one_step_resident(self):
        if self.reducer:             # gradient groups all-reduced on a side stream under the running step
            return self.reducer.step_overlapped()
        return self.step.run_resident()

    def barrier(self):
        if self.dist:
            self.dist.barrier()
        torch.cuda.synchronize()

    def launches_per_step(self):
        st = self.step
        st.use_cuda_graph = False
        n0 = self.lib.alexcap_launch_count(); st.run_resident(); torch.cuda.synchronize()
        n = int(self.lib.alexcap_launch_count() - n0)
        st.use_cuda_graph = self.graph
        return n

    def time_resident(self, steps, warmup, local=None):
        for _ in range(warmup):
            self.one_step_resident()
        self.barrier()
        clk = ClockSampler(local) if local is not None else None
        if clk:
            clk.__enter__()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            loss4 = self.one_step_resident()
        e.record()
        self.barrier()
        if clk:
            clk.__exit__()
        self.loss = float(loss4[0])
        return s.elapsed_time(e) / steps, (clk.summary() if clk else None)

    def time_e2e(self, steps):
        """Every step's inputs start in PINNED HOST memory, are copied H2D inside the timed region (prefetch(): on a side
        stream into the idle input set, overlapping the previous step's compute), the step runs (+ all-reduce), and the
        loss is read back D2H (a host sync) every step."""
        st = self.step
        st.prefetch(self.feat_h, self.gt_h)

        def one():
            st.swap()                                   # compute stream waits for this step's H2D copy
            st.prefetch(self.feat_h, self.gt_h)         # next step's inputs: H2D overlapped with this step
            return float(self.one_step_resident()[0].item())
        for _ in range(3):
            one()
        self.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            one()
        e.record()
        self.barrier()
        return s.elapsed_time(e) / steps

    def time_forward_train(self, steps):
        """The call INTEGRATION.md documents -- DecoderTrainStep.forward_train(features, gt_labels) with HOST tensors, exactly
        as the reference loop drives it (train_LSTMwAttention.py:78-96): zero_grad(set_to_none=True), forward_train, loss.item().
        No prefetch: the H2D copy is in line.  Single GPU only (the data-parallel caller is GradAllReducer)."""
        params = list(self.model.parameters())

        def one():
            for p in params:
                p.grad = None
            return float(self.step.forward_train(self.feat_h, self.gt_h).item())
        for _ in range(3):
            one()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            one()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / steps

    def close(self):
        if self.reducer:
            self.reducer.close()


def max_over_ranks(vals, dev, dist):
    t = torch.tensor(vals, device=dev, dtype=torch.float64)
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t]


def run_decode(args, c, dev, rank, world, local):
    """BASELINE.json configs[3]: greedy and beam-5 caption decoding, batch 1024, T_max = 20 (fp32, exact token ids are a
    parity test: tests/test_gpu_decode.py).  A 'step' is one decode of the whole batch; each token step streams Ws + feat once
    (B*L*(H+D)*4 bytes) -> the HBM figure.  Replicas only at N > 1 (no exchange step): every rank decodes its own batch."""
    from imagecaptioning_b200 import _lib
    from oracle.synth import synth_inputs
    model = make_model(c, dev)
    model.eval()
    with torch.no_grad():
        model.deep_output.bias[c['V'] + 2] += 6.0          # END reachable with random weights (as in the decode fixtures)
    feat, _ = synth_inputs(c['B'], c['L'], c['D'], c['V'], c['T'], seed=123 + rank, full_length=True)
    feat_h = feat.pin_memory()
    feat_d = feat_h.to(dev)
    empty = torch.empty(0, device=dev)
    lib = _lib.load()
    out = {}
    S = c['T'] + 1
    for name, beam in (('greedy', 0), ('beam5', 5)):
        model.use_beam = beam > 0
        model.beam_size = beam

        def one(f):
            ids, _ = model(f, empty)
            return ids
        for _ in range(max(args.warmup, 3)):
            ids = one(feat_d)
        torch.cuda.synchronize()
        n0 = lib.alexcap_launch_count()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(args.steps):
            ids = one(feat_d)
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / args.steps
        launches = int(lib.alexcap_launch_count() - n0) // args.steps
        # end to end: features from pinned host memory, ids back to the host
        s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s2.record()
        for _ in range(args.steps):
            ids_h = one(feat_h.to(dev, non_blocking=True)).cpu()
        e2.record(); torch.cuda.synchronize()
        ms2 = s2.elapsed_time(e2) / args.steps
        steps_run = S               # every token step's kernels run (the all-rows-emitted-END exit is evaluated on the device, no host sync)
        out[name] = dict(ms=ms, ms_e2e=ms2, launches=launches, token_steps=steps_run, d2h=int(ids_h.numel() * 8))
    if rank != 0:
        return
    pk = peaks()
    g, b5 = out['greedy'], out['beam5']
    step_bytes = c['B'] * c['L'] * (c['H'] + c['D']) * 4
    ach = step_bytes * g['token_steps'] / (g['ms'] * 1e-3) / 1e9
    line = dict(metric='captions/sec (attn-LSTM greedy decode, T_max=20)', value=round(c['B'] * world / (g['ms'] * 1e-3), 1),
                unit='captions/s', n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3), ms_per_step=round(g['ms'], 4),
                higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32', data='synthetic',
                config=dict(workload=f"LSTMwAttentionModel greedy / beam-5 decode (BASELINE.json configs[3]): B={c['B']}/GPU, L={c['L']}, "
                                     f"D={c['D']}, H={c['H']}, E={c['E']}, V={c['V']}, T_max={c['T']}, f32 (3xTF32 tensor-core engine, exact ids)",
                            global_batch=c['B'] * world, parallelism=f'replicas x{world}',
                            l2='inputs (Ws+feat = %d MB per token step) exceed the 126 MB L2' % (step_bytes >> 20)),
                e2e=dict(value=round(c['B'] * world / (g['ms_e2e'] * 1e-3), 1), unit='captions/s',
                         h2d_bytes_per_step=int(feat_h.numel() * 4), d2h_bytes_per_step=g['d2h'], ms_per_step=round(g['ms_e2e'], 4)),
                gpu_launches=g['launches'] * args.steps, gpu_launches_per_step=g['launches'],
                beam5=dict(value=round(c['B'] * world / (b5['ms'] * 1e-3), 1), unit='captions/s', ms_per_step=round(b5['ms'], 4),
                           e2e=round(c['B'] * world / (b5['ms_e2e'] * 1e-3), 1), gpu_launches_per_step=b5['launches']),
                roofline=dict(kernel='attn_fwd_stream_kernel (decode: one launch per token step)', bound='hbm', achieved=round(ach, 1),
                              peak=pk['hbm'], unit='GB/s', frac=round(ach / pk['hbm'], 4), traffic=None,
                              algorithmic_bytes_per_launch=step_bytes, token_steps=g['token_steps'],
                              note='whole greedy decode: token_steps x (Ws + feat) bytes over the decode time (GEMMs, argmax and '
                                   'bookkeeping included in the time, not in the bytes)', peak_source=pk['src']))
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--config', default='cfg2', choices=sorted(CONFIGS))
    ap.add_argument('--no-graph', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    c = CONFIGS[args.config]
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if args.impl == 'reference':
        run_reference(args, c, rank)
        return
    args.warmup = max(args.warmup, 3)

    assert torch.cuda.is_available(), 'bench.py (impl ours) needs a CUDA device: there is no CPU fallback'
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist = None
    saved_stdout = None
    if world > 1:
        # NCCL writes a "NCCL version ..." banner to stdout (fd 1) when the communicator is created: route fd 1 to stderr
        # until the result is ready, so that stdout carries exactly the one JSON line
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    if c.get('decode'):
        run_decode(args, c, dev, rank, world, local)
        if dist:
            dist.barrier(); sys.stdout.flush(); sys.stderr.flush(); os._exit(0)
        return

    run = TrainRun(c, dev, rank, world, dist, graph=not args.no_graph)
    step, model = run.step, run.model
    launches_per_step = run.launches_per_step()
    ms, clocks = run.time_resident(args.steps, args.warmup, local)
    loss_val = run.loss
    ms_e2e = run.time_e2e(args.steps)
    ms, ms_e2e = max_over_ranks([ms, ms_e2e], dev, dist)
    total = c['B'] * world
    extra = {}
    is_attn = c['model'] == 'LSTMwAttentionModel'

    if is_attn:
        # the reference's REAL training mode: nn.Dropout() p = 0.5 in train() (AttentionLanguageModule.py:54,79), mask drawn on
        # the device from a Philox counter that lives on the device -> same CUDA graph, fresh mask per replay
        drun = TrainRun(c, dev, rank, world, dist, graph=not args.no_graph, dropout_p=0.5)
        d_launch = drun.launches_per_step()
        d_ms, _ = drun.time_resident(args.steps, args.warmup)
        (d_ms,) = max_over_ranks([d_ms], dev, dist)
        extra['dropout_p05'] = dict(value=round(total / (d_ms * 1e-3), 1), unit='captions/s', ms_per_step=round(d_ms, 4),
                                    slowdown_vs_p0=round(d_ms / ms, 4), gpu_launches_per_step=d_launch, cuda_graph=not args.no_graph,
                                    loss=round(drun.loss, 5))
        drun.close(); del drun
    if world == 1 and is_attn:
        # north_star (3): the forward recurrence as ONE persistent weight-stationary kernel (time_loop.cu: LSTM pointwise in the
        # tcgen05 epilogue, next query GEMM behind a grid barrier).  Built, parity-tested, and measured here next to the default
        # PDL-chained per-step launches, which are faster on this shape -- so it is opt-in.
        from imagecaptioning_b200 import functional as F_
        prev = F_.set_time_loop(True)
        try:
            lrun = TrainRun(c, dev, rank, world, dist, graph=not args.no_graph)
            l_launch = lrun.launches_per_step()
            l_ms, _ = lrun.time_resident(args.steps, args.warmup)
            extra['time_loop_kernel'] = dict(ms_per_step=round(l_ms, 4), value=round(total / (l_ms * 1e-3), 1), unit='captions/s',
                                             gpu_launches_per_step=l_launch, loss=round(lrun.loss, 5),
                                             used=bool(l_launch < launches_per_step),
                                             what='alexcap_set_time_loop(1): persistent forward-recurrence kernel instead of 4 launches per time step')
            del lrun
        finally:
            F_.set_time_loop(prev)
    if world == 1:
        ft = run.time_forward_train(args.steps)
        extra['e2e_forward_train'] = dict(
            value=round(total / (ft * 1e-3), 1), unit='captions/s', ms_per_step=round(ft, 4),
            what='DecoderTrainStep.forward_train(features_host_pinned, gt_host) + loss.item() per step, zero_grad(set_to_none) '
                 'before each (the call INTEGRATION.md documents; in-line H2D copy, no prefetch)',
            h2d_bytes_per_step=int(run.feat_h.numel() * 4 + run.gt_h.numel() * 4), d2h_bytes_per_step=4)
    if world > 1 and is_attn:
        # correctness of the overlapped exchange ON THIS HARDWARE: ragged captions, every rank's averaged gradients vs the
        # single-process step on the concatenated batch; the all-reduced loss (tests/ddp_worker.py)
        from tests.ddp_worker import ddp_check
        cc = dict(c); cc['B'] = min(c['B'], 64)
        worst, loss_rel, spread, nv_ddp, nv_one = ddp_check(dev, rank, world, cc, graph=not args.no_graph)
        gl = run.reducer.global_loss()
        extra['ddp_check'] = dict(grad_rel_vs_single_process=float('%.3e' % worst), loss_rel=float('%.3e' % loss_rel),
                                  rank_spread=float('%.3e' % spread), n_valid_sum_over_ranks=nv_ddp, n_valid_single_process=nv_one,
                                  shape='B=%d/rank ragged captions, rest as the workload' % cc['B'],
                                  global_loss=round(float(gl[0]), 5))
        if args.config == 'cfg2':
            # the configuration BASELINE.json names for data parallelism (configs[2]): bf16 mode, B = 512 per GPU
            c3 = CONFIGS['cfg3']
            r3 = TrainRun(c3, dev, rank, world, dist, graph=not args.no_graph)
            m3, _ = r3.time_resident(max(5, args.steps // 2), args.warmup)
            m3e = r3.time_e2e(max(5, args.steps // 2))
            m3, m3e = max_over_ranks([m3, m3e], dev, dist)
            extra['cfg3'] = dict(value=round(c3['B'] * world / (m3 * 1e-3), 1), unit='captions/s', ms_per_step=round(m3, 4),
                                 e2e=round(c3['B'] * world / (m3e * 1e-3), 1), dtype='bf16', config=workload_config(c3, world),
                                 loss=round(r3.loss, 5))
            r3.close(); del r3

    if rank == 0:
        pk = peaks()
        line = dict(metric=METRIC, value=round(total / (ms * 1e-3), 1), unit='captions/s', n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=round(ms, 4), higher_is_better=True, scaling='weak', vs_baseline=None,
                    dtype=c['dtype'], data='synthetic', config=workload_config(c, world), loss=round(loss_val, 5),
                    cuda_graph=not args.no_graph,
                    e2e=dict(value=round(total / (ms_e2e * 1e-3), 1), unit='captions/s',
                             h2d_bytes_per_step=int(run.feat_h.numel() * 4 + run.gt_h.numel() * 4), d2h_bytes_per_step=4,
                             ms_per_step=round(ms_e2e, 4)),
                    gpu_launches=launches_per_step * args.steps, gpu_launches_per_step=launches_per_step,
                    clocks=clocks)
        line.update(extra)
        w = algorithmic_work(c)
        line['gemm_group'] = dict(algorithmic_gflop_per_step=round(w['gemm_flops_step'] / 1e9, 2),
                                  engine='tcgen05 bf16x3: hi/lo bf16 operand pairs, 3 MMAs per product, fp32 accumulate (fp32 in/out, 1e-4 parity)' if c['dtype'] == 'f32' else
                                  'tcgen05 bf16 operands, fp32 accumulate (2e-2 parity)')
        if is_attn:
            line['roofline'] = dominant_kernel_roofline(c, dev, pk)
            line['roofline_gemm'] = gemm_roofline(c, dev, pk)
        if world == 1 and is_attn:
            line['next_row_optimizer'] = optimizer_step_timing(model, step, dev)
        if world == 1 and not args.no_cpu_baseline:
            cb, _, _ = time_cpu_port(c, steps=3, warmup=1, budget_s=25.0)
            line['cpu_baseline'] = cb
        if saved_stdout is not None:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    reducer = run.reducer
    if dist:
        # the whole-step CUDA graphs hold captured NCCL kernels: drop them before the communicator, and do not let the
        # communicator's teardown outlive the measurement (every rank has finished its timed region by the barrier)
        reducer.close()
        torch.cuda.synchronize()
        dist.barrier()
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)


if __name__ == '__main__':
    main()
