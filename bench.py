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
    e = 4
    attn_fwd_step = B * L * (H + D) * e + B * L * 4 + B * (H + D) * e
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
    """The fused attention-step kernels (forward and backward: 34 launches per training step, each streaming Ws and
    feat once) timed alone at the bench shape.  S distinct q/bpre slices per pass over the same Ws/feat, exactly as
    inside a training step; launches replayed from a CUDA graph, CUDA events on the launching stream."""
    from imagecaptioning_b200 import functional as F_
    B, L, D, H, S = c['B'], c['L'], c['D'], c['H'], c['T'] + 1
    g = torch.Generator(device='cpu').manual_seed(1)
    Ws = torch.randn(B, L, H, generator=g).to(dev)
    feat = torch.randn(B, L, D, generator=g).relu().to(dev)
    q = torch.randn(S, B, H, generator=g).to(dev)
    bp = torch.randn(S, B, D, generator=g).to(dev)
    dg = torch.randn(S, B, D, generator=g).to(dev)
    wv = (torch.randn(H, generator=g) / H ** 0.5).to(dev)
    outs = [F_.attn_step_fwd(Ws, feat, q[t], bp[t], wv) for t in range(S)]
    state = {'t': 0}

    def fwd():
        t = state['t'] = (state['t'] + 1) % S
        F_.attn_step_fwd(Ws, feat, q[t], bp[t], wv)

    def bwd():
        t = state['t'] = (state['t'] + 1) % S
        a, cx, be, _ = outs[t]
        F_.attn_step_bwd(Ws, feat, q[t], wv, a, cx, be, dg[t])

    ms_f = time_graph_events(fwd, S, 12)
    ms_b = time_graph_events(bwd, S, 12)
    w = algorithmic_work(c)
    ach = w['attn_fwd_bytes_per_launch'] / (ms_f * 1e-3) / 1e9
    ach_b = w['attn_bwd_bytes_per_launch'] / (ms_b * 1e-3) / 1e9
    return dict(kernel='attn_fwd_stream_kernel', bound='hbm', achieved=round(ach, 1), peak=pk['hbm'], unit='GB/s',
                frac=round(ach / pk['hbm'], 4), traffic=NCU_TRAFFIC.get('attn_fwd') if c is CONFIGS['cfg2'] else None, ms_per_launch=round(ms_f, 5),
                algorithmic_bytes_per_launch=w['attn_fwd_bytes_per_launch'], peak_source=pk['src'],
                backward=dict(kernel='attn_bwd_stream_kernel', achieved=round(ach_b, 1), frac=round(ach_b / pk['hbm'], 4),
                              ms_per_launch=round(ms_b, 5), algorithmic_bytes_per_launch=w['attn_bwd_bytes_per_launch'],
                              traffic=NCU_TRAFFIC.get('attn_bwd') if c is CONFIGS['cfg2'] else None),
                note='Ws+feat = %.0f MB per launch, re-read by each of the 2*S launches of a step; with a 126 MB L2 part '
                     'of it can hit L2 between consecutive launches (ncu, cache flushed: dram__bytes_read = algorithmic '
                     'bytes, see profiles/)' % (2 * B * L * max(H, D) * 4 / 1e6))


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
    return dict(kernel='ps::gemm_ps_kernel', bound='tensor', achieved=round(tf, 1), peak=pk['bf16'], unit='TFLOP/s',
                frac=round(tf / pk['bf16'], 4), traffic=NCU_TRAFFIC.get('gemm_ps_logits') if (c is CONFIGS['cfg2']) else None,
                ms_per_launch=round(ms, 5), shape=[M, N, K],
                mma_per_product=1 if bf16 else 3, issued_frac=round((1 if bf16 else 3) * tf / pk['bf16'], 4),
                with_split_ms=round(ms_split, 5), convert_in_kernel_ms=round(ms_conv, 5), peak_source=pk['src'],
                note='operands pre-split into bf16 hi/lo copies in HBM, kernel = TMA -> smem -> tcgen05.mma only; bound by '
                     'L2 -> shared-memory operand traffic of 128-wide tiles (~10 TB/s), not by the tensor pipe; '
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


def workload_config(c, n):
    return dict(workload=f"{c['model']} decoder train step fwd+bwd ({c.get('ref', 'BASELINE.json configs[1]')}): B={c['B']}/GPU, "
                         f"L={c['L']}, D={c['D']}, H={c['H']}, E={c['E']}, V={c['V']}, T={c['T']}, {c['dtype']}, full-length captions, "
                         f"dropout p=0",
                global_batch=c['B'] * n, parallelism=f'dp{n}',
                l2='no explicit flush: per-step working set (Ws+feat+logits+saved state ~ 400 MB) exceeds the 126 MB L2')


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
    from imagecaptioning_b200 import _lib
    from imagecaptioning_b200.ddp import GradAllReducer
    from imagecaptioning_b200.model import DecoderTrainStep
    from oracle.synth import synth_inputs       # input generator shared with the oracle arm (not a compute path)

    model = make_model(c, dev)
    step = DecoderTrainStep(model, use_cuda_graph=not args.no_graph)
    feat, gt = synth_inputs(c['B'], c['L'], c['D'], c['V'], c['T'], seed=123 + rank, full_length=True)
    feat_h, gt_h = feat.pin_memory(), gt.pin_memory()
    step.stage(feat_h.to(dev), gt_h.to(dev))
    reducer = GradAllReducer(step, world) if world > 1 else None
    lib = _lib.load()

    def one_step_resident():
        if reducer:                  # all-reduce of the vocab-projection bucket overlapped with the reverse chain
            return reducer.step_overlapped()
        return step.run_resident()

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    # launches per step (counted on an un-graphed step)
    step.use_cuda_graph = False
    n0 = lib.alexcap_launch_count(); step.run_resident(); torch.cuda.synchronize()
    launches_per_step = int(lib.alexcap_launch_count() - n0)
    step.use_cuda_graph = not args.no_graph

    for _ in range(args.warmup):
        one_step_resident()
    barrier()
    with ClockSampler(local) as clk:
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(args.steps):
            loss4 = one_step_resident()
        e.record()
        barrier()
    ms = s.elapsed_time(e) / args.steps
    loss_val = float(loss4[0])

    # end to end through the public API: every step's inputs start in PINNED HOST memory, are copied H2D inside the
    # timed region (prefetch(): on a side stream into the idle input set, overlapping the previous step's compute),
    # the step runs (+ all-reduce), and the loss is read back D2H (a host sync) every step.
    step.prefetch(feat_h, gt_h)

    def one_step_e2e():
        step.swap()                          # compute stream waits for this step's H2D copy
        step.prefetch(feat_h, gt_h)          # next step's inputs: H2D overlapped with this step
        l4 = one_step_resident()
        return float(l4[0].item())           # D2H read of the loss
    for _ in range(3):
        one_step_e2e()
    barrier()
    s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s2.record()
    for _ in range(args.steps):
        one_step_e2e()
    e2.record()
    barrier()
    ms_e2e = s2.elapsed_time(e2) / args.steps

    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if dist:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])

    if rank == 0:
        pk = peaks()
        total = c['B'] * world
        line = dict(metric=METRIC, value=round(total / (ms * 1e-3), 1), unit='captions/s', n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=round(ms, 4), higher_is_better=True, scaling='weak', vs_baseline=None,
                    dtype=c['dtype'], data='synthetic', config=workload_config(c, world), loss=round(loss_val, 5),
                    cuda_graph=not args.no_graph,
                    e2e=dict(value=round(total / (ms_e2e * 1e-3), 1), unit='captions/s',
                             h2d_bytes_per_step=int(feat_h.numel() * 4 + gt_h.numel() * 4), d2h_bytes_per_step=4,
                             ms_per_step=round(ms_e2e, 4)),
                    gpu_launches=launches_per_step * args.steps, gpu_launches_per_step=launches_per_step,
                    clocks=clk.summary())
        w = algorithmic_work(c)
        line['gemm_group'] = dict(algorithmic_gflop_per_step=round(w['gemm_flops_step'] / 1e9, 2),
                                  engine='tcgen05 bf16x3: hi/lo bf16 operand pairs, 3 MMAs per product, fp32 accumulate (fp32 in/out, 1e-4 parity)' if c['dtype'] == 'f32' else
                                  'tcgen05 bf16 operands, fp32 accumulate (2e-2 parity)')
        if c['model'] == 'LSTMwAttentionModel':
            line['roofline'] = dominant_kernel_roofline(c, dev, pk)
            line['roofline_gemm'] = gemm_roofline(c, dev, pk)
        if world == 1 and c['model'] == 'LSTMwAttentionModel':
            line['next_row_optimizer'] = optimizer_step_timing(model, step, dev)
        if world == 1 and not args.no_cpu_baseline:
            cb, _, _ = time_cpu_port(c, steps=3, warmup=1, budget_s=25.0)
            line['cpu_baseline'] = cb
        if saved_stdout is not None:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
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
