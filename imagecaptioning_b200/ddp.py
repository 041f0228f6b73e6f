"""Data-parallel gradient exchange: one process per GPU (torchrun), NCCL all-reduce over NVLink 5 / NVSwitch.

The reference is single-GPU (AlexCap/LSTM_opts.py:12-13); data parallelism is new work required by BASELINE.json.
Every caption/image pair is independent, so the batch is sharded by rank with parameters replicated and the ONLY
exchange is the gradient all-reduce.  DecoderTrainStep keeps all gradients in one flat fp32
buffer laid out in gradient-finalisation order, so buckets are contiguous slices (no flatten/unflatten copies):
    group 0  deep_output.{weight,bias}   final after the vocab/CE backward, BEFORE the reverse-time chain starts (phase 1)
    group 1  lstm.*, embedding           final right after the chain (phase 4)
    group 2  everything else             final at the end (phase 8)
`allreduce()` averages the whole buffer; `step_overlapped()` all-reduces group g on a side stream while phase g+1 runs.

Ragged captions: CustomLoss.py:7 is a mean over the NON-IGNORED targets, so a rank's CE divisor must be the global
N_valid / world for the AVERAGED gradient to equal the single-process gradient on the concatenated batch.
`step_overlapped()` therefore counts the local N_valid from the staged captions (alexcap_count_valid_targets), averages that
one float over the ranks on the side stream while the forward recurrence (phase 16) runs, and hands it to phase 32
(vocabulary projection + loss) as `ce_denominator` -- all inside the captured whole-step graph.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


class GradAllReducer:
    def __init__(self, step, world_size: int, group=None):
        self.step, self.world, self.group = step, world_size, group
        self.side = torch.cuda.Stream() if torch.cuda.is_available() else None
        self._graphs = {}            # input set -> CUDA graph of the whole data-parallel step (kernels + NCCL)
        self._graph_ok = os.environ.get('ALEXCAP_DDP_GRAPH', '1') != '0'

    def close(self):
        """Destroy the captured whole-step graphs (they reference the NCCL communicator): call before
        dist.destroy_process_group()."""
        torch.cuda.synchronize()
        self._graphs.clear()

    def buckets(self):
        return self.step.grad_buckets()

    def allreduce(self):
        if self.world <= 1:
            return
        for b in self.buckets():
            dist.all_reduce(b, op=dist.ReduceOp.AVG if b.is_cuda else dist.ReduceOp.SUM, group=self.group)
            if not b.is_cuda:           # gloo (CPU tests) has no AVG
                b.div_(self.world)

    def _reduce(self, b, gi=None):
        if gi is not None and str(gi) in os.environ.get('ALEXCAP_DDP_SKIP', ''):     # timing experiments only
            return
        dist.all_reduce(b, op=dist.ReduceOp.AVG if b.is_cuda else dist.ReduceOp.SUM, group=self.group)
        if not b.is_cuda:               # gloo (CPU tests) has no AVG
            b.div_(self.world)

    def step_overlapped(self):
        """One data-parallel training step: group 0 (deep_output.*) is all-reduced on a side stream WHILE the reverse-time
        chain runs, group 1 (lstm.*, embedding) WHILE the remaining gradients are computed; only the small last group
        is exposed.  Returns the loss tensor."""
        step = self.step
        if self.world <= 1:
            return step.run_resident()
        if not step.is_attn:                                 # plain LSTM: one phase; the divisor exchange is exposed (1 float)
            from . import functional as F_
            F_.count_valid_targets(step._gt, out=step._slot.n_valid)
            self._reduce(step._slot.n_valid)
            step.ce_denominator = step._slot.n_valid
            loss = step.run_resident()
            self.allreduce()
            return loss
        step.ce_denominator = step._slot.n_valid            # (part of the graph key: set before any launch of this step)
        if self._graph_ok and step.use_cuda_graph:
            # the whole step -- the phase launches, the side-stream forks / joins and the NCCL all-reduces -- as ONE graph
            step._sync_dims()
            key = step._graph_key('ddp')
            graphs = self._graphs.setdefault(id(step._slot), {})
            if key in graphs:
                graphs[key].replay()
                return step._loss
            loss = self._run_phases(step, graphs=False)      # this call's step, eagerly (also warms up NCCL and the kernels)
            try:
                torch.cuda.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._run_phases(step, graphs=False)
                graphs[key] = g
            except Exception as e:                            # capture of the collectives not supported here: phase graphs
                self._graph_ok = False
                print(f'[imagecaptioning_b200.ddp] whole-step graph capture failed ({type(e).__name__}: {e}); '
                      'falling back to per-phase graphs', flush=True)
            return loss
        return self._run_phases(step, graphs=True)

    def _run_phases(self, step, graphs):
        from . import functional as F_
        main = torch.cuda.current_stream()
        groups = step.grad_groups()
        run = step.run_resident if graphs else (lambda phase: step._launch(phase))
        # N_valid_global / world -> phase 32's CE divisor; exchanged under the forward recurrence
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            F_.count_valid_targets(step._gt, out=step._slot.n_valid)
            self._reduce(step._slot.n_valid)
        run(phase=16)
        main.wait_stream(self.side)
        # vocabulary projection + loss + dH on the main stream (phase 128); d deep_output.* (phase 64: restricted to the SMs
        # the reverse loop leaves idle) and its all-reduce follow on the side stream, underneath phase 4
        run(phase=128)
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            run(phase=64)
            if groups[0].numel():
                self._reduce(groups[0], 0)
        run(phase=4)
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            if groups[1].numel():
                self._reduce(groups[1], 1)
        run(phase=8)
        if groups[2].numel():
            self._reduce(groups[2], 2)
        main.wait_stream(self.side)
        return step._loss

    def global_loss(self):
        """The data-parallel loss of the last step: CE terms sum to the global mean (each rank's CE is already divided by
        N_valid_global / world), the regulariser is a per-rank mean over equal B*L.  Returns a new 4-vector (AVG over ranks)."""
        t = self.step._loss.clone()
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.group)
        return t

    def global_ce_denominator(self, n_valid_local: torch.Tensor) -> torch.Tensor:
        """N_valid_global / world: with it as the CE divisor on every rank, the AVERAGED gradient equals the
        single-process gradient on the concatenated batch."""
        t = n_valid_local.clone().float()
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t / self.world
