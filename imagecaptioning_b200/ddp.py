"""Data-parallel gradient exchange: one process per GPU (torchrun), NCCL all-reduce over NVLink 5 / NVSwitch.

The reference is single-GPU (AlexCap/LSTM_opts.py:12-13); data parallelism is new work required by BASELINE.json.
Every caption/image pair is independent, so the batch is sharded by rank with parameters replicated and the ONLY
exchange is the gradient all-reduce.  DecoderTrainStep keeps all gradients in one flat fp32
buffer laid out in gradient-finalisation order, so buckets are contiguous slices (no flatten/unflatten copies):
    bucket 0  deep_output.{weight,bias}   final after the vocab/CE backward, BEFORE the reverse-time chain starts
    bucket 1+ everything else             final after the chain
`allreduce()` averages the whole buffer; `allreduce_bucket(i)` lets a caller overlap bucket 0 with the chain.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class GradAllReducer:
    def __init__(self, step, world_size: int, group=None):
        self.step, self.world, self.group = step, world_size, group
        self.side = torch.cuda.Stream() if torch.cuda.is_available() else None

    def buckets(self):
        return self.step.grad_buckets()

    def allreduce(self):
        if self.world <= 1:
            return
        for b in self.buckets():
            dist.all_reduce(b, op=dist.ReduceOp.AVG if b.is_cuda else dist.ReduceOp.SUM, group=self.group)
            if not b.is_cuda:           # gloo (CPU tests) has no AVG
                b.div_(self.world)

    def _reduce(self, b):
        dist.all_reduce(b, op=dist.ReduceOp.AVG if b.is_cuda else dist.ReduceOp.SUM, group=self.group)
        if not b.is_cuda:               # gloo (CPU tests) has no AVG
            b.div_(self.world)

    def step_overlapped(self):
        """One data-parallel training step with the first gradient bucket (deep_output.*: ~1/3 of the bytes, final
        after the vocabulary backward) all-reduced on a side stream WHILE the reverse-time chain runs; the remaining
        buckets follow the chain on the main stream.  Returns the loss tensor."""
        step = self.step
        if self.world <= 1 or not step.is_attn:
            loss = step.run_resident()
            self.allreduce()
            return loss
        main = torch.cuda.current_stream()
        step.run_resident(phase=1)
        ready = torch.cuda.Event()
        ready.record(main)
        bks = self.buckets()
        self.side.wait_event(ready)
        with torch.cuda.stream(self.side):
            self._reduce(bks[0])
            done = torch.cuda.Event()
            done.record(self.side)
        loss = step.run_resident(phase=2)
        for b in bks[1:]:
            self._reduce(b)
        main.wait_event(done)
        return loss

    def global_ce_denominator(self, n_valid_local: torch.Tensor) -> torch.Tensor:
        """N_valid_global / world: with it as the CE divisor on every rank, the AVERAGED gradient equals the
        single-process gradient on the concatenated batch."""
        t = n_valid_local.clone().float()
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t / self.world
