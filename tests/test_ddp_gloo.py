"""World-size-2 `gloo` test (CPU) of the data-parallel host logic (imagecaptioning_b200/ddp.py): bucketed gradient
all-reduce over the flat buffer + the global CE denominator.  The per-rank gradients come from the CPU oracle (there is
no CPU product path); the check is the property SURVEY.md section 8e states: the AVERAGED per-rank gradients, with
N_valid_global/world as every rank's CE divisor, equal the single-process gradient on the concatenated batch."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import restated as R
from oracle.synth import synth_inputs
from tests.gpu_util import random_attn_params

CFG = dict(V=40, E=8, D=12, H=16, T=5)
B, L = 6, 7


class _FakeStep:
    """What GradAllReducer needs from DecoderTrainStep: contiguous bucket slices of one flat gradient buffer."""

    def __init__(self, flat, first):
        self.flat, self.first, self.is_attn = flat, first, True

    def grad_buckets(self, bucket_elems=200):
        out, o = [self.flat[:self.first]], self.first
        while o < self.flat.numel():
            out.append(self.flat[o:o + bucket_elems])
            o += bucket_elems
        return out


def _rank_grads(p, feat, gt, ce_den=None):
    in_tok, tgt = R.make_targets(gt, CFG['V'])
    sv = R.attn_forward(p, feat, in_tok)
    Bq, S, Vp = sv['preds'].shape
    loss_ce, dlog, n_valid = R.ce_loss_and_grad(sv['preds'].reshape(Bq * S, Vp), tgt.reshape(-1))
    if ce_den is not None:
        dlog = dlog * (n_valid / float(ce_den))
    resid = 1.0 - sv['alphas'].sum(1)
    dal = (-2.0 * resid / (Bq * feat.shape[1])).unsqueeze(1).expand(Bq, S, feat.shape[1])
    grads, _ = R.attn_backward(p, sv, dlog.view(Bq, S, Vp), dal)
    return grads, n_valid


def _worker(rank, world, port, out_path):
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from imagecaptioning_b200.ddp import GradAllReducer
    p = random_attn_params(CFG, seed=11)
    feat, gt = synth_inputs(B, L, CFG['D'], CFG['V'], CFG['T'], seed=5)          # ragged captions: N_valid differs per rank
    sl = slice(rank * B // world, (rank + 1) * B // world)
    names = sorted(p, key=lambda k: (0 if k.startswith('deep_output.') else 1, k))  # vocab projection first
    red = GradAllReducer(_FakeStep(torch.zeros(1), 1), world)
    _, n_local = _rank_grads(p, feat[sl].double(), gt[sl])
    den = red.global_ce_denominator(torch.tensor(float(n_local)))
    g, _ = _rank_grads(p, feat[sl].double(), gt[sl], ce_den=float(den))
    flat = torch.cat([g[k].reshape(-1) for k in names])
    first = sum(g[k].numel() for k in names if k.startswith('deep_output.'))
    red = GradAllReducer(_FakeStep(flat, first), world)
    assert len(red.buckets()) > 3
    red.allreduce()
    if rank == 0:
        torch.save(dict(flat=flat, names=names, den=float(den)), out_path)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_allreduce_equals_single_process(tmp_path):
    s = socket.socket(); s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / 'rank0.pt')
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = torch.load(out)
    p = random_attn_params(CFG, seed=11)
    feat, gt = synth_inputs(B, L, CFG['D'], CFG['V'], CFG['T'], seed=5)
    ref, n_valid = _rank_grads(p, feat.double(), gt)
    assert abs(got['den'] * 2 - n_valid) < 1e-9
    want = torch.cat([ref[k].reshape(-1) for k in got['names']])
    # the regulariser is a mean over B*L, equal on both ranks: averaging it is exact as well
    assert float((got['flat'] - want).abs().max()) < 1e-12 * max(1.0, float(want.abs().max()))
