"""Worker of tests/test_gpu_ddp.py and of `bench.py`'s ddp_check: run under torchrun with >= 2 GPUs (one rank per GPU, NCCL).

Property (CustomLoss.py:7 is a mean over the NON-IGNORED targets): with ragged captions, the gradients
GradAllReducer.step_overlapped() leaves on every rank -- three phase groups all-reduced on a side stream under the running
step, N_valid exchanged under the forward recurrence, the whole thing replayed from one captured CUDA graph -- equal the
single-process gradients of the concatenated batch, and the averaged loss equals the single-process loss.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def ddp_check(dev, rank, world, cfg, graph=True, seed=31, steps=3):
    """Returns (max tensor-scale relative gradient difference, relative loss difference) on every rank."""
    from imagecaptioning_b200.ddp import GradAllReducer
    from imagecaptioning_b200.model import DecoderTrainStep
    from imagecaptioning_b200.modules import Decoder
    from oracle.synth import synth_inputs          # seeded input generator (not a compute path)
    B = cfg['B']
    feat, gt = synth_inputs(B * world, cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=seed)      # ragged: N_valid differs per rank
    torch.manual_seed(123)
    dec = Decoder(cfg['V'], cfg['E'], cfg['D'], cfg['H'], cfg['T'], {}, 0.0, dev).to(dev)
    dec.dropout.p = 0.0
    dec.train()
    step = DecoderTrainStep(dec, use_cuda_graph=graph)
    sl = slice(rank * B, (rank + 1) * B)
    step.stage(feat[sl].to(dev), gt[sl].to(dev))
    red = GradAllReducer(step, world)
    for _ in range(steps):                          # eager step, capture, replay
        red.step_overlapped()
    torch.cuda.synchronize()
    got = step._flat_grads.clone()
    loss_ddp = red.global_loss()
    # the same step, single process, concatenated batch (a second model with the same parameters)
    torch.manual_seed(123)
    dec1 = Decoder(cfg['V'], cfg['E'], cfg['D'], cfg['H'], cfg['T'], {}, 0.0, dev).to(dev)
    dec1.dropout.p = 0.0
    dec1.train()
    one = DecoderTrainStep(dec1)
    one.stage(feat.to(dev), gt.to(dev))
    loss_one = one.run_resident().clone()
    torch.cuda.synchronize()
    worst = 0.0
    from imagecaptioning_b200._lib import ATTN_STATE_NAMES
    detail = {}
    for name, a, b in zip(ATTN_STATE_NAMES, step._grads, one._grads):
        if name == 'attention.v.bias':          # identically zero (softmax shift invariance): no scale to be relative to
            continue
        scale = max(float(b.abs().max()), 1e-30)
        e = float((a - b).abs().max()) / scale
        detail[name] = e
        worst = max(worst, e)
    if rank == 0 and worst > 1e-4:
        sys.stderr.write('DDP_DETAIL graph=%s %s\n' % (graph, json.dumps({k: float('%.2e' % v) for k, v in detail.items()})))
    loss_rel = abs(float(loss_ddp[0]) - float(loss_one[0])) / abs(float(loss_one[0]))
    # every rank must hold the same averaged gradients
    chk = got.double().abs().sum().reshape(1)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    red.close()
    return worst, loss_rel, float(hi - lo) / max(float(hi), 1e-30), float(loss_ddp[3]) * world, float(loss_one[3])


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    os.dup2(2, 1)                                   # NCCL banner off stdout
    dist.init_process_group('nccl', device_id=dev)
    cfg = dict(B=48, L=49, D=256, H=192, E=96, V=2100, T=9)
    out = {}
    ok = True
    for graph in (False, True):
        worst, loss_rel, spread, nv_ddp, nv_one = ddp_check(dev, rank, world, cfg, graph=graph)
        out['graph' if graph else 'eager'] = dict(grad_rel=worst, loss_rel=loss_rel, rank_spread=spread,
                                                  n_valid_sum=nv_ddp, n_valid_single=nv_one)
        ok = ok and worst < 1e-4 and loss_rel < 1e-5 and spread < 1e-12
    if rank == 0:
        sys.stderr.write('DDP_CHECK ' + json.dumps(out) + '\n')
    dist.barrier()
    torch.cuda.synchronize()
    sys.stderr.flush()
    os._exit(0 if ok else 3)


if __name__ == '__main__':
    main()
