"""Run in a subprocess (ALEXCAP_BF16_MIRROR is read once per process): one bf16-mode training step whose per-step GEMMs are
large enough for the pre-split engine (M = B = 256, H = D = 1024), printing the loss, the flat gradient buffer's checksum per
parameter and the launch count -- tests/test_gpu_round2.py compares the producer-written operand mirrors with the split passes."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from oracle.synth import synth_inputs
from tests.gpu_util import DEV, build_decoder, random_attn_params


def main():
    from imagecaptioning_b200 import _lib
    from imagecaptioning_b200.model import DecoderTrainStep
    lib = _lib.load()
    cfg = dict(B=256, L=6, D=1024, H=1024, E=64, V=300, T=3)
    p = random_attn_params(cfg, seed=3, dtype=torch.float32)
    feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=4)
    dec = build_decoder(cfg, p)
    dec.compute_dtype = 'bf16'
    dec.train()
    st = DecoderTrainStep(dec, use_cuda_graph=False, need_dfeat=True)
    for _ in range(2):
        for q in dec.parameters():
            q.grad = None
        n0 = lib.alexcap_launch_count()
        loss = st.forward_train(feat.to(DEV), gt.to(DEV))
        torch.cuda.synchronize()
        n1 = lib.alexcap_launch_count()
    out_path = sys.argv[1]
    torch.save(dict(loss=float(loss), launches=int(n1 - n0), S=cfg['T'] + 1, flat=st._flat_grads.cpu(), dfeat=st.dfeat.cpu()), out_path)
    print('RESULT ' + json.dumps(dict(loss=float(loss), launches=int(n1 - n0))))


if __name__ == '__main__':
    main()
