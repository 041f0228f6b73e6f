"""Run in a subprocess with ALEXCAP_LSTM_Q_FUSED=1 (the switch is read once per process: the workspace layout depends on it):
the training step with the LSTM cell fused into the next step's query GEMM (gemm_tc_lstm_f32) against the fp64 oracle."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from oracle import restated as R
from oracle.synth import synth_inputs
from tests.gpu_util import DEV, build_decoder, random_attn_params
from tests.util import rel_err


def main():
    from imagecaptioning_b200 import _lib
    from imagecaptioning_b200.model import DecoderTrainStep
    lib = _lib.load()
    out = []
    for cfg, graph in ((dict(B=24, L=49, D=256, H=192, E=96, V=700, T=9), False),
                       (dict(B=130, L=12, D=128, H=256, E=64, V=300, T=4), True)):       # two row tiles, ragged M edge
        p = random_attn_params(cfg, seed=21)
        feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=22, dtype=torch.float64)
        ref = R.attn_train_step(p, feat, gt, cfg['V'], need_dfeat=True)
        dec = build_decoder(cfg, p)
        dec.train()
        st = DecoderTrainStep(dec, use_cuda_graph=graph, need_dfeat=True)
        for _ in range(2):
            for q in dec.parameters():
                q.grad = None
            n0 = lib.alexcap_launch_count()
            loss = st.forward_train(feat.float().to(DEV), gt.to(DEV))
            n1 = lib.alexcap_launch_count()
        worst = max(rel_err(q.grad, ref['grads'][k]) for k, q in dec.named_parameters() if k != 'attention.v.bias')
        out.append(dict(loss_rel=abs(float(loss) - float(ref['loss'])) / float(ref['loss']), grad_rel=worst,
                        dfeat_rel=rel_err(st.dfeat, ref['dfeat']), launches=int(n1 - n0), S=cfg['T'] + 1, graph=graph))
    print('RESULT ' + json.dumps(out))


if __name__ == '__main__':
    main()
