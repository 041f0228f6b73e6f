"""Per-tensor relative error of the fused train step vs the fp64 oracle, for both GEMM engines."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import restated as R
from oracle.synth import synth_inputs
from tests.gpu_util import build_decoder, random_attn_params
from tests.util import rel_err
from imagecaptioning_b200 import functional as F_
from imagecaptioning_b200.model import DecoderTrainStep
cfg = dict(B=24, L=49, D=256, H=192, E=96, V=700, T=9)
p = random_attn_params(cfg, seed=8)
feat, gt = synth_inputs(cfg['B'], cfg['L'], cfg['D'], cfg['V'], cfg['T'], seed=5, dtype=torch.float64)
ref = R.attn_train_step(p, feat, gt, cfg['V'], need_dfeat=True)
for eng in (0, 1, 3):
    F_.set_gemm_engine(eng)
    dec = build_decoder(cfg, p)
    step = DecoderTrainStep(dec, need_dfeat=True)
    loss = step.forward_train(feat.float().cuda(), gt.cuda())
    print('engine', eng, 'loss rel', abs(float(loss) - float(ref['loss'])) / float(ref['loss']))
    for k, q in dec.named_parameters():
        print(f'   {k:24s} {rel_err(q.grad, ref["grads"][k]):.2e}  max|ref| {float(ref["grads"][k].abs().max()):.2e}')
    print('   dfeat', rel_err(step.dfeat, ref['dfeat']))
