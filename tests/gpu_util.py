"""Helpers shared by the -m gpu parity tests (they call the CUDA path through the C ABI and check it against the
CPU oracle / committed golden vectors)."""
import torch

from imagecaptioning_b200.modules import Decoder, LanguageModel

DEV = 'cuda'


def build_decoder(cfg, params, dropout_p=0.0):
    dec = Decoder(cfg['V'], cfg['E'], cfg['D'], cfg['H'], cfg['T'], {}, 0.0, DEV)
    dec.load_state_dict({k: v.float() for k, v in params.items()})
    dec.dropout.p = dropout_p
    return dec.to(DEV)


def build_lstm(cfg, params, num_layers=1, dropout_p=0.0):
    lm = LanguageModel(cfg['V'], cfg['E'], cfg['D'], cfg['H'], cfg['T'], num_layers, {}, dropout_p, DEV)
    lm.load_state_dict({k: v.float() for k, v in params.items()})
    return lm.to(DEV)


def random_attn_params(cfg, seed=0, dtype=torch.float64, scale=0.3):
    g = torch.Generator().manual_seed(seed)
    V, E, D, H = cfg['V'], cfg['E'], cfg['D'], cfg['H']
    shp = {'init_h.weight': (H, D), 'init_h.bias': (H,), 'init_c.weight': (H, D), 'init_c.bias': (H,),
           'f_beta.weight': (D, H), 'f_beta.bias': (D,), 'deep_output.weight': (V + 3, H), 'deep_output.bias': (V + 3,),
           'attention.U.weight': (H, H), 'attention.U.bias': (H,), 'attention.W.weight': (H, D), 'attention.W.bias': (H,),
           'attention.v.weight': (1, H), 'attention.v.bias': (1,), 'embedding.weight': (V + 3, E),
           'lstm.weight_ih': (4 * H, E + D), 'lstm.weight_hh': (4 * H, H), 'lstm.bias_ih': (4 * H,), 'lstm.bias_hh': (4 * H,)}
    out = {}
    for k, s in shp.items():
        fan = s[-1] if len(s) > 1 else H
        w = (torch.rand(*s, generator=g, dtype=torch.float64) * 2 - 1) * (scale if k != 'embedding.weight' else 1.0)
        out[k] = (w / (fan ** 0.5 if k != 'embedding.weight' and len(s) > 1 else 1.0)).float().to(dtype)
    return out
