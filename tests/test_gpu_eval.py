"""-m gpu: eval side -- eval_split / beam_sweep over a dataset in the reference's
on-disk format, losses against the fp64 oracle, batched decoding against the reference's one-image-per-call protocol."""
import os

import pytest
import torch

from imagecaptioning_b200 import h5lite
from imagecaptioning_b200.data import AlexDataLoader
from imagecaptioning_b200.eval import beam_sweep, eval_split
from imagecaptioning_b200.model import DecoderTrainStep
from oracle import restated as R
from tests.gpu_util import DEV, build_decoder, random_attn_params

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
OPT = {'data_h5': os.path.join(GOLD, 'face2text_tiny.h5'), 'data_json': os.path.join(GOLD, 'face2text_tiny.json'),
       'device': DEV}
CFG = dict(B=1, L=6, D=32, H=24, E=16, V=50, T=16)


def _setup():
    ld = AlexDataLoader(OPT, source='features')
    p = random_attn_params(CFG, seed=5, scale=1.0)
    p['deep_output.bias'][CFG['V'] + 2] += 1.5
    dec = build_decoder(CFG, p)
    dec.idx_to_token = ld.info['idx_to_token']
    return ld, p, dec, DecoderTrainStep(dec)


def test_eval_split_losses_and_batched_captions():
    ld, p, dec, step = _setup()
    kw = {'model': step, 'loader': ld, 'split': 'val', 'max_images': -1, 'skip_incomplete': True}
    dec.use_beam, dec.beam_size = True, 3
    r1 = eval_split({**kw, 'val_batch_size': 1})
    r3 = eval_split({**kw, 'val_batch_size': 3})
    assert len(r1['records']) == len(r3['records']) == len(ld.val_ix) == 6
    with h5lite.File(OPT['data_h5']) as f:
        feats, labels = torch.from_numpy(f['features'][:]).double(), torch.from_numpy(f['labels'][:]).long()
    per_image = [float(R.attn_train_step(p, feats[i:i + 1], labels[i:i + 1], CFG['V'])['loss']) for i in ld.val_ix]
    assert abs(r1['loss_results'] - sum(per_image) / 6) < 1e-4 * sum(per_image) / 6        # eval_resnet.py:98
    per_batch = [float(R.attn_train_step(p, feats[ix], labels[ix], CFG['V'])['loss']) for ix in (ld.val_ix[:3], ld.val_ix[3:])]
    assert abs(r3['loss_results'] - 3 * sum(per_batch) / 6) < 1e-4 * r3['loss_results']
    # one device-side beam search per batch == the reference's one image per call
    assert [r['candidate'] for r in r1['records']] == [r['candidate'] for r in r3['records']]
    assert [r['imgid'] for r in r1['records']] == [ld.info['idx_to_filename'][str(i)] for i in ld.val_ix]
    assert [r['references'] for r in r1['records']] == dec.decode_sequence(labels[ld.val_ix])
    assert any(r['candidate'] for r in r1['records'])
    assert not dec.training and r1['ap_results'] is None
    scored = eval_split({**kw, 'val_batch_size': 2, 'max_images': 4}, scorer=lambda recs: {'n': len(recs)})
    assert scored['ap_results'] == {'n': 4}


def test_beam_sweep_restores_settings_and_beam1_is_greedy():
    ld, p, dec, step = _setup()
    dec.use_beam, dec.beam_size = False, 2
    out = beam_sweep({'model': step, 'loader': ld, 'split': 'test', 'val_batch_size': 5, 'skip_incomplete': True})
    assert sorted(out) == [1, 2, 3, 4, 5] and (dec.use_beam, dec.beam_size) == (False, 2)
    assert all(len(out[b]['records']) == len(ld.test_ix) == 5 for b in out)
    losses = {round(out[b]['loss_results'], 6) for b in out}
    assert len(losses) == 1                                   # the loss does not depend on the beam size
    greedy = eval_split({'model': step, 'loader': ld, 'split': 'test', 'val_batch_size': 1})
    for g, b1 in zip(greedy['records'], out[1]['records']):
        if b1['candidate']:
            assert g['candidate'] == b1['candidate']
