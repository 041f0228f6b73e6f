"""CPU: host logic of imagecaptioning_b200/eval.py (no CUDA: model and decoder are stand-ins that record the calls).
The loop has to reproduce AlexCap/eval/eval_resnet.py:39-107: counter arithmetic, the `split_bounds` stopping rule,
`loss_results = batch_size * all_losses / counter`, set_eval(True) around the loss and set_eval(False) around decoding."""
import os

import pytest
import torch

from imagecaptioning_b200.data import AlexDataLoader
from imagecaptioning_b200.eval import beam_sweep, eval_split

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
OPT = {'data_h5': os.path.join(GOLD, 'face2text_tiny.h5'), 'data_json': os.path.join(GOLD, 'face2text_tiny.json'),
       'device': 'cpu'}


class _Llm:
    use_beam, beam_size = False, 3

    def __init__(self):
        self.eval_called = 0

    def eval(self):
        self.eval_called += 1

    def decode_sequence(self, seq):
        return [' '.join(str(int(t)) for t in row if int(t) != 0) for row in seq]


class _Step:
    """forward_train returns the batch's mean label value as 'loss'; forward_test 'captions' each image by its first label."""

    def __init__(self, fail_on=()):
        self.llm, self.eval_mode, self.log, self.fail_on = _Llm(), None, [], set(fail_on)
        self._labels = None

    def set_eval(self, v):
        self.eval_mode = v

    def forward_train(self, features, gt):
        assert self.eval_mode is True
        self._labels = gt
        self.log.append(('loss', features.shape[0], self.llm.beam_size))
        return gt.float().mean()

    def forward_test(self, features):
        assert self.eval_mode is False
        if int(self._labels[0, 0]) in self.fail_on:
            raise ValueError('max() arg is an empty sequence')
        return [f'first={int(r[0])}' for r in self._labels], torch.zeros(features.shape[0], 1, 1)


def _labels(split):
    ld = AlexDataLoader(OPT, source='features')
    return ld, torch.from_numpy(ld.labels[{1: ld.val_ix, 2: ld.test_ix}[split]])


@pytest.mark.parametrize('bs', [1, 2, 4, 6, 7])
def test_eval_split_loop_matches_reference_arithmetic(bs):
    ld, lab = _labels(1)                                    # 6 validation images
    step = _Step()
    out = eval_split({'model': step, 'loader': ld, 'split': 'val', 'max_images': -1, 'val_batch_size': bs})
    # the reference's loop, restated on plain lists (eval_resnet.py:57-98 with MyDataLoader.py:67-73 underneath)
    counter, losses, seen, it = 0, 0.0, [], 0
    while True:
        counter += bs
        nxt = it + bs
        ix = list(range(it, min(it + bs, 6)))
        it = 0 if nxt >= 6 else nxt
        losses += float(lab[ix].float().mean())
        seen += ix
        counter = min(counter, 6)
        if counter >= 6:
            break
    assert [r['candidate'] for r in out['records']] == [f'first={int(lab[i, 0])}' for i in seen]
    assert out['loss_results'] == pytest.approx(bs * losses / counter)
    assert [r['imgid'] for r in out['records']] == [ld.info['idx_to_filename'][str(ld.val_ix[i])] for i in seen]
    assert step.llm.eval_called == 1 and step.eval_mode is False and out['ap_results'] is None
    assert ld.iterators[1] == it


def test_eval_split_max_images_scorer_and_incomplete_beams():
    ld, lab = _labels(2)                                    # 5 test images
    step = _Step()
    out = eval_split({'model': step, 'loader': ld, 'split': 'test', 'max_images': 3, 'val_batch_size': 1},
                     scorer=lambda recs: {'meteor': len(recs)})
    assert len(out['records']) == 3 and out['ap_results'] == {'meteor': 3}
    first = int(lab[1, 0])
    with pytest.raises(ValueError):                         # the reference aborts when caption() finds no completed beam
        eval_split({'model': _Step(fail_on=[first]), 'loader': ld, 'split': 'test', 'val_batch_size': 1})
    out = eval_split({'model': _Step(fail_on=[first]), 'loader': ld, 'split': 'test', 'val_batch_size': 1,
                      'skip_incomplete': True})
    assert out['incomplete_beams'] == 1 and out['records'][1]['candidate'] == '' and len(out['records']) == 5
    with pytest.raises(AssertionError):
        eval_split({'model': step, 'loader': ld, 'split': 'train'})
    with pytest.raises(ValueError):
        eval_split({'loader': ld})


def test_beam_sweep_sets_and_restores_decoder_attributes():
    ld, _ = _labels(2)
    step = _Step()
    step.llm.use_beam, step.llm.beam_size = False, 7
    out = beam_sweep({'model': step, 'loader': ld, 'split': 'test', 'val_batch_size': 5})
    assert sorted(out) == [1, 2, 3, 4, 5]
    assert [b for kind, n, b in step.log] == [1, 2, 3, 4, 5]            # train_LSTMwAttention.py:145-147
    assert (step.llm.use_beam, step.llm.beam_size) == (False, 7)
    step2 = _Step(fail_on=[int(torch.from_numpy(ld.labels[ld.test_ix])[0, 0])])
    step2.llm.beam_size = 9
    with pytest.raises(ValueError):
        beam_sweep({'model': step2, 'loader': ld, 'split': 'test', 'val_batch_size': 1})
    assert step2.llm.beam_size == 9                                      # restored on the error path too
