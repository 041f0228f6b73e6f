"""Eval side of the hot path: `eval_split` and the test-time beam sweep.

`eval_split(kwargs)` mirrors `AlexCap/eval/eval_resnet.py:39-107` -- same kwargs (`model`, `loader`, `split`,
`max_images`, `val_batch_size`), same loop (loss of the batch under `set_eval(True)`, captions under `set_eval(False)`,
`decode_sequence` of the ground truth, the `counter` / `split_bounds` stopping rule, `loss_results = batch_size *
all_losses / counter`) -- with `model` a `DecoderTrainStep` over precomputed feature grids and any `val_batch_size`:
the reference decodes one image per call with a host sync per token, here a batch is one device-side beam search
(`alexcap_attn_decode_beam`) whose rows equal the per-image runs (`tests/test_gpu_decode.py`).

Scoring is pluggable: METEOR / BLEU come from nltk in the reference (`eval_resnet.py:110-123`), which is not in this
image and is host-side text processing; pass `scorer=callable(records) -> dict` to get `ap_results`, else it is None and
the caller receives the `records` (candidate, references, imgid) the reference's evaluator would have scored.
"""
from .data import getopt


def eval_split(kwargs, scorer=None, verbose=False):
    model = getopt(kwargs, 'model')
    loader = getopt(kwargs, 'loader')
    split = getopt(kwargs, 'split', 'val')
    max_images = getopt(kwargs, 'max_images', -1)
    batch_size = getopt(kwargs, 'val_batch_size', 1)
    skip_incomplete = bool(kwargs.get('skip_incomplete', False))    # extension: record '' instead of aborting
    assert split == 'val' or split == 'test', 'split must be "val" or "test"'
    split = {'val': 1, 'test': 2}[split]
    model.llm.eval()
    loader.reset_iterator(split)
    records, counter, all_losses, no_beam = [], 0, 0.0, 0
    while True:
        counter += batch_size
        features, gt_labels, info, _ = loader.get_batch({'split': split, 'iterate': True}, batch_size)
        info = info[0]
        model.set_eval(True)
        all_losses += model.forward_train(features, gt_labels).item()
        model.set_eval(False)
        try:
            captions = model.forward_test(features)
            captions = captions[0] if isinstance(captions, tuple) else captions
        except ValueError:                     # the reference's caption() when no beam completed (B = 1 only): it aborts
            if not skip_incomplete:
                raise
            captions, no_beam = [''] * features.shape[0], no_beam + 1
        gt_captions = model.llm.decode_sequence(gt_labels)
        for i, t in enumerate(captions):       # DenseCaptioningEvaluator.addResult, eval_resnet.py:14-21
            records.append({'candidate': t, 'references': gt_captions[i], 'imgid': info['filename'][i]})
        num_images = info['split_bounds'][1]
        if max_images > 0:
            num_images = min(num_images, max_images)
            counter = min(counter, max_images)
        counter = min(counter, num_images)
        if verbose:
            print('Processed image %s (%d / %d) of split %d' % (info['filename'][0] + '...', counter, num_images, split))
        if max_images > 0 and counter >= max_images:
            break
        if counter >= num_images:
            break
    return {'loss_results': batch_size * all_losses / counter,
            'ap_results': scorer(records) if scorer is not None else None,
            'records': records, 'incomplete_beams': no_beam}


def beam_sweep(kwargs, beam_sizes=range(1, 6), scorer=None):
    """The test-time sweep of train_LSTMwAttention.py:143-155: `use_beam = True`, `beam_size = 1..5`, one `eval_split`
    each.  Returns {beam_size: eval_split result}; the decoder's `use_beam` / `beam_size` are restored afterwards."""
    llm = getopt(kwargs, 'model').llm
    saved = (llm.use_beam, llm.beam_size)
    out = {}
    try:
        llm.use_beam = True
        for b in beam_sizes:
            llm.beam_size = b
            out[b] = eval_split(kwargs, scorer=scorer)
    finally:
        llm.use_beam, llm.beam_size = saved
    return out
