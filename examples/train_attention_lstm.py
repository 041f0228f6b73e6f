"""The decoder side of `train_LSTMwAttention.py` on this package: same options, same loop, precomputed feature grids.

    python examples/train_attention_lstm.py --data_h5 data/face2text-feats.h5 --data_json data/face2text-dicts.json

What maps to what (reference train_LSTMwAttention.py / AlexCap/LSTMwAttention_opts.py):
  loader      AlexDataLoader(opt)                         -> imagecaptioning_b200.data.AlexDataLoader(opt, source='features')
                                                             (the HDF5 file carries a `/features [M, L, D]` dataset next to
                                                             `/labels`, `/split`, ...: the CNN is run once, offline)
  model       AlexCapModel(opt).llm / .crit / backward()  -> Decoder(...) + DecoderTrainStep(decoder).forward_train
  optimiser   clip_gradient(1) + Adam + LambdaLR  (:57-76) -> FusedClipAdam(max_norm=1) + the same LambdaLR lambda
  evaluation  eval_resnet.eval_split(kwargs)     (:118-124)-> imagecaptioning_b200.eval.eval_split (any val_batch_size)
  checkpoint  torch.save(model.state_dict())     (:127)    -> same key names under 'llm.' (loads back into the reference)
  test sweep  beam_size 1..5                     (:143-155)-> imagecaptioning_b200.eval.beam_sweep
Needs a CUDA device (there is no CPU path).  METEOR / BLEU need nltk, which the reference imports and this image lacks:
the best checkpoint is chosen by validation loss unless `--scorer nltk` finds it.
"""
import argparse
import json
import os
import sys
import time

import torch
from torch.optim.lr_scheduler import LambdaLR

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from imagecaptioning_b200.data import AlexDataLoader, load_reference_checkpoint          # noqa: E402
from imagecaptioning_b200.eval import beam_sweep, eval_split                              # noqa: E402
from imagecaptioning_b200.model import DecoderTrainStep                                   # noqa: E402
from imagecaptioning_b200.modules import Decoder                                          # noqa: E402
from imagecaptioning_b200.optim import FusedClipAdam, warmup_cosine_lambda                # noqa: E402


def get_config():
    """AlexCap/LSTMwAttention_opts.py:8-53 (the keys the decoder side reads), overridable from the command line."""
    cfg = dict(device='cuda:0', data_h5='data/face2text-feats.h5', data_json='data/face2text-dicts.json',
               use_scheduler=True, learning_rate=3e-4, embedding_size=1024, lstm_size=768, beta1=0.9, beta2=0.999, eps=1e-8,
               weight_decay=1e-6, min_lr=1e-6, num_epochs=50, save_checkpoint_every=8489,
               save_path='best_model_sch_LSTMwAttention.pth', batch_size=12, clip_grad=True, iterate=False,
               from_checkpoint=False, seed=123, val_batch_size=64, compute_dtype='f32', scorer='none')
    ap = argparse.ArgumentParser()
    for k, v in cfg.items():
        if isinstance(v, bool):
            ap.add_argument(f'--{k}', type=lambda s: s.lower() in ('1', 'true', 'yes'), default=v)
        else:
            ap.add_argument(f'--{k}', type=type(v), default=v)
    return vars(ap.parse_args())


def nltk_scorer(records):
    """eval_resnet.score_captions (:110-123): METEOR + BLEU-4 (method4 smoothing), averaged."""
    from nltk import word_tokenize
    from nltk.translate import bleu_score, meteor
    smooth = bleu_score.SmoothingFunction().method4
    ms = [round(meteor([word_tokenize(r['references'])], word_tokenize(r['candidate'])), 4) for r in records]
    bs = [round(bleu_score.sentence_bleu([word_tokenize(r['references'])], word_tokenize(r['candidate']),
                                         smoothing_function=smooth), 4) for r in records]
    return {'meteor': sum(ms) / len(ms), 'bleu': sum(bs) / len(bs)}


def main():
    opt = get_config()
    torch.manual_seed(opt['seed'])                                           # :20-23
    torch.cuda.manual_seed(opt['seed'])
    loader = AlexDataLoader(opt, source='features')
    feats = loader.h5_file['features']
    decoder = Decoder(loader.getVocabSize(), opt['embedding_size'], feats.shape[2], opt['lstm_size'], loader.getSeqLength(),
                      loader.info['idx_to_token'], 0.5, opt['device']).to(opt['device'])
    decoder.compute_dtype = opt['compute_dtype']
    if opt['from_checkpoint']:                                               # :31-39
        load_reference_checkpoint(decoder, opt['save_path'])
    step = DecoderTrainStep(decoder)
    scorer = nltk_scorer if opt['scorer'] == 'nltk' else None

    # setup_scheduler, :57-76
    optim = FusedClipAdam(decoder.parameters(), opt['learning_rate'], (opt['beta1'], opt['beta2']), opt['eps'],
                          weight_decay=opt['weight_decay'], max_norm=1.0 if opt['clip_grad'] else None)
    per_epoch = opt['save_checkpoint_every'] // opt['batch_size']
    max_iter = per_epoch * opt['num_epochs']
    warmup_steps = int(max_iter / opt['num_epochs'])
    lr_lambda = warmup_cosine_lambda(warmup_steps, max_iter, opt['min_lr'] / opt['learning_rate']) \
        if opt['use_scheduler'] else (lambda it: 1.0)
    scheduler = LambdaLR(optim, lr_lambda)

    best_val_loss, best_val_score, best_iter = float('inf'), -1.0, 0
    for it in range(max_iter):                                               # :99-135
        optim.zero_grad(set_to_none=True)
        decoder.train()
        step.set_eval(False)
        features, gt_labels, _, _ = loader.get_batch(opt, opt['batch_size'])
        t1 = time.time()
        loss = step.forward_train(features, gt_labels)                       # loss AND .grad (LSTMwAttentionModel.py:67-72)
        optim.step()                                                         # clip_grad_norm_(1) + Adam in three launches
        scheduler.step()
        if it % 50 == 0:
            print(f'iter : {it} -> captioning_loss: {loss.item():.5f}, epoch time in ms: {(time.time() - t1) * 1e3:.3f}')
        if it > 0 and ((it + 1) % per_epoch == 0 or it + 1 == max_iter):
            results = eval_split({'model': step, 'loader': loader, 'split': 'val', 'max_images': -1,
                                  'val_batch_size': opt['val_batch_size'], 'skip_incomplete': True}, scorer=scorer)
            better = (results['ap_results']['meteor'] > best_val_score) if scorer else (results['loss_results'] < best_val_loss)
            if better:
                torch.save({'llm.' + k: v for k, v in decoder.state_dict().items()}, opt['save_path'])   # :127
                best_iter = it
                if scorer:
                    best_val_score = results['ap_results']['meteor']
            best_val_loss = min(best_val_loss, results['loss_results'])
            print(json.dumps({'iter': it, 'val_loss': results['loss_results'], 'ap_results': results['ap_results'],
                              'best_iter': best_iter}))

    # MODEL EVALUATION ON TEST DATASET, :137-155
    sweep = beam_sweep({'model': step, 'loader': loader, 'split': 'test', 'max_images': -1,
                        'val_batch_size': opt['val_batch_size'], 'skip_incomplete': True}, scorer=scorer)
    for b, r in sweep.items():
        print(json.dumps({'beam_size': b, 'test_loss': r['loss_results'], 'ap_results': r['ap_results'],
                          'incomplete_beams': r['incomplete_beams']}))


if __name__ == '__main__':
    main()
