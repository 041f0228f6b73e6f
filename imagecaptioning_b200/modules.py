"""Drop-in mirrors of the reference's caption-decoder modules, computing on libalexcap_b200.so.

  LanguageModel          <- AlexCap/LanguageModule.py:6-198
  Attention, Decoder     <- AlexCap/AttentionLanguageModule.py:6-22, :23-268
  CustomCrossEntropyLoss <- AlexCap/CustomLoss.py:4-15

Same constructor signatures, same parameter names/shapes (reference checkpoints load with load_state_dict), same
forward(features, captions) / get_target / decode_sequence behaviour, so `from AlexCap.LanguageModule import
LanguageModel` in AlexCap/LSTMModel.py (and the Decoder import in LSTMwAttentionModel.py) can be pointed here.
nn.Linear / nn.Embedding / nn.LSTM(Cell) objects are used ONLY as parameter containers (native init, names);
their forward() is never called -- every FLOP runs in the sm_100a kernels.  There is no CPU fallback: calling
forward on non-CUDA tensors raises.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import functional as F_
from . import _lib
from ._lib import ATTN_STATE_NAMES, AlexcapError


def _i32(t):
    return t.to(torch.int32).contiguous()


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


# =============================================================================================== loss
def _compute_dtype(mod):
    """`mod.compute_dtype` ('f32' default, or 'bf16'): bf16 runs every GEMM of the training path on the tcgen05 bf16
    engine (operands rounded to bf16, fp32 accumulation; parameters/activations/gradients stay fp32) -- the
    mixed-precision mode of BASELINE.json configs 3 and 5.  The reference itself is fp32-only (train_LSTM.py:16)."""
    cd = getattr(mod, 'compute_dtype', 'f32')
    if cd not in ('f32', 'bf16'):
        raise AlexcapError("compute_dtype must be 'f32' or 'bf16'")
    return _lib.BF16 if cd == 'bf16' else _lib.F32


class _CEFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inp, target):
        N, T, Cc = inp.shape
        x = _f32c(inp).view(N * T, Cc)
        out, dl = F_.cross_entropy(x, _i32(target.reshape(-1)), 0.1, 0, want_grad=inp.requires_grad)
        ctx.dl = dl
        ctx.shape = inp.shape
        return out[0]

    @staticmethod
    def backward(ctx, g):
        return (ctx.dl * g).view(ctx.shape), None


class CustomCrossEntropyLoss(nn.Module):
    """CustomLoss.py:4-15 -- CrossEntropyLoss(reduction='mean', ignore_index=0, label_smoothing=0.1) over [N,T,C]."""

    def __init__(self):
        super().__init__()

    def forward(self, input, target):
        return _CEFunction.apply(input, target)


# =============================================================================================== shared helpers
class _DecodeMixin:
    def get_target(self, gt_sequence, make_target=False):
        """AttentionLanguageModule.py:197-212 / LanguageModule.py:113-128, one kernel instead of a Python double loop.
        Returned on gt_sequence's device with gt_sequence's dtype."""
        in_tok, tgt = F_.make_target(gt_sequence, self._V)
        return (tgt if make_target else in_tok).to(gt_sequence.dtype)

    def _decode_words(self, seq, start_as_sos):
        END, START = self.END_TOKEN, self.START_TOKEN

        def one(row, first):
            words = []
            for t in range(first, len(row)):
                idx = int(row[t])
                if idx == END or idx == 0:
                    break
                if start_as_sos and idx == START:
                    words.append('SOS')
                else:
                    words.append(self.idx_to_token[str(idx)])
            return ' '.join(words)

        if isinstance(seq, list):
            return [one(seq, 1)]
        if seq.dim() == 2:
            return [one(r, 0) for r in seq.tolist()]
        return [[one(r, 0) for r in s] for s in seq.tolist()]


# =============================================================================================== attention decoder
class Attention(nn.Module):
    """AttentionLanguageModule.py:6-22.  Parameter container for Decoder; its standalone forward (inference only)
    runs the fused attention-step kernel with the beta gate held open."""

    def __init__(self, encoder_dim, rnn_size):
        super().__init__()
        self.U = nn.Linear(rnn_size, rnn_size)
        self.W = nn.Linear(encoder_dim, rnn_size)
        self.v = nn.Linear(rnn_size, 1)
        self.tanh = nn.Tanh()
        self.softmax = nn.Softmax(1)

    @torch.no_grad()
    def forward(self, img_features, hidden_state):
        feat = _f32c(img_features)
        B, L, D = feat.shape
        H = self.U.weight.shape[0]
        Ws = F_.gemm(feat.view(B * L, D), _f32c(self.W.weight), bias=_f32c(self.W.bias)).view(B, L, H)
        q = F_.gemm(_f32c(hidden_state), _f32c(self.U.weight), bias=_f32c(self.U.bias))
        bpre = torch.zeros(B, D, device=feat.device)
        alpha, ctx, _, _ = F_.attn_step_fwd(Ws, feat, q, bpre, _f32c(self.v.weight).view(-1))
        return ctx, alpha


class _AttnDecoderFn(torch.autograd.Function):
    """Decoder.forward train branch as one autograd node: alexcap_attn_forward / alexcap_attn_backward."""

    @staticmethod
    def forward(ctx, mod, feat, gt, drop_mask, *params):
        dims = mod._dims(feat, gt, need_dfeat=feat.requires_grad)
        ws = F_.workspace(F_.attn_workspace_bytes(dims), feat.device)
        preds = torch.empty(dims.B, dims.S, dims.Vp, dtype=torch.float32, device=feat.device)
        alphas = torch.empty(dims.B, dims.S, dims.L, dtype=torch.float32, device=feat.device)
        p = [_f32c(t) for t in params]
        F_.attn_forward(dims, p, feat, gt, preds, alphas, ws, drop_mask)
        ctx.dims, ctx.ws, ctx.p, ctx.feat = dims, ws, p, feat
        ctx.need_dfeat = feat.requires_grad
        return preds, alphas

    @staticmethod
    def backward(ctx, dpreds, dalphas):
        grads = [torch.empty_like(t) for t in ctx.p]
        dfeat = torch.empty_like(ctx.feat) if ctx.need_dfeat else None
        F_.attn_backward(ctx.dims, ctx.p, ctx.feat, None if dpreds is None else dpreds.contiguous(),
                         None if dalphas is None else dalphas.contiguous(), grads, ctx.ws, dfeat)
        ctx.ws = None
        return (None, dfeat, None, None, *grads)


class Decoder(nn.Module, _DecodeMixin):
    """AttentionLanguageModule.py:23-268."""

    def __init__(self, vocab_size, embedding_size, image_vector_dim, rnn_size, seq_length, idx_to_token, dropout,
                 device):
        super().__init__()
        self.vocabulary_size = vocab_size
        self.embedding_size = embedding_size
        self.image_vector_dim = image_vector_dim
        self.rnn_size = rnn_size
        self.seq_length = seq_length
        self.idx_to_token = idx_to_token
        self.device = device
        self.START_TOKEN = vocab_size + 1
        self.END_TOKEN = vocab_size + 2
        self.init_h = nn.Linear(image_vector_dim, rnn_size)
        self.init_c = nn.Linear(image_vector_dim, rnn_size)
        self.tanh = nn.Tanh()
        self.f_beta = nn.Linear(rnn_size, image_vector_dim)
        self.sigmoid = nn.Sigmoid()
        self.deep_output = nn.Linear(rnn_size, vocab_size + 3)
        self.dropout = nn.Dropout()            # p=0.5 whatever the ctor argument says (reference :41 vs :54)
        self.attention = Attention(image_vector_dim, rnn_size)
        self.embedding = nn.Embedding(vocab_size + 3, embedding_size)
        self.lstm = nn.LSTMCell(embedding_size + image_vector_dim, rnn_size)
        self.use_beam = True
        self.beam_size = 5
        self._drop_mask_override = None        # tests inject an explicit [S,B,H] mask here

    @property
    def _V(self):
        return self.vocabulary_size

    def param_list(self):
        """Parameters in the C struct's order (== ATTN_STATE_NAMES)."""
        sd = dict(self.named_parameters())
        return [sd[n] for n in ATTN_STATE_NAMES]

    def _train_dropout_p(self):
        return self.dropout.p

    def _dims(self, feat, gt=None, need_dfeat=False, train_dropout=True):
        B, L, D = feat.shape
        if D != self.image_vector_dim:
            raise AlexcapError(f'feature dim {D} != image_vector_dim {self.image_vector_dim}')
        S = self.seq_length + 1
        if gt is not None and gt.shape[1] != self.seq_length:
            raise AlexcapError(f'captions have T={gt.shape[1]} but seq_length={self.seq_length}')
        p = float(self.dropout.p) if (self.training and train_dropout) else 0.0
        seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if p > 0 else 0
        return F_.make_dims(B, L, D, self.rnn_size, self.embedding_size, self.vocabulary_size + 3, S,
                            dtype=_compute_dtype(self), need_dfeat=need_dfeat, dropout_p=p, seed=seed)

    def forward(self, img_features, captions):
        feat = img_features if img_features.dtype == torch.float32 and img_features.is_contiguous() \
            else img_features.float().contiguous()
        F_.ensure_device(feat.device)
        if captions.nelement() > 0:                                   # :63 train branch
            gt = _i32(captions)
            return _AttnDecoderFn.apply(self, feat, gt, self._drop_mask_override, *self.param_list())
        if self.use_beam:                                             # :86-87
            return self.caption(feat, self.beam_size)
        with torch.no_grad():                                         # :88-115 greedy, batched
            dims = self._dims(feat, train_dropout=False)
            return F_.attn_decode_greedy(dims, [_f32c(t) for t in self.param_list()], feat.detach())

    def get_init_lstm_state(self, img_features):
        """:116-125 (inference helper): h0, c0."""
        with torch.no_grad():
            m = F_.mean_over_L(_f32c(img_features))             # mean_L_kernel, the kernel the train path uses
            h = F_.gemm(m, _f32c(self.init_h.weight), bias=_f32c(self.init_h.bias), act=1)
            c = F_.gemm(m, _f32c(self.init_c.weight), bias=_f32c(self.init_c.bias), act=1)
        return h, c

    @torch.no_grad()
    def caption(self, img_features, beam_size):
        """:128-195.  B=1: returns (sentence[1,n], alpha[1,n,L]) and raises ValueError when no beam completed, exactly
        like the reference.  B>1 (extension): returns (ids[B,S], alphas[B,S,L]) with 0 after END; rows whose beams never
        completed are all 0."""
        feat = _f32c(img_features)
        dims = self._dims(feat, train_dropout=False)
        ids, lens, scores, alphas = F_.attn_decode_beam(dims, [_f32c(t) for t in self.param_list()], feat, int(beam_size))
        if feat.shape[0] == 1:
            n = int(lens[0].item())
            if n == 0:
                raise ValueError('max() arg is an empty sequence')   # what the reference's max([]) raises (:191)
            return ids[:, :n], alphas[:, :n]
        self.last_beam_lens, self.last_beam_scores = lens, scores
        return ids, alphas

    def decode_sequence(self, seq):
        return self._decode_words(seq, start_as_sos=True)             # :214-268


# =============================================================================================== plain LSTM decoder
class _LstmDecoderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod, feat, gt, drop_mask, *params):
        dims = mod._dims(feat, gt, need_dfeat=feat.requires_grad)
        ws = F_.workspace(F_.lstm_workspace_bytes(dims), feat.device)
        preds = torch.empty(dims.B, dims.S, dims.Vp, dtype=torch.float32, device=feat.device)
        p = [_f32c(t) for t in params]
        P = mod._struct(p)
        F_.lstm_forward(dims, P, feat, gt, preds, ws, drop_mask)
        ctx.dims, ctx.ws, ctx.p, ctx.feat, ctx.mod = dims, ws, p, feat, mod
        ctx.need_dfeat = feat.requires_grad
        return preds

    @staticmethod
    def backward(ctx, dpreds):
        grads = [torch.empty_like(t) for t in ctx.p]
        dfeat = torch.empty_like(ctx.feat) if ctx.need_dfeat else None
        F_.lstm_backward(ctx.dims, ctx.mod._struct(ctx.p), ctx.feat, dpreds.contiguous(), ctx.mod._struct(grads), ctx.ws,
                         dfeat)
        ctx.ws = None
        return (None, dfeat, None, None, *grads)


class LanguageModel(nn.Module, _DecodeMixin):
    """LanguageModule.py:6-198."""

    def __init__(self, vocab_size, embedding_size, image_vector_dim, rnn_size, seq_length, num_layers, idx_to_token,
                 dropout, device):
        super().__init__()
        self.vocab_size = vocab_size
        self.embedding_size = embedding_size
        self.image_vector_dim = image_vector_dim
        self.rnn_size = rnn_size
        self.seq_length = seq_length
        self.num_layers = num_layers
        self.idx_to_token = idx_to_token
        self.dropout = dropout
        self.device = device
        W, D, V, H = embedding_size, image_vector_dim, vocab_size, rnn_size
        self.image_encoder = nn.Sequential()
        self.image_encoder.add_module('encode', nn.Linear(D, W))
        self.image_encoder.add_module('relu', nn.ReLU(inplace=True))
        self.START_TOKEN = V + 1
        self.END_TOKEN = V + 2
        self.lookup_table = nn.Embedding(V + 3, W)
        self.use_beam = False
        self.beam_size = 3
        self.rnn = nn.Sequential()
        self.lstm = nn.LSTM(W, H, num_layers, batch_first=True)
        self.rnn.add_module('drop', nn.Dropout(dropout))
        self.rnn.add_module('linear', nn.Linear(H, V + 3))
        self._drop_mask_override = None
        self.train()

    @property
    def _V(self):
        return self.vocab_size

    def param_names(self):
        names = ['image_encoder.encode.weight', 'image_encoder.encode.bias', 'lookup_table.weight',
                 'rnn.linear.weight', 'rnn.linear.bias']
        for l in range(self.num_layers):
            names += [f'lstm.weight_ih_l{l}', f'lstm.weight_hh_l{l}', f'lstm.bias_ih_l{l}', f'lstm.bias_hh_l{l}']
        return names

    def param_list(self):
        sd = dict(self.named_parameters())
        return [sd[n] for n in self.param_names()]

    def _struct(self, tensors):
        nl = self.num_layers
        rest = tensors[5:]
        return F_.lstm_param_struct(tensors[0], tensors[1], tensors[2], tensors[3], tensors[4],
                                    [rest[4 * l] for l in range(nl)], [rest[4 * l + 1] for l in range(nl)],
                                    [rest[4 * l + 2] for l in range(nl)], [rest[4 * l + 3] for l in range(nl)])

    def _train_dropout_p(self):
        return self.rnn.drop.p

    def _dims(self, feat, gt=None, need_dfeat=False, train_dropout=True):
        B, L, D = feat.shape
        if gt is not None and gt.shape[1] != self.seq_length:
            raise AlexcapError(f'captions have T={gt.shape[1]} but seq_length={self.seq_length}')
        p = float(self.rnn.drop.p) if (self.training and train_dropout) else 0.0
        seed = int(torch.randint(0, 2 ** 62, (1,)).item()) if p > 0 else 0
        return F_.make_dims(B, L, D, self.rnn_size, self.embedding_size, self.vocab_size + 3, self.seq_length + 1,
                            num_layers=self.num_layers, dtype=_compute_dtype(self), need_dfeat=need_dfeat, dropout_p=p,
                            seed=seed, reg_weight=0.0)

    def forward(self, image_vectors, gt_sequence):
        feat = image_vectors if image_vectors.dtype == torch.float32 and image_vectors.is_contiguous() \
            else image_vectors.float().contiguous()
        F_.ensure_device(feat.device)
        if gt_sequence.nelement() > 0:                                # :100-107
            gt = _i32(gt_sequence)
            self._gt_with_start = self.get_target(gt_sequence)
            self.output = _LstmDecoderFn.apply(self, feat, gt, self._drop_mask_override, *self.param_list())
            return self.output
        if self.use_beam:
            return self.caption(feat, self.beam_size)
        return self.predict_caption(feat)

    @torch.no_grad()
    def predict_caption(self, image_vectors):
        """:130-150 batched greedy."""
        feat = _f32c(image_vectors)
        dims = self._dims(feat, train_dropout=False)
        return F_.lstm_decode_greedy(dims, self._struct([_f32c(t) for t in self.param_list()]), feat)

    @torch.no_grad()
    def caption(self, img_features, beam_size):
        """:152-198 -- beam search for the plain LSTM.  The raw features go through the LSTM (no image_encoder, :160), so
        image_vector_dim must equal embedding_size.  B=1: returns the sentence as a Python list of ids INCLUDING the
        leading START, and raises ValueError when no beam completed, exactly like the reference.  B>1 (extension): returns
        ids[B, S+1] in predict_caption's form (no START, 0 after END; rows whose beams never completed are all 0), which
        is what decode_sequence takes for a 2-D tensor."""
        feat = _f32c(img_features)
        dims = self._dims(feat, train_dropout=False)
        ids, lens, scores = F_.lstm_decode_beam(dims, self._struct([_f32c(t) for t in self.param_list()]), feat,
                                                int(beam_size))
        if feat.shape[0] == 1:
            n = int(lens[0].item())
            if n == 0:
                raise ValueError('max() arg is an empty sequence')   # what the reference's max([]) raises (:194)
            return [self.START_TOKEN] + ids[0, :n].tolist()
        self.last_beam_lens, self.last_beam_scores = lens, scores
        return ids

    def decode_sequence(self, seq):
        return self._decode_words(seq, start_as_sos=False)            # :52-97
