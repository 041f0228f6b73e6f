"""ORACLE (test infrastructure, NOT product code) -- the reference's CPU execution path,
re-expressed with stock torch.nn ops + autograd so that it costs what the reference costs.

This is the thing bench.py times as `cpu_baseline` / `--impl reference` (kind "port"): the
reference itself is Python under /root/reference and cannot travel to the GPU box, so its
per-timestep structure is restated here op for op:
  * AlexCap/AttentionLanguageModule.py:62-84  -- S-step Python loop; W(features) is recomputed
    inside Attention.forward every step (:17); nn.LSTMCell; per-step vocab Linear written into
    a preallocated preds buffer (:81) -> CopySlices autograd nodes, as in the reference.
  * AlexCap/LanguageModule.py:99-107          -- nn.LSTM called twice (regions, then words).
  * AlexCap/CustomLoss.py:7-14                -- nn.CrossEntropyLoss(mean, ignore_index=0, ls=0.1).
  * AlexCap/LSTMwAttentionModel.py:67-72      -- loss + regulariser, loss.backward().
It is validated to be bit-identical to the imported reference modules on CPU in
tests/test_oracle_vs_reference.py (same ops, same order, same seed).
State-dict names equal the reference's so one set of weights loads into both.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _targets(gt, vocab_size, make_target):
    # get_target (AttentionLanguageModule.py:197-212): Python double loop kept on purpose --
    # it is part of what the reference pays per step.
    n, t = gt.shape
    out = torch.zeros(n, t + 1, dtype=gt.dtype)
    if make_target:
        out[:, :t] = gt
        for i in range(n):
            for j in range(1, t + 1):
                if out[i, j] == 0:
                    out[i, j] = vocab_size + 2
                    break
    else:
        out[:, 0] = vocab_size + 1
        out[:, 1:] = gt
    return out


class AttnDecoderPort(nn.Module):
    def __init__(self, vocab_size, embedding_size, image_vector_dim, rnn_size, seq_length, dropout_p=0.5):
        super().__init__()
        self.vocab_size, self.seq_length = vocab_size, seq_length
        self.init_h = nn.Linear(image_vector_dim, rnn_size)
        self.init_c = nn.Linear(image_vector_dim, rnn_size)
        self.f_beta = nn.Linear(rnn_size, image_vector_dim)
        self.deep_output = nn.Linear(rnn_size, vocab_size + 3)
        self.dropout = nn.Dropout(dropout_p)
        self.attention = nn.Module()
        self.attention.U = nn.Linear(rnn_size, rnn_size)
        self.attention.W = nn.Linear(image_vector_dim, rnn_size)
        self.attention.v = nn.Linear(rnn_size, 1)
        self.embedding = nn.Embedding(vocab_size + 3, embedding_size)
        self.lstm = nn.LSTMCell(embedding_size + image_vector_dim, rnn_size)

    def forward(self, feat, gt):
        b = feat.size(0)
        mean = feat.mean(dim=1)
        c = torch.tanh(self.init_c(mean))
        h = torch.tanh(self.init_h(mean))
        steps = self.seq_length + 1
        emb = self.embedding(_targets(gt, self.vocab_size, False))
        preds = torch.zeros(b, steps, self.vocab_size + 3)
        alphas = torch.zeros(b, steps, feat.size(1))
        for t in range(steps):
            uh = self.attention.U(h).unsqueeze(1)
            ws = self.attention.W(feat)                      # recomputed every step, as in the reference
            e = self.attention.v(torch.tanh(ws + uh)).squeeze(2)
            a = torch.softmax(e, dim=1)
            ctx = (feat * a.unsqueeze(2)).sum(1)
            gated = torch.sigmoid(self.f_beta(h)) * ctx
            h, c = self.lstm(torch.cat((emb[:, t], gated), dim=1), (h, c))
            preds[:, t] = self.deep_output(self.dropout(h))
            alphas[:, t] = a
        return preds, alphas

    def train_step(self, feat, gt):
        preds, alphas = self(feat, gt)
        tgt = _targets(gt, self.vocab_size, True).long()
        loss = F.cross_entropy(preds.view(-1, preds.size(2)), tgt.view(-1), ignore_index=0, label_smoothing=0.1)
        loss = loss + ((1 - alphas.sum(1)) ** 2).mean()
        loss.backward()
        return loss


class LstmDecoderPort(nn.Module):
    def __init__(self, vocab_size, embedding_size, image_vector_dim, rnn_size, seq_length, num_layers=1, dropout_p=0.0):
        super().__init__()
        self.vocab_size, self.seq_length = vocab_size, seq_length
        self.image_encoder = nn.Sequential()
        self.image_encoder.add_module('encode', nn.Linear(image_vector_dim, embedding_size))
        self.lookup_table = nn.Embedding(vocab_size + 3, embedding_size)
        self.lstm = nn.LSTM(embedding_size, rnn_size, num_layers, batch_first=True)
        self.rnn = nn.Sequential()
        self.rnn.add_module('drop', nn.Dropout(dropout_p))
        self.rnn.add_module('linear', nn.Linear(rnn_size, vocab_size + 3))

    def forward(self, feat, gt):
        words = self.lookup_table(_targets(gt, self.vocab_size, False))
        enc = torch.relu(self.image_encoder(feat))
        _, state = self.lstm(enc)
        out, _ = self.lstm(words, state)
        return self.rnn(out).contiguous()

    def train_step(self, feat, gt):
        preds = self(feat, gt)
        tgt = _targets(gt, self.vocab_size, True).long()
        loss = F.cross_entropy(preds.view(-1, preds.size(2)), tgt.view(-1), ignore_index=0, label_smoothing=0.1)
        loss.backward()
        return loss
