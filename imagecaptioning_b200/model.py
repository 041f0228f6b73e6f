"""DecoderTrainStep -- the decoder half of AlexCapModel.forward_train as ONE native call.

Reference contract (AlexCap/LSTMwAttentionModel.py:48-74, AlexCap/LSTMModel.py:53-74): forward_train computes
    loss = crit(llm(features, gt), llm.get_target(gt, make_target=True)) [+ ((1 - alphas.sum(1))**2).mean()]
calls loss.backward() itself and returns the loss; the caller then clips and steps (train_LSTMwAttention.py:85-109).
Here the same contract is served by alexcap_attn_train_fused / alexcap_lstm_train_fused: no logits or autograd
tape are materialised on the Python side, gradients land in `.grad`.

Optionally the whole call is captured once into a CUDA graph (static input buffers) and replayed: the step is a
chain of ~150 small launches whose launch latency otherwise rivals the kernel time at batch 128.
"""
from __future__ import annotations

import torch

from . import functional as F_
from .modules import Decoder, LanguageModel, _f32c


class DecoderTrainStep:
    def __init__(self, llm, use_cuda_graph: bool = False, need_dfeat: bool = False):
        if not isinstance(llm, (Decoder, LanguageModel)):
            raise TypeError('llm must be an imagecaptioning_b200 Decoder or LanguageModel')
        self.llm = llm
        self.is_attn = isinstance(llm, Decoder)
        self.use_cuda_graph = use_cuda_graph
        self.need_dfeat = need_dfeat
        self.eval_mode = False
        self.ce_denominator = None          # device float tensor: global N_valid / world (data parallel), or None
        self._key = None
        self._graphs = {}                   # (input set, phase) -> captured CUDA graph
        self._cur = 0
        self._copy_stream = None
        self._copied = [None, None]         # event: H2D into buffer i finished (copy stream)
        self._consumed = [None, None]       # event: last step reading buffer i finished (compute stream)
        self.last_alphas = None
        self.dfeat = None

    # -- reference API ---------------------------------------------------------------------------------
    def set_eval(self, value: bool):
        """AlexCapModel.set_eval (LSTMwAttentionModel.py:43-44): eval_mode => loss only, no backward."""
        self.eval_mode = value

    def clip_gradient(self, norm):
        torch.nn.utils.clip_grad_norm_(self.llm.parameters(), max_norm=norm)

    # -- internals --------------------------------------------------------------------------------------
    def _prepare(self, feat, gt, dev=None):
        dev = dev if dev is not None else feat.device
        key = (tuple(feat.shape), tuple(gt.shape), dev, self.llm.training, self.need_dfeat)
        if key == self._key:
            return
        self._key, self._graphs, self._cur = key, {}, 0
        self._copied, self._consumed = [None, None], [None, None]
        F_.ensure_device(dev)
        self._params = self.llm.param_list()
        for p in self._params:
            if p.dtype != torch.float32 or not p.is_contiguous() or p.device != dev:
                raise F_._lib.AlexcapError('parameters must be contiguous float32 tensors on the features device')
        # one flat fp32 gradient buffer, slices 16-byte aligned, in gradient-finalisation order: vocab projection (final
        # after phase 1), then lstm.* / embedding (final after phase 4), then everything else
        order = sorted(range(len(self._params)), key=self._group)
        sizes = [(self._params[i].numel() + 3) // 4 * 4 for i in order]
        self._flat_grads = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
        self._grads = [None] * len(self._params)
        off = 0
        self._vocab_elems = 0
        self._group_end = [0, 0, 0]
        for i, n in zip(order, sizes):
            self._grads[i] = self._flat_grads[off:off + self._params[i].numel()].view_as(self._params[i])
            off += n
            if self._is_vocab(i):
                self._vocab_elems = off
            for gi in range(self._group(i), 3):
                self._group_end[gi] = off
        self._dims = self.llm._dims(feat, gt, need_dfeat=self.need_dfeat)
        nbytes = F_.attn_workspace_bytes(self._dims) if self.is_attn else F_.lstm_workspace_bytes(self._dims)
        self._ws = F_.workspace(nbytes, dev)
        self._loss = torch.zeros(4, dtype=torch.float32, device=dev)
        # two static input buffer sets: the step reads set `_cur` while prefetch() fills the other one
        self._feats = [torch.empty(feat.shape, dtype=torch.float32, device=dev) for _ in range(2)]
        self._gts = [torch.empty(gt.shape, dtype=torch.int32, device=dev) for _ in range(2)]
        self.dfeat = torch.empty_like(self._feat) if self.need_dfeat else None
        self._alphas = None

    @property
    def _feat(self):
        return self._feats[self._cur]

    @property
    def _gt(self):
        return self._gts[self._cur]

    def _launch(self, phase=3):
        d = self._dims
        if d.dropout_p > 0 and phase in (1, 3):   # fresh mask every step (Philox counter advances)
            d.offset += 1 << 32
        mask = self.llm._drop_mask_override
        if self.is_attn:
            F_.attn_train_fused(d, self._params, self._feat, self._gt, self._grads, self._ws, dfeat=self.dfeat,
                                loss_out=self._loss, alphas_out=self._alphas, drop_mask=mask,
                                ce_denominator=self.ce_denominator, phase=phase)
        elif phase in (2, 4, 8):
            return                               # plain LSTM: one call does the whole step (phase 1)
        else:
            P = self.llm._struct(self._params)
            G = self.llm._struct(self._grads)
            F_.lstm_train_fused(d, P, self._feat, self._gt, G, self._ws, dfeat=self.dfeat, loss_out=self._loss,
                                drop_mask=mask, ce_denominator=self.ce_denominator)

    def run_resident(self, phase=3):
        """One step on the inputs already staged in the static device buffers (bench: inputs resident in HBM).
        phase=1 then phase=2 -- or 1, 4, 8 (the two halves of 2) -- run the same step in several launches (see
        GradAllReducer.step_overlapped)."""
        if self.use_cuda_graph and self._dims.dropout_p == 0:
            key = (self._cur, phase)
            if key not in self._graphs:
                s = torch.cuda.Stream()
                s.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s):
                    self._launch(phase)      # warm-up outside capture (cudaFuncSetAttribute, scratch allocation)
                torch.cuda.current_stream().wait_stream(s)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._launch(phase)
                self._graphs[key] = g
                # the warm-up consumed the saved state of the earlier phases (gates -> dgates in place, accumulations)
                for prev in {2: (1,), 4: (1,), 8: (1, 4)}.get(phase, ()):
                    self._launch(prev)
            self._graphs[key].replay()
        else:
            self._launch(phase)
        return self._loss

    def stage(self, features, gt_labels):
        """Copy one batch into the static device buffers.  Host tensors (ideally pinned) are accepted as the SOURCE of
        the H2D copy only; all compute happens on the parameters' CUDA device."""
        dev = next(self.llm.parameters()).device
        if dev.type != 'cuda':
            raise F_._lib.AlexcapError('imagecaptioning_b200 has no CPU path: move the decoder to a CUDA device')
        feat = features.detach()
        if feat.dtype != torch.float32:
            feat = feat.float()
        gt = gt_labels if gt_labels.dtype == torch.int32 else gt_labels.to(torch.int32)
        self._prepare(feat, gt, dev)
        self._feat.copy_(feat, non_blocking=True)
        self._gt.copy_(gt, non_blocking=True)

    def stage_nchw(self, cnn_out, gt_labels):
        """Stage one batch from the backbone's NCHW output [B, D, h, w] (device tensor): replaces the reference's
        permute(0, 2, 3, 1).view(B, -1, C) (LSTMwAttentionModel.py:64-66) AND the staging copy by one transposing pass
        into the static feature buffer.  Follow with run_resident() / forward_train_nchw()."""
        if not cnn_out.is_cuda:
            raise F_._lib.AlexcapError('stage_nchw takes the CNN output on the CUDA device')
        x = cnn_out.detach()
        if x.dtype != torch.float32:
            x = x.float()
        x = x.contiguous()
        B, D, h, w = x.shape
        gt = gt_labels if gt_labels.dtype == torch.int32 else gt_labels.to(torch.int32)
        self._prepare(torch.empty(B, h * w, D, dtype=torch.float32, device='meta'), gt, x.device)
        F_.features_from_nchw(x, out=self._feat)
        self._gt.copy_(gt, non_blocking=True)
        self._hw = (h, w)

    def forward_train_nchw(self, cnn_out, gt_labels):
        """forward_train on the backbone's NCHW output.  With need_dfeat=True, `dfeat_nchw()` afterwards returns the
        gradient w.r.t. `cnn_out` in its own layout (the fine-tune phase, train_LSTMwAttention.py:51-55)."""
        self.stage_nchw(cnn_out, gt_labels)
        loss4 = self.run_resident()
        for p, g in zip(self._params, self._grads):
            if p.requires_grad:
                if p.grad is None:
                    p.grad = g.clone()
                else:
                    p.grad.add_(g)
        return loss4[0]

    def dfeat_nchw(self, out=None, accumulate=False):
        if self.dfeat is None:
            raise F_._lib.AlexcapError('dfeat_nchw needs DecoderTrainStep(need_dfeat=True)')
        return F_.dfeat_to_nchw(self.dfeat, self._hw, out=out, accumulate=accumulate)

    def prefetch(self, features, gt_labels):
        """Start the H2D copy of the NEXT batch (pinned host tensors) into the idle input buffer set on a side stream,
        so that it overlaps the step running on the current set.  Pair with swap()+run_resident()."""
        dev = next(self.llm.parameters()).device
        if self._key is None:
            self.stage(features, gt_labels)     # first call sizes the buffers
        if features.dtype != torch.float32 or gt_labels.dtype != torch.int32:
            raise F_._lib.AlexcapError('prefetch() takes float32 features and int32 captions')
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        nxt = self._cur ^ 1
        cs = self._copy_stream
        if self._consumed[nxt] is not None:
            cs.wait_event(self._consumed[nxt])  # the step that last read this set must have finished
        with torch.cuda.stream(cs):
            self._feats[nxt].copy_(features, non_blocking=True)
            self._gts[nxt].copy_(gt_labels, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._copied[nxt] = ev

    def swap(self):
        """Make the prefetched input set current; the compute stream waits for its H2D copy."""
        cur = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(cur)
        self._consumed[self._cur] = ev
        self._cur ^= 1
        if self._copied[self._cur] is None:
            raise F_._lib.AlexcapError('swap() without a preceding prefetch()')
        cur.wait_event(self._copied[self._cur])
        self._copied[self._cur] = None

    def forward_train(self, features, gt_labels):
        """features [B,L,D] (the CNN grid after permute/view, LSTMwAttentionModel.py:64-66), gt_labels [B,T] int.
        Returns the scalar loss tensor (device); parameter .grad fields are ACCUMULATED into, like loss.backward()."""
        if self.eval_mode:
            return self.forward_eval(features, gt_labels)
        self.stage(features, gt_labels)
        loss4 = self.run_resident()
        for p, g in zip(self._params, self._grads):
            if p.requires_grad:
                if p.grad is None:
                    p.grad = g.clone()
                else:
                    p.grad.add_(g)
        return loss4[0]

    @torch.no_grad()
    def forward_eval(self, features, gt_labels):
        """eval_mode branch (LSTMwAttentionModel.py:50-60): loss only, through the materialising forward."""
        feat = _f32c(features)
        from .modules import CustomCrossEntropyLoss
        if self.is_attn:
            preds, alphas = self.llm.forward(feat, gt_labels)
            tgt = self.llm.get_target(gt_labels, make_target=True)
            loss = CustomCrossEntropyLoss()(preds, tgt)
            return loss + ((1 - alphas.sum(1)) ** 2).mean()
        preds = self.llm.forward(feat, gt_labels)
        return CustomCrossEntropyLoss()(preds, self.llm.get_target(gt_labels, make_target=True))

    @torch.no_grad()
    def forward_test(self, features):
        """Decoder half of AlexCapModel.forward_test (LSTMwAttentionModel.py:76-87 -> (captions, alphas);
        LSTMModel.py:76-86 -> captions): eval mode, decode branch of `llm.forward`, ids -> strings."""
        self.llm.eval()
        feat = _f32c(features)
        dev = next(self.llm.parameters()).device
        if feat.device != dev:
            feat = feat.to(dev)
        out = self.llm.forward(feat, feat.new())
        if self.is_attn:
            preds, alphas = out
            return self.llm.decode_sequence(preds), alphas
        return self.llm.decode_sequence(out)

    def _is_vocab(self, i):
        names = F_._lib.ATTN_STATE_NAMES if self.is_attn else self.llm.param_names()
        return names[i].startswith('deep_output.') or names[i].startswith('rnn.linear.')

    def _group(self, i):
        """0: vocab projection, 1: lstm.* and embedding (attention model), 2: the rest."""
        if self._is_vocab(i):
            return 0
        if self.is_attn:
            n = F_._lib.ATTN_STATE_NAMES[i]
            if n.startswith('lstm.') or n.startswith('embedding.'):
                return 1
        return 2

    def grad_groups(self):
        """The flat gradient buffer as contiguous slices in finalisation order (attention model: after phase 1 / 4 / 8)."""
        e = self._group_end
        return [self._flat_grads[:e[0]], self._flat_grads[e[0]:e[1]], self._flat_grads[e[1]:e[2]]]

    def grad_buckets(self, bucket_elems: int = 8 << 20):
        """Contiguous slices of the flat gradient buffer: [vocab projection] + ~32 MB chunks of the rest."""
        out = [self._flat_grads[:self._vocab_elems]]
        n = self._flat_grads.numel()
        o = self._vocab_elems
        while o < n:
            out.append(self._flat_grads[o:min(n, o + bucket_elems)])
            o += bucket_elems
        return out

    # convenience for tests
    def grads_by_name(self):
        names = F_._lib.ATTN_STATE_NAMES if self.is_attn else self.llm.param_names()
        return dict(zip(names, self._grads))
