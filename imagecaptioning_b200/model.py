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


class _Slot:
    """Everything that depends on the batch SHAPE: dims, workspace, loss cell, the two static input sets, captured graphs."""
    __slots__ = ('dims', 'ws', 'loss', 'feats', 'gts', 'dfeat', 'graphs', 'cur', 'copied', 'consumed', 'n_valid', 'ce_den')


class DecoderTrainStep:
    MAX_SLOTS = 4                            # batch shapes kept alive (a loader's partial last batch gets its own slot)

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
        self._slots = {}                    # shape key -> _Slot (insertion order = age)
        self._slot = None
        self._param_dev = None
        self._copy_stream = None
        self._philox = None                 # int64 {seed, offset} on the device: the dropout counter (advanced by the library)
        self.last_alphas = None

    # -- reference API ---------------------------------------------------------------------------------
    def set_eval(self, value: bool):
        """AlexCapModel.set_eval (LSTMwAttentionModel.py:43-44): eval_mode => loss only, no backward."""
        self.eval_mode = value

    def clip_gradient(self, norm):
        torch.nn.utils.clip_grad_norm_(self.llm.parameters(), max_norm=norm)

    # -- internals --------------------------------------------------------------------------------------
    def _drop_p(self):
        return float(self.llm._train_dropout_p()) if self.llm.training else 0.0

    def _prepare_params(self, dev):
        if self._param_dev == dev:
            return
        self._param_dev = dev
        self._slots, self._slot, self._key = {}, None, None
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
        # dropout counter on the device: seeded from torch's generator (torch.manual_seed reproducibility), offset 0
        self._philox = torch.tensor([int(torch.randint(0, 2 ** 62, (1,)).item()), 0], dtype=torch.int64, device=dev)

    def _prepare(self, feat, gt, dev=None):
        dev = dev if dev is not None else feat.device
        self._prepare_params(dev)
        # everything a workspace LAYOUT depends on; what only changes the captured work (compute dtype, ce_denominator,
        # an injected mask) is part of the graph key instead (see _graph_key)
        key = (tuple(feat.shape), tuple(gt.shape), self.need_dfeat, self._drop_p() > 0)
        if key == self._key:
            return
        self._key = key
        slot = self._slots.pop(key, None)
        if slot is None:
            while len(self._slots) >= self.MAX_SLOTS:           # drop the oldest shape (its graphs and buffers with it)
                self._slots.pop(next(iter(self._slots)))
            slot = _Slot()
            slot.dims = self.llm._dims(feat, gt, need_dfeat=self.need_dfeat)
            nbytes = F_.attn_workspace_bytes(slot.dims) if self.is_attn else F_.lstm_workspace_bytes(slot.dims)
            slot.ws = F_.workspace(nbytes, dev)
            slot.loss = torch.zeros(4, dtype=torch.float32, device=dev)
            # two static input buffer sets: the step reads set `cur` while prefetch() fills the other one
            slot.feats = [torch.empty(feat.shape, dtype=torch.float32, device=dev) for _ in range(2)]
            slot.gts = [torch.empty(gt.shape, dtype=torch.int32, device=dev) for _ in range(2)]
            slot.dfeat = torch.empty(feat.shape, dtype=torch.float32, device=dev) if self.need_dfeat else None
            slot.graphs, slot.cur = {}, 0
            slot.copied, slot.consumed = [None, None], [None, None]
            slot.n_valid = torch.zeros(1, dtype=torch.float32, device=dev)    # data parallel: local, then global / world
        self._slots[key] = slot                                  # (re)insert as the youngest
        self._slot = slot

    # the current slot's members under their historical names
    _dims = property(lambda self: self._slot.dims)
    _ws = property(lambda self: self._slot.ws)
    _loss = property(lambda self: self._slot.loss)
    _feats = property(lambda self: self._slot.feats)
    _gts = property(lambda self: self._slot.gts)
    _graphs = property(lambda self: self._slot.graphs)
    _copied = property(lambda self: self._slot.copied)
    _consumed = property(lambda self: self._slot.consumed)
    dfeat = property(lambda self: self._slot.dfeat if self._slot is not None else None)
    _alphas = None

    @property
    def _cur(self):
        return self._slot.cur

    @_cur.setter
    def _cur(self, v):
        self._slot.cur = v

    @property
    def _feat(self):
        return self._slot.feats[self._slot.cur]

    @property
    def _gt(self):
        return self._slot.gts[self._slot.cur]

    def _sync_dims(self):
        """Settings that may change between steps without changing the workspace layout are written into dims at every
        launch (and are part of the graph key): compute dtype, dropout p, the dropout counter."""
        d = self._slot.dims
        d.dtype = F_._lib.BF16 if getattr(self.llm, 'compute_dtype', 'f32') == 'bf16' else F_._lib.F32
        p = self._drop_p()
        if (p > 0) != (d.dropout_p > 0):
            raise F_._lib.AlexcapError('dropout was switched on/off without re-staging the batch (the workspace layout differs)')
        d.dropout_p = p
        d.philox_state = self._philox.data_ptr() if (p > 0 and self.llm._drop_mask_override is None) else None
        return d

    def _graph_key(self, phase):
        d = self._slot.dims
        mask = self.llm._drop_mask_override
        den = self.ce_denominator
        return (self._slot.cur, phase, int(d.dtype), float(d.dropout_p), None if mask is None else mask.data_ptr(),
                None if den is None else den.data_ptr())

    def _launch(self, phase=3):
        d = self._sync_dims()
        mask = self.llm._drop_mask_override
        if self.is_attn:
            F_.attn_train_fused(d, self._params, self._feat, self._gt, self._grads, self._ws, dfeat=self.dfeat,
                                loss_out=self._loss, alphas_out=self._alphas, drop_mask=mask,
                                ce_denominator=self.ce_denominator, phase=phase)
        elif phase in (2, 4, 8, 32, 64, 128):
            return                               # plain LSTM: one call does the whole step (phase 1 / 16)
        else:
            P = self.llm._struct(self._params)
            G = self.llm._struct(self._grads)
            F_.lstm_train_fused(d, P, self._feat, self._gt, G, self._ws, dfeat=self.dfeat, loss_out=self._loss,
                                drop_mask=mask, ce_denominator=self.ce_denominator)

    # phases whose saved state a warm-up run of `phase` consumes (gates -> dgates in place, accumulations): re-run after it
    _PREREQ = {2: (1,), 4: (1,), 8: (1, 4), 32: (16,), 128: (16,), 64: ()}

    def run_resident(self, phase=3):
        """One step on the inputs already staged in the static device buffers (bench: inputs resident in HBM).
        phase=1 then phase=2 -- or 16, 32, 4, 8 (the halves of 1 and of 2) -- run the same step in several launches (see
        GradAllReducer.step_overlapped).  With use_cuda_graph every (input set, phase, compute dtype, dropout p, injected
        mask, ce_denominator) combination is captured once and replayed; dropout p > 0 is captured too (the Philox counter
        lives on the device and advances at every replay)."""
        if self.use_cuda_graph:
            self._sync_dims()
            key = self._graph_key(phase)
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
                for prev in self._PREREQ.get(phase, ()):
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
        self._unalias_grads()
        loss4 = self.run_resident()
        self._publish_grads()
        return loss4[0]

    def _unalias_grads(self):
        """A .grad that still IS a slice of the flat gradient buffer (the caller did not zero_grad(set_to_none=True) since the
        previous step) is about to be overwritten: give it its own storage so that accumulation semantics hold."""
        for p, g in zip(self._params, self._grads):
            if p.grad is not None and p.grad.data_ptr() == g.data_ptr():
                p.grad = p.grad.clone()

    def _publish_grads(self):
        """loss.backward() semantics on .grad (LSTMwAttentionModel.py:72) without a copy in the common case: after
        optim.zero_grad(set_to_none=True) (train_LSTMwAttention.py:80) every .grad is None and simply BECOMES the slice of
        the flat gradient buffer the step wrote (valid until the next forward_train); otherwise one multi-tensor add."""
        acc_dst, acc_src = [], []
        for p, g in zip(self._params, self._grads):
            if not p.requires_grad:
                continue
            if p.grad is None:
                p.grad = g
            else:
                acc_dst.append(p.grad); acc_src.append(g)
        if acc_dst:
            torch._foreach_add_(acc_dst, acc_src)

    def dfeat_nchw(self, out=None, accumulate=False):
        if self.dfeat is None:
            raise F_._lib.AlexcapError('dfeat_nchw needs DecoderTrainStep(need_dfeat=True)')
        return F_.dfeat_to_nchw(self.dfeat, self._hw, out=out, accumulate=accumulate)

    def prefetch(self, features, gt_labels):
        """Start the H2D copy of the NEXT batch (pinned host tensors) into the idle input buffer set on a side stream,
        so that it overlaps the step running on the current set.  Pair with swap()+run_resident()."""
        dev = next(self.llm.parameters()).device
        if self._slot is None or tuple(features.shape) != tuple(self._feat.shape) or tuple(gt_labels.shape) != tuple(self._gt.shape):
            self.stage(features, gt_labels)     # first call of a shape sizes (or re-selects) the buffers
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
        self._unalias_grads()
        loss4 = self.run_resident()
        self._publish_grads()
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
