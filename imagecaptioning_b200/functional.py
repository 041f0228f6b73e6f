"""Tensor-level wrappers over the C ABI (include/alexcap.h).  torch is used for device memory and streams only;
every FLOP on these paths runs in libalexcap_b200.so.  No CPU path: tensors must live on a CUDA device."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import ATTN_FIELDS, AttnParams, Dims, LstmParams, check


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _req(*tensors):
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise _lib.AlexcapError('imagecaptioning_b200 has no CPU path: tensor is not on a CUDA device')
        if not t.is_contiguous():
            raise _lib.AlexcapError('non-contiguous tensor handed to the C ABI')


_checked_devices = set()


def ensure_device(device):
    """Fail loudly unless `device` is an sm_100 GPU and the native library is loadable."""
    lib = _lib.load()
    dev = torch.device(device)
    if dev.type != 'cuda':
        raise _lib.AlexcapError(f'imagecaptioning_b200 needs a CUDA (sm_100a) device, got {dev}')
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if idx not in _checked_devices:
        with torch.cuda.device(idx):
            check(lib.alexcap_device_check(), 'alexcap_device_check')
        _checked_devices.add(idx)
    return lib


def make_dims(B, L, D, H, E, Vp, S, num_layers=1, dtype=_lib.F32, need_dfeat=False, label_smoothing=0.1,
              ignore_index=0, reg_weight=1.0, dropout_p=0.0, seed=0, offset=0, philox_state=None) -> Dims:
    """philox_state: int64 CUDA tensor {seed, offset} -- the dropout counter lives on the device and is advanced there, so a
    captured CUDA graph draws a fresh mask at every replay (the caller keeps the tensor alive)."""
    if philox_state is not None and (not philox_state.is_cuda or philox_state.dtype != torch.int64 or philox_state.numel() < 2):
        raise _lib.AlexcapError('philox_state must be a CUDA int64 tensor with 2 elements {seed, offset}')
    return Dims(B, L, D, H, E, Vp, S, num_layers, dtype, int(bool(need_dfeat)), ignore_index, label_smoothing,
                reg_weight, dropout_p, seed, offset, None if philox_state is None else philox_state.data_ptr())


def attn_param_struct(tensors) -> AttnParams:
    """tensors: list in ATTN_FIELDS order (fp32, contiguous, CUDA)."""
    _req(*tensors)
    p = AttnParams()
    for name, t in zip(ATTN_FIELDS, tensors):
        if t.dtype != torch.float32:
            raise _lib.AlexcapError(f'parameter {name} must be float32')
        setattr(p, name, t.data_ptr())
    return p


def lstm_param_struct(enc_w, enc_b, emb_w, out_w, out_b, w_ih, w_hh, b_ih, b_hh) -> LstmParams:
    _req(enc_w, enc_b, emb_w, out_w, out_b, *w_ih, *w_hh, *b_ih, *b_hh)
    p = LstmParams()
    p.enc_w, p.enc_b, p.emb_w, p.out_w, p.out_b = (t.data_ptr() for t in (enc_w, enc_b, emb_w, out_w, out_b))
    for l in range(len(w_ih)):
        p.w_ih[l], p.w_hh[l], p.b_ih[l], p.b_hh[l] = (w_ih[l].data_ptr(), w_hh[l].data_ptr(), b_ih[l].data_ptr(),
                                                     b_hh[l].data_ptr())
    return p


def workspace(nbytes: int, device) -> torch.Tensor:
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


# ------------------------------------------------------------------------------------------------ kernels
def features_from_nchw(nchw: torch.Tensor, out: torch.Tensor = None) -> torch.Tensor:
    """[B, D, h, w] CNN output -> the decoder's [B, h*w, D] grid (LSTMwAttentionModel.py:64-66) in one transposing pass."""
    lib = ensure_device(nchw.device)
    if nchw.dim() != 4 or nchw.dtype != torch.float32 or not nchw.is_contiguous():
        raise _lib.AlexcapError('features_from_nchw takes a contiguous float32 [B, D, h, w] tensor')
    B, D, L = nchw.shape[0], nchw.shape[1], nchw.shape[2] * nchw.shape[3]
    if out is None:
        out = torch.empty(B, L, D, dtype=torch.float32, device=nchw.device)
    if tuple(out.shape) != (B, L, D) or not out.is_contiguous() or out.dtype != torch.float32:
        raise _lib.AlexcapError('features_from_nchw: out must be a contiguous float32 [B, h*w, D] tensor')
    check(lib.alexcap_features_from_nchw(B, D, L, _ptr(nchw), _ptr(out), _stream()), 'alexcap_features_from_nchw')
    return out


def dfeat_to_nchw(dfeat: torch.Tensor, hw, out: torch.Tensor = None, accumulate: bool = False) -> torch.Tensor:
    """Gradient w.r.t. the [B, L, D] grid -> [B, D, h, w] (what autograd would hand the CNN); accumulate adds into `out`."""
    lib = ensure_device(dfeat.device)
    B, L, D = dfeat.shape
    h, w = hw
    if h * w != L or dfeat.dtype != torch.float32 or not dfeat.is_contiguous():
        raise _lib.AlexcapError('dfeat_to_nchw takes a contiguous float32 [B, h*w, D] tensor')
    if out is None:
        if accumulate:
            raise _lib.AlexcapError('dfeat_to_nchw: accumulate needs an existing out tensor')
        out = torch.empty(B, D, h, w, dtype=torch.float32, device=dfeat.device)
    if tuple(out.shape) != (B, D, h, w) or not out.is_contiguous() or out.dtype != torch.float32:
        raise _lib.AlexcapError('dfeat_to_nchw: out must be a contiguous float32 [B, D, h, w] tensor')
    check(lib.alexcap_dfeat_to_nchw(B, D, L, _ptr(dfeat), _ptr(out), int(accumulate), _stream()), 'alexcap_dfeat_to_nchw')
    return out


def gemm(A, B, transa=False, transb=True, M=None, N=None, K=None, lda=None, ldb=None, out=None, ldc=None, bias=None,
         bias2=None, accumulate=False, act=0, alpha=1.0, engine=0):
    """C = act(alpha*op(A)op(B) + bias + bias2) (+C).  Shapes default from 2-D contiguous A, B."""
    lib = ensure_device(A.device)
    _req(A, B, out, bias, bias2)
    if M is None:
        M, K = (A.shape[1], A.shape[0]) if transa else (A.shape[0], A.shape[1])
        N = B.shape[0] if transb else B.shape[1]
    lda = lda if lda is not None else A.stride(0)
    ldb = ldb if ldb is not None else B.stride(0)
    if out is None:
        out = torch.empty(M, N, dtype=torch.float32, device=A.device)
    ldc = ldc if ldc is not None else out.stride(0)
    check(lib.alexcap_gemm_f32(int(transa), int(transb), M, N, K, alpha, _ptr(A), lda, _ptr(B), ldb, _ptr(out), ldc,
                               _ptr(bias), _ptr(bias2), int(accumulate), act, engine, _stream()), 'alexcap_gemm_f32')
    return out


def _rows_bf16(Ws, feat):
    if Ws.dtype != feat.dtype or Ws.dtype not in (torch.float32, torch.bfloat16):
        raise _lib.AlexcapError('Ws and feat must both be float32 or both bfloat16')
    return int(Ws.dtype == torch.bfloat16)


def ws_exp_form(Ws):
    """The form the training step keeps W(features) in for the streaming attention kernels: 2^(clamp(2 log2(e) Ws, +-60))
    = e^(2 Ws) (csrc/common.cuh: exp2c).  Pass the result with ws_exp=True to attn_step_fwd / attn_step_bwd."""
    return torch.exp2((2.8853900817779268 * Ws.float()).clamp(-60.0, 60.0)).to(Ws.dtype)


def attn_step_fwd(Ws, feat, q, bpre, w_v, ws_exp=False):
    """One fused attention step.  Ws[B,L,H], feat[B,L,D] (both fp32, or both bf16 = the bf16 storage mode of the streamed
    grids), q[B,H], bpre[B,D], w_v[H] -> (alpha[B,L], ctx[B,D], beta[B,D], gctx[B,D]).  ws_exp: Ws is ws_exp_form(Ws)."""
    lib = ensure_device(Ws.device)
    _req(Ws, feat, q, bpre, w_v)
    B, L, H = Ws.shape
    D = feat.shape[2]
    alpha = torch.empty(B, L, dtype=torch.float32, device=Ws.device)
    ctx = torch.empty(B, D, dtype=torch.float32, device=Ws.device)
    beta = torch.empty_like(ctx)
    gctx = torch.empty_like(ctx)
    check(lib.alexcap_attn_step_fwd(B, L, D, H, _ptr(Ws), _ptr(feat), _ptr(q), q.stride(0), _ptr(bpre), bpre.stride(0),
                                    _ptr(w_v), _ptr(alpha), L, _ptr(ctx), _ptr(beta), _ptr(gctx), D,
                                    _rows_bf16(Ws, feat) | (2 if ws_exp else 0), _stream()),
          'alexcap_attn_step_fwd')
    return alpha, ctx, beta, gctx


def attn_step_bwd(Ws, feat, q, w_v, alpha, ctx, beta, dgctx, dalpha_ext=None, reg_r=None, need_dfeat=False, ws_exp=False):
    """Backward of one attention step -> (dq[B,H], dbpre[B,D], de[B,L], dwv_part[B,H], dfeat or None)."""
    lib = ensure_device(Ws.device)
    _req(Ws, feat, q, w_v, alpha, ctx, beta, dgctx, dalpha_ext, reg_r)
    B, L, H = Ws.shape
    D = feat.shape[2]
    dev = Ws.device
    dq = torch.empty(B, H, dtype=torch.float32, device=dev)
    dbpre = torch.empty(B, D, dtype=torch.float32, device=dev)
    de = torch.empty(B, L, dtype=torch.float32, device=dev)
    dwv = torch.zeros(B, H, dtype=torch.float32, device=dev)
    dfeat = torch.zeros(feat.shape, dtype=torch.float32, device=dev) if need_dfeat else None
    check(lib.alexcap_attn_step_bwd(B, L, D, H, _ptr(Ws), _ptr(feat), _ptr(q), q.stride(0), _ptr(w_v), _ptr(alpha),
                                    alpha.stride(0), _ptr(ctx), _ptr(beta), _ptr(dgctx), dgctx.stride(0),
                                    _ptr(dalpha_ext), L, _ptr(reg_r), _ptr(dq), H, _ptr(dbpre), D, _ptr(de), L,
                                    _ptr(dwv), _ptr(dfeat), _rows_bf16(Ws, feat) | (2 if ws_exp else 0), _stream()),
          'alexcap_attn_step_bwd')
    return dq, dbpre, de, dwv, dfeat


def make_target(gt, vocab_size):
    """(in_tok[B,T+1], tgt[B,T+1]) int32 on gt's device."""
    lib = ensure_device(gt.device)
    gt = gt.to(torch.int32).contiguous()
    B, T = gt.shape
    in_tok = torch.empty(B, T + 1, dtype=torch.int32, device=gt.device)
    tgt = torch.empty_like(in_tok)
    check(lib.alexcap_make_target(_ptr(gt), _ptr(in_tok), _ptr(tgt), B, T, vocab_size + 1, vocab_size + 2, _stream()),
          'alexcap_make_target')
    return in_tok, tgt


def count_valid_targets(gt, out=None):
    """Number of non-ignored targets get_target(gt, make_target=True) yields (nnz(gt) + B) as a 1-element device float."""
    lib = ensure_device(gt.device)
    if gt.dtype != torch.int32:
        raise _lib.AlexcapError('count_valid_targets takes int32 captions')
    _req(gt, out)
    B, T = gt.shape
    if out is None:
        out = torch.empty(1, dtype=torch.float32, device=gt.device)
    check(lib.alexcap_count_valid_targets(_ptr(gt), B, T, _ptr(out), _stream()), 'alexcap_count_valid_targets')
    return out


def mean_over_L(feat):
    """[B, L, D] -> [B, D] mean over regions (Decoder.get_init_lstm_state, AttentionLanguageModule.py:117)."""
    lib = ensure_device(feat.device)
    _req(feat)
    B, L, D = feat.shape
    m = torch.empty(B, D, dtype=torch.float32, device=feat.device)
    check(lib.alexcap_mean_over_L(_ptr(feat), B, L, D, _ptr(m), _stream()), 'alexcap_mean_over_L')
    return m


def attn_workspace_region(dims: Dims, ws: torch.Tensor, name: str, shape) -> torch.Tensor:
    """fp32 view of a named region of an attention training workspace (test / introspection hook)."""
    off, n = C.c_size_t(0), C.c_size_t(0)
    check(_lib.load().alexcap_attn_workspace_region(C.byref(dims), name.encode(), C.byref(off), C.byref(n)),
          'alexcap_attn_workspace_region')
    return ws[off.value:off.value + n.value].view(torch.float32).view(*shape)


def cross_entropy(logits2d, tgt, label_smoothing=0.1, ignore_index=0, want_grad=True):
    """logits2d [N,C] fp32 contiguous, tgt [N] int32 -> (loss_out[2] = {loss, n_valid}, dlogits or None)."""
    lib = ensure_device(logits2d.device)
    _req(logits2d, tgt)
    N, Cc = logits2d.shape
    out = torch.empty(2, dtype=torch.float32, device=logits2d.device)
    scratch = torch.empty(2 * N + 8, dtype=torch.float32, device=logits2d.device)
    dl = torch.empty_like(logits2d) if want_grad else None
    check(lib.alexcap_cross_entropy(_ptr(logits2d), _ptr(tgt), N, Cc, logits2d.stride(0), label_smoothing, ignore_index,
                                    _ptr(out), _ptr(dl), _ptr(scratch), _stream()), 'alexcap_cross_entropy')
    return out, dl


def cross_entropy_static(logits2d, tgt, loss_out, dlogits=None, label_smoothing=0.1, ignore_index=0):
    """alexcap_cross_entropy_static: the same loss and gradient with the library's file-scope device scratch instead of a
    caller-provided one.  Allocates nothing (not here either: the caller passes loss_out [2] and, optionally, dlogits);
    N is limited to alexcap_cross_entropy_static_max_rows(); calls that overlap on the GPU race on the shared scratch."""
    lib = ensure_device(logits2d.device)
    _req(logits2d, tgt, loss_out, dlogits)
    N, Cc = logits2d.shape
    check(lib.alexcap_cross_entropy_static(_ptr(logits2d), _ptr(tgt), N, Cc, logits2d.stride(0), label_smoothing,
                                           ignore_index, _ptr(loss_out), _ptr(dlogits), _stream()),
          'alexcap_cross_entropy_static')
    return loss_out, dlogits


# ------------------------------------------------------------------------------------------------ attention decoder
def attn_workspace_bytes(dims: Dims) -> int:
    return int(_lib.load().alexcap_attn_workspace_bytes(C.byref(dims)))


def attn_train_fused(dims, params, feat, gt, grads, ws, dfeat=None, loss_out=None, alphas_out=None, drop_mask=None,
                     ce_denominator=None, phase=3):
    """phase 3: the whole step.  phase 2 may itself be run as 4 (chain + lstm/embedding gradients) then 8 (the rest).
    phase 1 (forward + loss + vocab backward) then phase 2 (reverse chain + the other
    weight gradients) is the same step in two calls (data-parallel overlap of the first gradient bucket)."""
    lib = ensure_device(feat.device)
    _req(feat, gt, dfeat, loss_out, alphas_out, drop_mask, ce_denominator, ws)
    if loss_out is None:
        loss_out = torch.empty(4, dtype=torch.float32, device=feat.device)
    P = attn_param_struct(params)
    G = attn_param_struct(grads)
    check(lib.alexcap_attn_train_phase(C.byref(dims), C.byref(P), _ptr(feat), _ptr(gt), C.byref(G), _ptr(dfeat),
                                       _ptr(loss_out), _ptr(alphas_out), _ptr(drop_mask), _ptr(ce_denominator),
                                       _ptr(ws), ws.numel(), int(phase), _stream()), 'alexcap_attn_train_phase')
    return loss_out


def attn_forward(dims, params, feat, gt, preds, alphas, ws, drop_mask=None):
    lib = ensure_device(feat.device)
    _req(feat, gt, preds, alphas, drop_mask, ws)
    P = attn_param_struct(params)
    check(lib.alexcap_attn_forward(C.byref(dims), C.byref(P), _ptr(feat), _ptr(gt), _ptr(preds), _ptr(alphas),
                                   _ptr(drop_mask), _ptr(ws), ws.numel(), _stream()), 'alexcap_attn_forward')


def attn_backward(dims, params, feat, dpreds, dalphas, grads, ws, dfeat=None):
    lib = ensure_device(feat.device)
    _req(feat, dpreds, dalphas, dfeat, ws)
    P = attn_param_struct(params)
    G = attn_param_struct(grads)
    check(lib.alexcap_attn_backward(C.byref(dims), C.byref(P), _ptr(feat), _ptr(dpreds), _ptr(dalphas), C.byref(G),
                                    _ptr(dfeat), _ptr(ws), ws.numel(), _stream()), 'alexcap_attn_backward')


def attn_decode_greedy(dims, params, feat):
    lib = ensure_device(feat.device)
    _req(feat)
    P = attn_param_struct(params)
    ids = torch.zeros(dims.B, dims.S, dtype=torch.int64, device=feat.device)
    alphas = torch.zeros(dims.B, dims.S, dims.L, dtype=torch.float32, device=feat.device)
    ws = workspace(lib.alexcap_attn_decode_workspace_bytes(C.byref(dims), 1), feat.device)
    check(lib.alexcap_attn_decode_greedy(C.byref(dims), C.byref(P), _ptr(feat), _ptr(ids), _ptr(alphas), _ptr(ws),
                                         ws.numel(), _stream()), 'alexcap_attn_decode_greedy')
    return ids, alphas


def attn_decode_beam(dims, params, feat, beam):
    lib = ensure_device(feat.device)
    _req(feat)
    P = attn_param_struct(params)
    dev = feat.device
    ids = torch.zeros(dims.B, dims.S, dtype=torch.int64, device=dev)
    lens = torch.zeros(dims.B, dtype=torch.int32, device=dev)
    scores = torch.zeros(dims.B, dtype=torch.float32, device=dev)
    alphas = torch.zeros(dims.B, dims.S, dims.L, dtype=torch.float32, device=dev)
    ws = workspace(lib.alexcap_attn_decode_workspace_bytes(C.byref(dims), beam), dev)
    check(lib.alexcap_attn_decode_beam(C.byref(dims), C.byref(P), _ptr(feat), beam, _ptr(ids), _ptr(lens), _ptr(scores),
                                       _ptr(alphas), _ptr(ws), ws.numel(), _stream()), 'alexcap_attn_decode_beam')
    return ids, lens, scores, alphas


# ------------------------------------------------------------------------------------------------ plain LSTM decoder
def lstm_workspace_bytes(dims: Dims) -> int:
    return int(_lib.load().alexcap_lstm_workspace_bytes(C.byref(dims)))


def lstm_train_fused(dims, P: LstmParams, feat, gt, G: LstmParams, ws, dfeat=None, loss_out=None, drop_mask=None,
                     ce_denominator=None):
    lib = ensure_device(feat.device)
    _req(feat, gt, dfeat, loss_out, drop_mask, ce_denominator, ws)
    if loss_out is None:
        loss_out = torch.empty(4, dtype=torch.float32, device=feat.device)
    check(lib.alexcap_lstm_train_fused(C.byref(dims), C.byref(P), _ptr(feat), _ptr(gt), C.byref(G), _ptr(dfeat),
                                       _ptr(loss_out), _ptr(drop_mask), _ptr(ce_denominator), _ptr(ws), ws.numel(),
                                       _stream()), 'alexcap_lstm_train_fused')
    return loss_out


def lstm_forward(dims, P: LstmParams, feat, gt, preds, ws, drop_mask=None):
    lib = ensure_device(feat.device)
    _req(feat, gt, preds, drop_mask, ws)
    check(lib.alexcap_lstm_forward(C.byref(dims), C.byref(P), _ptr(feat), _ptr(gt), _ptr(preds), _ptr(drop_mask),
                                   _ptr(ws), ws.numel(), _stream()), 'alexcap_lstm_forward')


def lstm_backward(dims, P: LstmParams, feat, dpreds, G: LstmParams, ws, dfeat=None):
    lib = ensure_device(feat.device)
    _req(feat, dpreds, dfeat, ws)
    check(lib.alexcap_lstm_backward(C.byref(dims), C.byref(P), _ptr(feat), _ptr(dpreds), C.byref(G), _ptr(dfeat),
                                    _ptr(ws), ws.numel(), _stream()), 'alexcap_lstm_backward')


def lstm_decode_greedy(dims, P: LstmParams, feat):
    lib = ensure_device(feat.device)
    _req(feat)
    ids = torch.zeros(dims.B, dims.S, dtype=torch.int64, device=feat.device)
    ws = workspace(lib.alexcap_lstm_decode_workspace_bytes(C.byref(dims)), feat.device)
    check(lib.alexcap_lstm_decode_greedy(C.byref(dims), C.byref(P), _ptr(feat), _ptr(ids), _ptr(ws), ws.numel(),
                                         _stream()), 'alexcap_lstm_decode_greedy')
    return ids


def lstm_decode_beam(dims, P: LstmParams, feat, beam):
    """ids [B, S+1] (no leading START, END included, 0 after), lens [B] (0 = no beam completed), scores [B]."""
    lib = ensure_device(feat.device)
    _req(feat)
    dev = feat.device
    ids = torch.zeros(dims.B, dims.S + 1, dtype=torch.int64, device=dev)
    lens = torch.zeros(dims.B, dtype=torch.int32, device=dev)
    scores = torch.zeros(dims.B, dtype=torch.float32, device=dev)
    ws = workspace(lib.alexcap_lstm_decode_beam_workspace_bytes(C.byref(dims), beam), dev)
    check(lib.alexcap_lstm_decode_beam(C.byref(dims), C.byref(P), _ptr(feat), beam, _ptr(ids), _ptr(lens), _ptr(scores),
                                       _ptr(ws), ws.numel(), _stream()), 'alexcap_lstm_decode_beam')
    return ids, lens, scores


# ------------------------------------------------------------------------------------------------ engine selection
def set_gemm_engine(engine: int):
    """0 = fp32 CUDA-core tiles, 1 = tcgen05/TMEM/TMA 3xTF32 tiles (process-wide)."""
    check(_lib.load().alexcap_set_gemm_engine(int(engine)), 'alexcap_set_gemm_engine')


def get_gemm_engine() -> int:
    return int(_lib.load().alexcap_get_gemm_engine())


def set_time_loop(on: bool) -> bool:
    """Forward recurrence as one persistent weight-stationary kernel (True) or as the PDL-chained per-step launches (False,
    the default: faster at cfg2 on B200).  Set before staging a batch (the workspace layout depends on it).  Returns the
    previous setting."""
    return bool(_lib.load().alexcap_set_time_loop(int(bool(on))))
