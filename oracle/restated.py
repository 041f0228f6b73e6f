"""ORACLE (test infrastructure, NOT product code) -- explicit-math restatement of the
AlexCap caption-decoder training step, forward AND hand-derived backward, on CPU.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
leg may import this package.  The product path (imagecaptioning_b200/) never does.

What it restates (all citations relative to /root/reference/):
  * AlexCap/AttentionLanguageModule.py:15-22   Attention.forward          -> attn_step()
  * AlexCap/AttentionLanguageModule.py:62-84   Decoder.forward (train)    -> attn_forward()
  * AlexCap/AttentionLanguageModule.py:116-125 get_init_lstm_state        -> attn_forward()
  * AlexCap/AttentionLanguageModule.py:197-212 get_target                 -> make_targets()
  * AlexCap/LanguageModule.py:99-107           LanguageModel.forward      -> lstm_forward()
  * AlexCap/CustomLoss.py:7-14                 masked label-smoothed CE   -> ce_loss_and_grad()
  * AlexCap/LSTMwAttentionModel.py:67-72       loss = CE + regulariser; backward -> attn_train_step()
  * AlexCap/LSTMModel.py:70-73                 loss = CE; backward        -> lstm_train_step()
The arithmetic of nn.LSTMCell / nn.LSTM / nn.CrossEntropyLoss lives in PyTorch (third
party, un-pinned by the reference; torch 2.11.0 installed here); their documented formulas
are restated explicitly (gate order i,f,g,o; two bias vectors; label smoothing spread over
all classes, mean over non-ignored rows).

Parity pinning: the reference ships no tests / golden vectors ("parity unpinned" by the
reference itself).  This restatement is pinned against the reference's OWN modules imported
from /root/reference in the build container (tests/test_oracle_vs_reference.py, skipped
where /root/reference is absent) and against committed fixtures generated from those
modules (tests/golden/*.npz, made by tests/golden/make_golden.py).

Everything here is plain torch-on-CPU tensor arithmetic with NO autograd: the backward is
written out, because it is exactly the chain the CUDA kernels implement.
"""
from __future__ import annotations

import torch

EPS_SMOOTH = 0.1  # CustomLoss.py:7


# --------------------------------------------------------------------------- targets
def make_targets(gt: torch.Tensor, vocab_size: int):
    """get_target(gt) and get_target(gt, make_target=True)  (AttentionLanguageModule.py:197-212,
    LanguageModule.py:113-128).  Returns (in_tok[B,S], tgt[B,S]) as int64."""
    gt = gt.long()
    B, T = gt.shape
    start, end = vocab_size + 1, vocab_size + 2
    in_tok = torch.zeros(B, T + 1, dtype=torch.long)
    in_tok[:, 0] = start
    in_tok[:, 1:] = gt
    tgt = torch.zeros(B, T + 1, dtype=torch.long)
    tgt[:, :T] = gt
    for b in range(B):
        for t in range(1, T + 1):  # t=0 is never replaced (reference loop starts at 1)
            if tgt[b, t] == 0:
                tgt[b, t] = end
                break
    return in_tok, tgt


# --------------------------------------------------------------------------- loss
def ce_loss_and_grad(logits: torch.Tensor, tgt: torch.Tensor, eps: float = EPS_SMOOTH):
    """nn.CrossEntropyLoss(reduction='mean', ignore_index=0, label_smoothing=eps) on
    logits[N,C], tgt[N]  (CustomLoss.py:7-14).  Returns (loss, dlogits, n_valid)."""
    N, C = logits.shape
    mx = logits.max(dim=1, keepdim=True).values
    lse = mx + (logits - mx).exp().sum(dim=1, keepdim=True).log()
    logp = logits - lse
    valid = (tgt != 0)
    n_valid = int(valid.sum())
    nll = -logp.gather(1, tgt.view(-1, 1)).squeeze(1)
    smooth = -logp.sum(dim=1) / C
    row = (1.0 - eps) * nll + eps * smooth
    denom = max(n_valid, 1)
    loss = (row * valid).sum() / denom
    if n_valid == 0:
        loss = loss * float('nan')  # torch returns nan for an all-ignored batch
    onehot = torch.zeros_like(logits)
    onehot.scatter_(1, tgt.view(-1, 1), 1.0)
    dlogits = (logp.exp() - (1.0 - eps) * onehot - eps / C) * valid.view(-1, 1).to(logits.dtype) / denom
    return loss, dlogits, n_valid


def _sigmoid(x):
    return 1.0 / (1.0 + (-x).exp())


# --------------------------------------------------------------------------- attention decoder
ATTN_PARAM_NAMES = [
    'init_h.weight', 'init_h.bias', 'init_c.weight', 'init_c.bias',
    'f_beta.weight', 'f_beta.bias', 'deep_output.weight', 'deep_output.bias',
    'attention.U.weight', 'attention.U.bias', 'attention.W.weight', 'attention.W.bias',
    'attention.v.weight', 'attention.v.bias', 'embedding.weight',
    'lstm.weight_ih', 'lstm.weight_hh', 'lstm.bias_ih', 'lstm.bias_hh',
]


def attn_forward(p: dict, feat: torch.Tensor, in_tok: torch.Tensor, drop_mask=None):
    """Teacher-forced forward of Decoder (AttentionLanguageModule.py:62-84) with the
    time-invariant W(feat) hoisted out of the loop (the reference recomputes it per step,
    :17 -- same value).  drop_mask[B,S,H] (already scaled by 1/(1-p)) or None.
    Returns dict of everything the backward needs + preds[B,S,V'], alphas[B,S,L]."""
    B, L, D = feat.shape
    S = in_tok.shape[1]
    H = p['lstm.weight_hh'].shape[1]
    E = p['embedding.weight'].shape[1]
    dt = feat.dtype
    m = feat.mean(dim=1)                                                    # :117
    h = torch.tanh(m @ p['init_h.weight'].T + p['init_h.bias'])              # :122-123
    c = torch.tanh(m @ p['init_c.weight'].T + p['init_c.bias'])              # :119-120
    Ws = feat @ p['attention.W.weight'].T + p['attention.W.bias']            # :17  [B,L,H]
    wv = p['attention.v.weight'].view(-1)
    emb = p['embedding.weight'][in_tok]                                      # :69  [B,S,E]
    sv = dict(m=m, h0=h, c0=c, Ws=Ws, emb=emb,
              q=[], att=[], alpha=[], ctx=[], beta=[], gctx=[], hprev=[], cprev=[],
              gi=[], gf=[], gg=[], go=[], c=[], h=[])
    for t in range(S):                                                       # :72
        q = h @ p['attention.U.weight'].T + p['attention.U.bias']            # :16
        att = torch.tanh(Ws + q.unsqueeze(1))                                # :18
        e = att @ wv + p['attention.v.bias']                                 # :19
        e = e - e.max(dim=1, keepdim=True).values
        a = e.exp()
        alpha = a / a.sum(dim=1, keepdim=True)                               # :20
        ctx = (feat * alpha.unsqueeze(2)).sum(1)                             # :21
        beta = _sigmoid(h @ p['f_beta.weight'].T + p['f_beta.bias'])         # :74
        gctx = beta * ctx                                                    # :75
        x = torch.cat([emb[:, t], gctx], dim=1)                              # :76
        g = x @ p['lstm.weight_ih'].T + p['lstm.bias_ih'] + h @ p['lstm.weight_hh'].T + p['lstm.bias_hh']
        gi, gf, gg, go = g.split(H, dim=1)                                   # torch LSTMCell order i,f,g,o
        gi, gf, gg, go = _sigmoid(gi), _sigmoid(gf), torch.tanh(gg), _sigmoid(go)
        sv['hprev'].append(h); sv['cprev'].append(c)
        c = gf * c + gi * gg
        h = go * torch.tanh(c)                                               # :78
        for k, v in (('q', q), ('att', att), ('alpha', alpha), ('ctx', ctx), ('beta', beta), ('gctx', gctx),
                     ('gi', gi), ('gf', gf), ('gg', gg), ('go', go), ('c', c), ('h', h)):
            sv[k].append(v)
    hs = torch.stack(sv['h'], dim=1)                                          # [B,S,H]
    hd = hs * drop_mask if drop_mask is not None else hs                     # :79 dropout
    preds = hd @ p['deep_output.weight'].T + p['deep_output.bias']           # :79,:81
    alphas = torch.stack(sv['alpha'], dim=1)                                 # :82
    sv.update(hs=hs, hd=hd, preds=preds, alphas=alphas, drop_mask=drop_mask, in_tok=in_tok, feat=feat)
    return sv


def attn_backward(p: dict, sv: dict, dpreds: torch.Tensor, dalphas: torch.Tensor, need_dfeat=False):
    """Hand-written BPTT for attn_forward.  dpreds[B,S,V'], dalphas[B,S,L] are the upstream
    gradients (what autograd hands the reference's Decoder.forward outputs).
    Returns (grads dict keyed like state_dict, dfeat or None)."""
    feat = sv['feat']
    B, L, D = feat.shape
    S = dpreds.shape[1]
    H = p['lstm.weight_hh'].shape[1]
    E = p['embedding.weight'].shape[1]
    wv = p['attention.v.weight'].view(-1)
    Wih, Whh = p['lstm.weight_ih'], p['lstm.weight_hh']
    g = {k: torch.zeros_like(v) for k, v in p.items()}
    # vocab projection
    g['deep_output.weight'] = dpreds.reshape(B * S, -1).T @ sv['hd'].reshape(B * S, H)
    g['deep_output.bias'] = dpreds.sum(dim=(0, 1))
    dhd = dpreds @ p['deep_output.weight']                                    # [B,S,H]
    dhs = dhd * sv['drop_mask'] if sv['drop_mask'] is not None else dhd
    dh_next = torch.zeros(B, H, dtype=feat.dtype)
    dc_next = torch.zeros(B, H, dtype=feat.dtype)
    dWs = torch.zeros_like(sv['Ws'])
    dfeat = torch.zeros_like(feat) if need_dfeat else None
    demb = torch.zeros(B, S, E, dtype=feat.dtype)
    for t in reversed(range(S)):
        dh = dhs[:, t] + dh_next
        gi, gf, gg, go = sv['gi'][t], sv['gf'][t], sv['gg'][t], sv['go'][t]
        c, cprev, hprev = sv['c'][t], sv['cprev'][t], sv['hprev'][t]
        tc = torch.tanh(c)
        do_ = dh * tc * go * (1 - go)
        dc = dc_next + dh * go * (1 - tc * tc)
        di = dc * gg * gi * (1 - gi)
        df = dc * cprev * gf * (1 - gf)
        dg = dc * gi * (1 - gg * gg)
        dc_next = dc * gf
        dgates = torch.cat([di, df, dg, do_], dim=1)                          # [B,4H]
        x = torch.cat([sv['emb'][:, t], sv['gctx'][t]], dim=1)
        g['lstm.weight_ih'] += dgates.T @ x
        g['lstm.weight_hh'] += dgates.T @ hprev
        g['lstm.bias_ih'] += dgates.sum(0)
        g['lstm.bias_hh'] += dgates.sum(0)
        dx = dgates @ Wih                                                     # [B,E+D]
        dhp = dgates @ Whh                                                    # [B,H]
        demb[:, t] = dx[:, :E]
        dgctx = dx[:, E:]
        beta, ctx, alpha, att = sv['beta'][t], sv['ctx'][t], sv['alpha'][t], sv['att'][t]
        dbeta = dgctx * ctx
        dbpre = dbeta * beta * (1 - beta)
        dctx = dgctx * beta
        g['f_beta.weight'] += dbpre.T @ hprev
        g['f_beta.bias'] += dbpre.sum(0)
        dhp = dhp + dbpre @ p['f_beta.weight']
        dalpha = (feat * dctx.unsqueeze(1)).sum(2) + dalphas[:, t]             # [B,L]
        if need_dfeat:
            dfeat += alpha.unsqueeze(2) * dctx.unsqueeze(1)
        de = alpha * (dalpha - (alpha * dalpha).sum(1, keepdim=True))         # softmax bwd
        g['attention.v.bias'] += de.sum().view(1)
        g['attention.v.weight'] += (de.unsqueeze(2) * att).sum(dim=(0, 1)).view(1, -1)
        dz = de.unsqueeze(2) * wv.view(1, 1, -1) * (1 - att * att)            # [B,L,H]
        dWs += dz
        dq = dz.sum(1)                                                        # [B,H]
        g['attention.U.weight'] += dq.T @ hprev
        g['attention.U.bias'] += dq.sum(0)
        dhp = dhp + dq @ p['attention.U.weight']
        dh_next = dhp
    # hoisted projection and init state
    g['attention.W.weight'] = dWs.reshape(B * L, H).T @ feat.reshape(B * L, D)
    g['attention.W.bias'] = dWs.sum(dim=(0, 1))
    if need_dfeat:
        dfeat += dWs @ p['attention.W.weight']
    dh0 = dh_next * (1 - sv['h0'] ** 2)
    dc0 = dc_next * (1 - sv['c0'] ** 2)
    g['init_h.weight'] = dh0.T @ sv['m']; g['init_h.bias'] = dh0.sum(0)
    g['init_c.weight'] = dc0.T @ sv['m']; g['init_c.bias'] = dc0.sum(0)
    if need_dfeat:
        dm = dh0 @ p['init_h.weight'] + dc0 @ p['init_c.weight']
        dfeat += (dm / L).unsqueeze(1)
    g['embedding.weight'].index_put_((sv['in_tok'].reshape(-1),), demb.reshape(B * S, E), accumulate=True)
    return g, dfeat


def attn_train_step(p: dict, feat: torch.Tensor, gt: torch.Tensor, vocab_size: int,
                    drop_mask=None, need_dfeat=False):
    """Decoder half of AlexCapModel.forward_train (LSTMwAttentionModel.py:67-72):
    loss = CE(preds, target) + mean((1 - sum_t alpha)^2); backward.
    Returns dict(loss, loss_ce, loss_reg, grads, dfeat, preds, alphas)."""
    in_tok, tgt = make_targets(gt, vocab_size)
    sv = attn_forward(p, feat, in_tok, drop_mask)
    B, S, Vp = sv['preds'].shape
    L = feat.shape[1]
    loss_ce, dlog, n_valid = ce_loss_and_grad(sv['preds'].reshape(B * S, Vp), tgt.reshape(-1))
    resid = 1.0 - sv['alphas'].sum(1)                                         # [B,L]
    loss_reg = (resid ** 2).mean()                                            # :70
    dal = (-2.0 * resid / (B * L)).unsqueeze(1).expand(B, S, L)               # same for every t
    grads, dfeat = attn_backward(p, sv, dlog.view(B, S, Vp), dal, need_dfeat)
    return dict(loss=loss_ce + loss_reg, loss_ce=loss_ce, loss_reg=loss_reg, grads=grads, dfeat=dfeat,
                preds=sv['preds'], alphas=sv['alphas'], n_valid=n_valid, in_tok=in_tok, tgt=tgt)


# --------------------------------------------------------------------------- plain LSTM decoder
def _lstm_layer_fwd(x_seq, h, c, Wih, Whh, bih, bhh):
    """One nn.LSTM layer over x_seq[B,N,I] from state (h,c); returns outputs + saved."""
    H = Whh.shape[1]
    sv = dict(x=x_seq, hprev=[], cprev=[], gi=[], gf=[], gg=[], go=[], c=[], h=[])
    for t in range(x_seq.shape[1]):
        g = x_seq[:, t] @ Wih.T + bih + h @ Whh.T + bhh
        gi, gf, gg, go = g.split(H, dim=1)
        gi, gf, gg, go = _sigmoid(gi), _sigmoid(gf), torch.tanh(gg), _sigmoid(go)
        sv['hprev'].append(h); sv['cprev'].append(c)
        c = gf * c + gi * gg
        h = go * torch.tanh(c)
        for k, v in (('gi', gi), ('gf', gf), ('gg', gg), ('go', go), ('c', c), ('h', h)):
            sv[k].append(v)
    out = torch.stack(sv['h'], dim=1)
    return out, h, c, sv


def _lstm_layer_bwd(sv, dout, dh_last, dc_last, Wih, Whh):
    """BPTT through one layer.  dout[B,N,H] grads on outputs; (dh_last,dc_last) on final state.
    Returns dx_seq, dh0, dc0, dWih, dWhh, db."""
    x = sv['x']
    B, N, _ = x.shape
    dWih = torch.zeros_like(Wih); dWhh = torch.zeros_like(Whh); db = torch.zeros(Wih.shape[0], dtype=x.dtype)
    dx = torch.zeros_like(x)
    dh_next, dc_next = dh_last, dc_last
    for t in reversed(range(N)):
        dh = dout[:, t] + dh_next
        gi, gf, gg, go = sv['gi'][t], sv['gf'][t], sv['gg'][t], sv['go'][t]
        tc = torch.tanh(sv['c'][t])
        do_ = dh * tc * go * (1 - go)
        dc = dc_next + dh * go * (1 - tc * tc)
        di = dc * gg * gi * (1 - gi)
        df = dc * sv['cprev'][t] * gf * (1 - gf)
        dg = dc * gi * (1 - gg * gg)
        dc_next = dc * gf
        dgates = torch.cat([di, df, dg, do_], dim=1)
        dWih += dgates.T @ x[:, t]
        dWhh += dgates.T @ sv['hprev'][t]
        db += dgates.sum(0)
        dx[:, t] = dgates @ Wih
        dh_next = dgates @ Whh
    return dx, dh_next, dc_next, dWih, dWhh, db


def lstm_forward(p: dict, feat: torch.Tensor, in_tok: torch.Tensor, num_layers: int = 1, drop_mask=None):
    """LanguageModel.forward train branch (LanguageModule.py:99-107): encode every region,
    run them as an L-step sequence to get the state, then the S word steps, then vocab Linear."""
    B, L, D = feat.shape
    pre = feat @ p['image_encoder.encode.weight'].T + p['image_encoder.encode.bias']
    enc = pre.clamp_min(0)                                                    # :36-37,:103
    emb = p['lookup_table.weight'][in_tok]                                    # :102
    H = p['lstm.weight_hh_l0'].shape[1]
    z = torch.zeros(B, H, dtype=feat.dtype)
    img_sv, word_sv = [], []
    x_img, x_word = enc, emb
    for l in range(num_layers):
        w = [p[f'lstm.weight_ih_l{l}'], p[f'lstm.weight_hh_l{l}'], p[f'lstm.bias_ih_l{l}'], p[f'lstm.bias_hh_l{l}']]
        x_img, h, c, s1 = _lstm_layer_fwd(x_img, z, z, *w)                    # :104
        x_word, _, _, s2 = _lstm_layer_fwd(x_word, h, c, *w)                  # :105
        img_sv.append(s1); word_sv.append(s2)
    hs = x_word
    hd = hs * drop_mask if drop_mask is not None else hs
    preds = hd @ p['rnn.linear.weight'].T + p['rnn.linear.bias']              # :106
    return dict(feat=feat, pre=pre, enc=enc, emb=emb, in_tok=in_tok, img_sv=img_sv, word_sv=word_sv,
                hs=hs, hd=hd, preds=preds, drop_mask=drop_mask, num_layers=num_layers)


def lstm_backward(p: dict, sv: dict, dpreds: torch.Tensor, need_dfeat=False):
    feat = sv['feat']
    B, L, D = feat.shape
    S = dpreds.shape[1]
    H = sv['hs'].shape[2]
    g = {k: torch.zeros_like(v) for k, v in p.items()}
    g['rnn.linear.weight'] = dpreds.reshape(B * S, -1).T @ sv['hd'].reshape(B * S, H)
    g['rnn.linear.bias'] = dpreds.sum(dim=(0, 1))
    dhd = dpreds @ p['rnn.linear.weight']
    dword = dhd * sv['drop_mask'] if sv['drop_mask'] is not None else dhd
    dimg = torch.zeros(B, L, H, dtype=feat.dtype)   # nothing reads the image-pass outputs of the top layer
    z = torch.zeros(B, H, dtype=feat.dtype)
    for l in reversed(range(sv['num_layers'])):
        Wih, Whh = p[f'lstm.weight_ih_l{l}'], p[f'lstm.weight_hh_l{l}']
        dxw, dh, dc, a1, b1, c1 = _lstm_layer_bwd(sv['word_sv'][l], dword, z, z, Wih, Whh)
        dxi, _, _, a2, b2, c2 = _lstm_layer_bwd(sv['img_sv'][l], dimg, dh, dc, Wih, Whh)
        g[f'lstm.weight_ih_l{l}'] = a1 + a2
        g[f'lstm.weight_hh_l{l}'] = b1 + b2
        g[f'lstm.bias_ih_l{l}'] = c1 + c2
        g[f'lstm.bias_hh_l{l}'] = c1 + c2
        dword, dimg = dxw, dxi
    g['lookup_table.weight'].index_put_((sv['in_tok'].reshape(-1),), dword.reshape(B * S, -1), accumulate=True)
    dpre = dimg * (sv['pre'] > 0).to(feat.dtype)
    g['image_encoder.encode.weight'] = dpre.reshape(B * L, -1).T @ feat.reshape(B * L, D)
    g['image_encoder.encode.bias'] = dpre.sum(dim=(0, 1))
    dfeat = dpre @ p['image_encoder.encode.weight'] if need_dfeat else None
    return g, dfeat


def lstm_train_step(p: dict, feat: torch.Tensor, gt: torch.Tensor, vocab_size: int, num_layers: int = 1,
                    drop_mask=None, need_dfeat=False):
    """Decoder half of AlexCapModel.forward_train for the plain LSTM (LSTMModel.py:70-73)."""
    in_tok, tgt = make_targets(gt, vocab_size)
    sv = lstm_forward(p, feat, in_tok, num_layers, drop_mask)
    B, S, Vp = sv['preds'].shape
    loss, dlog, n_valid = ce_loss_and_grad(sv['preds'].reshape(B * S, Vp), tgt.reshape(-1))
    grads, dfeat = lstm_backward(p, sv, dlog.view(B, S, Vp), need_dfeat)
    return dict(loss=loss, loss_ce=loss, grads=grads, dfeat=dfeat, preds=sv['preds'], n_valid=n_valid,
                in_tok=in_tok, tgt=tgt)


# --------------------------------------------------------------------------- decoding
def attn_greedy(p: dict, feat: torch.Tensor, vocab_size: int, seq_length: int):
    """Greedy branch of Decoder.forward (AttentionLanguageModule.py:88-115) with the B>1
    shape bug at :108-109 fixed (token.squeeze(1)); eval mode (no dropout).
    Returns (ids[B,S] int64, alphas[B,S,L], margins[B,S] top1-top2 logit gap)."""
    B, L, D = feat.shape
    S = seq_length + 1
    H = p['lstm.weight_hh'].shape[1]
    start, end = vocab_size + 1, vocab_size + 2
    m = feat.mean(dim=1)
    h = torch.tanh(m @ p['init_h.weight'].T + p['init_h.bias'])
    c = torch.tanh(m @ p['init_c.weight'].T + p['init_c.bias'])
    Ws = feat @ p['attention.W.weight'].T + p['attention.W.bias']
    wv = p['attention.v.weight'].view(-1)
    tok = torch.full((B,), start, dtype=torch.long)
    ids = torch.zeros(B, S, dtype=torch.long)
    alphas = torch.zeros(B, S, L, dtype=feat.dtype)
    margins = torch.full((B, S), float('inf'), dtype=feat.dtype)
    for t in range(S):
        q = h @ p['attention.U.weight'].T + p['attention.U.bias']
        e = torch.tanh(Ws + q.unsqueeze(1)) @ wv + p['attention.v.bias']
        alpha = torch.softmax(e, dim=1)
        ctx = (feat * alpha.unsqueeze(2)).sum(1)
        beta = _sigmoid(h @ p['f_beta.weight'].T + p['f_beta.bias'])
        x = torch.cat([p['embedding.weight'][tok], beta * ctx], dim=1)
        g = x @ p['lstm.weight_ih'].T + p['lstm.bias_ih'] + h @ p['lstm.weight_hh'].T + p['lstm.bias_hh']
        gi, gf, gg, go = g.split(H, dim=1)
        c = _sigmoid(gf) * c + _sigmoid(gi) * torch.tanh(gg)
        h = _sigmoid(go) * torch.tanh(c)
        out = h @ p['deep_output.weight'].T + p['deep_output.bias']
        top2 = out.topk(2, dim=1).values
        margins[:, t] = top2[:, 0] - top2[:, 1]
        tok = out.argmax(1)
        ids[:, t] = tok
        alphas[:, t] = alpha
        if bool((tok == end).all()):                                          # :111
            break
    return ids, alphas, margins


def attn_beam(p: dict, feat1: torch.Tensor, vocab_size: int, seq_length: int, beam_size: int):
    """Decoder.caption (AttentionLanguageModule.py:128-195) for ONE image feat1[1,L,D].
    Scores are cumulative RAW logits; step 1 takes top-k from row 0; finished beams shrink k;
    result = completed beam with the max score.  Raises ValueError when nothing completed
    (as the reference does at :191).  Returns (ids list without START, score)."""
    k = beam_size
    Vp = vocab_size + 3
    start, end = vocab_size + 1, vocab_size + 2
    H = p['lstm.weight_hh'].shape[1]
    feat = feat1.expand(k, -1, -1)
    wv = p['attention.v.weight'].view(-1)
    prev = torch.full((k,), start, dtype=torch.long)
    sentences = prev.view(k, 1)
    top = torch.zeros(k, 1, dtype=feat.dtype)
    m = feat.mean(dim=1)
    h = torch.tanh(m @ p['init_h.weight'].T + p['init_h.bias'])
    c = torch.tanh(m @ p['init_c.weight'].T + p['init_c.bias'])
    done, done_scores = [], []
    step = 1
    while True:
        Ws = feat @ p['attention.W.weight'].T + p['attention.W.bias']
        q = h @ p['attention.U.weight'].T + p['attention.U.bias']
        e = torch.tanh(Ws + q.unsqueeze(1)) @ wv + p['attention.v.bias']
        alpha = torch.softmax(e, dim=1)
        ctx = (feat * alpha.unsqueeze(2)).sum(1)
        beta = _sigmoid(h @ p['f_beta.weight'].T + p['f_beta.bias'])
        x = torch.cat([p['embedding.weight'][prev], beta * ctx], dim=1)
        g = x @ p['lstm.weight_ih'].T + p['lstm.bias_ih'] + h @ p['lstm.weight_hh'].T + p['lstm.bias_hh']
        gi, gf, gg, go = g.split(H, dim=1)
        c = _sigmoid(gf) * c + _sigmoid(gi) * torch.tanh(gg)
        h = _sigmoid(go) * torch.tanh(c)
        out = h @ p['deep_output.weight'].T + p['deep_output.bias']
        out = top.expand_as(out) + out                                        # :156
        if step == 1:
            top_s, top_w = out[0].topk(k, 0, True, True)                      # :158-159
        else:
            top_s, top_w = out.reshape(-1).topk(k, 0, True, True)             # :161
        prev_idx = top_w // Vp
        next_w = top_w % Vp
        sentences = torch.cat([sentences[prev_idx], next_w.view(-1, 1)], dim=1)
        inc = [i for i, w in enumerate(next_w.tolist()) if w != end]
        comp = [i for i, w in enumerate(next_w.tolist()) if w == end]
        for i in comp:
            done.append(sentences[i].tolist()); done_scores.append(float(top_s[i]))
        k -= len(comp)
        if k == 0:
            break
        sentences = sentences[inc]
        h = h[prev_idx[inc]]; c = c[prev_idx[inc]]
        feat = feat[prev_idx[inc]]
        top = top_s[inc].view(-1, 1)
        prev = next_w[inc]
        if step > seq_length:                                                 # :187-188
            break
        step += 1
    if not done:
        raise ValueError('no completed beam')
    best = done_scores.index(max(done_scores))                               # :191 first max wins
    return done[best][1:], done_scores[best]


def lstm_greedy(p: dict, feat: torch.Tensor, vocab_size: int, seq_length: int, num_layers: int = 1):
    """LanguageModel.predict_caption (LanguageModule.py:130-150).  Returns ids[B,S] int64."""
    B, L, D = feat.shape
    S = seq_length + 1
    start, end = vocab_size + 1, vocab_size + 2
    enc = (feat @ p['image_encoder.encode.weight'].T + p['image_encoder.encode.bias']).clamp_min(0)
    H = p['lstm.weight_hh_l0'].shape[1]
    z = torch.zeros(B, H, dtype=feat.dtype)
    states = []
    x = enc
    for l in range(num_layers):
        w = [p[f'lstm.weight_ih_l{l}'], p[f'lstm.weight_hh_l{l}'], p[f'lstm.bias_ih_l{l}'], p[f'lstm.bias_hh_l{l}']]
        x, h, c, _ = _lstm_layer_fwd(x, z, z, *w)
        states.append((h, c))
    ids = torch.zeros(B, S, dtype=torch.long)
    tok = torch.full((B,), start, dtype=torch.long)
    for i in range(S):
        x = p['lookup_table.weight'][tok].unsqueeze(1)
        for l in range(num_layers):
            w = [p[f'lstm.weight_ih_l{l}'], p[f'lstm.weight_hh_l{l}'], p[f'lstm.bias_ih_l{l}'], p[f'lstm.bias_hh_l{l}']]
            x, h, c, _ = _lstm_layer_fwd(x, states[l][0], states[l][1], *w)
            states[l] = (h, c)
        out = x[:, 0] @ p['rnn.linear.weight'].T + p['rnn.linear.bias']
        tok = out.argmax(1)                    # argmax(log_softmax) == argmax(logits), :141-142
        if bool((tok == end).all()):           # :145-147
            ids[:, i] = end
            break
        ids[:, i] = tok
    return ids


def lstm_beam(p: dict, feat1: torch.Tensor, vocab_size: int, seq_length: int, beam_size: int, num_layers: int = 1):
    """LanguageModel.caption (LanguageModule.py:152-198) for ONE image feat1[1,L,D].
    The RAW features go through the LSTM (no image_encoder, :160), so D must equal the embedding size; scores are
    cumulative raw logits; step 1 takes top-k from row 0 (:167-168); finished beams shrink k; the loop runs while
    `step <= seq_length+1` is true at the check AFTER a step, i.e. up to seq_length+2 steps (:190-192); result =
    completed beam with the max score, first on ties (:194).  Raises ValueError when nothing completed.
    Returns (ids list INCLUDING the leading START, as the reference does, score)."""
    k = beam_size
    Vp = vocab_size + 3
    start, end = vocab_size + 1, vocab_size + 2
    H = p['lstm.weight_hh_l0'].shape[1]
    W = [[p[f'lstm.weight_ih_l{l}'], p[f'lstm.weight_hh_l{l}'], p[f'lstm.bias_ih_l{l}'], p[f'lstm.bias_hh_l{l}']]
         for l in range(num_layers)]
    x = feat1.expand(k, -1, -1)
    z = torch.zeros(k, H, dtype=feat1.dtype)
    states = []
    for l in range(num_layers):                                                # :160
        x, h, c, _ = _lstm_layer_fwd(x, z, z, *W[l])
        states.append((h, c))
    prev = torch.full((k,), start, dtype=torch.long)
    sentences = prev.view(k, 1)
    top = torch.zeros(k, 1, dtype=feat1.dtype)
    done, done_scores = [], []
    step = 1
    while True:
        x = p['lookup_table.weight'][prev].unsqueeze(1)                        # :163
        for l in range(num_layers):
            x, h, c, _ = _lstm_layer_fwd(x, states[l][0], states[l][1], *W[l])
            states[l] = (h, c)
        out = x[:, 0] @ p['rnn.linear.weight'].T + p['rnn.linear.bias']        # :165 (eval: dropout is the identity)
        out = top.expand_as(out) + out                                         # :166
        if step == 1:
            top_s, top_w = out[0].topk(k, 0)                                   # :168-169
        else:
            top_s, top_w = out.reshape(-1).topk(k, 0)                          # :171
        prev_idx = top_w // Vp
        next_w = top_w % Vp
        sentences = torch.cat([sentences[prev_idx], next_w.view(-1, 1)], dim=1)
        inc = [i for i, w in enumerate(next_w.tolist()) if w != end]
        comp = [i for i, w in enumerate(next_w.tolist()) if w == end]
        for i in comp:
            done.append(sentences[i].tolist()); done_scores.append(float(top_s[i]))
        k -= len(comp)
        if k == 0:
            break
        sentences = sentences[inc]
        states = [(h[prev_idx[inc]], c[prev_idx[inc]]) for h, c in states]     # :186
        top = top_s[inc].view(-1, 1)
        prev = next_w[inc]
        if step > seq_length + 1:                                              # :190-191
            break
        step += 1
    if not done:
        raise ValueError('no completed beam')
    best = done_scores.index(max(done_scores))
    return done[best], done_scores[best]
