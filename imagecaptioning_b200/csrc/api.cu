// api.cu -- the C ABI of libalexcap_b200.so (include/alexcap.h) and the host-side orchestration of the
// caption-decoder training step: which kernels run, in which order, over which workspace slices.
//
// Structure of one attention training step (reference: AlexCap/LSTMwAttentionModel.py:67-72 driving
// AlexCap/AttentionLanguageModule.py:62-84 and AlexCap/CustomLoss.py:9-15):
//   hoisted (once):  targets, mean_L(feat), h0/c0, Ws = W(feat) [reference: recomputed S times], embedding gather,
//                    Xe = Emb W_ih[:, :E]^T + b_ih + b_hh for all S steps, weight concatenations
//   per step t:      [q | bpre] = h_{t-1} [W_U ; W_beta]^T      (1 GEMM)
//                    fused attention step                        (1 kernel)
//                    gates_t += [gctx_t | h_{t-1}] [W_ih[:,E:] | W_hh]^T   (1 GEMM)
//                    LSTM pointwise -> h_t, c_t                  (1 kernel)
//   after the loop:  logits = H_all W_o^T + b_o, CE + gradient, regulariser
//   reverse loop:    pointwise bwd, dgctx GEMM, fused attention bwd, dh_{t-1} GEMM
//   after the loop:  every weight gradient as ONE GEMM over K = S*B rows (the "fused dW accumulation"),
//                    dWs in one pass, embedding scatter-add, init-state gradients.
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

#include "api_util.h"

using namespace alexcap;

namespace alexcap {
char* err_buf() {
  static thread_local char buf[512] = "";
  return buf;
}
unsigned long long& launch_counter() {
  static unsigned long long n = 0;
  return n;
}
bool& pdl_enabled() {
  static bool on = [] { const char* v = getenv("ALEXCAP_PDL"); return !(v && v[0] == '0'); }();
  return on;
}
}  // namespace alexcap

namespace {

// ------------------------------------------------------------------------------------------ attention workspace
struct AttnWs {
  int32_t *in_bm, *tgt_bm, *in_tm, *tgt_tm, *bad_rows;
  float *m, *XH, *Call, *Ws, *EmbIn, *DGQ, *XG, *QB, *BETA, *CTX, *ALPHA, *DE;
  float *Wcat, *Wub, *bub, *Wback;
  float *MASK, *Hd;
  float *logits, *row_loss, *scal;   // scal: [0]=n_valid [1]=ce_sum [2]=reg
  float *dHout, *dh_next, *dc, *dgctx, *R, *dwv_part, *dWs, *dEmbIn, *dh0, *dc0, *dm;
  float *cs_scratch, *cs_scratch2, *reg_scratch;
  uint8_t* arena; size_t arena_bytes;   // bf16 hi/lo operand copies of the big GEMMs (training layout only)
  uint16_t* loop_bf16; unsigned* loop_bar;   // persistent time loop: bf16 operand copies of h / gctx / weight slices; grid barrier
  float4* ce_stats; float* row_lse;          // fused vocabulary CE: per-row partial statistics, log-sum-exp
  int64_t ldV;
  size_t bytes;
};

// ALEXCAP_LSTM_Q_FUSED=1: fuse the LSTM cell into the next step's query GEMM (see attn_fwd_core); read once per process,
// because the workspace layout (XG) depends on it.
static bool cell_fusion_enabled() {
  static const bool on = [] { const char* v = getenv("ALEXCAP_LSTM_Q_FUSED"); return v && v[0] == '1'; }();
  return on;
}

AttnWs attn_layout(const alexcap_dims* d, void* base, bool with_logits) {
  AttnWs w;
  Bump a(base);
  const int64_t B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, Vp = d->Vp, S = d->S, W5 = 5 * H + D;
  w.ldV = (Vp + 3) & ~int64_t(3);
  w.in_bm = a.take<int32_t>(B * S); w.tgt_bm = a.take<int32_t>(B * S);
  w.in_tm = a.take<int32_t>(B * S); w.tgt_tm = a.take<int32_t>(B * S);
  w.bad_rows = a.take<int32_t>(B);
  w.m = a.take<float>(B * D);
  w.XH = a.take<float>((S + 1) * B * (D + H));
  w.Call = a.take<float>((S + 1) * B * H);
  w.Ws = a.take<float>(B * L * H);
  w.EmbIn = a.take<float>(S * B * E);
  w.DGQ = a.take<float>(S * B * W5);
  w.XG = cell_fusion_enabled() ? a.take<float>(S * B * 4 * H) : nullptr;   // gate PRE-activations of the fused cell + query GEMM (attn_fwd_core)
  w.QB = a.take<float>(S * B * (H + D));
  w.BETA = a.take<float>(S * B * D);
  w.CTX = a.take<float>(S * B * D);
  w.ALPHA = a.take<float>(B * S * L);
  w.DE = a.take<float>(B * S * L);
  w.Wcat = a.take<float>(4 * H * (D + H));
  w.Wub = a.take<float>((H + D) * H);
  w.bub = a.take<float>(H + D);
  w.Wback = a.take<float>(W5 * H);
  const bool drop = d->dropout_p > 0.f;
  w.MASK = drop ? a.take<float>(S * B * H) : nullptr;
  w.Hd = drop ? a.take<float>(S * B * H) : nullptr;
  const bool ce_fused = with_logits && vocab_ce_fused_ok((int)(S * B), (int)Vp, (int)H);
  w.logits = (with_logits && !ce_fused) ? a.take<float>((int64_t)vocab_chunk_rows((int)(S * B), w.ldV) * w.ldV) : nullptr;   // ONE chunk
  w.ce_stats = ce_fused ? a.take<float4>(S * B * gemm_ps_ce_partials((int)Vp)) : nullptr;   // fused CE: no logits anywhere
  w.row_lse = ce_fused ? a.take<float>(S * B) : nullptr;
  w.row_loss = a.take<float>(S * B);
  w.scal = a.take<float>(64);
  w.dHout = a.take<float>(S * B * H);
  w.dh_next = a.take<float>(S * B * H);       // slot t: dh_{t-1} produced by reverse step t (zeroed once per step)
  w.dc = a.take<float>(B * H);
  w.dgctx = a.take<float>(S * B * D);         // slot t: d gctx_t (zeroed once per step: split-K GEMMs red.add into it)
  w.R = a.take<float>(B * L);
  w.dwv_part = a.take<float>(B * H);
  w.dWs = a.take<float>(B * L * H);
  w.dEmbIn = a.take<float>(S * B * E);
  w.dh0 = a.take<float>(2 * B * H);      // [B, 2H] = [dh0 | dc0]: one column-sum pass gives both init biases
  w.dc0 = w.dh0 ? w.dh0 + H : nullptr;
  w.dm = a.take<float>(B * D);
  int64_t maxn = Vp;
  if (W5 > maxn) maxn = W5;
  w.cs_scratch = a.take<float>(64 * maxn);
  w.cs_scratch2 = a.take<float>(64 * Vp);        // column-sum scratch of the deferred dW_o / db_o work (runs on a side stream)
  w.reg_scratch = a.take<float>(256);
  w.loop_bar = a.take<unsigned>(64);
  {
    const size_t n = attn_loop_workspace_elems_bf16((int)B, (int)S, (int)D, (int)H);
    w.loop_bf16 = n ? a.take<uint16_t>(n) : nullptr;
  }
  w.arena = nullptr; w.arena_bytes = 0;
  if (with_logits) {
    w.arena_bytes = split_arena_bytes(B, L, D, H, E, Vp, S, vocab_chunk_rows((int)(S * B), w.ldV));
    a.off = (a.off + 1023) & ~size_t(1023);
    w.arena = reinterpret_cast<uint8_t*>(a.take<char>(w.arena_bytes));
  }
  w.bytes = a.off + 256;
  return w;
}

// ALEXCAP_BF16 storage of the attention streams (BASELINE.json configs[2] / [4]): the per-step attention kernels read Ws and
// feat 2 * S times per training step; in bf16 mode they read bf16 copies (half the bytes per launch).  feat's copy is the one
// the hoisted Ws GEMM made of its A operand; Ws gets one split pass per training step.  NULL pointers = stream fp32.
// bf16 configurations (engine 2) whose per-step GEMMs run on the pre-split engine (M = B >= 256): the kernels that PRODUCE the
// per-step GEMM operands -- attention (gctx, dq, dbpre) and the LSTM pointwise kernels (h, d gates) -- also write them as
// bf16 into a mirror of XH / DGQ in the operand arena, so none of the 4 split passes per time step is needed (they sat on
// the serial chain: 8.6 % of the cfg3 step) and the post-loop dW GEMMs find their operands ready as well.
// ALEXCAP_BF16_MIRROR=0 disables (A/B timing).
static bool bf16_mirror_ok(int B, int D, int H) {
  static const bool off = [] { const char* v = getenv("ALEXCAP_BF16_MIRROR"); return v && v[0] == '0'; }();
  return !off && gemm_engine() == 2 && split_arena() && attn_engine() == 1 && attn_stream_supported(D, H) && !(D & 7) && !(H & 7) &&
         gemm_ps_wanted(B, H + D, H) && gemm_ps_wanted(B, 4 * H, D + H);
}
struct AttnRows { const float* Ws; const float* feat; bool bf16; };
// The streaming attention kernels take the hoisted Ws = W(features) in exponential form (ACT_EXP2C, common.cuh): one MUFU op
// per score element instead of two in the forward, backward and dWs kernels.  Engine 0 (direct-load kernels) keeps raw Ws.
// ALEXCAP_ATT_WS_EXP=0 switches the factoring off (A/B timing).
static bool attn_ws_exp(int D, int H) {
  static const bool off = [] { const char* v = getenv("ALEXCAP_ATT_WS_EXP"); return v && v[0] == '0'; }();
  return !off && attn_engine() == 1 && attn_stream_supported(D, H);
}
AttnRows attn_rows(const alexcap_dims* d, const float* feat, const AttnWs& w, cudaStream_t st) {
  AttnRows r{w.Ws, feat, false};
  static const bool off = [] { const char* v = getenv("ALEXCAP_ATT_BF16_ROWS"); return v && v[0] == '0'; }();
  SplitArena* arena = split_arena();
  if (off || d->dtype != ALEXCAP_BF16 || !arena || attn_engine() != 1 || !attn_stream_supported(d->D, d->H) || (d->D & 7) || (d->H & 7))
    return r;
  const uint16_t *fh = nullptr, *wh = nullptr;
  int64_t fld = 0, wld = 0;
  const size_t mark = arena->used;
  if (split_arena_get_hi(st, arena, feat, d->B * d->L, d->D, d->D, &fh, &fld) != cudaSuccess ||
      split_arena_get_hi(st, arena, w.Ws, d->B * d->L, d->H, d->H, &wh, &wld) != cudaSuccess || fld != d->D || wld != d->H) {
    split_arena_release(arena, mark);
    return r;
  }
  r.Ws = reinterpret_cast<const float*>(wh); r.feat = reinterpret_cast<const float*>(fh); r.bf16 = true;
  return r;
}

// ------------------------------------------------------------------------------------------ attention forward core
int attn_fwd_core(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, const int32_t* gt,
                  const float* drop_mask, AttnWs& w, cudaStream_t st) {
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, Vp = d->Vp, S = d->S;
  const int64_t DH = D + H, W5 = 5 * (int64_t)H + D, HD = H + D;
  const int V = Vp - 3;
  API_CK(make_target(st, gt, B, S - 1, V + 1, V + 2, w.in_bm, w.tgt_bm, w.in_tm, w.tgt_tm, Vp, w.bad_rows));
  // initial state (AttentionLanguageModule.py:116-125)
  API_CK(mean_over_L(st, feat, B, L, D, w.m));
  API_CK(gemm_f32(st, false, true, B, H, D, 1.f, w.m, D, P->init_h_w, D, w.XH + D, DH, P->init_h_b, nullptr, false, ACT_TANH));
  API_CK(gemm_f32(st, false, true, B, H, D, 1.f, w.m, D, P->init_c_w, D, w.Call, H, P->init_c_b, nullptr, false, ACT_TANH));
  // hoisted projections
  const bool ws_exp = attn_ws_exp(D, H);
  API_CK(gemm_f32(st, false, true, B * L, H, D, 1.f, feat, D, P->att_W_w, D, w.Ws, H, P->att_W_b, nullptr, false,
                  ws_exp ? ACT_EXP2C : ACT_NONE));
  API_CK(gather_rows(st, P->emb_w, w.in_tm, S * B, E, w.EmbIn, E));
  // The recurrence.  Default (tcgen05 engines, shapes whose recurrent weights fit in shared memory): ONE persistent kernel
  // for all S steps (time_loop.cu).  Otherwise the per-step chain below.
  const bool loop_ok = w.loop_bf16 && gemm_engine() >= 2 && ws_exp && attn_loop_supported(B, L, D, H);   // (the loop's attention body reads EXP2C(Ws))
  // Per-step chain, north_star (3), opt-in (ALEXCAP_LSTM_Q_FUSED=1): the LSTM cell of step t is computed by the converter
  // warps of step t+1's query / beta GEMM (gemm_tc_lstm_f32) -- three launches per step instead of four.  The gate
  // PRE-activations then live in XG (the hoisted embedding part + the gates GEMM), the activations the backward pass reads
  // go to DGQ as before.  Measured at cfg2: 2.006 ms vs 1.993 ms for the separate launches -- every one of the 8 column
  // tiles has to repeat the cell math of its 128 x 32 k-block (4.2 us of gate loads + MUFU on the GEMM's critical path)
  // which costs what the removed kernel boundary saves -- so the default stays the four-launch chain.
  const bool cell_fused = !loop_ok && cell_fusion_enabled() && (gemm_engine() == 2 || gemm_engine() == 3) && pdl_enabled() &&
                          gemm_tc_lstm_supported(B, H + D, H);
  const int64_t GL = cell_fused ? 4 * (int64_t)H : W5;                 // row pitch of the gate pre-activations
  float* const gates_pre = cell_fused ? w.XG : w.DGQ;
  API_CK(gemm_f32(st, false, true, S * B, 4 * H, E, 1.f, w.EmbIn, E, P->lstm_w_ih, E + D, gates_pre, GL, P->lstm_b_ih,
                  P->lstm_b_hh, false, ACT_NONE));
  // concatenated weights: Wcat = [W_ih[:,E:] | W_hh] (4H x (D+H));  Wub = [W_U ; W_beta] ((H+D) x H)
  if (!loop_ok) {
    CopySegments cs;
    cs.add(P->lstm_w_ih + E, E + D, w.Wcat, DH, 4 * H, D);
    cs.add(P->lstm_w_hh, H, w.Wcat + D, DH, 4 * H, H);
    cs.add(P->att_U_w, H, w.Wub, H, H, H);
    cs.add(P->f_beta_w, H, w.Wub + (int64_t)H * H, H, D, H);
    cs.add(P->att_U_b, H, w.bub, H, 1, H);
    cs.add(P->f_beta_b, D, w.bub + H, D, 1, D);
    API_CK(copy_segments(st, cs));
  }

  if (loop_ok) {
    API_CK(attn_loop_fwd(st, B, L, D, H, E, S, feat, w.Ws, P->lstm_w_hh, P->lstm_w_ih, P->att_U_w, P->att_U_b, P->f_beta_w,
                         P->f_beta_b, P->att_v_w, w.QB, w.XH, w.DGQ, w.Call, w.ALPHA, w.CTX, w.BETA, w.loop_bf16, w.loop_bar));
  } else {
  // the per-step split-K GEMMs accumulate with red.add: their outputs are zeroed here, once, instead of by a memset
    // node in front of every GEMM (which would also cut the programmatic-dependent-launch chain of the time loop)
    API_CK(cudaMemsetAsync(w.QB, 0, sizeof(float) * (size_t)S * B * HD, st));
    const AttnRows rows = attn_rows(d, feat, w, st);
    uint16_t *xhb = nullptr, *xhb_lo = nullptr;
    int64_t xld = 0;
    const bool mirror = !cell_fused && (rows.bf16 || ws_exp) && bf16_mirror_ok(B, D, H) &&
                        split_arena_reserve_mirror(split_arena(), w.XH, (S + 1) * B, (int)DH, DH, false, &xhb, &xhb_lo, &xld);
    if (mirror) API_CK(split_arena_fill_view(st, split_arena(), w.XH + D, B, H, DH));     // h_0 (written by the init GEMM)
    for (int t = 0; t < S; ++t) {                                            // AttentionLanguageModule.py:72
      float* xh = w.XH + (int64_t)t * B * DH;
      float* qb = w.QB + (int64_t)t * B * HD;
      float* g = w.DGQ + (int64_t)t * B * W5;
      float* gp = gates_pre + (int64_t)t * B * GL;
      if (t == 0 || !cell_fused)     // (fused: q_t / bpre_t were produced together with the cell of step t-1)
      { WeightBScope wb; API_CK(gemm_f32(st, false, true, B, H + D, H, 1.f, xh + D, DH, w.Wub, H, qb, HD, w.bub, nullptr, true, ACT_NONE)); }
      if (rows.bf16 || ws_exp)
        API_CK(attn_step_fwd_stream(st, B, L, D, H, rows.Ws, rows.feat, qb, HD, qb + H, HD, P->att_v_w, w.ALPHA + (int64_t)t * L,
                                    (int64_t)S * L, w.CTX + (int64_t)t * B * D, w.BETA + (int64_t)t * B * D, xh, DH, 1, rows.bf16,
                                    ws_exp, true, mirror ? xhb + (int64_t)t * B * xld : nullptr, xld));
      else
        API_CK(attn_step_fwd(st, B, L, D, H, w.Ws, feat, qb, HD, qb + H, HD, P->att_v_w, w.ALPHA + (int64_t)t * L,
                             (int64_t)S * L, w.CTX + (int64_t)t * B * D, w.BETA + (int64_t)t * B * D, xh, DH));
      { WeightBScope wb; API_CK(gemm_f32(st, false, true, B, 4 * H, D + H, 1.f, xh, DH, w.Wcat, DH, gp, GL, nullptr, nullptr, true, ACT_NONE)); }
      float* c_t = w.Call + (int64_t)t * B * H;
      float* h_next = xh + (int64_t)B * DH + D;                              // the h columns of step t+1's [gctx | h] row
      if (cell_fused && t + 1 < S) {
        const LstmA la{gp, GL, c_t, c_t + (int64_t)B * H, h_next, DH, g, W5};
        API_CK(gemm_tc_lstm_f32(st, B, H + D, H, la, w.Wub, H, qb + (int64_t)B * HD, HD, w.bub, gemm_engine() == 2 ? 1 : 2));
      } else {
        API_CK(lstm_pointwise_fwd(st, B, H, gp, GL, c_t, c_t + (int64_t)B * H, h_next, DH, nullptr, 0,
                                  cell_fused ? g : nullptr, W5, mirror ? xhb + (int64_t)(t + 1) * B * xld + D : nullptr, xld));
      }
    }
  }
  if (d->dropout_p > 0.f) {                                                 // :79 dropout(h)
    if (drop_mask) {
      API_CK(cudaMemcpyAsync(w.MASK, drop_mask, sizeof(float) * S * B * H, cudaMemcpyDeviceToDevice, st));
      API_CK(mul_2d(st, S * B, H, w.XH + (int64_t)B * DH + D, DH, w.MASK, H, w.Hd, H));
    } else {              // mask drawn and applied in one pass; counter on the device when the caller gave one (graph replays)
      API_CK(dropout_mask(st, w.MASK, (int64_t)S * B * H, d->dropout_p, d->seed, d->offset, d->philox_state,
                          w.XH + (int64_t)B * DH + D, DH, H, w.Hd));
    }
  }
  return 0;
}

// A second stream + fork / join events per (thread, device) for the one place where two independent kernels of different
// pipes are worth running side by side; inside a stream capture the event pair becomes a fork / join of the graph.
struct SideStream { cudaStream_t s = nullptr; cudaEvent_t fork = nullptr, join = nullptr; int dev = -1; bool ok = false; };
SideStream& side_stream() {
  static thread_local SideStream ss;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { ss.ok = false; return ss; }
  if (ss.dev != dev) {
    ss.ok = cudaStreamCreateWithFlags(&ss.s, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming) == cudaSuccess &&
            cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming) == cudaSuccess;
    ss.dev = dev;
  }
  return ss;
}
bool dws_overlap_enabled() {
  // measured at cfg2: 2.08-2.17 ms/step with the fork (any CTA cap) vs 2.06 without -- the MUFU kernel slows the co-resident
  // persistent GEMM CTAs more than it hides -- so the fork is OFF unless ALEXCAP_DWS_OVERLAP=1
  static const bool on = [] { const char* v = getenv("ALEXCAP_DWS_OVERLAP"); return v && v[0] == '1'; }();
  return on;
}

// Second fork (its own stream and events): dW_o / db_o of the vocabulary projection underneath the reverse-time loop.
SideStream& side_stream2() {
  static thread_local SideStream ss;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { ss.ok = false; return ss; }
  if (ss.dev != dev) {
    ss.ok = cudaStreamCreateWithFlags(&ss.s, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming) == cudaSuccess &&
            cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming) == cudaSuccess;
    ss.dev = dev;
  }
  return ss;
}
struct CtaCapScope {
  int saved;
  explicit CtaCapScope(int cap) : saved(gemm_cta_cap()) { gemm_cta_cap() = cap; }
  ~CtaCapScope() { gemm_cta_cap() = saved; }
};
// SMs the reverse-time loop leaves idle: its attention kernel runs one image per CTA (B CTAs) and its per-step GEMMs are
// sized for <= 128 CTAs, so with B <= #SMs - 16 a persistent GEMM capped to the remaining SMs runs beside the whole loop
// without taking a slot from it.  0 = no room (or ALEXCAP_DWO_DEFER=0).
int dwo_defer_ctas(int B) {
  static const bool off = [] { const char* v = getenv("ALEXCAP_DWO_DEFER"); return v && v[0] == '0'; }();
  int nsm = 0, dev = 0;
  if (off || cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
    return 0;
  const int busy = B > 128 ? B : 128;
  const int spare = (nsm - busy) & ~1;                 // (an even count: CTA pairs)
  static const int want = [] { const char* v = getenv("ALEXCAP_DWO_CAP"); return v ? atoi(v) & ~1 : 0; }();   // (tuning)
  if (spare < 16) return 0;
  return (want >= 2 && want < spare) ? want : spare;
}

// ------------------------------------------------------------------------------------------ attention backward core
// dHout [S*B,H] (time-major) = d loss / d (h_t before dropout is applied => masked inside if dropout is on).
// part bit 1: the reverse-time chain, then lstm.* and embedding gradients (the largest tensors after deep_output);
// part bit 2: every remaining gradient (attention.*, f_beta, init_*, dfeat).  3 = everything.
int attn_bwd_core(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat,
                  const float* dalphas, const float* reg_r, const alexcap_attn_params* G, float* dfeat,
                  AttnWs& w, cudaStream_t st, int part = 3) {
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, S = d->S;
  const int64_t DH = D + H, W5 = 5 * (int64_t)H + D, HD = H + D;
  const bool need_dfeat = d->need_dfeat && dfeat;
  const int SB = S * B;
  if (part & 1) {
  // Wback = [W_hh ; W_U ; W_beta]  ((4H + H + D) x H): dh_{t-1} = [dgates | dq | dbpre] Wback in one GEMM
  {
    CopySegments cs;
    cs.add(P->lstm_w_hh, H, w.Wback, H, 4 * H, H);
    cs.add(P->att_U_w, H, w.Wback + (int64_t)4 * H * H, H, H, H);
    cs.add(P->f_beta_w, H, w.Wback + (int64_t)5 * H * H, H, D, H);
    API_CK(copy_segments(st, cs));
  }
  API_CK(fill_f32(st, w.dc, (int64_t)B * H, 0.f));
  API_CK(fill_f32(st, w.dwv_part, (int64_t)B * H, 0.f));
  if (need_dfeat) API_CK(fill_f32(st, dfeat, (int64_t)B * L * D, 0.f));
  API_CK(cudaMemsetAsync(w.dgctx, 0, sizeof(float) * (size_t)S * B * D, st));
  API_CK(cudaMemsetAsync(w.dh_next, 0, sizeof(float) * (size_t)S * B * H, st));

  const AttnRows rows = attn_rows(d, feat, w, st);      // (arena hits: the copies were made by the forward phase)
  const bool ws_exp = attn_ws_exp(D, H);                // (the same decision the forward phase took for the Ws GEMM)
  uint16_t *gb = nullptr, *gb_lo = nullptr;
  int64_t gld = 0;
  const bool mirror = (rows.bf16 || ws_exp) && bf16_mirror_ok(B, D, H) &&
                      split_arena_reserve_mirror(split_arena(), w.DGQ, S * B, (int)W5, W5, false, &gb, &gb_lo, &gld);
  for (int t = S - 1; t >= 0; --t) {
    float* dgc = w.dgctx + (int64_t)t * B * D;
    float* dhn = w.dh_next + (int64_t)t * B * H;                  // dh_{t-1}, consumed by reverse step t-1
    float* g = w.DGQ + (int64_t)t * B * W5;
    const float* qb = w.QB + (int64_t)t * B * HD;
    API_CK(lstm_pointwise_bwd(st, B, H, g, W5, w.Call + (int64_t)(t + 1) * B * H, w.Call + (int64_t)t * B * H,
                              w.dHout + (int64_t)t * B * H, H, w.MASK ? w.MASK + (int64_t)t * B * H : nullptr,
                              (t == S - 1) ? nullptr : dhn + (int64_t)B * H, H, w.dc,
                              mirror ? gb + (int64_t)t * B * gld : nullptr, gld));
    // d gctx = dgates W_ih[:, E:]
    { WeightBScope wb; API_CK(gemm_f32(st, false, false, B, D, 4 * H, 1.f, g, W5, P->lstm_w_ih + E, E + D, dgc, D, nullptr, nullptr,
                    true, ACT_NONE)); }
    if (rows.bf16 || ws_exp)
      API_CK(attn_step_bwd_stream(st, B, L, D, H, rows.Ws, rows.feat, qb, HD, P->att_v_w, w.ALPHA + (int64_t)t * L, (int64_t)S * L,
                                  w.CTX + (int64_t)t * B * D, w.BETA + (int64_t)t * B * D, dgc, D,
                                  dalphas ? dalphas + (int64_t)t * L : nullptr, (int64_t)S * L, reg_r,
                                  g + 4 * H, W5, g + 5 * H, W5, w.DE + (int64_t)t * L, (int64_t)S * L, w.dwv_part,
                                  need_dfeat ? dfeat : nullptr, rows.bf16, ws_exp, true,
                                  mirror ? gb + (int64_t)t * B * gld + 4 * H : nullptr,
                                  mirror ? gb + (int64_t)t * B * gld + 5 * H : nullptr, gld));
    else
      API_CK(attn_step_bwd(st, B, L, D, H, w.Ws, feat, qb, HD, P->att_v_w, w.ALPHA + (int64_t)t * L, (int64_t)S * L,
                           w.CTX + (int64_t)t * B * D, w.BETA + (int64_t)t * B * D, dgc, D,
                           dalphas ? dalphas + (int64_t)t * L : nullptr, (int64_t)S * L, reg_r,
                           g + 4 * H, W5, g + 5 * H, W5, w.DE + (int64_t)t * L, (int64_t)S * L, w.dwv_part,
                           need_dfeat ? dfeat : nullptr));
    // dh_{t-1} = [dgates | dq | dbpre] [W_hh ; W_U ; W_beta]
    { WeightBScope wb; API_CK(gemm_f32(st, false, false, B, H, (int)W5, 1.f, g, W5, w.Wback, H, dhn, H, nullptr, nullptr, true, ACT_NONE)); }
  }

  // dWs in one pass over Ws (MUFU-bound, 85 us at cfg2), FORKED onto a side stream: it only needs de / q, which are final
  // here, and its only readers are the dW_W / dfeat GEMMs of part 2 -- so it runs concurrently with the tensor-pipe-bound
  // lstm / embedding GEMMs below (different pipes, its CTAs need no shared memory and co-reside with the GEMM CTAs).
  SideStream& side = side_stream();
  const bool overlap = side.ok && dws_overlap_enabled();
  cudaStream_t st_dws = overlap ? side.s : st;
  if (overlap) {
    API_CK(cudaEventRecord(side.fork, st));
    API_CK(cudaStreamWaitEvent(side.s, side.fork, 0));
  }
  {
    // when dW_W (and dfeat) run on the pre-split path the kernel writes the bf16 copies itself and the fp32 matrix is
    // never materialised
    uint16_t *d_hi = nullptr, *d_lo = nullptr;
    int64_t d_ld = 0;
    SplitArena* arena = split_arena();
    const bool direct = arena && gemm_engine() >= 2 && gemm_ps_wanted(H, D, B * L) && (!need_dfeat || gemm_ps_wanted(B * L, D, H)) &&
                        split_arena_reserve(arena, w.dWs, B * L, H, H, gemm_engine() == 3, &d_hi, &d_lo, &d_ld);
    static const int per_sm = [] { const char* v = getenv("ALEXCAP_DWS_CTAS_PER_SM"); return v ? atoi(v) : 4; }();
    API_CK(attn_dws(st_dws, B, L, H, S, w.Ws, w.QB, HD, P->att_v_w, w.DE, direct ? nullptr : w.dWs, d_hi, d_lo, d_ld,
                    overlap ? 148 * per_sm : 0, ws_exp));
  }
  if (overlap) API_CK(cudaEventRecord(side.join, side.s));
  // lstm weights: dW = dgates_all^T [EmbIn | gctx | h_prev], K = S*B
  API_CK(gemm_f32(st, true, false, 4 * H, E, SB, 1.f, w.DGQ, W5, w.EmbIn, E, G->lstm_w_ih, E + D, nullptr, nullptr, false, ACT_NONE));
  API_CK(gemm_f32(st, true, false, 4 * H, D, SB, 1.f, w.DGQ, W5, w.XH, DH, G->lstm_w_ih + E, E + D, nullptr, nullptr, false, ACT_NONE));
  API_CK(gemm_f32(st, true, false, 4 * H, H, SB, 1.f, w.DGQ, W5, w.XH + D, DH, G->lstm_w_hh, H, nullptr, nullptr, false, ACT_NONE));
  // bias gradients = column sums of [dgates | dq | dbpre] (all final after the chain): lstm.bias_*, attention.U.bias, f_beta.bias
  // in ONE pass over DGQ
  API_CK(col_sum3(st, w.DGQ, SB, W5, G->lstm_b_ih, 4 * H, G->att_U_b, H, G->f_beta_b, D, w.cs_scratch));
  API_CK(cudaMemcpyAsync(G->lstm_b_hh, G->lstm_b_ih, sizeof(float) * 4 * H, cudaMemcpyDeviceToDevice, st));
  // embedding: d EmbIn = dgates_all W_ih[:, :E], scatter-add by input token (row 0 / pad included: no padding_idx)
  API_CK(gemm_f32(st, false, false, SB, E, 4 * H, 1.f, w.DGQ, W5, P->lstm_w_ih, E + D, w.dEmbIn, E, nullptr, nullptr, false, ACT_NONE));
  API_CK(fill_f32(st, G->emb_w, (int64_t)d->Vp * E, 0.f));
  API_CK(scatter_add_rows(st, w.dEmbIn, E, w.in_tm, SB, E, G->emb_w));
  if (overlap) API_CK(cudaStreamWaitEvent(st, side.join, 0));     // join: dWs is complete for whatever follows on `st`
  }   // part & 1
  if (!(part & 2)) return 0;

  // attention.W : dWs (made in part 1, concurrently with the lstm / embedding GEMMs), then one K = B*L GEMM
  API_CK(gemm_f32(st, true, false, H, D, B * L, 1.f, w.dWs, H, feat, D, G->att_W_w, D, nullptr, nullptr, false, ACT_NONE));
  if (need_dfeat)
    API_CK(gemm_f32(st, false, false, B * L, D, H, 1.f, w.dWs, H, P->att_W_w, D, dfeat, D, nullptr, nullptr, true, ACT_NONE));
  // attention.U, f_beta
  API_CK(gemm_f32(st, true, false, H, H, SB, 1.f, w.DGQ + 4 * H, W5, w.XH + D, DH, G->att_U_w, H, nullptr, nullptr, false, ACT_NONE));
  API_CK(gemm_f32(st, true, false, D, H, SB, 1.f, w.DGQ + 5 * H, W5, w.XH + D, DH, G->f_beta_w, H, nullptr, nullptr, false, ACT_NONE));
  // d b_W = sum_{b,l} dWs = sum_{b,t,l} dz = sum_{b,t} dq = d b_U: the same number; taken from the dq path, whose sums are
  // formed without the cancellation that a column sum of dWs would re-introduce (see attention_stream.cu, pass D)
  API_CK(cudaMemcpyAsync(G->att_W_b, G->att_U_b, sizeof(float) * H, cudaMemcpyDeviceToDevice, st));
  // attention.v
  API_CK(col_sum(st, w.dwv_part, B, H, H, G->att_v_w, false, w.cs_scratch));
  API_CK(reduce_sum(st, w.DE, B * S * L, G->att_v_b, w.reg_scratch));
  // init_h / init_c (AttentionLanguageModule.py:116-125)
  const int64_t H2 = 2 * (int64_t)H;
  API_CK(tanh_bwd(st, B, H, w.dh_next, H, w.XH + D, DH, w.dh0, H2));
  API_CK(tanh_bwd(st, B, H, w.dc, H, w.Call, H, w.dc0, H2));
  API_CK(gemm_f32(st, true, false, H, D, B, 1.f, w.dh0, H2, w.m, D, G->init_h_w, D, nullptr, nullptr, false, ACT_NONE));
  API_CK(gemm_f32(st, true, false, H, D, B, 1.f, w.dc0, H2, w.m, D, G->init_c_w, D, nullptr, nullptr, false, ACT_NONE));
  API_CK(col_sum3(st, w.dh0, B, H2, G->init_h_b, H, G->init_c_b, H, nullptr, 0, w.cs_scratch));
  if (need_dfeat) {
    API_CK(gemm_f32(st, false, false, B, D, H, 1.f, w.dh0, H2, P->init_h_w, D, w.dm, D, nullptr, nullptr, false, ACT_NONE));
    API_CK(gemm_f32(st, false, false, B, D, H, 1.f, w.dc0, H2, P->init_c_w, D, w.dm, D, nullptr, nullptr, true, ACT_NONE));
    API_CK(add_broadcast_L(st, B, L, D, w.dm, 1.f / (float)L, dfeat));
  }
  return 0;
}

// Host-side bookkeeping of the training step a workspace holds (which bf16 operand copies exist in its arena after which
// phase).  Keyed by the workspace pointer; entries live for the process (a few hundred bytes + the entry table each).
struct StepState { std::mutex mu; SplitArena arena; size_t mark_after_16 = 0, mark_after_1 = 0, mark_after_4 = 0; };
StepState& step_state(void* workspace) {
  static std::mutex mu;
  static std::unordered_map<void*, StepState*> table;
  std::lock_guard<std::mutex> lock(mu);
  StepState*& e = table[workspace];
  if (!e) e = new StepState();
  return *e;
}

int check_attn_args(const alexcap_dims* d, const alexcap_attn_params* P, const void* feat, const void* ws, size_t ws_bytes,
                    size_t need) {
  int rc = check_dims_common(d);
  if (rc) return rc;
  if (!P || !feat || !ws) return fail(ALEXCAP_E_BADARG, "NULL params / feat / workspace");
  if (ws_bytes < need) return fail(ALEXCAP_E_WORKSPACE, "workspace too small (see alexcap_attn_workspace_bytes)");
  if (!aligned16(feat) || !aligned16(ws)) return fail(ALEXCAP_E_BADARG, "feat / workspace must be 16-byte aligned");
  return 0;
}

}  // namespace

// =========================================================================================== exported
extern "C" {

int alexcap_abi_version(void) { return ALEXCAP_ABI_VERSION; }
const char* alexcap_last_error(void) { return err_buf(); }

unsigned long long alexcap_launch_count(void) { return launch_counter(); }

int alexcap_set_gemm_engine(int engine) {
  if (engine < 0 || engine > 3) return fail(ALEXCAP_E_BADARG, "engine must be 0 (fp32 CUDA cores), 1 (tcgen05 3xTF32) or 2 (tcgen05 bf16)");
  gemm_engine_set_default(engine);
  gemm_engine() = engine;
  return 0;
}
int alexcap_get_gemm_engine(void) { return gemm_engine(); }

int alexcap_device_check(void) {
  int dev = 0;
  API_CK(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  API_CK(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) return fail(ALEXCAP_E_UNSUPPORTED, "libalexcap_b200 needs an sm_100 (B200) device; no fallback exists");
  return 0;
}

size_t alexcap_attn_workspace_bytes(const alexcap_dims* d) {
  if (!d) return 0;
  return attn_layout(d, nullptr, true).bytes;
}

// phase bit 1: forward + loss + vocabulary backward (deep_output.{weight,bias} gradients final afterwards);
// phase bit 2: reverse-time chain + every other weight gradient.  3 = the whole step.
// phase 4 / phase 8 are the two halves of bit 2: 4 = chain + lstm.* and embedding gradients, 8 = all remaining gradients
// (so a data-parallel caller can all-reduce the lstm/embedding bucket while the rest is still being computed).
static int attn_train_phases(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, const int32_t* gt,
                             const alexcap_attn_params* G, float* dfeat, float* loss_out, float* alphas_out,
                             const float* drop_mask, const float* ce_denominator, void* workspace,
                             size_t workspace_bytes, int phase, void* stream) {
  if (int rc = check_dims_common(d)) return rc;
  DtypeScope dtype_scope(d);
  AttnWs w = attn_layout(d, nullptr, true);
  if (int rc = check_attn_args(d, P, feat, workspace, workspace_bytes, w.bytes)) return rc;
  if (!gt || !G || !loss_out) return fail(ALEXCAP_E_BADARG, "NULL gt / grads / loss_out");
  if (phase != 1 && phase != 2 && phase != 3 && phase != 4 && phase != 8 && phase != 16 && phase != 32 && phase != 64 && phase != 128)
    return fail(ALEXCAP_E_BADARG, "phase must be 1, 2, 3 (= 1|2), 4 or 8 (the halves of 2), 16 or 32 (the halves of 1), or 128 + 64 "
                                  "(32 with dW_o / db_o split off)");
  w = attn_layout(d, workspace, true);
  // phases 2 / 4 / 8 / 32 continue the step an earlier phase started (same arguments and workspace): feat, W_W, dgates ... stay
  // split -- but only what the EARLIER phases made: a repeated phase 4 (say) must not find the copies of its previous run.
  // The bookkeeping belongs to the WORKSPACE (not to the thread): interleaved steps of two models, or a validation step
  // between the phases of a training step, each keep their own.
  StepState& ss = step_state(workspace);
  std::lock_guard<std::mutex> step_lock(ss.mu);
  SplitArena& step_arena = ss.arena;
  const bool fresh = (phase & 1) || phase == 16;
  if (!fresh && (step_arena.base != w.arena || step_arena.bytes != w.arena_bytes))
    return fail(ALEXCAP_E_BADARG, "phase 2/4/8/32 on a workspace whose step was not started by phase 1/16 with the same dims");
  ArenaScope arena_scope(w.arena, w.arena_bytes, &step_arena, fresh);
  if (!fresh && step_arena.base && phase != 64) {      // (64 runs BESIDE phase 4 and only reads copies phase 128 made)
    size_t mark = (phase == 8) ? ss.mark_after_4 : (phase == 32 || phase == 128) ? ss.mark_after_16 : ss.mark_after_1;
    if (mark > step_arena.used) mark = step_arena.used;
    split_arena_release(&step_arena, mark);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, D = d->D, H = d->H, Vp = d->Vp, S = d->S, L = d->L, SB = S * B;
  const int64_t DH = D + H;
  const bool reg = d->reg_weight != 0.f;
  bool deferred_dwo = false;
  if ((phase & 1) || phase == 16) {
    if (int rc = attn_fwd_core(d, P, feat, gt, drop_mask, w, st)) return rc;
    ss.mark_after_16 = ss.mark_after_1 = ss.mark_after_4 = step_arena.used;
  }
  const float* hout = w.Hd ? w.Hd : w.XH + (int64_t)B * DH + D;
  const int64_t hld = w.Hd ? H : DH;
  if (phase == 64) {
    // dW_o / db_o split off phase 128: on the caller's stream (a data-parallel caller runs it on its side stream, followed by
    // the all-reduce of deep_output.*, while phase 4 runs on the main stream), capped to the SMs the reverse loop leaves idle
    // (the same pure predicate phase 128 evaluates: when it is false, 128 computed the two gradients itself)
    const int spare = (w.ce_stats && attn_engine() == 1) ? dwo_defer_ctas(B) : 0;
    if (spare <= 0) return 0;
    CtaCapScope cap(spare);
    return vocab_ce_dwo(st, SB, Vp, H, hout, hld, w.ldV, w.ce_stats, G->deep_w, G->deep_b, w.cs_scratch2);
  }
  if ((phase & 1) || phase == 32 || phase == 128) {
    // vocabulary projection (AttentionLanguageModule.py:79) + CustomCrossEntropyLoss + vocab backward, in L2-sized row chunks
    API_CK(count_valid(st, w.tgt_tm, SB, d->ignore_index, w.scal));
    if (w.ce_stats) {
      // Whole step in one call: dW_o and db_o, which nothing else in the step reads, are taken off the critical path -- they
      // run on a side stream, capped to the SMs the reverse-time loop leaves idle, underneath that loop (joined at the end).
      SideStream& s2 = side_stream2();
      const int spare = ((phase == 3 || phase == 128) && attn_engine() == 1) ? dwo_defer_ctas(B) : 0;
      deferred_dwo = spare > 0 && (s2.ok || phase == 128);        // (phase 128: phase 64, on the caller's stream, does them)
      if (int rc = vocab_ce_fused(st, SB, Vp, H, hout, hld, P->deep_w, P->deep_b, w.tgt_tm, w.ldV, w.ce_stats, w.row_lse,
                                  d->label_smoothing, d->ignore_index, ce_denominator ? ce_denominator : w.scal, w.row_loss,
                                  G->deep_w, G->deep_b, w.dHout, w.cs_scratch, deferred_dwo)) return rc;
      if (deferred_dwo && phase == 128) deferred_dwo = false;      // nothing to fork or join inside this call
      if (deferred_dwo) {
        API_CK(cudaEventRecord(s2.fork, st));
        API_CK(cudaStreamWaitEvent(s2.s, s2.fork, 0));
        {
          CtaCapScope cap(spare);
          if (int rc = vocab_ce_dwo(s2.s, SB, Vp, H, hout, hld, w.ldV, w.ce_stats, G->deep_w, G->deep_b, w.cs_scratch2)) return rc;
        }
        API_CK(cudaEventRecord(s2.join, s2.s));
      }
    } else if (int rc = vocab_ce_chunked(st, SB, Vp, H, hout, hld, P->deep_w, P->deep_b, w.tgt_tm, w.ldV, w.logits, d->label_smoothing,
                                         d->ignore_index, ce_denominator ? ce_denominator : w.scal, w.row_loss, G->deep_w, G->deep_b,
                                         w.dHout, w.cs_scratch)) return rc;
    API_CK(reduce_sum(st, w.row_loss, SB, w.scal + 1));
    if (reg) API_CK(attn_regulariser(st, B, S, L, d->reg_weight, w.ALPHA, w.R, w.scal + 2, w.reg_scratch));
    API_CK(finalize_loss(st, w.scal + 1, reg ? w.scal + 2 : nullptr, w.scal, loss_out, w.bad_rows, B));
    if (alphas_out)
      API_CK(cudaMemcpyAsync(alphas_out, w.ALPHA, sizeof(float) * (size_t)B * S * L, cudaMemcpyDeviceToDevice, st));
    ss.mark_after_1 = ss.mark_after_4 = step_arena.used;
  }
  int rc = 0;
  if (phase & 2) rc = attn_bwd_core(d, P, feat, nullptr, reg ? w.R : nullptr, G, dfeat, w, st, 3);
  if (deferred_dwo) API_CK(cudaStreamWaitEvent(st, side_stream2().join, 0));     // join: dW_o / db_o are complete on `st`
  else if (phase == 4 || phase == 8) rc = attn_bwd_core(d, P, feat, nullptr, reg ? w.R : nullptr, G, dfeat, w, st, phase == 4 ? 1 : 2);
  if (phase == 4) ss.mark_after_4 = step_arena.used;
  return rc;
}

int alexcap_attn_train_fused(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, const int32_t* gt,
                             const alexcap_attn_params* G, float* dfeat, float* loss_out, float* alphas_out,
                             const float* drop_mask, const float* ce_denominator, void* workspace,
                             size_t workspace_bytes, void* stream) {
  return attn_train_phases(d, P, feat, gt, G, dfeat, loss_out, alphas_out, drop_mask, ce_denominator, workspace,
                           workspace_bytes, 3, stream);
}

int alexcap_attn_train_phase(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, const int32_t* gt,
                             const alexcap_attn_params* G, float* dfeat, float* loss_out, float* alphas_out,
                             const float* drop_mask, const float* ce_denominator, void* workspace,
                             size_t workspace_bytes, int32_t phase, void* stream) {
  return attn_train_phases(d, P, feat, gt, G, dfeat, loss_out, alphas_out, drop_mask, ce_denominator, workspace,
                           workspace_bytes, phase, stream);
}

int alexcap_attn_forward(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, const int32_t* gt,
                         float* preds, float* alphas, const float* drop_mask, void* workspace, size_t workspace_bytes,
                         void* stream) {
  if (int rc = check_dims_common(d)) return rc;
  DtypeScope dtype_scope(d);
  AttnWs w = attn_layout(d, nullptr, false);
  if (int rc = check_attn_args(d, P, feat, workspace, workspace_bytes, w.bytes)) return rc;
  if (!gt || !preds) return fail(ALEXCAP_E_BADARG, "NULL gt / preds");
  w = attn_layout(d, workspace, false);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, D = d->D, H = d->H, Vp = d->Vp, S = d->S, L = d->L;
  const int64_t DH = D + H;
  if (int rc = attn_fwd_core(d, P, feat, gt, drop_mask, w, st)) return rc;
  for (int t = 0; t < S; ++t) {   // preds[:, t] = deep_output(dropout(h_t))  (AttentionLanguageModule.py:79-81)
    const float* h = w.Hd ? w.Hd + (int64_t)t * B * H : w.XH + (int64_t)(t + 1) * B * DH + D;
    API_CK(gemm_f32(st, false, true, B, Vp, H, 1.f, h, w.Hd ? H : DH, P->deep_w, H, preds + (int64_t)t * Vp, (int64_t)S * Vp,
                    P->deep_b, nullptr, false, ACT_NONE));
  }
  if (alphas)
    API_CK(cudaMemcpyAsync(alphas, w.ALPHA, sizeof(float) * (size_t)B * S * L, cudaMemcpyDeviceToDevice, st));
  return 0;
}

int alexcap_attn_backward(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, const float* dpreds,
                          const float* dalphas, const alexcap_attn_params* G, float* dfeat, void* workspace,
                          size_t workspace_bytes, void* stream) {
  if (int rc = check_dims_common(d)) return rc;
  DtypeScope dtype_scope(d);
  AttnWs w = attn_layout(d, nullptr, false);
  if (int rc = check_attn_args(d, P, feat, workspace, workspace_bytes, w.bytes)) return rc;
  if (!G) return fail(ALEXCAP_E_BADARG, "NULL grads");
  w = attn_layout(d, workspace, false);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, D = d->D, H = d->H, Vp = d->Vp, S = d->S;
  const int64_t DH = D + H;
  if (dpreds) {
    for (int t = 0; t < S; ++t) {
      const float* h = w.Hd ? w.Hd + (int64_t)t * B * H : w.XH + (int64_t)(t + 1) * B * DH + D;
      const float* dp = dpreds + (int64_t)t * Vp;
      API_CK(gemm_f32(st, false, false, B, H, Vp, 1.f, dp, (int64_t)S * Vp, P->deep_w, H, w.dHout + (int64_t)t * B * H, H,
                      nullptr, nullptr, false, ACT_NONE));
      API_CK(gemm_f32(st, true, false, Vp, H, B, 1.f, dp, (int64_t)S * Vp, h, w.Hd ? H : DH, G->deep_w, H, nullptr, nullptr,
                      t > 0, ACT_NONE));
    }
    API_CK(col_sum(st, dpreds, B * S, Vp, Vp, G->deep_b, false, w.cs_scratch));
  } else {
    API_CK(fill_f32(st, w.dHout, (int64_t)S * B * H, 0.f));
    API_CK(fill_f32(st, G->deep_w, (int64_t)Vp * H, 0.f));
    API_CK(fill_f32(st, G->deep_b, Vp, 0.f));
  }
  return attn_bwd_core(d, P, feat, dalphas, nullptr, G, dfeat, w, st);
}

int alexcap_make_target(const int32_t* gt, int32_t* in_tok, int32_t* tgt, int32_t B, int32_t T, int32_t start_token,
                        int32_t end_token, void* stream) {
  if (!gt || B <= 0 || T <= 0) return fail(ALEXCAP_E_BADARG, "alexcap_make_target: bad arguments");
  API_CK(make_target(static_cast<cudaStream_t>(stream), gt, B, T, start_token, end_token, in_tok, tgt, nullptr, nullptr));
  return 0;
}

int alexcap_count_valid_targets(const int32_t* gt, int32_t B, int32_t T, float* n_valid_out, void* stream) {
  if (!gt || !n_valid_out || B <= 0 || T <= 0) return fail(ALEXCAP_E_BADARG, "alexcap_count_valid_targets: bad arguments");
  API_CK(count_valid_targets(static_cast<cudaStream_t>(stream), gt, B, T, n_valid_out));
  return 0;
}

int alexcap_mean_over_L(const float* feat, int32_t B, int32_t L, int32_t D, float* m, void* stream) {
  if (!feat || !m || B <= 0 || L <= 0 || D <= 0 || (D & 3)) return fail(ALEXCAP_E_BADARG, "alexcap_mean_over_L: bad arguments (D % 4 == 0)");
  API_CK(mean_over_L(static_cast<cudaStream_t>(stream), feat, B, L, D, m));
  return 0;
}

int alexcap_attn_workspace_region(const alexcap_dims* d, const char* name, size_t* offset, size_t* bytes) {
  if (int rc = check_dims_common(d)) return rc;
  if (!name || !offset || !bytes) return fail(ALEXCAP_E_BADARG, "alexcap_attn_workspace_region: NULL argument");
  char* base = reinterpret_cast<char*>(uintptr_t(1) << 20);       // a fake, aligned base: only differences are used
  const AttnWs w = attn_layout(d, base, true);
  const size_t B = d->B, L = d->L, D = d->D, H = d->H, S = d->S;
  const void* p = nullptr; size_t n = 0;
  if (!strcmp(name, "MASK")) { p = w.MASK; n = S * B * H * 4; }
  else if (!strcmp(name, "ALPHA")) { p = w.ALPHA; n = B * S * L * 4; }
  else if (!strcmp(name, "H")) { p = w.XH; n = (S + 1) * B * (D + H) * 4; }
  else if (!strcmp(name, "LOGITS")) { p = w.logits; n = (size_t)vocab_chunk_rows((int)(S * B), w.ldV) * w.ldV * 4; }   // absent when the CE is fused
  if (!p) return fail(ALEXCAP_E_BADARG, "alexcap_attn_workspace_region: unknown region name, or a region these dims do not have");
  *offset = (size_t)(static_cast<const char*>(p) - base); *bytes = n;
  return 0;
}

int alexcap_set_time_loop(int32_t on) {
  const int prev = time_loop_enabled() ? 1 : 0;
  time_loop_enabled() = on != 0;
  return prev;
}

int alexcap_time_loop_profile(unsigned long long* host_out, int32_t max_entries) {
  if (!host_out || max_entries <= 0) return 0;
  return time_loop_profile_read(host_out, max_entries);
}

int alexcap_cross_entropy(const float* logits, const int32_t* tgt, int32_t N, int32_t C, int64_t ld, float label_smoothing,
                          int32_t ignore_index, float* loss_out, float* dlogits, float* scratch, void* stream) {
  if (!logits || !tgt || !loss_out || !scratch || N <= 0 || C <= 0 || ld < C)
    return fail(ALEXCAP_E_BADARG, "alexcap_cross_entropy: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // scratch: [0..N) row losses, [N] n_valid
  API_CK(count_valid(st, tgt, N, ignore_index, scratch + N));
  API_CK(ce_rows(st, logits, tgt, N, C, ld, label_smoothing, ignore_index, scratch + N, scratch, dlogits));
  API_CK(reduce_sum(st, scratch, N, loss_out));
  API_CK(cudaMemcpyAsync(loss_out + 1, scratch + N, sizeof(float), cudaMemcpyDeviceToDevice, st));
  return 0;
}

int alexcap_gemm_f32(int32_t transa, int32_t transb, int32_t M, int32_t N, int32_t K, float alpha, const float* A,
                     int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, const float* bias,
                     const float* bias2, int32_t accumulate, int32_t act, int32_t engine, void* stream) {
  if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0) return fail(ALEXCAP_E_BADARG, "alexcap_gemm_f32: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (engine == 0) {
    API_CK(gemm_simt_f32(st, transa != 0, transb != 0, M, N, K, alpha, A, lda, B, ldb, C, ldc, bias, bias2, accumulate != 0, act));
  } else {
    if (!gemm_tc_supported(M, N, K, A, lda, B, ldb, C))
      return fail(ALEXCAP_E_UNSUPPORTED, "alexcap_gemm_f32: tensor-core engine needs 16-byte aligned operands with ld % 4 == 0");
    // engine 1 = 3xTF32, heuristic split-K; 2..99 = 3xTF32 with split-K forced to `engine` (tests);
    // 100 = bf16 operands / fp32 accumulate, heuristic split-K; 100 + s = bf16 with split-K forced to s
    // 200 (+s) = bf16x3 (hi/lo bf16 pairs, three kind::f16 MMAs)
    // 300 (+s) = bf16x3 over operands pre-split in HBM (gemm_ps.cu), 400 (+s) = bf16 pre-split; the copies live in a
    // process-lifetime arena grown on demand (this entry point is the test / micro-benchmark hook: not capturable when it grows)
    // 500 (+s) / 600 (+s) = the same two, but the operand copies of the previous call are KEPT when the pointers match
    // (times the GEMM kernel alone, as inside a training step where copies are reused; the caller must not have changed the data)
    const bool keep = engine >= 500;
    if (keep) engine -= 200;
    if (engine >= 300) {
      static SplitArena arena;
      const size_t need = ((size_t)M * (K + 8) + (size_t)N * (K + 8) + (size_t)K * (M + N + 16)) * 4 + (1 << 16);
      if (arena.bytes < need) {
        API_CK(cudaDeviceSynchronize());
        if (arena.base) API_CK(cudaFree(arena.base));
        arena.base = nullptr; arena.bytes = 0;
        API_CK(cudaMalloc(reinterpret_cast<void**>(&arena.base), need));
        arena.bytes = need; arena.used = 0; arena.n = 0;
      }
      if (!keep) { arena.used = 0; arena.n = 0; }
      const int np = engine >= 400 ? 1 : 3;
      API_CK(gemm_ps_f32(st, transa != 0, transb != 0, M, N, K, alpha, A, lda, B, ldb, C, ldc, bias, bias2, accumulate != 0, act,
                         engine >= 400 ? engine - 400 : engine - 300, np, &arena));
      return 0;
    }
    const int mode = engine >= 200 ? 2 : engine >= 100 ? 1 : 0;
    const int force = engine >= 200 ? engine - 200 : engine >= 100 ? engine - 100 : (engine >= 2 ? engine : 0);
    API_CK(gemm_tc_f32(st, transa != 0, transb != 0, M, N, K, alpha, A, lda, B, ldb, C, ldc, bias, bias2, accumulate != 0, act,
                       force, mode));
  }
  return 0;
}

int alexcap_attn_step_fwd(int32_t B, int32_t L, int32_t D, int32_t H, const float* Ws, const float* feat, const float* q,
                          int64_t q_ld, const float* bpre, int64_t bpre_ld, const float* w_v, float* alpha,
                          int64_t alpha_ld, float* ctx, float* beta, float* gctx, int64_t gctx_ld, int32_t rows_bf16, void* stream) {
  if (rows_bf16) {       // bit 0: bf16 rows; bit 1: Ws holds 2^(clamp(2 log2(e) Ws)) (the training step's exponential form)
    API_CK(attn_step_fwd_stream(static_cast<cudaStream_t>(stream), B, L, D, H, Ws, feat, q, q_ld, bpre, bpre_ld, w_v, alpha, alpha_ld,
                                ctx, beta, gctx, gctx_ld, 1, (rows_bf16 & 1) != 0, (rows_bf16 & 2) != 0));
    return 0;
  }
  API_CK(attn_step_fwd(static_cast<cudaStream_t>(stream), B, L, D, H, Ws, feat, q, q_ld, bpre, bpre_ld, w_v, alpha, alpha_ld,
                       ctx, beta, gctx, gctx_ld));
  return 0;
}

int alexcap_attn_step_bwd(int32_t B, int32_t L, int32_t D, int32_t H, const float* Ws, const float* feat, const float* q,
                          int64_t q_ld, const float* w_v, const float* alpha, int64_t alpha_ld, const float* ctx,
                          const float* beta, const float* dgctx, int64_t dgctx_ld, const float* dalpha_ext, int64_t ext_ld,
                          const float* reg_r, float* dq, int64_t dq_ld, float* dbpre, int64_t dbpre_ld, float* de,
                          int64_t de_ld, float* dwv_part, float* dfeat, int32_t rows_bf16, void* stream) {
  if (rows_bf16) {
    API_CK(attn_step_bwd_stream(static_cast<cudaStream_t>(stream), B, L, D, H, Ws, feat, q, q_ld, w_v, alpha, alpha_ld, ctx, beta,
                                dgctx, dgctx_ld, dalpha_ext, ext_ld, reg_r, dq, dq_ld, dbpre, dbpre_ld, de, de_ld, dwv_part, dfeat,
                                (rows_bf16 & 1) != 0, (rows_bf16 & 2) != 0));
    return 0;
  }
  API_CK(attn_step_bwd(static_cast<cudaStream_t>(stream), B, L, D, H, Ws, feat, q, q_ld, w_v, alpha, alpha_ld, ctx, beta,
                       dgctx, dgctx_ld, dalpha_ext, ext_ld, reg_r, dq, dq_ld, dbpre, dbpre_ld, de, de_ld, dwv_part, dfeat));
  return 0;
}

}  // extern "C"
