// api_lstm.cu -- plain LSTM decoder (AlexCap/LanguageModule.py) behind the C ABI: teacher-forced train step
// (LanguageModule.py:99-107 driven by AlexCap/LSTMModel.py:70-73) and greedy decoding (:130-150).
//
// The reference calls nn.LSTM twice -- first over the L encoded regions to obtain a state, then over the S word
// embeddings from that state.  That is ONE LSTM over the concatenated (L+S)-step sequence from a zero state, which is
// how it is run here: the input projection X W_ih^T is hoisted into a single GEMM over all L+S steps and each
// recurrent step is one [B,H]x[H,4H] GEMM + one pointwise kernel.
#include "api_util.h"

namespace alexcap {
// beam.cu
cudaError_t beam_select(cudaStream_t st, int B, int K, int Vp, int S, int step, int end_tok, const float* logits, int64_t ldl,
                        float* top, int32_t* kact, int32_t* words, int32_t* back, int32_t* rowrank, int32_t* src_row,
                        int32_t* prev_word, int32_t* ncomp, float* comp_score, int32_t* comp_step, int32_t* comp_rank);
cudaError_t beam_gather_state(cudaStream_t st, int B, int K, int H, const int32_t* src_row, const float* h_in, int64_t h_in_ld,
                              const float* c_in, float* h_out, int64_t h_out_ld, float* c_out);
cudaError_t beam_finalize(cudaStream_t st, int B, int K, int S, int L, const int32_t* words, const int32_t* back,
                          const int32_t* rowrank, const int32_t* ncomp, const float* comp_score, const int32_t* comp_step,
                          const int32_t* comp_rank, const float* alpha_steps, int64_t* ids, int32_t* lens, float* scores,
                          float* alphas);
}  // namespace alexcap

using namespace alexcap;

namespace {

struct LstmWs {
  int32_t *in_bm, *tgt_bm, *in_tm, *tgt_tm, *bad_rows;
  float *enc_bm, *In0;
  float *G[ALEXCAP_MAX_LAYERS], *Hs[ALEXCAP_MAX_LAYERS], *Cs[ALEXCAP_MAX_LAYERS];
  float *MASK, *Hd, *logits, *row_loss, *scal, *dHout, *dA, *dB, *dh_next, *dc, *dpre_bm, *cs_scratch;
  uint8_t* arena; size_t arena_bytes;   // bf16 hi/lo operand copies of the big GEMMs (training layout only)
  float4* ce_stats; float* row_lse;          // fused vocabulary CE: per-row partial statistics, log-sum-exp
  int64_t ldV;
  size_t bytes;
};

LstmWs lstm_layout(const alexcap_dims* d, void* base, bool with_logits) {
  LstmWs w;
  Bump a(base);
  const int64_t B = d->B, L = d->L, H = d->H, E = d->E, Vp = d->Vp, S = d->S, LT = L + S;
  const int nl = d->num_layers;
  w.ldV = (Vp + 3) & ~int64_t(3);
  w.in_bm = a.take<int32_t>(B * S); w.tgt_bm = a.take<int32_t>(B * S);
  w.in_tm = a.take<int32_t>(B * S); w.tgt_tm = a.take<int32_t>(B * S);
  w.bad_rows = a.take<int32_t>(B);
  w.enc_bm = a.take<float>(B * L * E);
  w.In0 = a.take<float>(LT * B * E);
  for (int l = 0; l < nl; ++l) {
    w.G[l] = a.take<float>(LT * B * 4 * H);
    w.Hs[l] = a.take<float>((LT + 1) * B * H);
    w.Cs[l] = a.take<float>((LT + 1) * B * H);
  }
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
  const int64_t wmax = E > H ? E : H;
  w.dA = a.take<float>(LT * B * wmax);
  w.dB = a.take<float>(LT * B * wmax);
  w.dh_next = a.take<float>(B * H);
  w.dc = a.take<float>(B * H);
  w.dpre_bm = a.take<float>(B * L * E);
  int64_t maxn = Vp > 4 * H ? Vp : 4 * H;
  if (E > maxn) maxn = E;
  w.cs_scratch = a.take<float>(64 * maxn);
  w.arena = nullptr; w.arena_bytes = 0;
  if (with_logits) {
    const int64_t I = E > H ? E : H, D = d->D;
    const int64_t el = B * L * (D + E) + E * D + LT * B * E + nl * (2 * LT * B * 4 * H + 2 * (LT + 1) * B * H + 4 * H * (I + H) + 4 * H * H) +
                       S * B * H + Vp * H + (int64_t)vocab_chunk_rows((int)(S * B), w.ldV) * (Vp + 8);
    w.arena_bytes = (size_t)el * 4 + (size_t)SplitArena::MAX_ENTRIES * 2048 + (1 << 20);
    a.off = (a.off + 1023) & ~size_t(1023);
    w.arena = reinterpret_cast<uint8_t*>(a.take<char>(w.arena_bytes));
  }
  w.bytes = a.off + 256;
  return w;
}

int check_lstm_args(const alexcap_dims* d, const alexcap_lstm_params* P, const void* feat, const void* ws, size_t ws_bytes,
                    size_t need) {
  if (!P || !feat || !ws) return fail(ALEXCAP_E_BADARG, "NULL params / feat / workspace");
  if (ws_bytes < need) return fail(ALEXCAP_E_WORKSPACE, "workspace too small (see alexcap_lstm_workspace_bytes)");
  if (!aligned16(feat) || !aligned16(ws)) return fail(ALEXCAP_E_BADARG, "feat / workspace must be 16-byte aligned");
  return 0;
}

int check_lstm_dims(const alexcap_dims* d) {
  if (int rc = check_dims_common(d)) return rc;
  if (d->num_layers < 1 || d->num_layers > ALEXCAP_MAX_LAYERS) return fail(ALEXCAP_E_BADARG, "num_layers out of range");
  return 0;
}

// one LSTM layer over LT steps (hoisted input GEMM + recurrence)
int run_layer_fwd(const alexcap_dims* d, const float* in, int I, const float* w_ih, const float* w_hh, const float* b_ih,
                  const float* b_hh, float* G, float* Hs, float* Cs, int LT, cudaStream_t st) {
  const int B = d->B, H = d->H;
  API_CK(gemm_f32(st, false, true, LT * B, 4 * H, I, 1.f, in, I, w_ih, I, G, 4 * H, b_ih, b_hh, false, ACT_NONE));
  API_CK(fill_f32(st, Hs, (int64_t)B * H, 0.f));
  API_CK(fill_f32(st, Cs, (int64_t)B * H, 0.f));
  for (int t = 0; t < LT; ++t) {
    float* g = G + (int64_t)t * B * 4 * H;
    if (t > 0)   // h_{-1} = 0
      API_CK(gemm_f32(st, false, true, B, 4 * H, H, 1.f, Hs + (int64_t)t * B * H, H, w_hh, H, g, 4 * H, nullptr, nullptr, true,
                      ACT_NONE));
    API_CK(lstm_pointwise_fwd(st, B, H, g, 4 * H, Cs + (int64_t)t * B * H, Cs + (int64_t)(t + 1) * B * H,
                              Hs + (int64_t)(t + 1) * B * H, H, nullptr, 0));
  }
  return 0;
}

int lstm_fwd_core(const alexcap_dims* d, const alexcap_lstm_params* P, const float* feat, const int32_t* gt,
                  const float* drop_mask, LstmWs& w, cudaStream_t st) {
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, S = d->S, LT = L + S, V = d->Vp - 3;
  API_CK(make_target(st, gt, B, S - 1, V + 1, V + 2, w.in_bm, w.tgt_bm, w.in_tm, w.tgt_tm, d->Vp, w.bad_rows));
  // image_encoder: ReLU(Linear) on every region (LanguageModule.py:35-37,103), then to time-major
  API_CK(gemm_f32(st, false, true, B * L, E, D, 1.f, feat, D, P->enc_w, D, w.enc_bm, E, P->enc_b, nullptr, false, ACT_RELU));
  API_CK(permute_01(st, B, L, E, w.enc_bm, w.In0));
  API_CK(gather_rows(st, P->emb_w, w.in_tm, S * B, E, w.In0 + (int64_t)L * B * E, E));   // :102
  for (int l = 0; l < d->num_layers; ++l) {
    const float* in = (l == 0) ? w.In0 : w.Hs[l - 1] + (int64_t)B * H;
    if (int rc = run_layer_fwd(d, in, l == 0 ? E : H, P->w_ih[l], P->w_hh[l], P->b_ih[l], P->b_hh[l], w.G[l], w.Hs[l], w.Cs[l],
                               LT, st))
      return rc;
  }
  if (d->dropout_p > 0.f) {   // rnn.drop (:48,:106)
    const float* htop = w.Hs[d->num_layers - 1] + (int64_t)(1 + L) * B * H;
    if (drop_mask) {
      API_CK(cudaMemcpyAsync(w.MASK, drop_mask, sizeof(float) * S * B * H, cudaMemcpyDeviceToDevice, st));
      API_CK(mul_2d(st, S * B, H, htop, H, w.MASK, H, w.Hd, H));
    } else {
      API_CK(dropout_mask(st, w.MASK, (int64_t)S * B * H, d->dropout_p, d->seed, d->offset, d->philox_state, htop, H, H, w.Hd));
    }
  }
  return 0;
}

const float* top_out(const alexcap_dims* d, const LstmWs& w) {
  return w.Hd ? w.Hd : w.Hs[d->num_layers - 1] + (int64_t)(1 + d->L) * d->B * d->H;
}

int lstm_bwd_core(const alexcap_dims* d, const alexcap_lstm_params* P, const float* feat, const alexcap_lstm_params* G,
                  float* dfeat, LstmWs& w, cudaStream_t st) {
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, S = d->S, LT = L + S, nl = d->num_layers;
  float* dOut = nullptr;       // d loss / d (outputs of layer l), all LT steps, for l < top
  float* bufs[2] = {w.dA, w.dB};
  for (int l = nl - 1; l >= 0; --l) {
    const bool top = (l == nl - 1);
    const int I = (l == 0) ? E : H;
    const float* in = (l == 0) ? w.In0 : w.Hs[l - 1] + (int64_t)B * H;
    API_CK(fill_f32(st, w.dc, (int64_t)B * H, 0.f));
    for (int t = LT - 1; t >= 0; --t) {
      float* g = w.G[l] + (int64_t)t * B * 4 * H;
      const float* dh_a = nullptr; const float* mask = nullptr;
      if (top) {
        if (t >= L) {
          dh_a = w.dHout + (int64_t)(t - L) * B * H;
          mask = w.MASK ? w.MASK + (int64_t)(t - L) * B * H : nullptr;
        }
      } else {
        dh_a = dOut + (int64_t)t * B * H;
      }
      API_CK(lstm_pointwise_bwd(st, B, H, g, 4 * H, w.Cs[l] + (int64_t)(t + 1) * B * H, w.Cs[l] + (int64_t)t * B * H, dh_a, H,
                                mask, (t == LT - 1) ? nullptr : w.dh_next, H, w.dc));
      if (t > 0)
        API_CK(gemm_f32(st, false, false, B, H, 4 * H, 1.f, g, 4 * H, P->w_hh[l], H, w.dh_next, H, nullptr, nullptr, false,
                        ACT_NONE));
    }
    // weight gradients as single K = LT*B GEMMs
    API_CK(gemm_f32(st, true, false, 4 * H, I, LT * B, 1.f, w.G[l], 4 * H, in, I, G->w_ih[l], I, nullptr, nullptr, false, ACT_NONE));
    API_CK(gemm_f32(st, true, false, 4 * H, H, LT * B, 1.f, w.G[l], 4 * H, w.Hs[l], H, G->w_hh[l], H, nullptr, nullptr, false,
                    ACT_NONE));
    API_CK(col_sum(st, w.G[l], LT * B, 4 * H, 4 * H, G->b_ih[l], false, w.cs_scratch));
    API_CK(cudaMemcpyAsync(G->b_hh[l], G->b_ih[l], sizeof(float) * 4 * H, cudaMemcpyDeviceToDevice, st));
    float* dIn = bufs[l & 1];
    API_CK(gemm_f32(st, false, false, LT * B, I, 4 * H, 1.f, w.G[l], 4 * H, P->w_ih[l], I, dIn, I, nullptr, nullptr, false,
                    ACT_NONE));
    dOut = dIn;
  }
  // dOut = d In0 [LT,B,E]: words -> embedding rows, regions -> image encoder
  API_CK(fill_f32(st, G->emb_w, (int64_t)d->Vp * E, 0.f));
  API_CK(scatter_add_rows(st, dOut + (int64_t)L * B * E, E, w.in_tm, S * B, E, G->emb_w));
  API_CK(relu_bwd(st, (int64_t)L * B * E, dOut, w.In0, dOut));
  API_CK(permute_01(st, L, B, E, dOut, w.dpre_bm));
  API_CK(gemm_f32(st, true, false, E, D, B * L, 1.f, w.dpre_bm, E, feat, D, G->enc_w, D, nullptr, nullptr, false, ACT_NONE));
  API_CK(col_sum(st, w.dpre_bm, B * L, E, E, G->enc_b, false, w.cs_scratch));
  if (d->need_dfeat && dfeat)
    API_CK(gemm_f32(st, false, false, B * L, D, E, 1.f, w.dpre_bm, E, P->enc_w, D, dfeat, D, nullptr, nullptr, false, ACT_NONE));
  return 0;
}

}  // namespace

extern "C" {

size_t alexcap_lstm_workspace_bytes(const alexcap_dims* d) {
  if (!d || d->num_layers < 1 || d->num_layers > ALEXCAP_MAX_LAYERS) return 0;
  return lstm_layout(d, nullptr, true).bytes;
}

int alexcap_lstm_train_fused(const alexcap_dims* d, const alexcap_lstm_params* P, const float* feat, const int32_t* gt,
                             const alexcap_lstm_params* G, float* dfeat, float* loss_out, const float* drop_mask,
                             const float* ce_denominator, void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_lstm_dims(d)) return rc;
  DtypeScope dtype_scope(d);
  LstmWs w = lstm_layout(d, nullptr, true);
  if (int rc = check_lstm_args(d, P, feat, workspace, workspace_bytes, w.bytes)) return rc;
  if (!gt || !G || !loss_out) return fail(ALEXCAP_E_BADARG, "NULL gt / grads / loss_out");
  w = lstm_layout(d, workspace, true);
  ArenaScope arena_scope(w.arena, w.arena_bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, H = d->H, Vp = d->Vp, S = d->S, SB = S * B;
  if (int rc = lstm_fwd_core(d, P, feat, gt, drop_mask, w, st)) return rc;
  const float* hout = top_out(d, w);
  API_CK(count_valid(st, w.tgt_tm, SB, d->ignore_index, w.scal));
  if (w.ce_stats) {
    if (int rc = vocab_ce_fused(st, SB, Vp, H, hout, H, P->out_w, P->out_b, w.tgt_tm, w.ldV, w.ce_stats, w.row_lse,
                                d->label_smoothing, d->ignore_index, ce_denominator ? ce_denominator : w.scal, w.row_loss,
                                G->out_w, G->out_b, w.dHout, w.cs_scratch)) return rc;
  } else if (int rc = vocab_ce_chunked(st, SB, Vp, H, hout, H, P->out_w, P->out_b, w.tgt_tm, w.ldV, w.logits, d->label_smoothing,
                                       d->ignore_index, ce_denominator ? ce_denominator : w.scal, w.row_loss, G->out_w, G->out_b,
                                       w.dHout, w.cs_scratch)) return rc;
  API_CK(reduce_sum(st, w.row_loss, SB, w.scal + 1));
  API_CK(finalize_loss(st, w.scal + 1, nullptr, w.scal, loss_out, w.bad_rows, B));
  return lstm_bwd_core(d, P, feat, G, dfeat, w, st);
}

int alexcap_lstm_forward(const alexcap_dims* d, const alexcap_lstm_params* P, const float* feat, const int32_t* gt,
                         float* preds, const float* drop_mask, void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_lstm_dims(d)) return rc;
  DtypeScope dtype_scope(d);
  LstmWs w = lstm_layout(d, nullptr, false);
  if (int rc = check_lstm_args(d, P, feat, workspace, workspace_bytes, w.bytes)) return rc;
  if (!gt || !preds) return fail(ALEXCAP_E_BADARG, "NULL gt / preds");
  w = lstm_layout(d, workspace, false);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, H = d->H, Vp = d->Vp, S = d->S;
  if (int rc = lstm_fwd_core(d, P, feat, gt, drop_mask, w, st)) return rc;
  const float* hout = top_out(d, w);
  for (int t = 0; t < S; ++t)   // preds[B,S,Vp] batch-major like the reference (LanguageModule.py:106-107)
    API_CK(gemm_f32(st, false, true, B, Vp, H, 1.f, hout + (int64_t)t * B * H, H, P->out_w, H, preds + (int64_t)t * Vp,
                    (int64_t)S * Vp, P->out_b, nullptr, false, ACT_NONE));
  return 0;
}

int alexcap_lstm_backward(const alexcap_dims* d, const alexcap_lstm_params* P, const float* feat, const float* dpreds,
                          const alexcap_lstm_params* G, float* dfeat, void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_lstm_dims(d)) return rc;
  DtypeScope dtype_scope(d);
  LstmWs w = lstm_layout(d, nullptr, false);
  if (int rc = check_lstm_args(d, P, feat, workspace, workspace_bytes, w.bytes)) return rc;
  if (!G || !dpreds) return fail(ALEXCAP_E_BADARG, "NULL grads / dpreds");
  w = lstm_layout(d, workspace, false);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, H = d->H, Vp = d->Vp, S = d->S;
  const float* hout = top_out(d, w);
  for (int t = 0; t < S; ++t) {
    const float* dp = dpreds + (int64_t)t * Vp;
    API_CK(gemm_f32(st, false, false, B, H, Vp, 1.f, dp, (int64_t)S * Vp, P->out_w, H, w.dHout + (int64_t)t * B * H, H, nullptr,
                    nullptr, false, ACT_NONE));
    API_CK(gemm_f32(st, true, false, Vp, H, B, 1.f, dp, (int64_t)S * Vp, hout + (int64_t)t * B * H, H, G->out_w, H, nullptr, nullptr,
                    t > 0, ACT_NONE));
  }
  API_CK(col_sum(st, dpreds, B * S, Vp, Vp, G->out_b, false, w.cs_scratch));
  return lstm_bwd_core(d, P, feat, G, dfeat, w, st);
}

// ------------------------------------------------------------------------------------------ greedy decoding
struct LstmDecWs {
  float *enc_bm, *seqA, *seqB, *Gimg, *Himg, *Cimg, *hst, *cst, *x, *g, *logits;
  int32_t *tok, *done;
  size_t bytes;
};
static LstmDecWs lstm_dec_layout(const alexcap_dims* d, void* base) {
  LstmDecWs w;
  Bump a(base);
  const int64_t B = d->B, L = d->L, H = d->H, E = d->E, nl = d->num_layers;
  const int64_t ldV = (d->Vp + 3) & ~int64_t(3);
  w.enc_bm = a.take<float>(B * L * E);
  w.seqA = a.take<float>(L * B * E);                 // layer-0 input sequence (time-major)
  w.seqB = a.take<float>(L * B * H);                 // inputs of layers >= 1
  w.Gimg = a.take<float>(L * B * 4 * H);
  w.Himg = a.take<float>((L + 1) * B * H);
  w.Cimg = a.take<float>((L + 1) * B * H);
  w.hst = a.take<float>(nl * 2 * B * H);             // [layer][pingpong][B,H]
  w.cst = a.take<float>(nl * 2 * B * H);
  w.x = a.take<float>(B * E);
  w.g = a.take<float>(B * 4 * H);
  w.logits = a.take<float>(B * ldV);
  w.tok = a.take<int32_t>(B);
  w.done = a.take<int32_t>(16);
  w.bytes = a.off + 256;
  return w;
}

size_t alexcap_lstm_decode_workspace_bytes(const alexcap_dims* d) {
  if (!d || d->num_layers < 1 || d->num_layers > ALEXCAP_MAX_LAYERS) return 0;
  return lstm_dec_layout(d, nullptr).bytes;
}

int alexcap_lstm_decode_greedy(const alexcap_dims* d, const alexcap_lstm_params* P, const float* feat, int64_t* ids,
                               void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_lstm_dims(d)) return rc;
  ExactScope exact_scope;
  if (int rc = check_lstm_args(d, P, feat, workspace, workspace_bytes, lstm_dec_layout(d, nullptr).bytes)) return rc;
  if (!ids) return fail(ALEXCAP_E_BADARG, "NULL ids");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, Vp = d->Vp, S = d->S, nl = d->num_layers, V = Vp - 3;
  const int64_t ldV = (Vp + 3) & ~int64_t(3);
  LstmDecWs w = lstm_dec_layout(d, workspace);
  float *enc_bm = w.enc_bm, *seqA = w.seqA, *seqB = w.seqB, *Gimg = w.Gimg, *Himg = w.Himg, *Cimg = w.Cimg;
  float *hst = w.hst, *cst = w.cst, *x = w.x, *g = w.g, *logits = w.logits;
  int32_t *tok = w.tok, *done = w.done;
  // image pass (LanguageModule.py:133,135): state of every layer after the L region steps
  API_CK(gemm_f32(st, false, true, B * L, E, D, 1.f, feat, D, P->enc_w, D, enc_bm, E, P->enc_b, nullptr, false, ACT_RELU));
  API_CK(permute_01(st, B, L, E, enc_bm, seqA));
  for (int l = 0; l < nl; ++l) {
    const float* in = (l == 0) ? seqA : seqB;
    if (int rc = run_layer_fwd(d, in, l == 0 ? E : H, P->w_ih[l], P->w_hh[l], P->b_ih[l], P->b_hh[l], Gimg, Himg, Cimg, L, st))
      return rc;
    API_CK(cudaMemcpyAsync(hst + (int64_t)(l * 2) * B * H, Himg + (int64_t)L * B * H, sizeof(float) * B * H, cudaMemcpyDeviceToDevice, st));
    API_CK(cudaMemcpyAsync(cst + (int64_t)(l * 2) * B * H, Cimg + (int64_t)L * B * H, sizeof(float) * B * H, cudaMemcpyDeviceToDevice, st));
    if (l + 1 < nl)   // this layer's outputs are the next layer's inputs
      API_CK(cudaMemcpyAsync(seqB, Himg + (int64_t)B * H, sizeof(float) * L * B * H, cudaMemcpyDeviceToDevice, st));
  }
  API_CK(fill_i32(st, tok, B, V + 1));
  API_CK(fill_i32(st, done, 1, 0));
  API_CK(cudaMemsetAsync(ids, 0, sizeof(int64_t) * (size_t)B * S, st));
  for (int t = 0; t < S; ++t) {   // :138-149
    const int cur = t & 1, nxt = cur ^ 1;
    API_CK(gather_rows(st, P->emb_w, tok, B, E, x, E));
    const float* in = x; int I = E;
    for (int l = 0; l < nl; ++l) {
      float* hc = hst + (int64_t)(l * 2 + cur) * B * H; float* hn = hst + (int64_t)(l * 2 + nxt) * B * H;
      float* cc = cst + (int64_t)(l * 2 + cur) * B * H; float* cn = cst + (int64_t)(l * 2 + nxt) * B * H;
      API_CK(gemm_f32(st, false, true, B, 4 * H, I, 1.f, in, I, P->w_ih[l], I, g, 4 * H, P->b_ih[l], P->b_hh[l], false, ACT_NONE));
      API_CK(gemm_f32(st, false, true, B, 4 * H, H, 1.f, hc, H, P->w_hh[l], H, g, 4 * H, nullptr, nullptr, true, ACT_NONE));
      API_CK(lstm_pointwise_fwd(st, B, H, g, 4 * H, cc, cn, hn, H, nullptr, 0));
      in = hn; I = H;
    }
    API_CK(gemm_f32(st, false, true, B, Vp, H, 1.f, in, H, P->out_w, H, logits, ldV, P->out_b, nullptr, false, ACT_NONE));
    API_CK(argmax_rows(st, logits, B, Vp, ldV, tok));
    API_CK(greedy_commit(st, B, S, 0, t, V + 2, tok, nullptr, ids, nullptr, done));
  }
  return 0;
}

// ---- beam search for the plain LSTM: LanguageModel.caption (AlexCap/LanguageModule.py:152-198), batched over images ----
struct LstmBeamWs {
  float *seqA, *seqB, *Gimg, *Himg, *Cimg, *hst, *cst, *x, *g, *logits, *top, *comp_score;
  int32_t *tok, *kact, *words, *back, *rowrank, *src_row, *ncomp, *comp_step, *comp_rank;
  size_t bytes;
};
static LstmBeamWs lstm_beam_layout(const alexcap_dims* d, int K, void* base) {
  LstmBeamWs w;
  Bump a(base);
  const int64_t B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, nl = d->num_layers, R = B * K, S1 = d->S + 1;
  const int64_t ldV = (d->Vp + 3) & ~int64_t(3);
  w.seqA = a.take<float>(L * B * D);                 // raw features, time-major (the image pass runs on B rows)
  w.seqB = a.take<float>(L * B * H);
  w.Gimg = a.take<float>(L * B * 4 * H);
  w.Himg = a.take<float>((L + 1) * B * H);
  w.Cimg = a.take<float>((L + 1) * B * H);
  w.hst = a.take<float>(nl * 2 * R * H);             // [layer][0 = current, 1 = after the step][R,H]
  w.cst = a.take<float>(nl * 2 * R * H);
  w.x = a.take<float>(R * E);
  w.g = a.take<float>(R * 4 * H);
  w.logits = a.take<float>(R * ldV);
  w.top = a.take<float>(R);
  w.comp_score = a.take<float>(R);
  w.tok = a.take<int32_t>(R);
  w.kact = a.take<int32_t>(B);
  w.words = a.take<int32_t>(S1 * R);
  w.back = a.take<int32_t>(S1 * R);
  w.rowrank = a.take<int32_t>((S1 + 1) * R);
  w.src_row = a.take<int32_t>(R);
  w.ncomp = a.take<int32_t>(B);
  w.comp_step = a.take<int32_t>(R);
  w.comp_rank = a.take<int32_t>(R);
  w.bytes = a.off + 256;
  return w;
}

size_t alexcap_lstm_decode_beam_workspace_bytes(const alexcap_dims* d, int32_t beam) {
  if (!d || d->num_layers < 1 || d->num_layers > ALEXCAP_MAX_LAYERS || beam < 1 || beam > 32) return 0;
  return lstm_beam_layout(d, beam, nullptr).bytes;
}

int alexcap_lstm_decode_beam(const alexcap_dims* d, const alexcap_lstm_params* P, const float* feat, int32_t beam,
                             int64_t* ids, int32_t* lens, float* scores, void* workspace, size_t workspace_bytes,
                             void* stream) {
  if (int rc = check_lstm_dims(d)) return rc;
  if (beam < 1 || beam > 32) return fail(ALEXCAP_E_BADARG, "beam size must be in [1,32]");
  if (d->D != d->E)   // :160 feeds img_features straight into nn.LSTM(input_size = embedding size)
    return fail(ALEXCAP_E_BADARG, "LanguageModel.caption feeds the raw features to the LSTM: D must equal E");
  ExactScope exact_scope;
  if (int rc = check_lstm_args(d, P, feat, workspace, workspace_bytes, lstm_beam_layout(d, beam, nullptr).bytes)) return rc;
  if (!ids || !lens || !scores) return fail(ALEXCAP_E_BADARG, "NULL ids / lens / scores");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int K = beam;
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, Vp = d->Vp, S1 = d->S + 1, nl = d->num_layers, V = Vp - 3;
  const int R = B * K;
  const int64_t ldV = (Vp + 3) & ~int64_t(3);
  LstmBeamWs w = lstm_beam_layout(d, K, workspace);
  // image pass (:160): all K rows of an image are identical -> run it on B rows, then replicate the final states
  API_CK(permute_01(st, B, L, D, feat, w.seqA));
  for (int l = 0; l < nl; ++l) {
    const float* in = (l == 0) ? w.seqA : w.seqB;
    if (int rc = run_layer_fwd(d, in, l == 0 ? D : H, P->w_ih[l], P->w_hh[l], P->b_ih[l], P->b_hh[l], w.Gimg, w.Himg, w.Cimg, L, st))
      return rc;
    float* h0 = w.hst + (int64_t)(l * 2) * R * H;
    float* c0 = w.cst + (int64_t)(l * 2) * R * H;
    for (int j = 0; j < K; ++j) {
      API_CK(copy_2d(st, B, H, w.Himg + (int64_t)L * B * H, H, h0 + (int64_t)j * H, (int64_t)K * H));
      API_CK(copy_2d(st, B, H, w.Cimg + (int64_t)L * B * H, H, c0 + (int64_t)j * H, (int64_t)K * H));
    }
    if (l + 1 < nl)
      API_CK(cudaMemcpyAsync(w.seqB, w.Himg + (int64_t)B * H, sizeof(float) * L * B * H, cudaMemcpyDeviceToDevice, st));
  }
  API_CK(fill_i32(st, w.tok, R, V + 1));
  API_CK(fill_i32(st, w.kact, B, K));
  API_CK(fill_i32(st, w.ncomp, B, 0));
  API_CK(fill_i32(st, w.rowrank, (int64_t)(S1 + 1) * R, 0));
  API_CK(fill_f32(st, w.top, R, 0.f));
  for (int s = 0; s < S1; ++s) {   // the reference's `step` is s+1; it stops after step seq_length+2 = S+1 (:190-192)
    API_CK(gather_rows(st, P->emb_w, w.tok, R, E, w.x, E));
    const float* in = w.x; int I = E;
    for (int l = 0; l < nl; ++l) {
      float* hc = w.hst + (int64_t)(l * 2) * R * H; float* hn = hc + (int64_t)R * H;
      float* cc = w.cst + (int64_t)(l * 2) * R * H; float* cn = cc + (int64_t)R * H;
      API_CK(gemm_f32(st, false, true, R, 4 * H, I, 1.f, in, I, P->w_ih[l], I, w.g, 4 * H, P->b_ih[l], P->b_hh[l], false, ACT_NONE));
      API_CK(gemm_f32(st, false, true, R, 4 * H, H, 1.f, hc, H, P->w_hh[l], H, w.g, 4 * H, nullptr, nullptr, true, ACT_NONE));
      API_CK(lstm_pointwise_fwd(st, R, H, w.g, 4 * H, cc, cn, hn, H, nullptr, 0));
      in = hn; I = H;
    }
    // eval mode: rnn.drop is the identity
    API_CK(gemm_f32(st, false, true, R, Vp, H, 1.f, in, H, P->out_w, H, w.logits, ldV, P->out_b, nullptr, false, ACT_NONE));
    API_CK(beam_select(st, B, K, Vp, S1, s, V + 2, w.logits, ldV, w.top, w.kact, w.words + (int64_t)s * R,
                       w.back + (int64_t)s * R, w.rowrank + (int64_t)(s + 1) * R, w.src_row, w.tok, w.ncomp, w.comp_score,
                       w.comp_step, w.comp_rank));
    for (int l = 0; l < nl; ++l) {   // surviving beams continue from their parent's state of every layer (:186)
      float* hc = w.hst + (int64_t)(l * 2) * R * H; float* cc = w.cst + (int64_t)(l * 2) * R * H;
      API_CK(beam_gather_state(st, B, K, H, w.src_row, hc + (int64_t)R * H, H, cc + (int64_t)R * H, hc, H, cc));
    }
  }
  API_CK(beam_finalize(st, B, K, S1, 0, w.words, w.back, w.rowrank, w.ncomp, w.comp_score, w.comp_step, w.comp_rank, nullptr,
                       ids, lens, scores, nullptr));
  return 0;
}

}  // extern "C"
