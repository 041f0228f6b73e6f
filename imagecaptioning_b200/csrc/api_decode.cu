// api_decode.cu -- attention-LSTM decoding behind the C ABI: batched greedy
// (AlexCap/AttentionLanguageModule.py:88-115) and batched beam search (:128-195).
// Both run entirely on the device for a fixed S = seq_length+1 steps with device-side "done" bookkeeping instead
// of the reference's per-token host synchronisations.
#include "api_util.h"

using namespace alexcap;

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

namespace {

struct DecWs {
  float *m, *Ws, *xh[2], *c[2], *qb, *ctx, *beta, *alpha_steps, *x, *g, *logits, *Wcat, *Wub, *bub, *top, *comp_score;
  int32_t *tok, *done, *kact, *words, *back, *rowrank, *src_row, *ncomp, *comp_step, *comp_rank;
  int64_t ldV;
  size_t bytes;
};

// R = rows decoded at once (B for greedy, B*beam for beam search)
DecWs dec_layout(const alexcap_dims* d, int K, void* base) {
  DecWs w;
  Bump a(base);
  const int64_t B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, Vp = d->Vp, S = d->S, R = B * K;
  w.ldV = (Vp + 3) & ~int64_t(3);
  w.m = a.take<float>(R * D);
  w.Ws = a.take<float>(B * L * H);
  for (int i = 0; i < 2; ++i) { w.xh[i] = a.take<float>(R * (D + H)); w.c[i] = a.take<float>(R * H); }
  w.qb = a.take<float>(R * (H + D));
  w.ctx = a.take<float>(R * D);
  w.beta = a.take<float>(R * D);
  w.alpha_steps = a.take<float>((K > 1 ? S : 1) * R * L);
  w.x = a.take<float>(R * E);
  w.g = a.take<float>(R * 4 * H);
  w.logits = a.take<float>(R * w.ldV);
  w.Wcat = a.take<float>(4 * H * (D + H));
  w.Wub = a.take<float>((H + D) * H);
  w.bub = a.take<float>(H + D);
  w.top = a.take<float>(R);
  w.comp_score = a.take<float>(R);
  w.tok = a.take<int32_t>(R);
  w.done = a.take<int32_t>(16);
  w.kact = a.take<int32_t>(B);
  w.words = a.take<int32_t>(S * R);
  w.back = a.take<int32_t>(S * R);
  w.rowrank = a.take<int32_t>((S + 1) * R);
  w.src_row = a.take<int32_t>(R);
  w.ncomp = a.take<int32_t>(B);
  w.comp_step = a.take<int32_t>(R);
  w.comp_rank = a.take<int32_t>(R);
  w.bytes = a.off + 256;
  return w;
}

int check_dec(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, void* ws, size_t ws_bytes, int K) {
  if (int rc = check_dims_common(d)) return rc;
  if (K < 1 || K > 32) return fail(ALEXCAP_E_BADARG, "beam size must be in [1,32]");
  if (!P || !feat || !ws) return fail(ALEXCAP_E_BADARG, "NULL params / feat / workspace");
  if (ws_bytes < dec_layout(d, K, nullptr).bytes) return fail(ALEXCAP_E_WORKSPACE, "decode workspace too small");
  if (!aligned16(feat) || !aligned16(ws)) return fail(ALEXCAP_E_BADARG, "feat / workspace must be 16-byte aligned");
  return 0;
}

// shared prologue: h0/c0 (replicated K times per image), Ws, concatenated weights
int dec_prologue(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, int K, DecWs& w, cudaStream_t st) {
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E;
  const int64_t DH = D + H;
  API_CK(mean_over_L(st, feat, B, L, D, w.m));
  // rows b*K + j all start from image b's state: compute for B rows into row stride K*DH, then replicate
  API_CK(gemm_f32(st, false, true, B, H, D, 1.f, w.m, D, P->init_h_w, D, w.xh[0] + D, (int64_t)K * DH, P->init_h_b, nullptr, false,
                  ACT_TANH));
  API_CK(gemm_f32(st, false, true, B, H, D, 1.f, w.m, D, P->init_c_w, D, w.c[0], (int64_t)K * H, P->init_c_b, nullptr, false, ACT_TANH));
  for (int j = 1; j < K; ++j) {
    API_CK(copy_2d(st, B, H, w.xh[0] + D, (int64_t)K * DH, w.xh[0] + (int64_t)j * DH + D, (int64_t)K * DH));
    API_CK(copy_2d(st, B, H, w.c[0], (int64_t)K * H, w.c[0] + (int64_t)j * H, (int64_t)K * H));
  }
  API_CK(gemm_f32(st, false, true, B * L, H, D, 1.f, feat, D, P->att_W_w, D, w.Ws, H, P->att_W_b, nullptr, false, ACT_NONE));
  API_CK(copy_2d(st, 4 * H, D, P->lstm_w_ih + E, E + D, w.Wcat, DH));
  API_CK(copy_2d(st, 4 * H, H, P->lstm_w_hh, H, w.Wcat + D, DH));
  API_CK(cudaMemcpyAsync(w.Wub, P->att_U_w, sizeof(float) * H * H, cudaMemcpyDeviceToDevice, st));
  API_CK(cudaMemcpyAsync(w.Wub + (int64_t)H * H, P->f_beta_w, sizeof(float) * D * H, cudaMemcpyDeviceToDevice, st));
  API_CK(cudaMemcpyAsync(w.bub, P->att_U_b, sizeof(float) * H, cudaMemcpyDeviceToDevice, st));
  API_CK(cudaMemcpyAsync(w.bub + H, P->f_beta_b, sizeof(float) * D, cudaMemcpyDeviceToDevice, st));
  return 0;
}

// one decoder step for R rows: attention -> gate -> LSTMCell -> logits.  State slot `cur` -> slot cur^1.
int dec_step(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, int K, int cur, float* alpha_out,
             DecWs& w, cudaStream_t st) {
  const int B = d->B, L = d->L, D = d->D, H = d->H, E = d->E, Vp = d->Vp, R = B * K;
  const int64_t DH = D + H, HD = H + D;
  float* xh = w.xh[cur];
  API_CK(gemm_f32(st, false, true, R, H + D, H, 1.f, xh + D, DH, w.Wub, H, w.qb, HD, w.bub, nullptr, false, ACT_NONE));
  API_CK(attn_step_fwd(st, R, L, D, H, w.Ws, feat, w.qb, HD, w.qb + H, HD, P->att_v_w, alpha_out, L, w.ctx, w.beta, xh, DH, K));
  API_CK(gather_rows(st, P->emb_w, w.tok, R, E, w.x, E));
  API_CK(gemm_f32(st, false, true, R, 4 * H, E, 1.f, w.x, E, P->lstm_w_ih, E + D, w.g, 4 * H, P->lstm_b_ih, P->lstm_b_hh, false,
                  ACT_NONE));
  API_CK(gemm_f32(st, false, true, R, 4 * H, D + H, 1.f, xh, DH, w.Wcat, DH, w.g, 4 * H, nullptr, nullptr, true, ACT_NONE));
  API_CK(lstm_pointwise_fwd(st, R, H, w.g, 4 * H, w.c[cur], w.c[cur ^ 1], w.xh[cur ^ 1] + D, DH, nullptr, 0));
  // eval mode: dropout is the identity (LSTMwAttentionModel.py:77 calls self.eval())
  API_CK(gemm_f32(st, false, true, R, Vp, H, 1.f, w.xh[cur ^ 1] + D, DH, P->deep_w, H, w.logits, w.ldV, P->deep_b, nullptr, false,
                  ACT_NONE));
  return 0;
}

}  // namespace

extern "C" {

size_t alexcap_attn_decode_workspace_bytes(const alexcap_dims* d, int32_t beam) {
  if (!d || beam < 1 || beam > 32) return 0;
  return dec_layout(d, beam, nullptr).bytes;
}

int alexcap_attn_decode_greedy(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, int64_t* ids,
                               float* alphas, void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_dec(d, P, feat, workspace, workspace_bytes, 1)) return rc;
  ExactScope exact_scope;
  if (!ids) return fail(ALEXCAP_E_BADARG, "NULL ids");
  DecWs w = dec_layout(d, 1, workspace);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, L = d->L, S = d->S, V = d->Vp - 3;
  if (int rc = dec_prologue(d, P, feat, 1, w, st)) return rc;
  API_CK(fill_i32(st, w.tok, B, V + 1));
  API_CK(fill_i32(st, w.done, 1, 0));
  API_CK(cudaMemsetAsync(ids, 0, sizeof(int64_t) * (size_t)B * S, st));
  if (alphas) API_CK(cudaMemsetAsync(alphas, 0, sizeof(float) * (size_t)B * S * L, st));
  for (int t = 0; t < S; ++t) {
    if (int rc = dec_step(d, P, feat, 1, t & 1, w.alpha_steps, w, st)) return rc;
    API_CK(argmax_rows(st, w.logits, B, d->Vp, w.ldV, w.tok));                       // :108
    API_CK(greedy_commit(st, B, S, L, t, V + 2, w.tok, w.alpha_steps, ids, alphas, w.done));   // :109-112
  }
  return 0;
}

int alexcap_attn_decode_beam(const alexcap_dims* d, const alexcap_attn_params* P, const float* feat, int32_t beam,
                             int64_t* ids, int32_t* lens, float* scores, float* alphas, void* workspace,
                             size_t workspace_bytes, void* stream) {
  if (int rc = check_dec(d, P, feat, workspace, workspace_bytes, beam)) return rc;
  ExactScope exact_scope;
  if (!ids || !lens || !scores) return fail(ALEXCAP_E_BADARG, "NULL ids / lens / scores");
  const int K = beam;
  DecWs w = dec_layout(d, K, workspace);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int B = d->B, L = d->L, D = d->D, H = d->H, S = d->S, V = d->Vp - 3, R = B * K;
  const int64_t DH = D + H;
  if (int rc = dec_prologue(d, P, feat, K, w, st)) return rc;
  API_CK(fill_i32(st, w.tok, R, V + 1));
  API_CK(fill_i32(st, w.kact, B, K));
  API_CK(fill_i32(st, w.ncomp, B, 0));
  API_CK(fill_i32(st, w.rowrank, (int64_t)(S + 1) * R, 0));
  API_CK(fill_f32(st, w.top, R, 0.f));
  for (int s = 0; s < S; ++s) {     // the reference's `step` is s+1; it stops after step seq_length+1 = S (:187-189)
    const int cur = s & 1;
    if (int rc = dec_step(d, P, feat, K, cur, w.alpha_steps + (int64_t)s * R * L, w, st)) return rc;
    API_CK(beam_select(st, B, K, d->Vp, S, s, V + 2, w.logits, w.ldV, w.top, w.kact, w.words + (int64_t)s * R,
                       w.back + (int64_t)s * R, w.rowrank + (int64_t)(s + 1) * R, w.src_row, w.tok, w.ncomp, w.comp_score,
                       w.comp_step, w.comp_rank));
    // surviving beams continue from their parent's (h, c): gather slot cur^1 -> slot cur (free after the step)
    API_CK(beam_gather_state(st, B, K, H, w.src_row, w.xh[cur ^ 1] + D, DH, w.c[cur ^ 1], w.xh[cur] + D, DH, w.c[cur]));
    API_CK(copy_2d(st, R, H, w.xh[cur] + D, DH, w.xh[cur ^ 1] + D, DH));
    API_CK(cudaMemcpyAsync(w.c[cur ^ 1], w.c[cur], sizeof(float) * (size_t)R * H, cudaMemcpyDeviceToDevice, st));
  }
  API_CK(beam_finalize(st, B, K, S, L, w.words, w.back, w.rowrank, w.ncomp, w.comp_score, w.comp_step, w.comp_rank,
                       w.alpha_steps, ids, lens, scores, alphas));
  return 0;
}

}  // extern "C"
