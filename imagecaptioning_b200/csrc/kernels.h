// kernels.h -- host-side launchers of the sm_100a kernels (internal; the public surface is include/alexcap.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace alexcap {

enum Act { ACT_NONE = 0, ACT_TANH = 1, ACT_RELU = 2, ACT_EXP2C = 3 };   // (applied by act_apply, common.cuh)

// C[M,N] = act(alpha * op(A) op(B) + bias + bias2) (+ C).  See alexcap_gemm_f32 in alexcap.h.
cudaError_t gemm_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                     const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                     const float* bias, const float* bias2, bool accumulate, int act);

cudaError_t gemm_simt_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                          const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                          const float* bias, const float* bias2, bool accumulate, int act);
// tcgen05 / TMEM / TMA engine (gemm_tc.cu); force_splits > 0 overrides the split-K heuristic (tests)
cudaError_t gemm_tc_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                        const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                        const float* bias, const float* bias2, bool accumulate, int act, int force_splits, int mode = 0);
// Fused LSTM cell -> next GEMM (gemm_tc.cu, north_star 3): C[M,N] += cell(g, c_prev) B^T where the A operand h = o * tanh(c)
// is computed by the kernel's converter warps from the gate pre-activations instead of being loaded: the LSTM pointwise pass
// and the GEMM that consumes its h (the next step's attention query / beta projection) are ONE launch.
//   g [M, 4H] (row stride g_ld): gate pre-activations i|f|g|o (read only);  c_prev [M, H]
//   written by the CTAs of the first column tile: acts [M, 4H] (row stride acts_ld: the gate ACTIVATIONS the backward pass
//   reads -- a different buffer than g, other CTAs still read g), c_out [M, H], h_out [M, H] (row stride h_ld)
struct LstmA { const float* g; int64_t g_ld; const float* c_prev; float* c_out; float* h_out; int64_t h_ld; float* acts; int64_t acts_ld; };
// K = H (H % 8 == 0), B = weights [N, H] row-major (ldb), C pre-initialised (accumulated with red.add), bias added once.
// Returns cudaErrorNotSupported when the shape would make too many CTAs repeat the cell math (caller: separate launches).
bool gemm_tc_lstm_supported(int M, int N, int H);
cudaError_t gemm_tc_lstm_f32(cudaStream_t st, int M, int N, int H, const LstmA& la, const float* B, int64_t ldb, float* C,
                             int64_t ldc, const float* bias, int mode);
// Pre-split engine (gemm_ps.cu): operands split once in HBM into bf16 hi (+ lo) copies that live in a bump arena (cached
// per source pointer until released), then a persistent TMA -> tcgen05 kernel.  nprod: 1 = bf16, 3 = bf16x3.
// Returns cudaErrorMemoryAllocation when the arena cannot hold the copies (the caller falls back to gemm_tc_f32).
struct SplitArena {
  static constexpr int MAX_ENTRIES = 256;
  struct Entry { const float* src; int rows, cols; int64_t ld, ldo; uint16_t* hi; uint16_t* lo; size_t end; bool produced; bool mirror; };
  // produced: the bf16 copies are written by the matrix's producer kernel, not by a split pass.  mirror: ... beside the fp32
  // matrix, which exists as well (a GEMM operand that is a row / column sub-block of `src` uses the matching sub-view).
  uint8_t* base = nullptr; size_t bytes = 0, used = 0;
  Entry e[MAX_ENTRIES]; int n = 0;
};
SplitArena*& split_arena();                       // the arena big GEMMs of the current call may use (nullptr: none)
void split_arena_release(SplitArena* a, size_t mark);   // forget every copy made after `mark` (= a->used at that time)
bool gemm_ps_wanted(int M, int N, int K);
// true when `p` is the key of copies written directly by their producer: the fp32 matrix behind it does NOT hold the data
bool split_arena_is_produced(const SplitArena* a, const float* p);
cudaError_t split_arena_get_hi(cudaStream_t st, SplitArena* a, const float* X, int rows, int cols, int64_t ld, const uint16_t** hi,
                               int64_t* ldo);
bool split_arena_reserve(SplitArena* a, const float* key, int rows, int cols, int64_t ld, bool with_lo, uint16_t** hi,
                         uint16_t** lo, int64_t* ldo_out);
bool split_arena_find(const SplitArena* a, const float* key, uint16_t** hi, uint16_t** lo, int64_t* ldo_out);
// bf16 (+ lo) MIRROR of the fp32 matrix `src` [rows, cols] (pitch ld): the kernels that write `src` write the same values into
// the mirror (row pitch *ldo_out elements), so that no split pass is needed when `src` -- or any sub-block of it with a column
// offset that is a multiple of 8 -- is a GEMM operand later in the step.  The caller must make sure EVERY element a GEMM will
// read has been written by a producer (or by split_arena_fill_view).
bool split_arena_reserve_mirror(SplitArena* a, const float* src, int rows, int cols, int64_t ld, bool with_lo, uint16_t** hi,
                                uint16_t** lo, int64_t* ldo_out);
// split pass over the sub-block X [rows, cols] (pitch ld) of a mirrored matrix into the matching part of its mirror
cudaError_t split_arena_fill_view(cudaStream_t st, SplitArena* a, const float* X, int rows, int cols, int64_t ld);
cudaError_t gemm_ps_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                        const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                        const float* bias, const float* bias2, bool accumulate, int act, int force_splits, int nprod,
                        SplitArena* arena);
// Fused vocabulary projection + masked label-smoothed cross-entropy on the pre-split engine (gemm_ps.cu): the logits tile is
// reduced / differentiated in the tcgen05 epilogue, straight from TMEM; logits are never written.  Two passes over the same
// product (the second recomputes it): statistics -> (row_lse, row_loss), then the gradient as bf16 hi (+ lo) copies.
bool gemm_ps_ce_wanted(int M, int N, int K);
int gemm_ps_ce_partials(int N);          // float4 statistics per row the first pass writes
cudaError_t gemm_ps_ce(cudaStream_t st, bool want_grad, int M, int N, int K, const float* A, int64_t lda, const float* B, int64_t ldb,
                       const float* bias, const int32_t* tgt, float eps, int ignore_index, const float* denom, float4* stats,
                       float* row_lse, float* row_loss, uint16_t* dl_hi, uint16_t* dl_lo, int64_t dl_ld, int nprod,
                       SplitArena* arena);
bool gemm_tc_supported(int M, int N, int K, const float* A, int64_t lda, const float* B, int64_t ldb, const float* C);
int& gemm_engine();                      // this thread's engine (scoped overrides write it)
void gemm_engine_set_default(int engine); // process-wide default, picked up by every thread
// Hint for the calls made while it is true: the B operand holds weights that were final before the PREVIOUS kernel of the
// stream started, so the kernel may fetch B tiles ahead of its programmatic-dependent-launch wait (api.cu: WeightBScope).
bool& gemm_b_is_weight();
// > 0: the pre-split engine's persistent kernels launched while it is set use at most this many CTAs (SMs), whatever the
// device has -- for a GEMM forked onto a side stream next to kernels that need the other SMs (api.cu: CtaCapScope).
int& gemm_cta_cap();

// ---- attention (attention.cu)
cudaError_t attn_step_fwd(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                          const float* q, int64_t q_ld, const float* bpre, int64_t bpre_ld, const float* wv,
                          float* alpha, int64_t alpha_ld, float* ctx, float* beta, float* gctx, int64_t gctx_ld,
                          int rows_per_image = 1);
cudaError_t attn_step_bwd(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                          const float* q, int64_t q_ld, const float* wv, const float* alpha, int64_t alpha_ld,
                          const float* ctx, const float* beta, const float* dgctx, int64_t dgctx_ld,
                          const float* dalpha_ext, int64_t ext_ld, const float* reg_r,
                          float* dq, int64_t dq_ld, float* dbpre, int64_t dbpre_ld, float* de, int64_t de_ld,
                          float* dwv_part, float* dfeat);
// TMA-fed streaming variants (attention_stream.cu): the default engine; attention.cu keeps the direct-load kernels
// (engine 0: ALEXCAP_ATT_ENGINE=0, or shapes the streaming kernels do not cover).
bool attn_stream_supported(int D, int H);
int& attn_engine();
cudaError_t attn_step_fwd_stream(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                                 const float* q, int64_t q_ld, const float* bpre, int64_t bpre_ld, const float* wv,
                                 float* alpha, int64_t alpha_ld, float* ctx, float* beta, float* gctx, int64_t gctx_ld,
                                 int rows_per_image, bool rows_bf16 = false, bool ws_exp = false, bool static_ready = false,
                                 uint16_t* gctx_b16 = nullptr, int64_t gb_ld = 0);   // gctx_b16: bf16 mirror of gctx
// static_ready: Ws / feat were complete before the kernel that precedes this launch in the stream started (true inside the
// time loops): the producer may fetch row pieces before griddepcontrol.wait; only q / bpre (dgctx) wait for the predecessor
// ws_exp: Ws holds EXP2C(W(features)) (GEMM activation ACT_EXP2C; common.cuh) -- one MUFU op per score element instead of two
// rows_bf16: Ws and feat point at bf16 rows of the same shapes (ALEXCAP_BF16 storage: half the streamed bytes; D, H % 8 == 0)
cudaError_t attn_step_bwd_stream(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                                 const float* q, int64_t q_ld, const float* wv, const float* alpha, int64_t alpha_ld,
                                 const float* ctx, const float* beta, const float* dgctx, int64_t dgctx_ld,
                                 const float* dalpha_ext, int64_t ext_ld, const float* reg_r,
                                 float* dq, int64_t dq_ld, float* dbpre, int64_t dbpre_ld, float* de, int64_t de_ld,
                                 float* dwv_part, float* dfeat, bool rows_bf16 = false, bool ws_exp = false, bool static_ready = false,
                                 uint16_t* dq_b16 = nullptr, uint16_t* dbpre_b16 = nullptr, int64_t db_ld = 0);   // bf16 mirrors of dq / dbpre
// dWs[b,l,h] = wv[h] * sum_t de[b,t,l] * (1 - tanh^2(Ws[b,l,h] + q[t,b,h]));  q: [S][B] rows of stride q_ld,
// de: [B][S][L].
cudaError_t attn_dws(cudaStream_t st, int B, int L, int H, int S, const float* Ws, const float* q, int64_t q_ld,
                     const float* wv, const float* de, float* dWs, uint16_t* d_hi = nullptr, uint16_t* d_lo = nullptr,
                     int64_t d_ld = 0, int max_ctas = 0, bool ws_exp = false);
// d_hi / d_lo: bf16 copies [B*L, d_ld] written as well as / instead of dWs (may be NULL); max_ctas > 0 caps the grid
// resid = 1 - sum_t alpha[b,t,l]; r[b,l] = -2*w*resid/(B*L); loss_reg = w*mean(resid^2) -> *loss_reg
cudaError_t attn_regulariser(cudaStream_t st, int B, int S, int L, float weight, const float* alphas, float* r,
                             float* loss_reg, float* scratch);

// ---- persistent time loop (time_loop.cu): the whole forward recurrence in one cooperative launch, recurrent weights
// resident in shared memory, LSTM pointwise in the tcgen05 epilogue.  Supported when the weight slices fit (B <= 128,
// H <= 512, H/4 <= #SMs ...); the caller falls back to the per-step chain otherwise.
bool& time_loop_enabled();     // process-wide switch (default: ALEXCAP_TIME_LOOP=1 in the environment, else off)
bool attn_loop_supported(int B, int L, int D, int H);
size_t attn_loop_workspace_elems_bf16(int B, int S, int D, int H);      // 0 when unsupported
cudaError_t attn_loop_fwd(cudaStream_t st, int B, int L, int D, int H, int E, int S, const float* feat, const float* Ws,
                          const float* w_hh, const float* w_ih, const float* w_u, const float* b_u, const float* w_beta,
                          const float* b_beta, const float* wv, float* QB, float* XH, float* DGQ, float* Call, float* ALPHA,
                          float* CTX, float* BETA, uint16_t* bf16ws, unsigned* bar);

int time_loop_profile_read(unsigned long long* host_out, int max_entries);

// ---- LSTM pointwise (lstm_pointwise.cu)
// gates [B,4H] (row stride g_ld) pre-activations -> activations in place; c_prev [B,H] -> c_out [B,H];
// h written to h_out (row stride h_ld) and optionally h_out2 (row stride h2_ld).
cudaError_t lstm_pointwise_fwd(cudaStream_t st, int B, int H, float* gates, int64_t g_ld, const float* c_prev,
                               float* c_out, float* h_out, int64_t h_ld, float* h_out2, int64_t h2_ld, float* acts_out = nullptr,
                               int64_t acts_ld = 0, uint16_t* h_b16 = nullptr, int64_t hb_ld = 0);
// acts_out: write the gate activations there instead of over `gates`; h_b16: bf16 mirror of h_out (row pitch hb_ld elements)
// gates hold activations (i,f,g,o) -> overwritten with d(pre-activations).  dh = dh_a + dh_b (either may be NULL;
// dh_a scaled by mask_a if given).  dc_io [B,H]: in = dc from step t+1, out = dc for step t-1.
cudaError_t lstm_pointwise_bwd(cudaStream_t st, int B, int H, float* gates, int64_t g_ld, const float* c_t,
                               const float* c_prev, const float* dh_a, int64_t dha_ld, const float* mask_a,
                               const float* dh_b, int64_t dhb_ld, float* dc_io, uint16_t* g_b16 = nullptr, int64_t gb_ld = 0);

// ---- loss (ce_loss.cu)
cudaError_t count_valid(cudaStream_t st, const int32_t* tgt, int N, int ignore_index, float* n_valid_out);
// per-row masked label-smoothed CE; denom (device float) = divisor; row_loss[N]; dlogits may alias logits.
// column sums of X[M, n0+n1+n2] written to three destinations by column range (n1 / n2 may be 0)
cudaError_t col_sum3(cudaStream_t st, const float* X, int M, int64_t ld, float* o0, int n0, float* o1, int n1, float* o2, int n2,
                     float* scratch);
// out[n] (+)= sum_m (hi[m, n] + lo[m, n]) over bf16 pair copies (lo may be NULL)
cudaError_t col_sum_bf16(cudaStream_t st, const uint16_t* hi, const uint16_t* lo, int M, int N, int64_t ld, float* out,
                         bool accumulate, float* scratch);
cudaError_t ce_rows(cudaStream_t st, const float* logits, const int32_t* tgt, int N, int C, int64_t ld, float eps,
                    int ignore_index, const float* denom, float* row_loss, float* dlogits,
                    uint16_t* dl_hi = nullptr, uint16_t* dl_lo = nullptr, int64_t dl_ld = 0);
// (dl_hi / dl_lo: the gradient is ALSO/INSTEAD written as bf16 hi (+ lo residual) copies [N, dl_ld] for the pre-split GEMMs)
// out[0] = sum(x[0..n)) (deterministic single-block tree)
cudaError_t reduce_sum(cudaStream_t st, const float* x, int n, float* out, float* scratch = nullptr);

// ---- misc (misc.cu)
// n_classes > 0: tokens outside [0, n_classes) are replaced by 0 and the row is flagged in bad_rows[B] (may be NULL)
cudaError_t make_target(cudaStream_t st, const int32_t* gt, int B, int T, int start_tok, int end_tok,
                        int32_t* in_bm, int32_t* tgt_bm, int32_t* in_tm, int32_t* tgt_tm, int n_classes = 0,
                        int32_t* bad_rows = nullptr);
// out[0] = nnz(gt) + B = number of non-ignored targets of get_target(gt, make_target=True)
cudaError_t count_valid_targets(cudaStream_t st, const int32_t* gt, int B, int T, float* out);
cudaError_t mean_over_L(cudaStream_t st, const float* feat, int B, int L, int D, float* m);
// out[r,:] = table[idx[r],:]   (rows of `width` floats; out row stride out_ld)
cudaError_t gather_rows(cudaStream_t st, const float* table, const int32_t* idx, int rows, int width, float* out,
                        int64_t out_ld);
// table[idx[r],:] += src[r,:]   (atomic)
cudaError_t scatter_add_rows(cudaStream_t st, const float* src, int64_t src_ld, const int32_t* idx, int rows,
                             int width, float* table);
// out[n] (+)= sum_m X[m*ld + n]     scratch: >= 64*N floats
cudaError_t col_sum(cudaStream_t st, const float* X, int M, int N, int64_t ld, float* out, bool accumulate,
                    float* scratch);
// y = dy * (1 - t*t)  elementwise (tanh backward), n elements, strides for 2-D views
cudaError_t tanh_bwd(cudaStream_t st, int rows, int cols, const float* dy, int64_t dy_ld, const float* t, int64_t t_ld,
                     float* out, int64_t out_ld);
cudaError_t relu_bwd(cudaStream_t st, int64_t n, const float* dy, const float* pre, float* out);
// dst[r, :cols] = src[r, :cols] with strides
struct CopySegments {
  static constexpr int MAX = 8;
  const float* src[MAX]; float* dst[MAX]; int64_t src_ld[MAX], dst_ld[MAX]; int rows[MAX], cols[MAX]; int n = 0;
  void add(const float* s, int64_t sld, float* d, int64_t dld, int r, int c) {
    src[n] = s; dst[n] = d; src_ld[n] = sld; dst_ld[n] = dld; rows[n] = r; cols[n] = c; ++n;
  }
};
// all segments in one launch; columns, pitches multiples of 4 floats, 16-byte aligned bases
cudaError_t copy_segments(cudaStream_t st, const CopySegments& t);
cudaError_t copy_2d(cudaStream_t st, int rows, int cols, const float* src, int64_t src_ld, float* dst, int64_t dst_ld);
cudaError_t fill_f32(cudaStream_t st, float* p, int64_t n, float v);
// dfeat[b,l,:] += dm[b,:] * scale
cudaError_t add_broadcast_L(cudaStream_t st, int B, int L, int D, const float* dm, float scale, float* dfeat);
// out[i] = a[i] * b[i]
cudaError_t mul_2d(cudaStream_t st, int rows, int cols, const float* a, int64_t a_ld, const float* b, int64_t b_ld,
                   float* out, int64_t out_ld);
// mask[i] = (philox(seed, offset+i) >= p) ? 1/(1-p) : 0.  state (device {seed, offset}) overrides seed / offset and is
// advanced on the device afterwards.  y != NULL: also y[i] = mask[i] * x[i / cols][i % cols] (x row stride x_ld).
cudaError_t dropout_mask(cudaStream_t st, float* mask, int64_t n, float p, uint64_t seed, uint64_t offset,
                         uint64_t* state = nullptr, const float* x = nullptr, int64_t x_ld = 0, int cols = 1,
                         float* y = nullptr);
// loss_out = {a+b, a, b, n_valid}; a = NaN when n_valid == 0 or any bad_rows[i] != 0
cudaError_t finalize_loss(cudaStream_t st, const float* ce_sum, const float* reg, const float* n_valid, float* loss_out,
                          const int32_t* bad_rows = nullptr, int n_bad = 0);

// ---- decode (decode.cu)
// per row argmax over logits[B,C] (row stride ld); first max wins (torch.argmax semantics)
cudaError_t argmax_rows(cudaStream_t st, const float* logits, int B, int C, int64_t ld, int32_t* out_tok);
cudaError_t fill_i32(cudaStream_t st, int32_t* p, int64_t n, int32_t v);
// greedy bookkeeping (AttentionLanguageModule.py:108-112): unless *done, ids[b,t] = tok[b] and (optionally)
// alphas[b,t,:] = alpha_step[b,:]; then *done |= all(tok == end_tok).  No host sync.
cudaError_t greedy_commit(cudaStream_t st, int B, int S, int L, int t, int end_tok, const int32_t* tok,
                          const float* alpha_step, int64_t* ids, float* alphas, int32_t* done);
// dst[a,b,:] = src[b,a,:]  (src [nb,na,w] -> dst [na,nb,w])
cudaError_t permute_01(cudaStream_t st, int nb, int na, int w, const float* src, float* dst);

}  // namespace alexcap
