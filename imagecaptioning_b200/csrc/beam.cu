// beam.cu -- device-side beam search bookkeeping for Decoder.caption (AlexCap/AttentionLanguageModule.py:128-195),
// batched over images.  Faithful to the reference's quirks:
//   * scores are cumulative RAW logits (no log-softmax, no length normalisation)           (:155-156)
//   * the first step expands beam row 0 only                                                (:158-159)
//   * later steps take top-k over the flattened [k, Vp] scores                              (:161-163)
//   * a beam that emits END is moved to the completed list and k shrinks                    (:171-176)
//   * the result is the completed beam with the highest score, first one on ties            (:191)
// Instead of copying sentences/alpha histories around every step (torch.cat in the reference) each step records
// (word, parent row) and the winner is recovered by backtracking.
#include "common.cuh"
#include "kernels.h"

namespace alexcap {

constexpr int BEAM_MAX = 32;
constexpr int BS_THREADS = 256;

__global__ void __launch_bounds__(BS_THREADS)
beam_select_kernel(int K, int Vp, int step, int end_tok, const float* __restrict__ logits, int64_t ldl,
                   float* __restrict__ top, int32_t* __restrict__ kact, int32_t* __restrict__ words,
                   int32_t* __restrict__ back, int32_t* __restrict__ rowrank_next, int32_t* __restrict__ src_row,
                   int32_t* __restrict__ prev_word, int32_t* __restrict__ ncomp, float* __restrict__ comp_score,
                   int32_t* __restrict__ comp_step, int32_t* __restrict__ comp_rank) {
  __shared__ float sv[BS_THREADS];
  __shared__ int si[BS_THREADS];
  __shared__ int sel_idx[BEAM_MAX];
  __shared__ float sel_val[BEAM_MAX];
  __shared__ float row_top[BEAM_MAX];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int k = kact[b];
  if (k == 0) {   // image finished: keep its (unused) rows where they are
    for (int j = tid; j < K; j += BS_THREADS) src_row[b * K + j] = j;
    return;
  }
  const int nrows = (step == 0) ? 1 : k;
  if (tid < K) row_top[tid] = top[b * K + tid];
  __syncthreads();
  const int ncand = nrows * Vp;
  for (int r = 0; r < k; ++r) {
    float bv = -INFINITY; int bi = 0x7fffffff;
    for (int i = tid; i < ncand; i += BS_THREADS) {
      bool taken = false;
      for (int q = 0; q < r; ++q) taken |= (sel_idx[q] == i);
      if (taken) continue;
      const int j = i / Vp, v = i - j * Vp;
      const float val = row_top[j] + logits[(int64_t)(b * K + j) * ldl + v];
      if (val > bv || (val == bv && i < bi)) { bv = val; bi = i; }
    }
    sv[tid] = bv; si[tid] = bi;
    __syncthreads();
    for (int s = BS_THREADS / 2; s > 0; s >>= 1) {
      if (tid < s) {
        const float v2 = sv[tid + s]; const int i2 = si[tid + s];
        if (v2 > sv[tid] || (v2 == sv[tid] && i2 < si[tid])) { sv[tid] = v2; si[tid] = i2; }
      }
      __syncthreads();
    }
    if (tid == 0) { sel_idx[r] = si[0]; sel_val[r] = sv[0]; }
    __syncthreads();
  }
  if (tid == 0) {
    int ninc = 0, nc = ncomp[b];
    for (int r = 0; r < k; ++r) {
      const int flat = sel_idx[r];
      const int prev = flat / Vp, word = flat - prev * Vp;
      words[b * K + r] = word;
      back[b * K + r] = prev;
      if (word == end_tok) {
        comp_score[b * K + nc] = sel_val[r];
        comp_step[b * K + nc] = step;
        comp_rank[b * K + nc] = r;
        ++nc;
      } else {
        src_row[b * K + ninc] = prev;
        top[b * K + ninc] = sel_val[r];
        prev_word[b * K + ninc] = word;
        rowrank_next[b * K + ninc] = r;
        ++ninc;
      }
    }
    for (int j = ninc; j < K; ++j) src_row[b * K + j] = j;
    kact[b] = ninc;
    ncomp[b] = nc;
  }
}

cudaError_t beam_select(cudaStream_t st, int B, int K, int Vp, int S, int step, int end_tok, const float* logits, int64_t ldl,
                        float* top, int32_t* kact, int32_t* words, int32_t* back, int32_t* rowrank, int32_t* src_row,
                        int32_t* prev_word, int32_t* ncomp, float* comp_score, int32_t* comp_step, int32_t* comp_rank) {
  (void)S;
  if (K > BEAM_MAX) return cudaErrorInvalidValue;
  ALEXCAP_NOTE_LAUNCH(); beam_select_kernel<<<B, BS_THREADS, 0, st>>>(K, Vp, step, end_tok, logits, ldl, top, kact, words, back, rowrank, src_row,
                                               prev_word, ncomp, comp_score, comp_step, comp_rank);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) beam_gather_kernel(int K, int H, const int32_t* __restrict__ src_row,
                                                          const float* __restrict__ h_in, int64_t h_in_ld,
                                                          const float* __restrict__ c_in, float* __restrict__ h_out,
                                                          int64_t h_out_ld, float* __restrict__ c_out) {
  const int row = blockIdx.x;                 // b*K + j
  const int b = row / K;
  const int src = b * K + src_row[row];
  for (int i = threadIdx.x; i < H; i += blockDim.x) {
    h_out[(int64_t)row * h_out_ld + i] = h_in[(int64_t)src * h_in_ld + i];
    c_out[(int64_t)row * H + i] = c_in[(int64_t)src * H + i];
  }
}
cudaError_t beam_gather_state(cudaStream_t st, int B, int K, int H, const int32_t* src_row, const float* h_in, int64_t h_in_ld,
                              const float* c_in, float* h_out, int64_t h_out_ld, float* c_out) {
  ALEXCAP_NOTE_LAUNCH(); beam_gather_kernel<<<B * K, 256, 0, st>>>(K, H, src_row, h_in, h_in_ld, c_in, h_out, h_out_ld, c_out);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(128) beam_finalize_kernel(int B, int K, int S, int L, const int32_t* __restrict__ words,
                                                            const int32_t* __restrict__ back,
                                                            const int32_t* __restrict__ rowrank,
                                                            const int32_t* __restrict__ ncomp,
                                                            const float* __restrict__ comp_score,
                                                            const int32_t* __restrict__ comp_step,
                                                            const int32_t* __restrict__ comp_rank,
                                                            const float* __restrict__ alpha_steps, int64_t* __restrict__ ids,
                                                            int32_t* __restrict__ lens, float* __restrict__ scores,
                                                            float* __restrict__ alphas) {
  __shared__ int path_row[256];   // parent row per step (S <= 256)
  const int b = blockIdx.x, R = B * K;
  const int nc = ncomp[b];
  for (int t = threadIdx.x; t < S; t += blockDim.x) ids[(int64_t)b * S + t] = 0;
  if (alphas) for (int i = threadIdx.x; i < S * L; i += blockDim.x) alphas[(int64_t)b * S * L + i] = 0.f;
  __syncthreads();
  if (nc == 0) {   // the reference raises ValueError here (:191)
    if (threadIdx.x == 0) { lens[b] = 0; scores[b] = 0.f; }
    return;
  }
  int best = 0;
  for (int i = 1; i < nc; ++i)
    if (comp_score[b * K + i] > comp_score[b * K + best]) best = i;
  const int s_end = comp_step[b * K + best];
  if (threadIdx.x == 0) {
    int r = comp_rank[b * K + best];
    for (int s = s_end; s >= 0; --s) {
      ids[(int64_t)b * S + s] = words[(int64_t)s * R + b * K + r];
      const int p = back[(int64_t)s * R + b * K + r];
      path_row[s] = p;
      if (s > 0) r = rowrank[(int64_t)s * R + b * K + p];
    }
    lens[b] = s_end + 1;
    scores[b] = comp_score[b * K + best];
  }
  __syncthreads();
  if (alphas) {
    for (int i = threadIdx.x; i < (s_end + 1) * L; i += blockDim.x) {
      const int s = i / L, l = i - s * L;
      alphas[((int64_t)b * S + s) * L + l] = alpha_steps[((int64_t)s * R + b * K + path_row[s]) * L + l];
    }
  }
}
cudaError_t beam_finalize(cudaStream_t st, int B, int K, int S, int L, const int32_t* words, const int32_t* back,
                          const int32_t* rowrank, const int32_t* ncomp, const float* comp_score, const int32_t* comp_step,
                          const int32_t* comp_rank, const float* alpha_steps, int64_t* ids, int32_t* lens, float* scores,
                          float* alphas) {
  if (S > 256) return cudaErrorInvalidValue;
  ALEXCAP_NOTE_LAUNCH(); beam_finalize_kernel<<<B, 128, 0, st>>>(B, K, S, L, words, back, rowrank, ncomp, comp_score, comp_step, comp_rank,
                                          alpha_steps, ids, lens, scores, alphas);
  return cudaGetLastError();
}

}  // namespace alexcap
