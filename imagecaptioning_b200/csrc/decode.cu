// decode.cu -- token selection kernels for greedy / beam decoding
// (AlexCap/AttentionLanguageModule.py:108,158-163 ; AlexCap/LanguageModule.py:141-142).
#include "common.cuh"
#include "kernels.h"

namespace alexcap {

// argmax over each row; ties -> lowest index (torch.argmax returns the first maximal element).
__global__ void __launch_bounds__(256) argmax_kernel(const float* __restrict__ logits, int C, int64_t ld,
                                                     int32_t* __restrict__ out) {
  __shared__ float sv[256];
  __shared__ int si[256];
  const float* x = logits + (int64_t)blockIdx.x * ld;
  float bv = -INFINITY; int bi = 0x7fffffff;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const float v = x[c];
    if (v > bv || (v == bv && c < bi)) { bv = v; bi = c; }
  }
  sv[threadIdx.x] = bv; si[threadIdx.x] = bi;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      const float v = sv[threadIdx.x + s]; const int i = si[threadIdx.x + s];
      if (v > sv[threadIdx.x] || (v == sv[threadIdx.x] && i < si[threadIdx.x])) { sv[threadIdx.x] = v; si[threadIdx.x] = i; }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) out[blockIdx.x] = si[0];
}
cudaError_t argmax_rows(cudaStream_t st, const float* logits, int B, int C, int64_t ld, int32_t* out_tok) {
  if (B <= 0) return cudaSuccess;
  ALEXCAP_NOTE_LAUNCH(); argmax_kernel<<<B, 256, 0, st>>>(logits, C, ld, out_tok);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) fill_i32_kernel(int32_t* __restrict__ p, int64_t n, int32_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
cudaError_t fill_i32(cudaStream_t st, int32_t* p, int64_t n, int32_t v) {
  if (n <= 0) return cudaSuccess;
  const int blocks = (int)(((n + 255) / 256 < 1024) ? (n + 255) / 256 : 1024);
  ALEXCAP_NOTE_LAUNCH(); fill_i32_kernel<<<blocks, 256, 0, st>>>(p, n, v);
  return cudaGetLastError();
}

// The reference stops its Python loop when every row emitted END (one host sync per step).  Here the loop always
// runs S steps on the device and a `done` flag freezes the outputs after that point instead.
__global__ void __launch_bounds__(256) commit_alpha_kernel(int B, int S, int L, int t, const float* __restrict__ alpha_step,
                                                           float* __restrict__ alphas, const int32_t* __restrict__ done) {
  if (*done) return;
  const int64_t n = (int64_t)B * L;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / L, l = i % L;
    alphas[(b * S + t) * L + l] = alpha_step[i];
  }
}
__global__ void __launch_bounds__(1024) commit_ids_kernel(int B, int S, int t, int end_tok, const int32_t* __restrict__ tok,
                                                          int64_t* __restrict__ ids, int32_t* done) {
  if (*done) return;   // uniform across the (single) block
  int all_end = 1;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const int32_t v = tok[b];
    ids[(int64_t)b * S + t] = v;
    all_end &= (v == end_tok);
  }
  all_end = __syncthreads_and(all_end);
  if (threadIdx.x == 0 && all_end) *done = 1;
}
cudaError_t greedy_commit(cudaStream_t st, int B, int S, int L, int t, int end_tok, const int32_t* tok,
                          const float* alpha_step, int64_t* ids, float* alphas, int32_t* done) {
  if (alphas && alpha_step) {
    const int64_t n = (int64_t)B * L;
    const int blocks = (int)(((n + 255) / 256 < 2048) ? (n + 255) / 256 : 2048);
    ALEXCAP_NOTE_LAUNCH(); commit_alpha_kernel<<<blocks, 256, 0, st>>>(B, S, L, t, alpha_step, alphas, done);
    ALEXCAP_LAUNCH_CHECK();
  }
  ALEXCAP_NOTE_LAUNCH(); commit_ids_kernel<<<1, 1024, 0, st>>>(B, S, t, end_tok, tok, ids, done);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) permute_01_kernel(int nb, int na, int w, const float* __restrict__ src,
                                                         float* __restrict__ dst) {
  const int64_t n = (int64_t)nb * na * w;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % w);
    const int64_t r = i / w;
    const int a = (int)(r % na), b = (int)(r / na);          // src index (b, a, c)
    dst[((int64_t)a * nb + b) * w + c] = src[i];
  }
}
cudaError_t permute_01(cudaStream_t st, int nb, int na, int w, const float* src, float* dst) {
  const int64_t n = (int64_t)nb * na * w;
  if (n <= 0) return cudaSuccess;
  const int blocks = (int)(((n + 255) / 256 < 8192) ? (n + 255) / 256 : 8192);
  ALEXCAP_NOTE_LAUNCH(); permute_01_kernel<<<blocks, 256, 0, st>>>(nb, na, w, src, dst);
  return cudaGetLastError();
}

}  // namespace alexcap
