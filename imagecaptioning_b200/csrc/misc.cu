// misc.cu -- the small kernels around the hot path: get_target, embedding gather / scatter-add, reductions,
// dropout mask, elementwise backward helpers.
#include "common.cuh"
#include "kernels.h"

namespace alexcap {

// get_target (AlexCap/AttentionLanguageModule.py:197-212, AlexCap/LanguageModule.py:113-128): both variants, written
// batch-major [B,S] (the reference layout) and time-major [S,B] (what the per-step kernels index).
// The reference runs a Python double loop with one device sync per element; this is one thread per caption.
__global__ void make_target_kernel(const int32_t* __restrict__ gt, int B, int T, int start_tok, int end_tok, int n_classes,
                                   int32_t* in_bm, int32_t* tgt_bm, int32_t* in_tm, int32_t* tgt_tm, int32_t* bad_rows) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int S = T + 1;
  bool placed = false;
  int bad = 0;
  for (int t = 0; t < S; ++t) {
    int in_v = (t == 0) ? start_tok : gt[(int64_t)b * T + t - 1];
    int tg = (t < T) ? gt[(int64_t)b * T + t] : 0;
    // A token outside [0, n_classes) would index the embedding / logits out of bounds (torch raises a device-side assert):
    // it is replaced by 0 here so that no kernel reads or writes out of range, and the row is reported in bad_rows, which
    // makes finalize_loss return NaN.
    if (n_classes > 0) {
      if ((unsigned)in_v >= (unsigned)n_classes) { in_v = 0; bad = 1; }
      if ((unsigned)tg >= (unsigned)n_classes) { tg = 0; bad = 1; }
    }
    if (t >= 1 && !placed && tg == 0) { tg = end_tok; placed = true; }   // first 0 at index >= 1 becomes END
    if (in_bm) in_bm[(int64_t)b * S + t] = in_v;
    if (tgt_bm) tgt_bm[(int64_t)b * S + t] = tg;
    if (in_tm) in_tm[(int64_t)t * B + b] = in_v;
    if (tgt_tm) tgt_tm[(int64_t)t * B + b] = tg;
  }
  if (bad_rows) bad_rows[b] = bad;
}
cudaError_t make_target(cudaStream_t st, const int32_t* gt, int B, int T, int start_tok, int end_tok,
                        int32_t* in_bm, int32_t* tgt_bm, int32_t* in_tm, int32_t* tgt_tm, int n_classes, int32_t* bad_rows) {
  if (B <= 0) return cudaSuccess;
  ALEXCAP_NOTE_LAUNCH(); make_target_kernel<<<ceil_div(B, 128), 128, 0, st>>>(gt, B, T, start_tok, end_tok, n_classes, in_bm, tgt_bm, in_tm, tgt_tm, bad_rows);
  return cudaGetLastError();
}

// N_valid of get_target(gt, make_target=True) straight from the captions: every row keeps its non-zero tokens and gets
// exactly one END (the first 0 at index >= 1 of [gt..., 0]), so N_valid = nnz(gt) + B.
__global__ void __launch_bounds__(256) count_valid_targets_kernel(const int32_t* __restrict__ gt, int B, int T, float* __restrict__ out) {
  __shared__ float scratch[40];
  float c = 0.f;
  for (int64_t i = threadIdx.x; i < (int64_t)B * T; i += blockDim.x) c += (gt[i] != 0) ? 1.f : 0.f;
  c = block_sum(c, scratch);
  if (threadIdx.x == 0) out[0] = c + (float)B;
}
cudaError_t count_valid_targets(cudaStream_t st, const int32_t* gt, int B, int T, float* out) {
  ALEXCAP_NOTE_LAUNCH(); count_valid_targets_kernel<<<1, 256, 0, st>>>(gt, B, T, out);
  return cudaGetLastError();
}

// mean over regions (AttentionLanguageModule.py:117): one CTA per image, D/4 float4 columns x row groups
constexpr int MEAN_THREADS = 1024;     // one CTA per image: 32 warps x 8 loads in flight keep the SM's HBM share busy
__global__ void __launch_bounds__(MEAN_THREADS) mean_L_kernel(const float* __restrict__ feat, int L, int D, float* __restrict__ m) {
  __shared__ float4 red[MEAN_THREADS];
  const int b = blockIdx.x, D4 = D >> 2, tid = threadIdx.x;
  const float4* fb = reinterpret_cast<const float4*>(feat) + (int64_t)b * L * D4;
  const bool wide = D4 >= MEAN_THREADS;
  const int G = wide ? 1 : MEAN_THREADS / D4;
  const float inv = 1.f / (float)L;
  for (int c0 = 0; c0 < D4; c0 += MEAN_THREADS) {
    const int c = wide ? c0 + tid : tid % D4;
    const int g = wide ? 0 : tid / D4;
    float4 acc = f4_zero();
    if (c < D4 && g < G) {
#pragma unroll 8
      for (int l = g; l < L; l += G) f4_add(acc, ldg4(fb + (int64_t)l * D4 + c));
    }
    red[tid] = acc;
    __syncthreads();
    if (g == 0 && c < D4) {
      for (int gg = 1; gg < G; ++gg) f4_add(acc, red[gg * D4 + c]);
      reinterpret_cast<float4*>(m)[(int64_t)b * D4 + c] = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
    }
    __syncthreads();
  }
}
cudaError_t mean_over_L(cudaStream_t st, const float* feat, int B, int L, int D, float* m) {
  if (D & 3) return cudaErrorInvalidValue;
  ALEXCAP_NOTE_LAUNCH(); mean_L_kernel<<<B, MEAN_THREADS, 0, st>>>(feat, L, D, m);
  return cudaGetLastError();
}

// embedding lookup (AttentionLanguageModule.py:69, LanguageModule.py:102)
__global__ void __launch_bounds__(128) gather_rows_kernel(const float* __restrict__ table, const int32_t* __restrict__ idx,
                                                          int width4, float* __restrict__ out, int64_t out_ld) {
  const int r = blockIdx.x;
  const float4* src = reinterpret_cast<const float4*>(table) + (int64_t)idx[r] * width4;
  float4* dst = reinterpret_cast<float4*>(out + (int64_t)r * out_ld);
  for (int c = threadIdx.x; c < width4; c += blockDim.x) dst[c] = ldg4(src + c);
}
cudaError_t gather_rows(cudaStream_t st, const float* table, const int32_t* idx, int rows, int width, float* out,
                        int64_t out_ld) {
  if ((width & 3) || (out_ld & 3)) return cudaErrorInvalidValue;
  if (rows <= 0) return cudaSuccess;
  ALEXCAP_NOTE_LAUNCH(); gather_rows_kernel<<<rows, 128, 0, st>>>(table, idx, width / 4, out, out_ld);
  return cudaGetLastError();
}

// embedding backward: index_put_(accumulate=True)
__global__ void __launch_bounds__(128) scatter_add_kernel(const float* __restrict__ src, int64_t src_ld,
                                                          const int32_t* __restrict__ idx, int width,
                                                          float* __restrict__ table) {
  const int r = blockIdx.x;
  float* dst = table + (int64_t)idx[r] * width;
  const float* s = src + (int64_t)r * src_ld;
  for (int c = threadIdx.x; c < width; c += blockDim.x) atomicAdd(dst + c, s[c]);
}
cudaError_t scatter_add_rows(cudaStream_t st, const float* src, int64_t src_ld, const int32_t* idx, int rows,
                             int width, float* table) {
  if (rows <= 0) return cudaSuccess;
  ALEXCAP_NOTE_LAUNCH(); scatter_add_kernel<<<rows, 128, 0, st>>>(src, src_ld, idx, width, table);
  return cudaGetLastError();
}

// column sums (bias gradients): two deterministic stages through scratch[nsplit][N]
constexpr int CS_SPLIT = 64;
__global__ void __launch_bounds__(256) col_sum_part_kernel(const float* __restrict__ X, int M, int N, int64_t ld,
                                                           float* __restrict__ part) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int s = blockIdx.y;
  const int rows_per = (M + CS_SPLIT - 1) / CS_SPLIT;
  const int m0 = s * rows_per, m1 = min(M, m0 + rows_per);
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;     // four independent chains: loads of several rows in flight
  int m = m0;
  for (; m + 3 < m1; m += 4) {
    a0 += __ldg(X + (int64_t)m * ld + n); a1 += __ldg(X + (int64_t)(m + 1) * ld + n);
    a2 += __ldg(X + (int64_t)(m + 2) * ld + n); a3 += __ldg(X + (int64_t)(m + 3) * ld + n);
  }
  for (; m < m1; ++m) a0 += __ldg(X + (int64_t)m * ld + n);
  part[(int64_t)s * N + n] = (a0 + a1) + (a2 + a3);
}
__global__ void __launch_bounds__(256) col_sum_final_kernel(const float* __restrict__ part, int N, float* __restrict__ out,
                                                            int accumulate) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float acc = 0.f;
  for (int s = 0; s < CS_SPLIT; ++s) acc += part[(int64_t)s * N + n];
  out[n] = accumulate ? out[n] + acc : acc;
}
cudaError_t col_sum(cudaStream_t st, const float* X, int M, int N, int64_t ld, float* out, bool accumulate,
                    float* scratch) {
  if (N <= 0) return cudaSuccess;
  dim3 grid(ceil_div(N, 256), CS_SPLIT);
  ALEXCAP_NOTE_LAUNCH(); col_sum_part_kernel<<<grid, 256, 0, st>>>(X, M, N, ld, scratch);
  ALEXCAP_LAUNCH_CHECK();
  ALEXCAP_NOTE_LAUNCH(); col_sum_final_kernel<<<ceil_div(N, 256), 256, 0, st>>>(scratch, N, out, accumulate ? 1 : 0);
  return cudaGetLastError();
}

// column sums of ONE matrix whose column ranges [0,n0) [n0,n0+n1) [n0+n1,n0+n1+n2) are different bias gradients
__global__ void __launch_bounds__(256) col_sum_final3_kernel(const float* __restrict__ part, int N, float* __restrict__ o0, int n0,
                                                             float* __restrict__ o1, int n1, float* __restrict__ o2) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float acc = 0.f;
  for (int s = 0; s < CS_SPLIT; ++s) acc += part[(int64_t)s * N + n];
  if (n < n0) o0[n] = acc;
  else if (n < n0 + n1) o1[n - n0] = acc;
  else o2[n - n0 - n1] = acc;
}
cudaError_t col_sum3(cudaStream_t st, const float* X, int M, int64_t ld, float* o0, int n0, float* o1, int n1, float* o2, int n2,
                     float* scratch) {
  const int N = n0 + n1 + n2;
  if (N <= 0) return cudaSuccess;
  dim3 grid(ceil_div(N, 256), CS_SPLIT);
  ALEXCAP_NOTE_LAUNCH(); col_sum_part_kernel<<<grid, 256, 0, st>>>(X, M, N, ld, scratch);
  ALEXCAP_LAUNCH_CHECK();
  ALEXCAP_NOTE_LAUNCH(); col_sum_final3_kernel<<<ceil_div(N, 256), 256, 0, st>>>(scratch, N, o0, n0, o1, n1, o2);
  return cudaGetLastError();
}

// the same over bf16 (hi, lo) pair copies: thread = two adjacent columns (one 4-byte load per copy and row)
__global__ void __launch_bounds__(256) col_sum_bf16_part_kernel(const uint16_t* __restrict__ hi, const uint16_t* __restrict__ lo,
                                                                int M, int N, int64_t ld, float* __restrict__ part) {
  const int n = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
  if (n >= N) return;
  const int s = blockIdx.y;
  const int rows_per = (M + CS_SPLIT - 1) / CS_SPLIT;
  const int m0 = s * rows_per, m1 = min(M, m0 + rows_per);
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  for (int m = m0; m < m1; ++m) {
    const uint32_t h = __ldg(reinterpret_cast<const uint32_t*>(hi + (int64_t)m * ld + n));
    a0 += __uint_as_float(h << 16); a1 += __uint_as_float(h & 0xffff0000u);
    if (lo) {
      const uint32_t l = __ldg(reinterpret_cast<const uint32_t*>(lo + (int64_t)m * ld + n));
      b0 += __uint_as_float(l << 16); b1 += __uint_as_float(l & 0xffff0000u);
    }
  }
  part[(int64_t)s * N + n] = a0 + b0;
  if (n + 1 < N) part[(int64_t)s * N + n + 1] = a1 + b1;
}
cudaError_t col_sum_bf16(cudaStream_t st, const uint16_t* hi, const uint16_t* lo, int M, int N, int64_t ld, float* out,
                         bool accumulate, float* scratch) {
  if (N <= 0) return cudaSuccess;
  if (ld & 1) return cudaErrorInvalidValue;
  dim3 grid(ceil_div(ceil_div(N, 2), 256), CS_SPLIT);
  ALEXCAP_NOTE_LAUNCH(); col_sum_bf16_part_kernel<<<grid, 256, 0, st>>>(hi, lo, M, N, ld, scratch);
  ALEXCAP_LAUNCH_CHECK();
  ALEXCAP_NOTE_LAUNCH(); col_sum_final_kernel<<<ceil_div(N, 256), 256, 0, st>>>(scratch, N, out, accumulate ? 1 : 0);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) tanh_bwd_kernel(int rows, int cols, const float* __restrict__ dy, int64_t dy_ld,
                                                       const float* __restrict__ t, int64_t t_ld, float* __restrict__ out,
                                                       int64_t out_ld) {
  const int64_t n = (int64_t)rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i % cols);
    const float tv = t[(int64_t)r * t_ld + c];
    out[(int64_t)r * out_ld + c] = dy[(int64_t)r * dy_ld + c] * (1.f - tv * tv);
  }
}
cudaError_t tanh_bwd(cudaStream_t st, int rows, int cols, const float* dy, int64_t dy_ld, const float* t, int64_t t_ld,
                     float* out, int64_t out_ld) {
  const int64_t n = (int64_t)rows * cols;
  if (n <= 0) return cudaSuccess;
  ALEXCAP_NOTE_LAUNCH(); tanh_bwd_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(rows, cols, dy, dy_ld, t, t_ld, out, out_ld);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) relu_bwd_kernel(int64_t n, const float* __restrict__ dy,
                                                       const float* __restrict__ y, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = (y[i] > 0.f) ? dy[i] : 0.f;
}
cudaError_t relu_bwd(cudaStream_t st, int64_t n, const float* dy, const float* y, float* out) {
  if (n <= 0) return cudaSuccess;
  const int blocks = (int)(((n + 255) / 256 < 4096) ? (n + 255) / 256 : 4096);
  ALEXCAP_NOTE_LAUNCH(); relu_bwd_kernel<<<blocks, 256, 0, st>>>(n, dy, y, out);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) copy_2d_kernel(int rows, int cols, const float* __restrict__ src, int64_t src_ld,
                                                      float* __restrict__ dst, int64_t dst_ld) {
  const int64_t n = (int64_t)rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i % cols);
    dst[(int64_t)r * dst_ld + c] = src[(int64_t)r * src_ld + c];
  }
}
cudaError_t copy_2d(cudaStream_t st, int rows, int cols, const float* src, int64_t src_ld, float* dst, int64_t dst_ld) {
  const int64_t n = (int64_t)rows * cols;
  if (n <= 0) return cudaSuccess;
  const int blocks = (int)(((n + 255) / 256 < 8192) ? (n + 255) / 256 : 8192);
  ALEXCAP_NOTE_LAUNCH(); copy_2d_kernel<<<blocks, 256, 0, st>>>(rows, cols, src, src_ld, dst, dst_ld);
  return cudaGetLastError();
}

// several strided 2-D copies in ONE launch (the per-step weight concatenations: 7 memcpy / copy nodes -> 1)
__global__ void __launch_bounds__(256) copy_segments_kernel(const CopySegments t) {
  const int sgm = blockIdx.y;
  const float* __restrict__ src = t.src[sgm];
  float* __restrict__ dst = t.dst[sgm];
  const int64_t sld = t.src_ld[sgm], dld = t.dst_ld[sgm];
  const int cols4 = t.cols[sgm] >> 2;
  const int64_t n = (int64_t)t.rows[sgm] * cols4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols4), c = (int)(i % cols4) * 4;
    *reinterpret_cast<float4*>(dst + (int64_t)r * dld + c) = __ldg(reinterpret_cast<const float4*>(src + (int64_t)r * sld + c));
  }
}
cudaError_t copy_segments(cudaStream_t st, const CopySegments& t) {
  if (t.n <= 0) return cudaSuccess;
  if (t.n > CopySegments::MAX) return cudaErrorInvalidValue;
  int64_t most = 0;
  for (int i = 0; i < t.n; ++i) {
    if ((t.cols[i] & 3) || (t.src_ld[i] & 3) || (t.dst_ld[i] & 3) || !aligned16(t.src[i]) || !aligned16(t.dst[i])) return cudaErrorInvalidValue;
    const int64_t n = (int64_t)t.rows[i] * (t.cols[i] >> 2);
    if (n > most) most = n;
  }
  int bx = (int)((most + 255) / 256);
  if (bx > 592) bx = 592;
  if (bx < 1) bx = 1;
  ALEXCAP_NOTE_LAUNCH(); copy_segments_kernel<<<dim3(bx, t.n), 256, 0, st>>>(t);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) fill_kernel(float* __restrict__ p, int64_t n, float v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
cudaError_t fill_f32(cudaStream_t st, float* p, int64_t n, float v) {
  if (n <= 0) return cudaSuccess;
  if (v == 0.f) return cudaMemsetAsync(p, 0, n * sizeof(float), st);
  const int blocks = (int)(((n + 255) / 256 < 8192) ? (n + 255) / 256 : 8192);
  ALEXCAP_NOTE_LAUNCH(); fill_kernel<<<blocks, 256, 0, st>>>(p, n, v);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) add_bcast_L_kernel(int L, int D, const float* __restrict__ dm, float scale,
                                                          float* __restrict__ dfeat, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int d = (int)(i % D);
    const int64_t b = i / ((int64_t)L * D);
    dfeat[i] += dm[b * D + d] * scale;
  }
}
cudaError_t add_broadcast_L(cudaStream_t st, int B, int L, int D, const float* dm, float scale, float* dfeat) {
  const int64_t n = (int64_t)B * L * D;
  if (n <= 0) return cudaSuccess;
  const int blocks = (int)(((n + 255) / 256 < 8192) ? (n + 255) / 256 : 8192);
  ALEXCAP_NOTE_LAUNCH(); add_bcast_L_kernel<<<blocks, 256, 0, st>>>(L, D, dm, scale, dfeat, n);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) mul_2d_kernel(int rows, int cols, const float* __restrict__ a, int64_t a_ld,
                                                     const float* __restrict__ b, int64_t b_ld, float* __restrict__ out,
                                                     int64_t out_ld) {
  const int64_t n = (int64_t)rows * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i % cols);
    out[(int64_t)r * out_ld + c] = a[(int64_t)r * a_ld + c] * b[(int64_t)r * b_ld + c];
  }
}
cudaError_t mul_2d(cudaStream_t st, int rows, int cols, const float* a, int64_t a_ld, const float* b, int64_t b_ld,
                   float* out, int64_t out_ld) {
  const int64_t n = (int64_t)rows * cols;
  if (n <= 0) return cudaSuccess;
  const int blocks = (int)(((n + 255) / 256 < 8192) ? (n + 255) / 256 : 8192);
  ALEXCAP_NOTE_LAUNCH(); mul_2d_kernel<<<blocks, 256, 0, st>>>(rows, cols, a, a_ld, b, b_ld, out, out_ld);
  return cudaGetLastError();
}

// Philox4x32-10 counter-based generator (Salmon et al. 2011) for the dropout mask of AttentionLanguageModule.py:79.
// Bit-exact agreement with torch's own Philox stream is not a goal; tests inject explicit masks.
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0; key.y += W1;
  }
  return ctr;
}
// state != NULL: {seed, offset} are read from the device (graph replays draw fresh masks; philox_advance bumps offset)
__global__ void __launch_bounds__(256) dropout_mask_kernel(float* __restrict__ mask, int64_t n, float p, float scale,
                                                           uint64_t seed, uint64_t offset, const uint64_t* __restrict__ state,
                                                           const float* __restrict__ x, int64_t x_ld, int cols,
                                                           float* __restrict__ y) {
  if (state) { seed = state[0]; offset = state[1]; }
  const int64_t n4 = (n + 3) / 4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const uint64_t c = offset + (uint64_t)i;
    const uint4 r = philox4x32_10(make_uint4((uint32_t)c, (uint32_t)(c >> 32), 0u, 0u),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t e = i * 4 + j;
      if (e < n) {
        const float u = (float)(rr[j] >> 8) * (1.0f / 16777216.0f);   // [0,1)
        const float mk = (u >= p) ? scale : 0.f;
        mask[e] = mk;
        if (y) y[e] = mk * x[(e / cols) * x_ld + (e % cols)];           // fused dropout(h): y = mask * x
      }
    }
  }
}
__global__ void philox_advance_kernel(uint64_t* state, uint64_t by) { state[1] += by; }
cudaError_t dropout_mask(cudaStream_t st, float* mask, int64_t n, float p, uint64_t seed, uint64_t offset, uint64_t* state,
                         const float* x, int64_t x_ld, int cols, float* y) {
  if (n <= 0) return cudaSuccess;
  const int64_t n4 = (n + 3) / 4;
  const int blocks = (int)(((n4 + 255) / 256 < 8192) ? (n4 + 255) / 256 : 8192);
  ALEXCAP_NOTE_LAUNCH(); dropout_mask_kernel<<<blocks, 256, 0, st>>>(mask, n, p, 1.f / (1.f - p), seed, offset, state, x, x_ld, cols, y);
  ACK(cudaGetLastError());
  if (state) {            // same stream: ordered after every reader of the old counter
    ALEXCAP_NOTE_LAUNCH(); philox_advance_kernel<<<1, 1, 0, st>>>(state, (uint64_t)1 << 32);
  }
  return cudaGetLastError();
}

// loss_out = {ce + reg, ce, reg, n_valid}.  ce is NaN when no target is valid (the reference's mean over zero rows,
// CustomLoss.py:7) or when make_target flagged an out-of-range token (bad_rows[0..n_bad)).
__global__ void finalize_loss_kernel(const float* ce_sum, const float* reg, const float* n_valid, float* out,
                                     const int32_t* bad_rows, int n_bad) {
  float a = ce_sum ? ce_sum[0] : 0.f;
  const float b = reg ? reg[0] : 0.f;
  const float nv = n_valid ? n_valid[0] : 0.f;
  bool bad = n_valid && nv == 0.f;
  for (int i = 0; bad_rows && i < n_bad; ++i) bad |= bad_rows[i] != 0;
  if (bad) a = __int_as_float(0x7fc00000);
  out[0] = a + b; out[1] = a; out[2] = b; out[3] = nv;
}
cudaError_t finalize_loss(cudaStream_t st, const float* ce_sum, const float* reg, const float* n_valid, float* loss_out,
                          const int32_t* bad_rows, int n_bad) {
  ALEXCAP_NOTE_LAUNCH(); finalize_loss_kernel<<<1, 1, 0, st>>>(ce_sum, reg, n_valid, loss_out, bad_rows, n_bad);
  return cudaGetLastError();
}

}  // namespace alexcap
