// ce_loss.cu -- CustomCrossEntropyLoss (AlexCap/CustomLoss.py:7-14):
//   nn.CrossEntropyLoss(reduction='mean', ignore_index=0, label_smoothing=0.1) over [N, C] logits,
// value and gradient in one pass over each row:
//   loss_row = (1-eps) * (lse - x[y]) + eps * (lse - mean_c x[c]),   only rows with y != ignore_index
//   dlogits  = (softmax(x) - (1-eps) * onehot(y) - eps/C) * [y != ignore] / denom
#include "common.cuh"
#include "kernels.h"

namespace alexcap {

__global__ void __launch_bounds__(256) count_valid_kernel(const int32_t* __restrict__ tgt, int N, int ignore_index,
                                                          float* __restrict__ out) {
  __shared__ float scratch[40];
  float c = 0.f;
  for (int i = threadIdx.x; i < N; i += blockDim.x) c += (tgt[i] != ignore_index) ? 1.f : 0.f;
  c = block_sum(c, scratch);
  if (threadIdx.x == 0) out[0] = c;
}
cudaError_t count_valid(cudaStream_t st, const int32_t* tgt, int N, int ignore_index, float* n_valid_out) {
  ALEXCAP_NOTE_LAUNCH(); count_valid_kernel<<<1, 256, 0, st>>>(tgt, N, ignore_index, n_valid_out);
  return cudaGetLastError();
}

constexpr int CE_THREADS = 512;
__global__ void __launch_bounds__(CE_THREADS) ce_rows_kernel(const float* __restrict__ logits,
                                                             const int32_t* __restrict__ tgt, int C, int64_t ld,
                                                             float eps, int ignore_index,
                                                             const float* __restrict__ denom,
                                                             float* __restrict__ row_loss, float* dlogits,
                                                             uint16_t* __restrict__ dl_hi, uint16_t* __restrict__ dl_lo,
                                                             int64_t dl_ld) {
  __shared__ float scratch[40];
  const int row = blockIdx.x;
  const float* x = logits + (int64_t)row * ld;
  float* dx = dlogits ? dlogits + (int64_t)row * ld : nullptr;
  const int y = tgt[row];
  uint32_t* hi2 = dl_hi ? reinterpret_cast<uint32_t*>(dl_hi + (int64_t)row * dl_ld) : nullptr;   // bf16 pairs (dl_ld even)
  uint32_t* lo2 = dl_lo ? reinterpret_cast<uint32_t*>(dl_lo + (int64_t)row * dl_ld) : nullptr;
  if (y == ignore_index) {   // ignored row: no loss, zero gradient
    if (threadIdx.x == 0) row_loss[row] = 0.f;
    if (dx) for (int c = threadIdx.x; c < C; c += CE_THREADS) dx[c] = 0.f;
    if (hi2) for (int c = threadIdx.x; 2 * c < C; c += CE_THREADS) { hi2[c] = 0u; if (lo2) lo2[c] = 0u; }
    return;
  }
  const float xy = x[y];   // read before any (possibly aliasing) gradient write
  float mx = -INFINITY, sx = 0.f;
  for (int c = threadIdx.x; c < C; c += CE_THREADS) {
    const float v = x[c];
    mx = fmaxf(mx, v);
    sx += v;
  }
  mx = block_max(mx, scratch);
  sx = block_sum(sx, scratch);
  float se = 0.f;
  for (int c = threadIdx.x; c < C; c += CE_THREADS) se += expf(x[c] - mx);
  se = block_sum(se, scratch);
  const float lse = mx + logf(se);
  const float dn = denom[0];
  const float inv_d = dn > 0.f ? 1.f / dn : 0.f;     // (a data-parallel divisor N_valid_global / world may be < 1)
  if (threadIdx.x == 0) {
    const float nll = lse - xy;
    const float smooth = lse - sx / (float)C;
    row_loss[row] = ((1.f - eps) * nll + eps * smooth) * inv_d;
  }
  const float u = eps / (float)C;
  if (hi2) {                 // two columns per thread: packed bf16 pairs, 4-byte stores
    for (int c = 2 * threadIdx.x; c < C; c += 2 * CE_THREADS) {
      float g0 = expf(x[c] - lse) - u, g1 = (c + 1 < C) ? expf(x[c + 1] - lse) - u : 0.f;
      if (c == y) g0 -= (1.f - eps);
      if (c + 1 == y) g1 -= (1.f - eps);
      g0 *= inv_d; g1 *= inv_d;
      if (c + 1 >= C) g1 = 0.f;
      uint32_t h;
      asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(g1), "f"(g0));
      hi2[c >> 1] = h;
      if (lo2) {
        uint32_t l;
        const float r0 = g0 - __uint_as_float(h << 16), r1 = g1 - __uint_as_float(h & 0xffff0000u);
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(l) : "f"(r1), "f"(r0));
        lo2[c >> 1] = l;
      }
      if (dx) { dx[c] = g0; if (c + 1 < C) dx[c + 1] = g1; }
    }
  } else if (dx) {
    for (int c = threadIdx.x; c < C; c += CE_THREADS) {
      float g = expf(x[c] - lse) - u;
      if (c == y) g -= (1.f - eps);
      dx[c] = g * inv_d;
    }
  }
}

cudaError_t ce_rows(cudaStream_t st, const float* logits, const int32_t* tgt, int N, int C, int64_t ld, float eps,
                    int ignore_index, const float* denom, float* row_loss, float* dlogits, uint16_t* dl_hi, uint16_t* dl_lo,
                    int64_t dl_ld) {
  if (N <= 0) return cudaSuccess;
  if (dl_hi && (dl_ld & 1)) return cudaErrorInvalidValue;
  ALEXCAP_NOTE_LAUNCH();
  ce_rows_kernel<<<N, CE_THREADS, 0, st>>>(logits, tgt, C, ld, eps, ignore_index, denom, row_loss, dlogits, dl_hi, dl_lo, dl_ld);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(1024) reduce_sum_kernel(const float* __restrict__ x, int n, float* __restrict__ out) {
  __shared__ float scratch[40];
  float acc = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) acc += x[i];
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) out[0] = acc;
}
// stage 1 of the large-n path: one partial per block (deterministic: fixed block -> element mapping)
__global__ void __launch_bounds__(256) reduce_partial_kernel(const float* __restrict__ x, int n, float* __restrict__ part) {
  __shared__ float scratch[40];
  float acc = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) acc += x[i];
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) part[blockIdx.x] = acc;
}
cudaError_t reduce_sum(cudaStream_t st, const float* x, int n, float* out, float* scratch) {
  if (scratch && n > 16384) {   // scratch: >= 256 floats
    ALEXCAP_NOTE_LAUNCH(); reduce_partial_kernel<<<256, 256, 0, st>>>(x, n, scratch);
    ALEXCAP_LAUNCH_CHECK();
    ALEXCAP_NOTE_LAUNCH(); reduce_sum_kernel<<<1, 1024, 0, st>>>(scratch, 256, out);
    return cudaGetLastError();
  }
  ALEXCAP_NOTE_LAUNCH(); reduce_sum_kernel<<<1, 1024, 0, st>>>(x, n, out);
  return cudaGetLastError();
}

}  // namespace alexcap
