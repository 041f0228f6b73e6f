// layout.cu -- the CNN hand-off: the backbone's output arrives NCHW [B, D, h, w]; the
// reference turns it into the decoder's [B, L = h*w, D] grid with `permute(0, 2, 3, 1).view(B, -1, C)`
// (AlexCap/LSTMwAttentionModel.py:64-66, forward_test :79-80), i.e. a per-image [D, L] -> [L, D] transposition that
// PyTorch materialises as a strided-copy kernel when the view is made contiguous.  Here the transposition writes straight
// into the training step's static feature buffer (one pass instead of permute-copy + staging copy), and the gradient
// w.r.t. the grid goes back to NCHW for the fine-tune phase (train_LSTMwAttention.py:51-55, :102-106) the same way.
// Both are pure HBM streams: B*L*D*4 bytes read + written, 32x32 tiles through shared memory, 128-byte rows both ways.
#include "api_util.h"

using namespace alexcap;

namespace {

// per image: src [R, C] (row pitch C) -> dst [C, R] (row pitch R);  dst = src^T (+ dst if ACC)
template <bool ACC>
__global__ void __launch_bounds__(256) transpose_batched_kernel(const float* __restrict__ src, float* __restrict__ dst, int R, int C) {
  __shared__ float tile[32][33];
  const int64_t img = (int64_t)blockIdx.z * R * C;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int r = r0 + ty + j, c = c0 + tx;
    tile[ty + j][tx] = (r < R && c < C) ? __ldg(src + img + (int64_t)r * C + c) : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const int c = c0 + ty + j, r = r0 + tx;
    if (c < C && r < R) {
      float* o = dst + img + (int64_t)c * R + r;
      const float v = tile[tx][ty + j];
      *o = ACC ? *o + v : v;
    }
  }
}

cudaError_t transpose_batched(cudaStream_t st, int B, int R, int C, const float* src, float* dst, bool acc) {
  if (B <= 0 || R <= 0 || C <= 0) return cudaSuccess;
  if (B > 65535) return cudaErrorInvalidValue;
  dim3 grid((C + 31) / 32, (R + 31) / 32, B);
  ALEXCAP_NOTE_LAUNCH();
  if (acc) transpose_batched_kernel<true><<<grid, 256, 0, st>>>(src, dst, R, C);
  else transpose_batched_kernel<false><<<grid, 256, 0, st>>>(src, dst, R, C);
  return cudaGetLastError();
}

}  // namespace

extern "C" {

int alexcap_features_from_nchw(int32_t B, int32_t D, int32_t L, const float* nchw, float* feat, void* stream) {
  if (B <= 0 || D <= 0 || L <= 0 || !nchw || !feat) return fail(ALEXCAP_E_BADARG, "alexcap_features_from_nchw: bad arguments");
  API_CK(transpose_batched(static_cast<cudaStream_t>(stream), B, D, L, nchw, feat, false));     // [D, L] -> [L, D]
  return 0;
}

int alexcap_dfeat_to_nchw(int32_t B, int32_t D, int32_t L, const float* dfeat, float* dnchw, int32_t accumulate, void* stream) {
  if (B <= 0 || D <= 0 || L <= 0 || !dfeat || !dnchw) return fail(ALEXCAP_E_BADARG, "alexcap_dfeat_to_nchw: bad arguments");
  API_CK(transpose_batched(static_cast<cudaStream_t>(stream), B, L, D, dfeat, dnchw, accumulate != 0)); // [L, D] -> [D, L]
  return 0;
}

}  // extern "C"
