// optimizer.cu -- global-norm gradient clipping + Adam as ONE multi-tensor pass (the step
// that immediately follows the hot path).
//
// Reference: train_LSTM.py:86-87 / train_LSTMwAttention.py:87-88 -> AlexCapModel.clip_gradient (LSTMModel.py:50-51:
// torch.nn.utils.clip_grad_norm_(parameters, max_norm)), then optim.step() with torch.optim.Adam(params, lr, (beta1, beta2),
// eps, weight_decay) (train_LSTMwAttention.py:60, :109).  Stock PyTorch launches a handful of kernels per parameter
// tensor (19 tensors for the attention decoder); here: one deterministic two-stage norm reduction and one update kernel
// over all tensors, the clip coefficient applied on the fly.
//
// torch semantics reproduced exactly (non-amsgrad Adam, L2 weight decay folded into the gradient):
//   total_norm = || concat(grads) ||_2 ;  coef = min(1, max_norm / (total_norm + 1e-6)) ;  g <- g * coef   (written back)
//   g' = g + wd * p ;  m <- b1 m + (1-b1) g' ;  v <- b2 v + (1-b2) g'^2
//   p <- p - (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
#include "api_util.h"

using namespace alexcap;

namespace {

constexpr int OPT_MAX_TENSORS = ALEXCAP_OPT_MAX_TENSORS;
constexpr int OPT_THREADS = 256;

struct TensorTable {
  float* p[OPT_MAX_TENSORS]; float* g[OPT_MAX_TENSORS]; float* m[OPT_MAX_TENSORS]; float* v[OPT_MAX_TENSORS];
  int64_t n[OPT_MAX_TENSORS];
  int count;
};

// stage 1: per-block partial sum of squares over every tensor (grid-stride inside each tensor, fixed order)
__global__ void __launch_bounds__(OPT_THREADS) grad_sqnorm_part_kernel(const TensorTable tt, float* __restrict__ partial) {
  __shared__ float scratch[40];
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (int t = 0; t < tt.count; ++t) {
    const float* g = tt.g[t];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tt.n[t]; i += stride) {
      const float x = g[i];
      acc = fmaf(x, x, acc);
    }
  }
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// stage 2: total norm and clip coefficient.  out[0] = total_norm, out[1] = coef
__global__ void __launch_bounds__(OPT_THREADS) grad_norm_final_kernel(const float* __restrict__ partial, int nblk, float max_norm,
                                                                      float* __restrict__ out) {
  __shared__ float scratch[40];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nblk; i += blockDim.x) acc += partial[i];
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) {
    const float norm = sqrtf(acc);
    float coef = 1.f;
    if (max_norm > 0.f) coef = fminf(1.f, max_norm / (norm + 1e-6f));
    out[0] = norm; out[1] = coef;
  }
}

__global__ void __launch_bounds__(OPT_THREADS) clip_adam_kernel(const TensorTable tt, const float* __restrict__ norm_coef,
                                                                float lr_over_bc1, float inv_sqrt_bc2, float beta1, float beta2,
                                                                float eps, float weight_decay) {
  const float coef = norm_coef ? norm_coef[1] : 1.f;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int t = 0; t < tt.count; ++t) {
    float* p = tt.p[t]; float* g = tt.g[t]; float* m = tt.m[t]; float* v = tt.v[t];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < tt.n[t]; i += stride) {
      const float gc = g[i] * coef;
      g[i] = gc;                                    // clip_grad_norm_ scales .grad in place
      const float pi = p[i];
      const float gd = fmaf(weight_decay, pi, gc);
      const float mi = fmaf(beta1, m[i], (1.f - beta1) * gd);
      const float vi = fmaf(beta2, v[i], (1.f - beta2) * gd * gd);
      m[i] = mi; v[i] = vi;
      p[i] = pi - lr_over_bc1 * (mi / (sqrtf(vi) * inv_sqrt_bc2 + eps));
    }
  }
}

}  // namespace

extern "C" {

size_t alexcap_clip_adam_scratch_floats(void) { return 2048 + 8; }

int alexcap_clip_adam_step(int32_t n_tensors, float* const* params, float* const* grads, float* const* exp_avg,
                           float* const* exp_avg_sq, const int64_t* numel, float max_norm, float lr, float beta1,
                           float beta2, float eps, float weight_decay, int64_t step, float* norm_out, float* scratch,
                           void* stream) {
  if (n_tensors <= 0 || n_tensors > OPT_MAX_TENSORS) return fail(ALEXCAP_E_BADARG, "alexcap_clip_adam_step: 1..ALEXCAP_OPT_MAX_TENSORS tensors");
  if (!params || !grads || !exp_avg || !exp_avg_sq || !numel || !scratch) return fail(ALEXCAP_E_BADARG, "alexcap_clip_adam_step: NULL table / scratch");
  if (step < 1) return fail(ALEXCAP_E_BADARG, "alexcap_clip_adam_step: step is 1-based");
  if (!(beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f)) return fail(ALEXCAP_E_BADARG, "alexcap_clip_adam_step: betas must be in [0,1)");
  TensorTable tt;
  tt.count = n_tensors;
  int64_t total = 0;
  for (int i = 0; i < n_tensors; ++i) {
    if (!params[i] || !grads[i] || !exp_avg[i] || !exp_avg_sq[i] || numel[i] < 0)
      return fail(ALEXCAP_E_BADARG, "alexcap_clip_adam_step: NULL tensor or negative numel");
    tt.p[i] = params[i]; tt.g[i] = grads[i]; tt.m[i] = exp_avg[i]; tt.v[i] = exp_avg_sq[i]; tt.n[i] = numel[i];
    total += numel[i];
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int blocks = (int)((total + OPT_THREADS * 8 - 1) / (OPT_THREADS * 8));
  if (blocks < 1) blocks = 1;
  if (blocks > 2048) blocks = 2048;
  float* partial = scratch;                 // [<= 2048]
  float* nc = scratch + 2048;               // [0] = norm, [1] = coef
  ALEXCAP_NOTE_LAUNCH(); grad_sqnorm_part_kernel<<<blocks, OPT_THREADS, 0, st>>>(tt, partial);
  API_CK(cudaGetLastError());
  ALEXCAP_NOTE_LAUNCH(); grad_norm_final_kernel<<<1, OPT_THREADS, 0, st>>>(partial, blocks, max_norm, nc);
  API_CK(cudaGetLastError());
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  ALEXCAP_NOTE_LAUNCH();
  clip_adam_kernel<<<blocks, OPT_THREADS, 0, st>>>(tt, nc, (float)(lr / bc1), (float)(1.0 / sqrt(bc2)), beta1, beta2, eps,
                                                   weight_decay);
  API_CK(cudaGetLastError());
  if (norm_out) API_CK(cudaMemcpyAsync(norm_out, nc, sizeof(float), cudaMemcpyDeviceToDevice, st));
  return 0;
}

}  // extern "C"
