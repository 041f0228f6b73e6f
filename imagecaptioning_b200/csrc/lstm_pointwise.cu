// lstm_pointwise.cu -- LSTM cell pointwise math, forward and backward (gate order i,f,g,o as torch.nn.LSTMCell /
// nn.LSTM: AlexCap/AttentionLanguageModule.py:58,78 ; AlexCap/LanguageModule.py:46,104-105).
// The gate pre-activations come from the GEMMs; these kernels are elementwise and negligible in bytes -- but they sit on the
// serial chain of every time step, so sigmoid / tanh are the two-MUFU forms of common.cuh (~2e-7 absolute error) rather than
// libm's expf / tanhf / IEEE division (4 cells per thread of those are ~450 dependent instructions).
#include <cuda_bf16.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace alexcap {

__global__ void __launch_bounds__(256) lstm_fwd_kernel(int B, int H, float* __restrict__ gates, int64_t g_ld,
                                                       const float* __restrict__ c_prev, float* __restrict__ c_out,
                                                       float* __restrict__ h_out, int64_t h_ld,
                                                       float* __restrict__ h_out2, int64_t h2_ld,
                                                       float* __restrict__ acts_out, int64_t acts_ld,
                                                       uint16_t* __restrict__ h_b16, int64_t hb_ld) {
  const int H4 = H >> 2;
  const int64_t n = (int64_t)B * H4;
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / H4), c = (int)(i % H4);
    float4* g = reinterpret_cast<float4*>(gates + (int64_t)b * g_ld);
    float4 gi = g[c], gf = g[H4 + c], gg = g[2 * H4 + c], go = g[3 * H4 + c];
    const float4 cp = reinterpret_cast<const float4*>(c_prev + (int64_t)b * H)[c];
    float4 cn, hn;
#define CELL(x)                                                        \
    gi.x = sigmoid_fast(gi.x); gf.x = sigmoid_fast(gf.x); gg.x = tanh_fast2(gg.x); go.x = sigmoid_fast(go.x); \
    cn.x = fmaf(gf.x, cp.x, gi.x * gg.x); hn.x = go.x * tanh_fast2(cn.x);
    CELL(x) CELL(y) CELL(z) CELL(w)
#undef CELL
    float4* ga = acts_out ? reinterpret_cast<float4*>(acts_out + (int64_t)b * acts_ld) : g;   // (default: activations in place)
    ga[c] = gi; ga[H4 + c] = gf; ga[2 * H4 + c] = gg; ga[3 * H4 + c] = go;
    reinterpret_cast<float4*>(c_out + (int64_t)b * H)[c] = cn;
    reinterpret_cast<float4*>(h_out + (int64_t)b * h_ld)[c] = hn;
    if (h_out2) reinterpret_cast<float4*>(h_out2 + (int64_t)b * h2_ld)[c] = hn;
    if (h_b16) {                 // bf16 mirror of h (operand copy of the next GEMMs: no split pass)
      uint2 o;
      asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o.x) : "f"(hn.y), "f"(hn.x));
      asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o.y) : "f"(hn.w), "f"(hn.z));
      reinterpret_cast<uint2*>(h_b16 + (int64_t)b * hb_ld)[c] = o;
    }
  }
}

// One hidden unit per thread: 4x the threads of the float4 kernel, a quarter of the work each.  The kernel is a link of the
// serial per-step chain (pure latency): spreading B*H = 64 K cells over 256 blocks instead of 64 is 0.45 us per step faster at
// cfg2 (1.8965 -> 1.8889 ms); used while all threads are resident at once, the float4 kernel beyond.
__global__ void __launch_bounds__(256) lstm_fwd1_kernel(int B, int H, float* __restrict__ gates, int64_t g_ld,
                                                        const float* __restrict__ c_prev, float* __restrict__ c_out,
                                                        float* __restrict__ h_out, int64_t h_ld,
                                                        float* __restrict__ h_out2, int64_t h2_ld,
                                                        float* __restrict__ acts_out, int64_t acts_ld,
                                                        uint16_t* __restrict__ h_b16, int64_t hb_ld) {
  const int64_t n = (int64_t)B * H;
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / H), h = (int)(i % H);
    const float* g = gates + (int64_t)b * g_ld;
    const float gi = sigmoid_fast(g[h]), gf = sigmoid_fast(g[H + h]), gg = tanh_fast2(g[2 * H + h]), go = sigmoid_fast(g[3 * H + h]);
    const float cn = fmaf(gf, c_prev[(int64_t)b * H + h], gi * gg);
    const float hn = go * tanh_fast2(cn);
    float* ga = acts_out ? acts_out + (int64_t)b * acts_ld : gates + (int64_t)b * g_ld;
    ga[h] = gi; ga[H + h] = gf; ga[2 * H + h] = gg; ga[3 * H + h] = go;
    c_out[(int64_t)b * H + h] = cn;
    h_out[(int64_t)b * h_ld + h] = hn;
    if (h_out2) h_out2[(int64_t)b * h2_ld + h] = hn;
    if (h_b16) h_b16[(int64_t)b * hb_ld + h] = __bfloat16_as_ushort(__float2bfloat16_rn(hn));
  }
}

cudaError_t lstm_pointwise_fwd(cudaStream_t st, int B, int H, float* gates, int64_t g_ld, const float* c_prev,
                               float* c_out, float* h_out, int64_t h_ld, float* h_out2, int64_t h2_ld, float* acts_out,
                               int64_t acts_ld, uint16_t* h_b16, int64_t hb_ld) {
  if ((H & 3) || (g_ld & 3) || (h_ld & 3) || (h_out2 && (h2_ld & 3)) || (acts_out && (acts_ld & 3))) return cudaErrorInvalidValue;
  const int64_t n1 = (int64_t)B * H;
  if (n1 <= (int64_t)148 * 2048) {            // every thread resident at once: one cell per thread
    ALEXCAP_NOTE_LAUNCH();
    return launch_pdl(lstm_fwd1_kernel, dim3((int)((n1 + 255) / 256)), dim3(256), 0, st, B, H, gates, g_ld, c_prev, c_out, h_out, h_ld,
                      h_out2, h2_ld, acts_out, acts_ld, h_b16, hb_ld);
  }
  const int64_t n = (int64_t)B * (H / 4);
  const int blocks = (int)((n + 255) / 256);
  ALEXCAP_NOTE_LAUNCH();
  return launch_pdl(lstm_fwd_kernel, dim3(blocks), dim3(256), 0, st, B, H, gates, g_ld, c_prev, c_out, h_out, h_ld, h_out2, h2_ld,
                    acts_out, acts_ld, h_b16, hb_ld);
}

__global__ void __launch_bounds__(256) lstm_bwd_kernel(int B, int H, float* __restrict__ gates, int64_t g_ld,
                                                       const float* __restrict__ c_t, const float* __restrict__ c_prev,
                                                       const float* __restrict__ dh_a, int64_t dha_ld,
                                                       const float* __restrict__ mask_a,
                                                       const float* __restrict__ dh_b, int64_t dhb_ld,
                                                       float* __restrict__ dc_io, uint16_t* __restrict__ g_b16, int64_t gb_ld) {
  const int64_t n = (int64_t)B * H;
  pdl_trigger();
  pdl_wait();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / H), h = (int)(i % H);
    float* g = gates + (int64_t)b * g_ld;
    const float gi = g[h], gf = g[H + h], gg = g[2 * H + h], go = g[3 * H + h];
    float dh = 0.f;
    if (dh_a) {
      float v = dh_a[(int64_t)b * dha_ld + h];
      if (mask_a) v *= mask_a[(int64_t)b * H + h];
      dh += v;
    }
    if (dh_b) dh += dh_b[(int64_t)b * dhb_ld + h];
    const float tc = tanh_fast2(c_t[(int64_t)b * H + h]);
    const float d_o = dh * tc * go * (1.f - go);
    const float dc = dc_io[(int64_t)b * H + h] + dh * go * (1.f - tc * tc);
    const float d_i = dc * gg * gi * (1.f - gi);
    const float d_f = dc * c_prev[(int64_t)b * H + h] * gf * (1.f - gf);
    const float d_g = dc * gi * (1.f - gg * gg);
    dc_io[(int64_t)b * H + h] = dc * gf;
    g[h] = d_i; g[H + h] = d_f; g[2 * H + h] = d_g; g[3 * H + h] = d_o;
    if (g_b16) {                 // bf16 mirror of d gates
      uint16_t* gb = g_b16 + (int64_t)b * gb_ld;
      gb[h] = __bfloat16_as_ushort(__float2bfloat16_rn(d_i)); gb[H + h] = __bfloat16_as_ushort(__float2bfloat16_rn(d_f));
      gb[2 * H + h] = __bfloat16_as_ushort(__float2bfloat16_rn(d_g)); gb[3 * H + h] = __bfloat16_as_ushort(__float2bfloat16_rn(d_o));
    }
  }
}

cudaError_t lstm_pointwise_bwd(cudaStream_t st, int B, int H, float* gates, int64_t g_ld, const float* c_t,
                               const float* c_prev, const float* dh_a, int64_t dha_ld, const float* mask_a,
                               const float* dh_b, int64_t dhb_ld, float* dc_io, uint16_t* g_b16, int64_t gb_ld) {
  const int64_t n = (int64_t)B * H;
  const int blocks = (int)((n + 255) / 256);
  ALEXCAP_NOTE_LAUNCH();
  return launch_pdl(lstm_bwd_kernel, dim3(blocks), dim3(256), 0, st, B, H, gates, g_ld, c_t, c_prev, dh_a, dha_ld, mask_a, dh_b, dhb_ld, dc_io,
                    g_b16, gb_ld);
}

}  // namespace alexcap
