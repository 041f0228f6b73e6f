// attention.cu -- fused additive soft-attention step (forward, backward) and the hoisted dWs pass.
//
// Replaces, per timestep, Attention.forward + the beta gate of Decoder.forward
// (AlexCap/AttentionLanguageModule.py:15-22, :73-75): ~8 ATen kernels and a GEMV in the reference.
// W_s = W(features) is time-invariant and computed ONCE per training step by the caller (the reference
// recomputes it every step, :17).
//
// HBM-bound: one CTA per image streams W_s[b] (L x H) and feat[b] (L x D) exactly once per step with 128-bit
// coalesced loads; scores, softmax (warp-shuffle reductions) and alpha live in shared memory.
// Algorithmic bytes per step: B*L*(H+D)*4 (+ small vectors).
#include "common.cuh"
#include "kernels.h"

namespace alexcap {

constexpr int ATT_THREADS = 512;
constexpr int ATT_WARPS = ATT_THREADS / 32;

struct AttnFwdP {
  const float* Ws; const float* feat; const float* q; int64_t q_ld; const float* bpre; int64_t bpre_ld;
  const float* wv; float* alpha; int64_t alpha_ld; float* ctx; float* beta; float* gctx; int64_t gctx_ld;
  int L, D, H, rpi;
};

// NJ: number of float4 per lane along H (H <= 128*NJ).
template <int NJ>
__global__ void __launch_bounds__(ATT_THREADS) attn_fwd_kernel(const AttnFwdP a) {
  extern __shared__ __align__(16) float sm[];
  float4* red = reinterpret_cast<float4*>(sm);              // [ATT_THREADS]
  float* es = sm + 4 * ATT_THREADS;                         // [L]
  float* scratch = es + ((a.L + 3) & ~3);                   // [33]

  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int L = a.L, H4 = a.H >> 2, D4 = a.D >> 2;

  // ---- phase 1: e[l] = w_v . tanh(Ws[b,l,:] + q[b,:])           (AttentionLanguageModule.py:18-19)
  float4 q4[NJ], v4[NJ];
  {
    const float4* qp = reinterpret_cast<const float4*>(a.q + (int64_t)b * a.q_ld);
    const float4* vp = reinterpret_cast<const float4*>(a.wv);
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int f = lane + 32 * j;
      q4[j] = (f < H4) ? __ldg(qp + f) : f4_zero();
      v4[j] = (f < H4) ? __ldg(vp + f) : f4_zero();
    }
  }
  const int img = b / a.rpi;   // beam rows of one image share its feature grid
  const float4* wsb = reinterpret_cast<const float4*>(a.Ws) + (int64_t)img * L * H4;
  for (int l = warp; l < L; l += ATT_WARPS) {
    const float4* row = wsb + (int64_t)l * H4;
    float4 w[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int f = lane + 32 * j;
      w[j] = (f < H4) ? ldg4(row + f) : f4_zero();
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      s = fmaf(v4[j].x, tanhf(w[j].x + q4[j].x), s);
      s = fmaf(v4[j].y, tanhf(w[j].y + q4[j].y), s);
      s = fmaf(v4[j].z, tanhf(w[j].z + q4[j].z), s);
      s = fmaf(v4[j].w, tanhf(w[j].w + q4[j].w), s);
    }
    s = warp_sum(s);
    if (lane == 0) es[l] = s;       // + b_v omitted: softmax is shift invariant (alpha identical)
  }
  __syncthreads();

  // ---- phase 2: alpha = softmax_l(e)                               (:20)
  float mx = -INFINITY;
  for (int l = tid; l < L; l += ATT_THREADS) mx = fmaxf(mx, es[l]);
  mx = block_max(mx, scratch);
  float sum = 0.f;
  for (int l = tid; l < L; l += ATT_THREADS) {
    const float e = expf(es[l] - mx);
    es[l] = e;
    sum += e;
  }
  sum = block_sum(sum, scratch);
  const float inv = 1.f / sum;
  for (int l = tid; l < L; l += ATT_THREADS) {
    const float al = es[l] * inv;
    es[l] = al;
    a.alpha[(int64_t)b * a.alpha_ld + l] = al;
  }
  __syncthreads();

  // ---- phase 3: ctx = sum_l alpha[l] * feat[b,l,:];  beta = sigmoid(bpre);  gctx = beta*ctx   (:21, :74-75)
  const float4* fb = reinterpret_cast<const float4*>(a.feat) + (int64_t)img * L * D4;
  const bool wide = D4 >= ATT_THREADS;
  const int G = wide ? 1 : ATT_THREADS / D4;
  for (int c0 = 0; c0 < D4; c0 += ATT_THREADS) {
    const int c = wide ? c0 + tid : tid % D4;
    const int g = wide ? 0 : tid / D4;
    float4 acc = f4_zero();
    if (c < D4 && g < G) {
      const float4* col = fb + c;
#pragma unroll 4
      for (int l = g; l < L; l += G) {
        const float4 f = ldg4(col + (int64_t)l * D4);
        f4_fma(acc, es[l], f);
      }
    }
    red[tid] = acc;
    __syncthreads();
    if (g == 0 && c < D4) {
      for (int gg = 1; gg < G; ++gg) f4_add(acc, red[gg * D4 + c]);
      const float4 bp = __ldg(reinterpret_cast<const float4*>(a.bpre + (int64_t)b * a.bpre_ld) + c);
      float4 be;
      be.x = sigmoid_f(bp.x); be.y = sigmoid_f(bp.y); be.z = sigmoid_f(bp.z); be.w = sigmoid_f(bp.w);
      reinterpret_cast<float4*>(a.ctx + (int64_t)b * a.D)[c] = acc;
      reinterpret_cast<float4*>(a.beta + (int64_t)b * a.D)[c] = be;
      reinterpret_cast<float4*>(a.gctx + (int64_t)b * a.gctx_ld)[c] =
          make_float4(be.x * acc.x, be.y * acc.y, be.z * acc.z, be.w * acc.w);
    }
    __syncthreads();
  }
}

static size_t attn_fwd_smem(int L) { return (size_t)(4 * ATT_THREADS + ((L + 3) & ~3) + 40) * sizeof(float); }

// Opt in to > 48 KB dynamic shared memory, once per kernel instantiation and size (the attribute call is kept out of
// steady-state launches so the step can be captured into a CUDA graph).
template <typename K>
static cudaError_t set_smem(K kernel, size_t bytes) {
  static size_t granted = 48 * 1024;   // one static per template instantiation K... but K is a type: key by pointer too
  static K granted_for = nullptr;
  if (bytes <= 48 * 1024) return cudaSuccess;
  if (granted_for == kernel && bytes <= granted) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) { granted = bytes; granted_for = kernel; }
  return e;
}

cudaError_t attn_step_fwd(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                          const float* q, int64_t q_ld, const float* bpre, int64_t bpre_ld, const float* wv,
                          float* alpha, int64_t alpha_ld, float* ctx, float* beta, float* gctx, int64_t gctx_ld,
                          int rows_per_image) {
  if ((H & 3) || (D & 3) || H > 1024 || (q_ld & 3) || (bpre_ld & 3) || (gctx_ld & 3)) return cudaErrorInvalidValue;
  AttnFwdP a{Ws, feat, q, q_ld, bpre, bpre_ld, wv, alpha, alpha_ld, ctx, beta, gctx, gctx_ld, L, D, H,
             rows_per_image < 1 ? 1 : rows_per_image};
  const size_t smem = attn_fwd_smem(L);
  const int nj = ceil_div(H, 128);
#define LAUNCH_FWD(NJ)                                              \
  do {                                                              \
    ACK(set_smem(attn_fwd_kernel<NJ>, smem));                       \
    ALEXCAP_NOTE_LAUNCH(); attn_fwd_kernel<NJ><<<B, ATT_THREADS, smem, st>>>(a);           \
  } while (0)
  if (nj <= 1) LAUNCH_FWD(1);
  else if (nj <= 2) LAUNCH_FWD(2);
  else if (nj <= 4) LAUNCH_FWD(4);
  else LAUNCH_FWD(8);
#undef LAUNCH_FWD
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ backward step
struct AttnBwdP {
  const float* Ws; const float* feat; const float* q; int64_t q_ld; const float* wv;
  const float* alpha; int64_t alpha_ld; const float* ctx; const float* beta; const float* dgctx; int64_t dgctx_ld;
  const float* ext; int64_t ext_ld; const float* reg_r;
  float* dq; int64_t dq_ld; float* dbpre; int64_t dbpre_ld; float* de; int64_t de_ld; float* dwv_part; float* dfeat;
  int L, D, H;
};

template <int NJ>
__global__ void __launch_bounds__(ATT_THREADS) attn_bwd_kernel(const AttnBwdP a) {
  extern __shared__ __align__(16) float sm[];
  const int L = a.L, D = a.D, H = a.H, H4 = H >> 2, D4 = D >> 2;
  const int Lp = (L + 3) & ~3;
  float* dctx_s = sm;                    // [D]
  float* al_s = dctx_s + D;              // [Lp]
  float* da_s = al_s + Lp;               // [Lp]   d alpha, then d e
  float* scratch = da_s + Lp;            // [40]
  float* red = scratch + 40;             // [ATT_WARPS][H]

  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- A: gate backward.  gctx = beta*ctx, beta = sigmoid(bpre)          (AttentionLanguageModule.py:74-75)
  for (int c = tid; c < D4; c += ATT_THREADS) {
    const float4 dg = __ldg(reinterpret_cast<const float4*>(a.dgctx + (int64_t)b * a.dgctx_ld) + c);
    const float4 be = __ldg(reinterpret_cast<const float4*>(a.beta + (int64_t)b * D) + c);
    const float4 cx = __ldg(reinterpret_cast<const float4*>(a.ctx + (int64_t)b * D) + c);
    float4 dc, db;
    dc.x = dg.x * be.x; dc.y = dg.y * be.y; dc.z = dg.z * be.z; dc.w = dg.w * be.w;
    db.x = dg.x * cx.x * be.x * (1.f - be.x); db.y = dg.y * cx.y * be.y * (1.f - be.y);
    db.z = dg.z * cx.z * be.z * (1.f - be.z); db.w = dg.w * cx.w * be.w * (1.f - be.w);
    reinterpret_cast<float4*>(dctx_s)[c] = dc;
    reinterpret_cast<float4*>(a.dbpre + (int64_t)b * a.dbpre_ld)[c] = db;
  }
  for (int l = tid; l < L; l += ATT_THREADS) al_s[l] = __ldg(a.alpha + (int64_t)b * a.alpha_ld + l);
  __syncthreads();

  // ---- B: d alpha[l] = dctx . feat[b,l,:] (+ upstream terms);  dfeat += alpha[l] * dctx        (:21)
  const float4* fb = reinterpret_cast<const float4*>(a.feat) + (int64_t)b * L * D4;
  float4* dfb = a.dfeat ? reinterpret_cast<float4*>(a.dfeat) + (int64_t)b * L * D4 : nullptr;
  for (int l = warp; l < L; l += ATT_WARPS) {
    const float4* row = fb + (int64_t)l * D4;
    float s = 0.f;
    const float al = al_s[l];
#pragma unroll 4
    for (int f = lane; f < D4; f += 32) {
      const float4 x = ldg4(row + f);
      const float4 dc = reinterpret_cast<const float4*>(dctx_s)[f];
      s = fmaf(x.x, dc.x, s); s = fmaf(x.y, dc.y, s); s = fmaf(x.z, dc.z, s); s = fmaf(x.w, dc.w, s);
      if (dfb) {
        float4 o = dfb[(int64_t)l * D4 + f];
        f4_fma(o, al, dc);
        dfb[(int64_t)l * D4 + f] = o;
      }
    }
    s = warp_sum(s);
    if (lane == 0) {
      if (a.ext) s += __ldg(a.ext + (int64_t)b * a.ext_ld + l);
      if (a.reg_r) s += __ldg(a.reg_r + (int64_t)b * L + l);
      da_s[l] = s;
    }
  }
  __syncthreads();

  // ---- C: softmax backward  de = alpha * (dalpha - sum alpha*dalpha)                         (:20)
  float part = 0.f;
  for (int l = tid; l < L; l += ATT_THREADS) part = fmaf(al_s[l], da_s[l], part);
  const float dot = block_sum(part, scratch);
  for (int l = tid; l < L; l += ATT_THREADS) {
    const float de = al_s[l] * (da_s[l] - dot);
    da_s[l] = de;
    a.de[(int64_t)b * a.de_ld + l] = de;
  }
  __syncthreads();

  // ---- D: score backward.  dz = de[l]*w_v*(1-att^2); dq = sum_l dz; dw_v += sum_l de[l]*att   (:18-19)
  float4 q4[NJ], v4[NJ], dq4[NJ], dv4[NJ];
  {
    const float4* qp = reinterpret_cast<const float4*>(a.q + (int64_t)b * a.q_ld);
    const float4* vp = reinterpret_cast<const float4*>(a.wv);
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int f = lane + 32 * j;
      q4[j] = (f < H4) ? __ldg(qp + f) : f4_zero();
      v4[j] = (f < H4) ? __ldg(vp + f) : f4_zero();
      dq4[j] = f4_zero();
      dv4[j] = f4_zero();
    }
  }
  const float4* wsb = reinterpret_cast<const float4*>(a.Ws) + (int64_t)b * L * H4;
  for (int l = warp; l < L; l += ATT_WARPS) {
    const float4* row = wsb + (int64_t)l * H4;
    const float de = da_s[l];
    float4 w[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int f = lane + 32 * j;
      w[j] = (f < H4) ? ldg4(row + f) : f4_zero();
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      float t;
      t = tanhf(w[j].x + q4[j].x); dv4[j].x = fmaf(de, t, dv4[j].x); dq4[j].x = fmaf(de * v4[j].x, 1.f - t * t, dq4[j].x);
      t = tanhf(w[j].y + q4[j].y); dv4[j].y = fmaf(de, t, dv4[j].y); dq4[j].y = fmaf(de * v4[j].y, 1.f - t * t, dq4[j].y);
      t = tanhf(w[j].z + q4[j].z); dv4[j].z = fmaf(de, t, dv4[j].z); dq4[j].z = fmaf(de * v4[j].z, 1.f - t * t, dq4[j].z);
      t = tanhf(w[j].w + q4[j].w); dv4[j].w = fmaf(de, t, dv4[j].w); dq4[j].w = fmaf(de * v4[j].w, 1.f - t * t, dq4[j].w);
    }
  }
  // cross-warp reduction of dq, then of dw_v, through red[ATT_WARPS][H]
  float4* red4 = reinterpret_cast<float4*>(red);
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const int f = lane + 32 * j;
    if (f < H4) red4[warp * H4 + f] = dq4[j];
  }
  __syncthreads();
  for (int f = tid; f < H4; f += ATT_THREADS) {
    float4 s = red4[f];
    for (int w = 1; w < ATT_WARPS; ++w) f4_add(s, red4[w * H4 + f]);
    reinterpret_cast<float4*>(a.dq + (int64_t)b * a.dq_ld)[f] = s;
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const int f = lane + 32 * j;
    if (f < H4) red4[warp * H4 + f] = dv4[j];
  }
  __syncthreads();
  for (int f = tid; f < H4; f += ATT_THREADS) {
    float4 s = red4[f];
    for (int w = 1; w < ATT_WARPS; ++w) f4_add(s, red4[w * H4 + f]);
    float4* out = reinterpret_cast<float4*>(a.dwv_part + (int64_t)b * H) + f;
    float4 o = *out;
    f4_add(o, s);
    *out = o;
  }
}

cudaError_t attn_step_bwd(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                          const float* q, int64_t q_ld, const float* wv, const float* alpha, int64_t alpha_ld,
                          const float* ctx, const float* beta, const float* dgctx, int64_t dgctx_ld,
                          const float* dalpha_ext, int64_t ext_ld, const float* reg_r,
                          float* dq, int64_t dq_ld, float* dbpre, int64_t dbpre_ld, float* de, int64_t de_ld,
                          float* dwv_part, float* dfeat) {
  if ((H & 3) || (D & 3) || H > 1024 || (q_ld & 3) || (dgctx_ld & 3) || (dq_ld & 3) || (dbpre_ld & 3))
    return cudaErrorInvalidValue;
  AttnBwdP a{Ws, feat, q, q_ld, wv, alpha, alpha_ld, ctx, beta, dgctx, dgctx_ld, dalpha_ext, ext_ld, reg_r,
             dq, dq_ld, dbpre, dbpre_ld, de, de_ld, dwv_part, dfeat, L, D, H};
  const int Lp = (L + 3) & ~3;
  const size_t smem = (size_t)(D + 2 * Lp + 40 + ATT_WARPS * H) * sizeof(float);
  const int nj = ceil_div(H, 128);
#define LAUNCH_BWD(NJ)                                              \
  do {                                                              \
    ACK(set_smem(attn_bwd_kernel<NJ>, smem));                       \
    ALEXCAP_NOTE_LAUNCH(); attn_bwd_kernel<NJ><<<B, ATT_THREADS, smem, st>>>(a);           \
  } while (0)
  if (nj <= 1) LAUNCH_BWD(1);
  else if (nj <= 2) LAUNCH_BWD(2);
  else if (nj <= 4) LAUNCH_BWD(4);
  else LAUNCH_BWD(8);
#undef LAUNCH_BWD
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ hoisted dWs
// dWs[b,l,:] = w_v * sum_t de[b,t,l] * (1 - tanh^2(Ws[b,l,:] + q_t[b,:])) : one read of Ws, one write of dWs,
// instead of a read-modify-write of the [B,L,H] accumulator on every reverse step.
constexpr int DWS_ROWS = 4;
__global__ void __launch_bounds__(256) attn_dws_kernel(int L, int H, int S, int B, const float* __restrict__ Ws,
                                                       const float* __restrict__ q, int64_t q_ld,
                                                       const float* __restrict__ wv, const float* __restrict__ de,
                                                       float* __restrict__ dWs) {
  const int b = blockIdx.y, l0 = blockIdx.x * DWS_ROWS, H4 = H >> 2;
  const float* deb = de + (int64_t)b * S * L;
  for (int c = threadIdx.x; c < H4; c += blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(wv) + c);
    float4 w[DWS_ROWS], acc[DWS_ROWS];
#pragma unroll
    for (int r = 0; r < DWS_ROWS; ++r) {
      const int l = l0 + r;
      w[r] = (l < L) ? ldg4(reinterpret_cast<const float4*>(Ws) + ((int64_t)b * L + l) * H4 + c) : f4_zero();
      acc[r] = f4_zero();
    }
    for (int t = 0; t < S; ++t) {
      const float4 qt = __ldg(reinterpret_cast<const float4*>(q + ((int64_t)t * B + b) * q_ld) + c);
#pragma unroll
      for (int r = 0; r < DWS_ROWS; ++r) {
        const int l = l0 + r;
        const float d = (l < L) ? __ldg(deb + (int64_t)t * L + l) : 0.f;
        float a;
        a = tanhf(w[r].x + qt.x); acc[r].x = fmaf(d, 1.f - a * a, acc[r].x);
        a = tanhf(w[r].y + qt.y); acc[r].y = fmaf(d, 1.f - a * a, acc[r].y);
        a = tanhf(w[r].z + qt.z); acc[r].z = fmaf(d, 1.f - a * a, acc[r].z);
        a = tanhf(w[r].w + qt.w); acc[r].w = fmaf(d, 1.f - a * a, acc[r].w);
      }
    }
#pragma unroll
    for (int r = 0; r < DWS_ROWS; ++r) {
      const int l = l0 + r;
      if (l < L)
        reinterpret_cast<float4*>(dWs)[((int64_t)b * L + l) * H4 + c] =
            make_float4(v.x * acc[r].x, v.y * acc[r].y, v.z * acc[r].z, v.w * acc[r].w);
    }
  }
}

cudaError_t attn_dws(cudaStream_t st, int B, int L, int H, int S, const float* Ws, const float* q, int64_t q_ld,
                     const float* wv, const float* de, float* dWs) {
  if ((H & 3) || (q_ld & 3)) return cudaErrorInvalidValue;
  dim3 grid(ceil_div(L, DWS_ROWS), B);
  const int threads = (H / 4 >= 256) ? 256 : ((H / 4 + 31) / 32) * 32;
  ALEXCAP_NOTE_LAUNCH(); attn_dws_kernel<<<grid, threads, 0, st>>>(L, H, S, B, Ws, q, q_ld, wv, de, dWs);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------ regulariser
// LSTMwAttentionModel.py:70 : ((1 - alphas.sum(1))**2).mean(); its gradient wrt alpha[b,t,l] is the same for all t.
__global__ void __launch_bounds__(256) reg_kernel(int B, int S, int L, float weight, const float* __restrict__ alphas,
                                                  float* __restrict__ r, float* __restrict__ partial) {
  __shared__ float scratch[40];
  const int64_t n = (int64_t)B * L;
  float acc = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / L), l = (int)(i % L);
    float s = 0.f;
    for (int t = 0; t < S; ++t) s += __ldg(alphas + ((int64_t)b * S + t) * L + l);
    const float resid = 1.f - s;
    r[i] = -2.f * weight * resid / (float)n;
    acc = fmaf(resid, resid, acc);
  }
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}
__global__ void __launch_bounds__(256) reg_final_kernel(int nblk, float scale, const float* __restrict__ partial,
                                                        float* __restrict__ out) {
  __shared__ float scratch[40];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nblk; i += blockDim.x) acc += partial[i];
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) out[0] = acc * scale;
}

cudaError_t attn_regulariser(cudaStream_t st, int B, int S, int L, float weight, const float* alphas, float* r,
                             float* loss_reg, float* scratch) {
  const int64_t n = (int64_t)B * L;
  const int nblk = (int)((n + 255) / 256 < 256 ? (n + 255) / 256 : 256);
  ALEXCAP_NOTE_LAUNCH(); reg_kernel<<<nblk, 256, 0, st>>>(B, S, L, weight, alphas, r, scratch);
  ALEXCAP_LAUNCH_CHECK();
  ALEXCAP_NOTE_LAUNCH(); reg_final_kernel<<<1, 256, 0, st>>>(nblk, weight / (float)n, scratch, loss_reg);
  return cudaGetLastError();
}

}  // namespace alexcap
