// attention.cu -- fused additive soft-attention step, forward and backward.
//
// Replaces, per timestep, Attention.forward + the beta gate of Decoder.forward
// (AlexCap/AttentionLanguageModule.py:15-22, :73-75): ~8 ATen kernels and a GEMV in the reference.
// W_s = W(features) is time-invariant and computed ONCE per training step by the caller (the reference
// recomputes it every step, :17).
//
// HBM/L2-bandwidth bound: W_s[b] (L x H) and feat[b] (L x D) are streamed exactly once per step with 128-bit coalesced
// loads.  One image is handled by a thread-block CLUSTER of CL CTAs that split the L regions: every CTA computes the
// scores of its rows, a local (max, sum) and an unnormalised partial context; the pieces are combined through
// distributed shared memory (flash-decoding style split softmax), so a batch of 128 images still puts 512 CTAs
// (~3.5 per SM) in flight and enough bytes to cover the memory latency.
// Algorithmic bytes per step: B*L*(H+D)*4 (+ small vectors).
#include <cooperative_groups.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"
#include "pipe.cuh"

namespace cg = cooperative_groups;

namespace alexcap {

int& attn_engine() {
  static int e = [] { const char* v = getenv("ALEXCAP_ATT_ENGINE"); return (v && *v) ? atoi(v) : 1; }();
  return e;
}

constexpr int ATT_THREADS = 256;
constexpr int ATT_WARPS = ATT_THREADS / 32;
constexpr int ATT_MAX_CL = 8;

// cluster size along L: enough CTAs to fill 148 SMs a couple of times, at least 8 rows per CTA
static int pick_cluster(int B, int L) {
  int cl = 1;
  while (cl < ATT_MAX_CL && (int64_t)B * cl < 296 && L / (cl * 2) >= 8) cl *= 2;
  return cl;
}

template <typename Kern, typename P>
static cudaError_t launch_cluster(Kern kern, int grid, size_t smem, int cl, cudaStream_t st, const P& p) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(ATT_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  ALEXCAP_NOTE_LAUNCH();
  return cudaLaunchKernelEx(&cfg, kern, p);
}

// Opt in to > 48 KB dynamic shared memory once per kernel (kept out of steady-state launches: graph capture).
template <typename K>
static cudaError_t set_smem(K kernel, size_t bytes) {
  static size_t granted = 48 * 1024;
  static K granted_for = nullptr;
  if (bytes <= 48 * 1024) return cudaSuccess;
  if (granted_for == kernel && bytes <= granted) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) { granted = bytes; granted_for = kernel; }
  return e;
}

// ------------------------------------------------------------------------------------------ forward step
struct AttnFwdP {
  const float* Ws; const float* feat; const float* q; int64_t q_ld; const float* bpre; int64_t bpre_ld;
  const float* wv; float* alpha; int64_t alpha_ld; float* ctx; float* beta; float* gctx; int64_t gctx_ld;
  int L, D, H, rpi, CL, Lc;
};

// NJ: number of float4 per lane along H (H <= 128*NJ).
template <int NJ>
__global__ void __launch_bounds__(ATT_THREADS) attn_fwd_kernel(const AttnFwdP a) {
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) float sm[];
  const int L = a.L, D = a.D, H4 = a.H >> 2, D4 = D >> 2, CL = a.CL;
  float4* red = reinterpret_cast<float4*>(sm);              // [ATT_THREADS]
  float4* pctx = red + ATT_THREADS;                         // [D4]  unnormalised partial context of this CTA
  float* es = reinterpret_cast<float*>(pctx + D4);          // [Lc]  scores, then p = exp(e - m_local)
  float* stats = es + ((a.Lc + 3) & ~3);                    // [2]   local max, local sum
  float* scratch = stats + 4;                               // [40]

  const int r = (int)cluster.block_rank();
  const int b = blockIdx.x / CL, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int l0 = r * a.Lc, l1 = min(L, l0 + a.Lc), nl = max(l1 - l0, 0);
  const int img = b / a.rpi;   // beam rows of one image share its feature grid

  // ---- phase 1: e[l] = w_v . tanh(Ws[b,l,:] + q[b,:]) for this CTA's rows     (AttentionLanguageModule.py:18-19)
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
  const float4* wsb = reinterpret_cast<const float4*>(a.Ws) + ((int64_t)img * L + l0) * H4;
  for (int i = warp; i < nl; i += 2 * ATT_WARPS) {          // two rows per iteration: 2*NJ 128-bit loads in flight
    const int i2 = i + ATT_WARPS;
    const bool has2 = i2 < nl;
    float4 w0[NJ], w1[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int f = lane + 32 * j;
      w0[j] = (f < H4) ? ldg4(wsb + (int64_t)i * H4 + f) : f4_zero();
      w1[j] = (f < H4 && has2) ? ldg4(wsb + (int64_t)i2 * H4 + f) : f4_zero();
    }
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      s0 = fmaf(v4[j].x, tanhf(w0[j].x + q4[j].x), s0); s0 = fmaf(v4[j].y, tanhf(w0[j].y + q4[j].y), s0);
      s0 = fmaf(v4[j].z, tanhf(w0[j].z + q4[j].z), s0); s0 = fmaf(v4[j].w, tanhf(w0[j].w + q4[j].w), s0);
      s1 = fmaf(v4[j].x, tanhf(w1[j].x + q4[j].x), s1); s1 = fmaf(v4[j].y, tanhf(w1[j].y + q4[j].y), s1);
      s1 = fmaf(v4[j].z, tanhf(w1[j].z + q4[j].z), s1); s1 = fmaf(v4[j].w, tanhf(w1[j].w + q4[j].w), s1);
    }
    s0 = warp_sum(s0); s1 = warp_sum(s1);
    if (lane == 0) {          // + b_v omitted: softmax is shift invariant (alpha identical)
      es[i] = s0;
      if (has2) es[i2] = s1;
    }
  }
  __syncthreads();

  // ---- phase 2: local softmax statistics                                        (:20, split over the cluster)
  float mx = -INFINITY;
  for (int i = tid; i < nl; i += ATT_THREADS) mx = fmaxf(mx, es[i]);
  mx = block_max(mx, scratch);
  float sum = 0.f;
  for (int i = tid; i < nl; i += ATT_THREADS) {
    const float e = expf(es[i] - mx);
    es[i] = e;
    sum += e;
  }
  sum = block_sum(sum, scratch);
  if (tid == 0) { stats[0] = mx; stats[1] = sum; }

  // ---- phase 3: unnormalised partial context  pctx = sum_{own l} p[l] * feat[b,l,:]            (:21)
  const float4* fb = reinterpret_cast<const float4*>(a.feat) + ((int64_t)img * L + l0) * D4;
  {
    const bool wide = D4 >= ATT_THREADS;
    const int G = wide ? 1 : ATT_THREADS / D4;
    for (int c0 = 0; c0 < D4; c0 += ATT_THREADS) {
      const int c = wide ? c0 + tid : tid % D4;
      const int g = wide ? 0 : tid / D4;
      float4 acc = f4_zero();
      if (c < D4 && g < G) {
        const float4* col = fb + c;
#pragma unroll 8
        for (int i = g; i < nl; i += G) f4_fma(acc, es[i], ldg4(col + (int64_t)i * D4));
      }
      red[tid] = acc;
      __syncthreads();
      if (g == 0 && c < D4) {
        for (int gg = 1; gg < G; ++gg) f4_add(acc, red[gg * D4 + c]);
        pctx[c] = acc;
      }
      __syncthreads();
    }
  }
  cluster.sync();                                            // partials + stats of every CTA visible cluster-wide

  // ---- phase 4: combine.  M = max m_r, S = sum s_r e^{m_r-M};  alpha = p * e^{m_own-M}/S;  ctx = sum_r (e^{m_r-M}/S) pctx_r
  float scale[ATT_MAX_CL];
  {
    float mr[ATT_MAX_CL], sr[ATT_MAX_CL], M = -INFINITY, S = 0.f;
#pragma unroll
    for (int rr = 0; rr < ATT_MAX_CL; ++rr) {
      if (rr < CL) {
        const float* st = cluster.map_shared_rank(stats, rr);
        mr[rr] = st[0]; sr[rr] = st[1];
        M = fmaxf(M, mr[rr]);
      }
    }
#pragma unroll
    for (int rr = 0; rr < ATT_MAX_CL; ++rr)
      if (rr < CL) { mr[rr] = (sr[rr] > 0.f) ? expf(mr[rr] - M) : 0.f; S = fmaf(sr[rr], mr[rr], S); }
    const float inv = 1.f / S;
#pragma unroll
    for (int rr = 0; rr < ATT_MAX_CL; ++rr) scale[rr] = (rr < CL) ? mr[rr] * inv : 0.f;
  }
  for (int i = tid; i < nl; i += ATT_THREADS)
    a.alpha[(int64_t)b * a.alpha_ld + l0 + i] = es[i] * scale[r];
  // this CTA finalises the float4 columns [r*per, (r+1)*per) of ctx / beta / gctx           (:21, :74-75)
  const int per = (D4 + CL - 1) / CL;
  for (int cc = tid; cc < per; cc += ATT_THREADS) {
    const int c = r * per + cc;
    if (c >= D4) break;
    float4 acc = f4_zero();
#pragma unroll
    for (int rr = 0; rr < ATT_MAX_CL; ++rr)
      if (rr < CL) f4_fma(acc, scale[rr], cluster.map_shared_rank(pctx, rr)[c]);
    const float4 bp = __ldg(reinterpret_cast<const float4*>(a.bpre + (int64_t)b * a.bpre_ld) + c);
    float4 be;
    be.x = sigmoid_f(bp.x); be.y = sigmoid_f(bp.y); be.z = sigmoid_f(bp.z); be.w = sigmoid_f(bp.w);
    reinterpret_cast<float4*>(a.ctx + (int64_t)b * D)[c] = acc;
    reinterpret_cast<float4*>(a.beta + (int64_t)b * D)[c] = be;
    reinterpret_cast<float4*>(a.gctx + (int64_t)b * a.gctx_ld)[c] =
        make_float4(be.x * acc.x, be.y * acc.y, be.z * acc.z, be.w * acc.w);
  }
  cluster.sync();                                            // keep shared memory alive until every peer has read it
}

cudaError_t attn_step_fwd(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                          const float* q, int64_t q_ld, const float* bpre, int64_t bpre_ld, const float* wv,
                          float* alpha, int64_t alpha_ld, float* ctx, float* beta, float* gctx, int64_t gctx_ld,
                          int rows_per_image) {
  if (attn_engine() == 1 && attn_stream_supported(D, H))
    return attn_step_fwd_stream(st, B, L, D, H, Ws, feat, q, q_ld, bpre, bpre_ld, wv, alpha, alpha_ld, ctx, beta, gctx,
                                gctx_ld, rows_per_image);
  if ((H & 3) || (D & 3) || H > 1024 || (q_ld & 3) || (bpre_ld & 3) || (gctx_ld & 3)) return cudaErrorInvalidValue;
  const int CL = pick_cluster(B, L);
  const int Lc = ceil_div(L, CL);
  AttnFwdP a{Ws, feat, q, q_ld, bpre, bpre_ld, wv, alpha, alpha_ld, ctx, beta, gctx, gctx_ld, L, D, H,
             rows_per_image < 1 ? 1 : rows_per_image, CL, Lc};
  const size_t smem = (size_t)(4 * ATT_THREADS + D + ((Lc + 3) & ~3) + 4 + 40) * sizeof(float);
  const int nj = ceil_div(H, 128);
#define LAUNCH_FWD(NJ)                                                             \
  do {                                                                             \
    ACK(set_smem(attn_fwd_kernel<NJ>, smem));                                      \
    return launch_cluster(attn_fwd_kernel<NJ>, B * CL, smem, CL, st, a);           \
  } while (0)
  if (nj <= 1) LAUNCH_FWD(1);
  else if (nj <= 2) LAUNCH_FWD(2);
  else if (nj <= 4) LAUNCH_FWD(4);
  else LAUNCH_FWD(8);
#undef LAUNCH_FWD
}

// ------------------------------------------------------------------------------------------ backward step
struct AttnBwdP {
  const float* Ws; const float* feat; const float* q; int64_t q_ld; const float* wv;
  const float* alpha; int64_t alpha_ld; const float* ctx; const float* beta; const float* dgctx; int64_t dgctx_ld;
  const float* ext; int64_t ext_ld; const float* reg_r;
  float* dq; int64_t dq_ld; float* dbpre; int64_t dbpre_ld; float* de; int64_t de_ld; float* dwv_part; float* dfeat;
  int L, D, H, CL, Lc;
};

template <int NJ>
__global__ void __launch_bounds__(ATT_THREADS) attn_bwd_kernel(const AttnBwdP a) {
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) float sm[];
  const int L = a.L, D = a.D, H = a.H, H4 = H >> 2, D4 = D >> 2, CL = a.CL;
  const int Lp = (a.Lc + 3) & ~3;
  float* dctx_s = sm;                    // [D]
  float* pdq = dctx_s + D;               // [H]  partial d q   of this CTA's rows
  float* pdv = pdq + H;                  // [H]  partial d w_v of this CTA's rows
  float* red = pdv + H;                  // [ATT_WARPS][H]
  float* al_s = red + ATT_WARPS * H;     // [Lp]
  float* da_s = al_s + Lp;               // [Lp]  d alpha, then d e
  float* stats = da_s + Lp;              // [4]   partial sum_l alpha*dalpha
  float* scratch = stats + 4;            // [40]

  const int r = (int)cluster.block_rank();
  const int b = blockIdx.x / CL, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int l0 = r * a.Lc, l1 = min(L, l0 + a.Lc), nl = max(l1 - l0, 0);

  // ---- A: gate backward.  gctx = beta*ctx, beta = sigmoid(bpre)          (AttentionLanguageModule.py:74-75)
  for (int c = tid; c < D4; c += ATT_THREADS) {
    const float4 dg = __ldg(reinterpret_cast<const float4*>(a.dgctx + (int64_t)b * a.dgctx_ld) + c);
    const float4 be = __ldg(reinterpret_cast<const float4*>(a.beta + (int64_t)b * D) + c);
    float4 dc;
    dc.x = dg.x * be.x; dc.y = dg.y * be.y; dc.z = dg.z * be.z; dc.w = dg.w * be.w;
    reinterpret_cast<float4*>(dctx_s)[c] = dc;
    if (r == 0) {
      const float4 cx = __ldg(reinterpret_cast<const float4*>(a.ctx + (int64_t)b * D) + c);
      float4 db;
      db.x = dg.x * cx.x * be.x * (1.f - be.x); db.y = dg.y * cx.y * be.y * (1.f - be.y);
      db.z = dg.z * cx.z * be.z * (1.f - be.z); db.w = dg.w * cx.w * be.w * (1.f - be.w);
      reinterpret_cast<float4*>(a.dbpre + (int64_t)b * a.dbpre_ld)[c] = db;
    }
  }
  for (int i = tid; i < nl; i += ATT_THREADS) al_s[i] = __ldg(a.alpha + (int64_t)b * a.alpha_ld + l0 + i);
  __syncthreads();

  // ---- B: d alpha[l] = dctx . feat[b,l,:] (+ upstream terms);  dfeat += alpha[l] * dctx        (:21)
  const float4* fb = reinterpret_cast<const float4*>(a.feat) + ((int64_t)b * L + l0) * D4;
  float4* dfb = a.dfeat ? reinterpret_cast<float4*>(a.dfeat) + ((int64_t)b * L + l0) * D4 : nullptr;
  for (int i = warp; i < nl; i += ATT_WARPS) {
    const float4* row = fb + (int64_t)i * D4;
    float s = 0.f;
    const float al = al_s[i];
#pragma unroll 4
    for (int f = lane; f < D4; f += 32) {
      const float4 x = ldg4(row + f);
      const float4 dc = reinterpret_cast<const float4*>(dctx_s)[f];
      s = fmaf(x.x, dc.x, s); s = fmaf(x.y, dc.y, s); s = fmaf(x.z, dc.z, s); s = fmaf(x.w, dc.w, s);
      if (dfb) {
        float4 o = dfb[(int64_t)i * D4 + f];
        f4_fma(o, al, dc);
        dfb[(int64_t)i * D4 + f] = o;
      }
    }
    s = warp_sum(s);
    if (lane == 0) {
      if (a.ext) s += __ldg(a.ext + (int64_t)b * a.ext_ld + l0 + i);
      if (a.reg_r) s += __ldg(a.reg_r + (int64_t)b * L + l0 + i);
      da_s[i] = s;
    }
  }
  __syncthreads();

  // ---- C: softmax backward  de = alpha * (dalpha - sum_L alpha*dalpha), the sum taken over the whole cluster    (:20)
  float part = 0.f;
  for (int i = tid; i < nl; i += ATT_THREADS) part = fmaf(al_s[i], da_s[i], part);
  part = block_sum(part, scratch);
  if (tid == 0) stats[0] = part;
  cluster.sync();
  float dot = 0.f;
#pragma unroll
  for (int rr = 0; rr < ATT_MAX_CL; ++rr)
    if (rr < CL) dot += cluster.map_shared_rank(stats, rr)[0];
  for (int i = tid; i < nl; i += ATT_THREADS) {
    const float de = al_s[i] * (da_s[i] - dot);
    da_s[i] = de;
    a.de[(int64_t)b * a.de_ld + l0 + i] = de;
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
  const float4* wsb = reinterpret_cast<const float4*>(a.Ws) + ((int64_t)b * L + l0) * H4;
  for (int i = warp; i < nl; i += 2 * ATT_WARPS) {
    const int i2 = i + ATT_WARPS;
    const bool has2 = i2 < nl;
    const float de0 = da_s[i], de1 = has2 ? da_s[i2] : 0.f;
    float4 w0[NJ], w1[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int f = lane + 32 * j;
      w0[j] = (f < H4) ? ldg4(wsb + (int64_t)i * H4 + f) : f4_zero();
      w1[j] = (f < H4 && has2) ? ldg4(wsb + (int64_t)i2 * H4 + f) : f4_zero();
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      float t;
#define ATT_BWD_ELEM(W, DE, X)                                         \
  t = tanhf(W[j].X + q4[j].X);                                         \
  dv4[j].X = fmaf(DE, t, dv4[j].X);                                    \
  dq4[j].X = fmaf(DE * v4[j].X, 1.f - t * t, dq4[j].X);
      ATT_BWD_ELEM(w0, de0, x) ATT_BWD_ELEM(w0, de0, y) ATT_BWD_ELEM(w0, de0, z) ATT_BWD_ELEM(w0, de0, w)
      ATT_BWD_ELEM(w1, de1, x) ATT_BWD_ELEM(w1, de1, y) ATT_BWD_ELEM(w1, de1, z) ATT_BWD_ELEM(w1, de1, w)
#undef ATT_BWD_ELEM
    }
  }
  // cross-warp reduction of dq then dw_v through red[ATT_WARPS][H] -> per-CTA partials pdq / pdv
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
    reinterpret_cast<float4*>(pdq)[f] = s;
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
    reinterpret_cast<float4*>(pdv)[f] = s;
  }
  cluster.sync();
  // this CTA finalises the float4 columns [r*per, (r+1)*per) of dq and dw_v (summing the cluster's partials via DSMEM)
  const int per = (H4 + CL - 1) / CL;
  for (int cc = tid; cc < per; cc += ATT_THREADS) {
    const int f = r * per + cc;
    if (f >= H4) break;
    float4 sq = f4_zero(), sv = f4_zero();
#pragma unroll
    for (int rr = 0; rr < ATT_MAX_CL; ++rr) {
      if (rr < CL) {
        f4_add(sq, reinterpret_cast<const float4*>(cluster.map_shared_rank(pdq, rr))[f]);
        f4_add(sv, reinterpret_cast<const float4*>(cluster.map_shared_rank(pdv, rr))[f]);
      }
    }
    reinterpret_cast<float4*>(a.dq + (int64_t)b * a.dq_ld)[f] = sq;
    float4* out = reinterpret_cast<float4*>(a.dwv_part + (int64_t)b * H) + f;
    float4 o = *out;
    f4_add(o, sv);
    *out = o;
  }
  cluster.sync();
}

cudaError_t attn_step_bwd(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                          const float* q, int64_t q_ld, const float* wv, const float* alpha, int64_t alpha_ld,
                          const float* ctx, const float* beta, const float* dgctx, int64_t dgctx_ld,
                          const float* dalpha_ext, int64_t ext_ld, const float* reg_r,
                          float* dq, int64_t dq_ld, float* dbpre, int64_t dbpre_ld, float* de, int64_t de_ld,
                          float* dwv_part, float* dfeat) {
  if (attn_engine() == 1 && attn_stream_supported(D, H))
    return attn_step_bwd_stream(st, B, L, D, H, Ws, feat, q, q_ld, wv, alpha, alpha_ld, ctx, beta, dgctx, dgctx_ld,
                                dalpha_ext, ext_ld, reg_r, dq, dq_ld, dbpre, dbpre_ld, de, de_ld, dwv_part, dfeat);
  if ((H & 3) || (D & 3) || H > 1024 || (q_ld & 3) || (dgctx_ld & 3) || (dq_ld & 3) || (dbpre_ld & 3))
    return cudaErrorInvalidValue;
  const int CL = pick_cluster(B, L);
  const int Lc = ceil_div(L, CL);
  AttnBwdP a{Ws, feat, q, q_ld, wv, alpha, alpha_ld, ctx, beta, dgctx, dgctx_ld, dalpha_ext, ext_ld, reg_r,
             dq, dq_ld, dbpre, dbpre_ld, de, de_ld, dwv_part, dfeat, L, D, H, CL, Lc};
  const int Lp = (Lc + 3) & ~3;
  const size_t smem = (size_t)(D + 2 * H + ATT_WARPS * H + 2 * Lp + 4 + 40) * sizeof(float);
  const int nj = ceil_div(H, 128);
#define LAUNCH_BWD(NJ)                                                             \
  do {                                                                             \
    ACK(set_smem(attn_bwd_kernel<NJ>, smem));                                      \
    return launch_cluster(attn_bwd_kernel<NJ>, B * CL, smem, CL, st, a);           \
  } while (0)
  if (nj <= 1) LAUNCH_BWD(1);
  else if (nj <= 2) LAUNCH_BWD(2);
  else if (nj <= 4) LAUNCH_BWD(4);
  else LAUNCH_BWD(8);
#undef LAUNCH_BWD
}

}  // namespace alexcap
