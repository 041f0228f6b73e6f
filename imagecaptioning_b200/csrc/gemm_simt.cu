// gemm_simt.cu -- fp32 CUDA-core GEMM tiles (exact fp32 FMA: the engine the 1e-4 fp32 parity bar is met with).
//   C[M,N] = act(alpha * op(A) op(B) + bias[N] + bias2[N]) (+ C)
// Replaces the cuBLAS sgemm calls behind nn.Linear / nn.LSTMCell / nn.LSTM in the reference
// (AlexCap/AttentionLanguageModule.py:16-17,74,78-79,119-123; AlexCap/LanguageModule.py:36,46,49) and the
// dX / dW GEMMs autograd issues for them.
// Layout: BMxBN output tile per CTA, BK=16 k-slab double-buffered in shared memory (k-major, so the inner loop
// reads float4 along m and n), 128-bit global loads when the operand's leading dimension allows it,
// (4*RM)x(4*RN) accumulators per thread.
#include <atomic>

#include "common.cuh"
#include "kernels.h"

namespace alexcap {

struct GemmP {
  const float* A; const float* B; float* C; const float* bias; const float* bias2;
  int M, N, K; int64_t lda, ldb, ldc; float alpha; int accumulate, act; int vecA, vecB, vecC;
};

constexpr int BK = 16;

// Load one BX x BK operand tile into registers.  KCONTIG: element (mn,k) at X[mn*ld + k]; else X[k*ld + mn].
template <int BX, int NT, bool KCONTIG>
struct TileLoader {
  static constexpr int NV = (BX * BK / 4) / NT;   // float4 per thread
  static_assert((BX * BK / 4) % NT == 0 && NV >= 1, "tile/thread mismatch");
  float4 r[NV];

  __device__ __forceinline__ void load(const float* __restrict__ X, int64_t ld, int mn0, int k0, int MN, int K, bool vec) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = threadIdx.x + i * NT;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (KCONTIG) {
        const int row = idx >> 2, kq = idx & 3;
        const int mn = mn0 + row, k = k0 + kq * 4;
        if (mn < MN && k < K) {
          const float* p = X + (int64_t)mn * ld + k;
          if (vec && k + 3 < K) {
            v = __ldg(reinterpret_cast<const float4*>(p));
          } else {
            v.x = __ldg(p);
            if (k + 1 < K) v.y = __ldg(p + 1);
            if (k + 2 < K) v.z = __ldg(p + 2);
            if (k + 3 < K) v.w = __ldg(p + 3);
          }
        }
      } else {
        constexpr int C4 = BX / 4;
        const int krow = idx / C4, c4 = idx % C4;
        const int k = k0 + krow, mn = mn0 + c4 * 4;
        if (k < K && mn < MN) {
          const float* p = X + (int64_t)k * ld + mn;
          if (vec && mn + 3 < MN) {
            v = __ldg(reinterpret_cast<const float4*>(p));
          } else {
            v.x = __ldg(p);
            if (mn + 1 < MN) v.y = __ldg(p + 1);
            if (mn + 2 < MN) v.z = __ldg(p + 2);
            if (mn + 3 < MN) v.w = __ldg(p + 3);
          }
        }
      }
      r[i] = v;
    }
  }
  // smem tile is [BK][BX+4]
  __device__ __forceinline__ void store(float* __restrict__ s) const {
    constexpr int LDS = BX + 4;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int idx = threadIdx.x + i * NT;
      if (KCONTIG) {
        const int row = idx >> 2, kq = idx & 3;
        s[(kq * 4 + 0) * LDS + row] = r[i].x;
        s[(kq * 4 + 1) * LDS + row] = r[i].y;
        s[(kq * 4 + 2) * LDS + row] = r[i].z;
        s[(kq * 4 + 3) * LDS + row] = r[i].w;
      } else {
        constexpr int C4 = BX / 4;
        const int krow = idx / C4, c4 = idx % C4;
        *reinterpret_cast<float4*>(&s[krow * LDS + c4 * 4]) = r[i];
      }
    }
  }
};

template <int BM, int BN, int RM, int RN, bool TA, bool TB>
__global__ void __launch_bounds__((BM / (4 * RM)) * (BN / (4 * RN)))
sgemm_kernel(const GemmP p) {
  constexpr int TXN = BN / (4 * RN), TYN = BM / (4 * RM), NT = TXN * TYN;
  constexpr int LDA = BM + 4, LDB = BN + 4;
  __shared__ __align__(16) float As[2][BK * LDA];
  __shared__ __align__(16) float Bs[2][BK * LDB];

  const int tx = threadIdx.x % TXN, ty = threadIdx.x / TXN;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

  // A: TA=false -> [M,K] k-contiguous; TA=true -> stored [K,M].  B: TB=true -> [N,K] k-contiguous; TB=false -> [K,N].
  TileLoader<BM, NT, !TA> la;
  TileLoader<BN, NT, TB> lb;

  float acc[4 * RM][4 * RN];
#pragma unroll
  for (int i = 0; i < 4 * RM; ++i)
#pragma unroll
    for (int j = 0; j < 4 * RN; ++j) acc[i][j] = 0.f;

  const int nk = (p.K + BK - 1) / BK;
  la.load(p.A, p.lda, m0, 0, p.M, p.K, p.vecA);
  lb.load(p.B, p.ldb, n0, 0, p.N, p.K, p.vecB);
  la.store(As[0]);
  lb.store(Bs[0]);
  __syncthreads();

  int buf = 0;
  for (int kt = 0; kt < nk; ++kt) {
    const bool more = (kt + 1) < nk;
    if (more) {
      la.load(p.A, p.lda, m0, (kt + 1) * BK, p.M, p.K, p.vecA);
      lb.load(p.B, p.ldb, n0, (kt + 1) * BK, p.N, p.K, p.vecB);
    }
    const float* as = As[buf];
    const float* bs = Bs[buf];
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[4 * RM], b[4 * RN];
#pragma unroll
      for (int r = 0; r < RM; ++r) {
        const float4 v = *reinterpret_cast<const float4*>(&as[k * LDA + r * (BM / RM) + ty * 4]);
        a[4 * r] = v.x; a[4 * r + 1] = v.y; a[4 * r + 2] = v.z; a[4 * r + 3] = v.w;
      }
#pragma unroll
      for (int r = 0; r < RN; ++r) {
        const float4 v = *reinterpret_cast<const float4*>(&bs[k * LDB + r * (BN / RN) + tx * 4]);
        b[4 * r] = v.x; b[4 * r + 1] = v.y; b[4 * r + 2] = v.z; b[4 * r + 3] = v.w;
      }
#pragma unroll
      for (int i = 0; i < 4 * RM; ++i)
#pragma unroll
        for (int j = 0; j < 4 * RN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (more) {
      la.store(As[buf ^ 1]);
      lb.store(Bs[buf ^ 1]);
    }
    __syncthreads();
    buf ^= 1;
  }

  // epilogue
#pragma unroll
  for (int rm = 0; rm < RM; ++rm) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + rm * (BM / RM) + ty * 4 + i;
      if (m >= p.M) continue;
#pragma unroll
      for (int rn = 0; rn < RN; ++rn) {
        const int n = n0 + rn * (BN / RN) + tx * 4;
        if (n >= p.N) continue;
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float x = p.alpha * acc[rm * 4 + i][rn * 4 + j];
          if (n + j < p.N) {
            if (p.bias) x += __ldg(p.bias + n + j);
            if (p.bias2) x += __ldg(p.bias2 + n + j);
          }
          v[j] = x;
        }
        float* cp = p.C + (int64_t)m * p.ldc + n;
        if (p.vecC && n + 3 < p.N) {
          if (p.accumulate) {
            const float4 c = *reinterpret_cast<const float4*>(cp);
            v[0] += c.x; v[1] += c.y; v[2] += c.z; v[3] += c.w;
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            v[j] = act_apply(v[j], p.act);
          }
          *reinterpret_cast<float4*>(cp) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (n + j < p.N) {
              float x = v[j];
              if (p.accumulate) x += cp[j];
              x = act_apply(x, p.act);
              cp[j] = x;
            }
          }
        }
      }
    }
  }
}

template <int BM, int BN, int RM, int RN>
static cudaError_t launch_cfg(cudaStream_t st, bool ta, bool tb, const GemmP& p) {
  constexpr int NT = (BM / (4 * RM)) * (BN / (4 * RN));
  dim3 grid(ceil_div(p.N, BN), ceil_div(p.M, BM));
  ALEXCAP_NOTE_LAUNCH();
  if (!ta && tb) sgemm_kernel<BM, BN, RM, RN, false, true><<<grid, NT, 0, st>>>(p);
  else if (!ta && !tb) sgemm_kernel<BM, BN, RM, RN, false, false><<<grid, NT, 0, st>>>(p);
  else if (ta && !tb) sgemm_kernel<BM, BN, RM, RN, true, false><<<grid, NT, 0, st>>>(p);
  else sgemm_kernel<BM, BN, RM, RN, true, true><<<grid, NT, 0, st>>>(p);
  return cudaGetLastError();
}

int& gemm_cta_cap() {
  static thread_local int cap = 0;
  return cap;
}
bool& gemm_b_is_weight() {
  static thread_local bool on = false;
  return on;
}
// 0 = fp32 CUDA-core tiles; 1 = tcgen05 3xTF32 (2^-21: decoding always uses it); 2 = tcgen05 bf16 (mixed precision);
// 3 = tcgen05 bf16x3 (hi/lo bf16 pairs, 2^-17, 1.3x faster than 3xTF32: the training default).  Per-call fallback to
// engine 0 when TMA cannot map the operands.  The process-wide default (alexcap_set_gemm_engine) is mirrored into a
// thread-local value, which is what the scoped per-call overrides (DtypeScope, ExactScope) change: concurrent calls on
// different host threads do not see each other's overrides.
static std::atomic<int> g_engine_default{3};
static std::atomic<unsigned> g_engine_gen{0};
int& gemm_engine() {
  static thread_local int eng = 3;
  static thread_local unsigned gen = ~0u;
  const unsigned g = g_engine_gen.load(std::memory_order_acquire);
  if (gen != g) { eng = g_engine_default.load(std::memory_order_relaxed); gen = g; }
  return eng;
}
void gemm_engine_set_default(int engine) {
  g_engine_default.store(engine, std::memory_order_relaxed);
  g_engine_gen.fetch_add(1, std::memory_order_release);
}

// Engine dispatch used by every GEMM on the hot path.
cudaError_t gemm_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                     const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                     const float* bias, const float* bias2, bool accumulate, int act) {
  // engine 1: tcgen05 3xTF32 (fp32-grade); engine 2: tcgen05 bf16 operands / fp32 accumulate (mixed-precision configs)
  if (gemm_engine() >= 1 && gemm_tc_supported(M, N, K, A, lda, B, ldb, C))
    return gemm_tc_f32(st, ta, tb, M, N, K, alpha, A, lda, B, ldb, C, ldc, bias, bias2, accumulate, act, 0,
                       gemm_engine() == 2 ? 1 : gemm_engine() == 3 ? 2 : 0);
  // an operand that exists only as producer-written bf16 copies cannot be read as fp32: loud, never silent garbage
  if (split_arena_is_produced(split_arena(), A) || split_arena_is_produced(split_arena(), B)) return cudaErrorInvalidValue;
  return gemm_simt_f32(st, ta, tb, M, N, K, alpha, A, lda, B, ldb, C, ldc, bias, bias2, accumulate, act);
}

cudaError_t gemm_simt_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                          const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                          const float* bias, const float* bias2, bool accumulate, int act) {
  if (M <= 0 || N <= 0) return cudaSuccess;
  GemmP p;
  p.A = A; p.B = B; p.C = C; p.bias = bias; p.bias2 = bias2;
  p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = ldb; p.ldc = ldc; p.alpha = alpha;
  p.accumulate = accumulate ? 1 : 0; p.act = act;
  p.vecA = (lda % 4 == 0) && aligned16(A);
  p.vecB = (ldb % 4 == 0) && aligned16(B);
  p.vecC = (ldc % 4 == 0) && aligned16(C);
  // pick the largest tile that still fills the 148 SMs
  auto ctas = [&](int bm, int bn) { return (int64_t)ceil_div(M, bm) * ceil_div(N, bn); };
  if (ctas(128, 128) >= 120) return launch_cfg<128, 128, 2, 2>(st, ta, tb, p);
  if (ctas(64, 64) >= 120) return launch_cfg<64, 64, 1, 1>(st, ta, tb, p);
  if (ctas(32, 64) >= 120) return launch_cfg<32, 64, 1, 1>(st, ta, tb, p);
  return launch_cfg<32, 32, 1, 1>(st, ta, tb, p);
}

}  // namespace alexcap
