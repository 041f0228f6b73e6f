// gemm_tc.cu -- tcgen05 / TMEM / TMA GEMM for sm_100a with fp32 in, fp32 out and fp32-grade accuracy (3xTF32).
//
//   C[M,N] = act(alpha * op(A) op(B) + bias + bias2) (+ C)         same contract as gemm_f32 (gemm_simt.cu)
//
// Why 3xTF32: BASELINE.json's fp32 config demands 1e-4 relative parity on losses AND gradients; a single
// kind::tf32 MMA keeps 10 mantissa bits (~1e-3).  Each fp32 operand x is split on the fly in shared memory into
// hi = tf32_round(x) and lo = tf32_round(x - hi), and every k-step issues
//     D += A_lo*B_hi;  D += A_hi*B_lo;  D += A_hi*B_hi          (fp32 accumulate in TMEM)
// which leaves a ~2^-21 relative error.  Reported FLOPs are the algorithmic 2MNK, not 3x.
//
// Structure (one 128 x BN output tile per CTA, optional split-K over gridDim.z):
//   warp 8      : TMA producer  -- cp.async.bulk.tensor.2d, 128B-swizzled tiles, STAGES-deep mbarrier ring
//   warps 0..7  : hi/lo splitter (generic-proxy smem pass, fence.proxy.async) and, at the end, the epilogue:
//                 tcgen05.ld 32x32b -> registers -> smem transpose -> coalesced global stores / red.add
//   warp 9      : TMEM allocator + single-thread tcgen05.mma issuer; tcgen05.commit frees smem stages
// Operands may be K-major or MN-major in global memory (NT / NN / TN GEMMs without any transposition pass):
//   K-major  tile: one TMA box {32 k, rows}, swizzle 128B   -> UMMA K-major SWIZZLE_128B layout (SBO = 1024 B)
//   MN-major tile: rows/32 TMA boxes {32 mn, 32 k}, swizzle 128B_ATOM_32B -> UMMA MN-major SWIZZLE_128B_BASE32B layout,
//                  the only MN-major layout tf32 operands have (LBO = 4096 B, SBO = 512 B)
// TMA zero-fills out-of-bounds elements, so M, N and K need no padding.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"
#include "tc_ptx.cuh"

namespace alexcap {

namespace tc {

constexpr int BM = 128;
constexpr int BK = 32;            // fp32 elements per k-block = one 128-byte swizzle row
constexpr int UMMA_K = 8;         // kind::tf32
constexpr int CONV_WARPS = 16;                // hi/lo splitter + epilogue warps
constexpr int CONV_THREADS = CONV_WARPS * 32;
constexpr int TMA_WARP = CONV_WARPS, MMA_WARP = CONV_WARPS + 1;
constexpr int NUM_THREADS = (CONV_WARPS + 2) * 32;

struct Params {
  float* C; const float* bias; const float* bias2;
  int M, N, K; int64_t ldc; float alpha; int accumulate, act;
  int kb_per_split, num_kb;
  int dbg_noepi;   // debug (ALEXCAP_DBG_NOEPI=1): skip the split-K red.add stores (timing experiments only; results are wrong)
  int b_const;     // B was written long before the previous kernel (weights): its first tiles may be fetched before pdl_wait
  int red_v4;      // accumulate with red.global.add.v4.f32 (16-byte vector atomics): a quarter of the red instructions
  int bulk_red;    // split-K / accumulate epilogue through shared memory + cp.reduce.async.bulk (.add.f32) instead of red.add
  unsigned long long* prof;   // debug (ALEXCAP_TC_PROF_PTR): phase time stamps of CTA (0,0,0) of the LA kernels
  LstmA la;        // LA kernels: the A operand is the LSTM cell output computed from these (see kernels.h)
};
// MN-major fp32/tf32 tiles (probed on B200, scripts/debug_tc_mn.py): boxes of {32 mn, 32 k} = 32 rows of 128 bytes,
// TMA swizzle 128B_ATOM_32B, UMMA layout SWIZZLE_128B_BASE32B: LBO = 4096 (next 32-wide mn block), SBO = 512 (next group
// of 4 k-rows); one K=8 MMA consumes 8 k-rows = 1024 bytes.
constexpr uint32_t MN_LBO = 4096, MN_SBO = 512, MN_KSTEP = 1024;

// fp32 tile [ROWS][32 k] as TMA delivered it -> bf16 tile, k-chunk-major core matrices: (r, k) at (k/8)*ROWS*16 + r*16 + (k%8)*2.
// K-major source: 128-byte rows, 16-byte chunk c of row r stored at chunk c ^ (r & 7) (SWIZZLE_128B).
// MN-major source: boxes of {32 mn, 32 k}, un-swizzled: (k, mn) at (mn/32)*4096 + k*128 + (mn%32)*4.
template <int ROWS, bool MN, bool WITH_LO>
__device__ __forceinline__ void convert_bf16(const uint8_t* src, uint8_t* dst, uint8_t* dst_lo, int t) {
  for (int idx = t; idx < ROWS * 4; idx += CONV_THREADS) {
    const int r = idx % ROWS, pch = idx / ROWS;          // row, 8-wide k chunk
    float v[8];
    if (!MN) {
      const uint8_t* row = src + r * 128;
      const float4 x0 = *reinterpret_cast<const float4*>(row + (((2 * pch) ^ (r & 7)) << 4));
      const float4 x1 = *reinterpret_cast<const float4*>(row + (((2 * pch + 1) ^ (r & 7)) << 4));
      v[0] = x0.x; v[1] = x0.y; v[2] = x0.z; v[3] = x0.w; v[4] = x1.x; v[5] = x1.y; v[6] = x1.z; v[7] = x1.w;
    } else {
      const float* col = reinterpret_cast<const float*>(src + (r >> 5) * 4096 + (r & 31) * 4) + pch * 8 * 32;
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = col[j * 32];
    }
    uint4 o;
    o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]); o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
    if (WITH_LO) {            // lo = bf16(x - float(hi)): x = hi + lo up to 2^-17 |x|
      uint4 l;
#define LO_PAIR(W, A, B) l.W = pack_bf16(A - __uint_as_float(o.W << 16), B - __uint_as_float(o.W & 0xffff0000u))
      LO_PAIR(x, v[0], v[1]); LO_PAIR(y, v[2], v[3]); LO_PAIR(z, v[4], v[5]); LO_PAIR(w, v[6], v[7]);
#undef LO_PAIR
      // the source chunk pair is dead once v[] is loaded: the lo tile overwrites the fp32 tile's first half in place?  No --
      // other threads still read it; the lo tile goes to its own region.
      *reinterpret_cast<uint4*>(dst_lo + pch * (ROWS * 16) + r * 16) = make_uint4(l.x, l.y, l.z, l.w);
    }
    *reinterpret_cast<uint4*>(dst + pch * (ROWS * 16) + r * 16) = o;
  }
}
// fp32 -> (hi, lo) with both parts exactly representable in TF32 (low 13 mantissa bits zero), round-to-nearest at both
// levels so the residual error (~2^-22 |x|) is unbiased; the tensor core would otherwise TRUNCATE the low bits.
__device__ __forceinline__ void split_tf32(uint32_t x, uint32_t& hi, uint32_t& lo) {
  hi = (x + 0x1000u) & 0xffffe000u;
  const float r = __uint_as_float(x) - __uint_as_float(hi);      // exact in fp32
  lo = (__float_as_uint(r) + 0x1000u) & 0xffffe000u;
}

// A tile of the fused LSTM -> GEMM kernels: thread t owns (row r = t / 4, 8 hidden units pch = t % 4) of the 128 x 32
// k-block at (m0, k0); `ld` phase = global loads into registers, `st` phase = cell math, optional global stores (the CTAs of
// the first column tile), bf16 hi (+ lo) chunks in the converter's k-chunk-major layout.
struct CellRegs { float4 v[10]; };     // i, f, g, o, c_prev: two float4 each
__device__ __forceinline__ void lstm_a_load(const Params& p, int m0, int k0, int t, CellRegs& x) {
  const int r = t >> 2, pch = t & 3, m = m0 + r, u = k0 + pch * 8;      // 4 lanes cover one 128-byte row segment
  const int H = p.K;
  if (m < p.M && u < H) {
    const float* g = p.la.g + (int64_t)m * p.la.g_ld + u;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      x.v[2 * q] = *reinterpret_cast<const float4*>(g + (int64_t)q * H);
      x.v[2 * q + 1] = *reinterpret_cast<const float4*>(g + (int64_t)q * H + 4);
    }
    const float* c = p.la.c_prev + (int64_t)m * H + u;
    x.v[8] = *reinterpret_cast<const float4*>(c);
    x.v[9] = *reinterpret_cast<const float4*>(c + 4);
  }
}
template <bool WITH_LO>
__device__ __forceinline__ void lstm_a_store(const Params& p, int m0, int k0, int t, CellRegs& x, uint8_t* dst, uint8_t* dst_lo,
                                             bool writer) {
  const int r = t >> 2, pch = t & 3, m = m0 + r, u = k0 + pch * 8;      // 4 lanes cover one 128-byte row segment
  const int H = p.K;
  float h[8];
  if (m < p.M && u < H) {
    float* gi = &x.v[0].x; float* gf = &x.v[2].x; float* gg = &x.v[4].x; float* go = &x.v[6].x; float* cp = &x.v[8].x;
#pragma unroll
    for (int j = 0; j < 8; ++j) {          // the cell of lstm_fwd_kernel (lstm_pointwise.cu) with two-MUFU sigmoid / tanh: this
      // CTA's 4096 cells sit on the critical path of the GEMM (libm tanhf / expf / division: 8 us; these: ~2 us)
      gi[j] = sigmoid_fast(gi[j]); gf[j] = sigmoid_fast(gf[j]); gg[j] = tanh_fast2(gg[j]); go[j] = sigmoid_fast(go[j]);
      cp[j] = fmaf(gf[j], cp[j], gi[j] * gg[j]);
      h[j] = go[j] * tanh_fast2(cp[j]);
    }
    if (writer) {
      float* a = p.la.acts + (int64_t)m * p.la.acts_ld + u;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        *reinterpret_cast<float4*>(a + (int64_t)q * H) = x.v[2 * q];
        *reinterpret_cast<float4*>(a + (int64_t)q * H + 4) = x.v[2 * q + 1];
      }
      float* c = p.la.c_out + (int64_t)m * H + u;
      *reinterpret_cast<float4*>(c) = x.v[8];
      *reinterpret_cast<float4*>(c + 4) = x.v[9];
      float* ho = p.la.h_out + (int64_t)m * p.la.h_ld + u;
      *reinterpret_cast<float4*>(ho) = make_float4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<float4*>(ho + 4) = make_float4(h[4], h[5], h[6], h[7]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) h[j] = 0.f;
  }
  uint4 o;
  o.x = pack_bf16(h[0], h[1]); o.y = pack_bf16(h[2], h[3]); o.z = pack_bf16(h[4], h[5]); o.w = pack_bf16(h[6], h[7]);
  if (WITH_LO) {
    uint4 l;
#define LO_PAIR(W, A, B) l.W = pack_bf16(A - __uint_as_float(o.W << 16), B - __uint_as_float(o.W & 0xffff0000u))
    LO_PAIR(x, h[0], h[1]); LO_PAIR(y, h[2], h[3]); LO_PAIR(z, h[4], h[5]); LO_PAIR(w, h[6], h[7]);
#undef LO_PAIR
    *reinterpret_cast<uint4*>(dst_lo + pch * (BM * 16) + r * 16) = l;
  }
  *reinterpret_cast<uint4*>(dst + pch * (BM * 16) + r * 16) = o;
}

// ------------------------------------------------------------------------------------------------ kernel
// MODE 0: 3xTF32 (fp32-grade accuracy).  MODE 1: bf16 operands, fp32 accumulate (the mixed-precision configs): the
// converter warps round the fp32 tile to bf16 and re-lay it out K-major / no-swizzle (core matrices of 8 rows x 16 bytes)
// in the slot the 3xTF32 path uses for the `lo` tile; one kind::f16 MMA per 16 k replaces three kind::tf32 per 8 k.
// LA (MODE >= 1, K-major A): no A tiles are loaded; the converter warps compute them (lstm_a_load / lstm_a_store).
template <int BN, bool A_MN, bool B_MN, int STAGES, int MODE, bool LA = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  constexpr uint32_t A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
  constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;    // A_hi, A_lo, B_hi, B_lo
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // manual 1024-byte alignment (swizzle atoms); the launcher over-allocates by 1 KB
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* conv_bar = full_bar + STAGES;
  uint64_t* empty_bar = conv_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
#define TC_STAMP(slot)                                                                                              \
  do {                                                                                                              \
    if (LA && p.prof && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {                \
      unsigned long long t_;                                                                                        \
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_));                                                         \
      p.prof[slot] = t_;                                                                                            \
    }                                                                                                               \
  } while (0)
  TC_STAMP(0);
  const int kb0 = blockIdx.z * p.kb_per_split;
  const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
  const int nkb = kb1 - kb0;     // >= 1 by construction

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&conv_bar[s], CONV_THREADS);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == MMA_WARP) {   // TMEM allocation (power of two >= 32 columns); the same warp frees it
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(BN));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (warp == TMA_WARP && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();      // the next kernel in the stream may start its own prologue
  // Weight operand (per-step GEMMs): the B tiles of the first ring pass do not depend on the previous kernel, so their
  // L2 -> smem latency is put under the previous kernel's tail as well; the A tiles follow after the wait.
  const int npre = p.b_const ? min(nkb, STAGES) : 0;
  if (npre > 0 && warp == TMA_WARP && lane == 0) {
    for (int i = 0; i < npre; ++i) {
      uint8_t* sb = smem + i * STAGE_BYTES + 2 * A_BYTES;
      mbar_expect_tx(&full_bar[i], (LA ? 0u : A_BYTES) + B_BYTES);
      const int k0 = (kb0 + i) * BK;
      if (B_MN) {
#pragma unroll
        for (int j = 0; j < BN / 32; ++j) tma_load_2d(sb + j * 4096, &tmB, n0 + 32 * j, k0, &full_bar[i]);
      } else {
        tma_load_2d(sb, &tmB, k0, n0, &full_bar[i]);
      }
    }
  }
  pdl_wait();         // everything above overlapped the previous kernel's tail; its results are visible from here on
  TC_STAMP(1);

  if (warp == TMA_WARP) {
    // ================================================================ TMA producer
    if (lane == 0) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        const bool pre = i < npre;                       // B tile already in flight, barrier already armed
        if (!pre) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          mbar_expect_tx(&full_bar[s], (LA ? 0u : A_BYTES) + B_BYTES);
        }
        uint8_t* st = smem + s * STAGE_BYTES;
        const int k0 = (kb0 + i) * BK;
        if (LA) {
        } else if (A_MN) {
#pragma unroll
          for (int j = 0; j < BM / 32; ++j) tma_load_2d(st + j * 4096, &tmA, m0 + 32 * j, k0, &full_bar[s]);
        } else {
          tma_load_2d(st, &tmA, k0, m0, &full_bar[s]);
        }
        uint8_t* sb = st + 2 * A_BYTES;
        if (pre) continue;
        if (B_MN) {
#pragma unroll
          for (int j = 0; j < BN / 32; ++j) tma_load_2d(sb + j * 4096, &tmB, n0 + 32 * j, k0, &full_bar[s]);
        } else {
          tma_load_2d(sb, &tmB, k0, n0, &full_bar[s]);
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ================================================================ MMA issuer (one elected thread)
    // instruction descriptor: D=f32, A=B=tf32, majors, N>>3, M>>4   (cute::UMMA::InstrDescriptor bit layout)
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    for (int i = 0; i < nkb; ++i) {
      const int s = i % STAGES;
      const uint32_t ph = (i / STAGES) & 1;
      mbar_wait(&conv_bar[s], ph);
      tcgen05_fence_after();
      if (lane == 0) {
        const uint32_t a_hi = smem_u32(smem + s * STAGE_BYTES), a_lo = a_hi + A_BYTES;
        const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + B_BYTES;
        if (MODE >= 1) {
          // bf16 tiles: k-chunk-major core matrices.  LBO = stride between 16-byte k chunks = rows*16, SBO = 128 (next 8 rows)
          const uint32_t idesc16 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = make_desc(a_lo + k * 2 * (BM * 16), BM * 16, 128, 0);
            const uint64_t db = make_desc(b_lo + k * 2 * (BN * 16), BN * 16, 128, 0);
            if (MODE == 2) {     // bf16x3: x = hi + lo (16 significand bits); lo*hi + hi*lo + hi*hi, fp32 accumulate
              const uint64_t dal = make_desc(a_lo + A_BYTES / 2 + k * 2 * (BM * 16), BM * 16, 128, 0);
              const uint64_t dbl = make_desc(b_lo + B_BYTES / 2 + k * 2 * (BN * 16), BN * 16, 128, 0);
              umma_bf16(tmem_base, dal, db, idesc16, (i | k) != 0);
              umma_bf16(tmem_base, da, dbl, idesc16, 1);
              umma_bf16(tmem_base, da, db, idesc16, 1);
            } else {
              umma_bf16(tmem_base, da, db, idesc16, (i | k) != 0);
            }
          }
        } else {
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // K-major: +32 bytes inside the swizzled 128-byte row; MN-major: next group of 8 k-rows (+1024 bytes)
          const uint32_t a_off = A_MN ? k * MN_KSTEP : k * 32, b_off = B_MN ? k * MN_KSTEP : k * 32;
          // MN-major fp32/tf32 operands only exist in the SWIZZLE_128B_BASE32B layout (32-byte swizzle atoms, 4 k-rows
          // of 128 bytes per atom): LBO = stride between 32-element mn blocks, SBO = stride between 4-row k groups.
          const uint64_t dah = A_MN ? make_desc(a_hi + a_off, MN_LBO, MN_SBO, 1) : make_desc(a_hi + a_off, 0, 1024);
          const uint64_t dal = A_MN ? make_desc(a_lo + a_off, MN_LBO, MN_SBO, 1) : make_desc(a_lo + a_off, 0, 1024);
          const uint64_t dbh = B_MN ? make_desc(b_hi + b_off, MN_LBO, MN_SBO, 1) : make_desc(b_hi + b_off, 0, 1024);
          const uint64_t dbl = B_MN ? make_desc(b_lo + b_off, MN_LBO, MN_SBO, 1) : make_desc(b_lo + b_off, 0, 1024);
          umma_tf32(tmem_base, dal, dbh, idesc, (i | k) != 0);
          umma_tf32(tmem_base, dah, dbl, idesc, 1);
          umma_tf32(tmem_base, dah, dbh, idesc, 1);
        }
        }
        umma_commit(&empty_bar[s]);                       // smem stage reusable once these MMAs retire
        if (i == nkb - 1) umma_commit(tmem_full_bar);     // accumulator complete
      }
      __syncwarp();
    }
  } else {
    // ================================================================ warps 0..3: splitter, then epilogue
    const int t = threadIdx.x;   // 0..CONV_THREADS-1
    for (int i = 0; i < nkb; ++i) {
      const int s = i % STAGES;
      const uint32_t ph = (i / STAGES) & 1;
      CellRegs cell;
      if (LA) lstm_a_load(p, m0, (kb0 + i) * BK, t, cell);      // (in flight while the weight tile lands)
      mbar_wait(&full_bar[s], ph);
      TC_STAMP(2);
      uint8_t* st = smem + s * STAGE_BYTES;
      if (MODE >= 1) {
        // bf16 tiles live in the `lo` slot: hi in its first half, (MODE 2) the bf16 residual in its second half
        if (LA) lstm_a_store<MODE == 2>(p, m0, (kb0 + i) * BK, t, cell, st + A_BYTES, st + A_BYTES + A_BYTES / 2, blockIdx.x == 0);
        else convert_bf16<BM, A_MN, MODE == 2>(st, st + A_BYTES, st + A_BYTES + A_BYTES / 2, t);
        convert_bf16<BN, B_MN, MODE == 2>(st + 2 * A_BYTES, st + 2 * A_BYTES + B_BYTES, st + 2 * A_BYTES + B_BYTES + B_BYTES / 2, t);
      } else {
      // elementwise, so it is oblivious to the swizzle: same byte offset in the hi and lo tiles
#pragma unroll
      for (int c = t; c < (int)(A_BYTES / 16); c += CONV_THREADS) {
        const uint4 x = *reinterpret_cast<uint4*>(st + c * 16);
        uint4 h, l;
        split_tf32(x.x, h.x, l.x); split_tf32(x.y, h.y, l.y); split_tf32(x.z, h.z, l.z); split_tf32(x.w, h.w, l.w);
        *reinterpret_cast<uint4*>(st + c * 16) = h;
        *reinterpret_cast<uint4*>(st + A_BYTES + c * 16) = l;
      }
      uint8_t* sb = st + 2 * A_BYTES;
#pragma unroll
      for (int c = t; c < (int)(B_BYTES / 16); c += CONV_THREADS) {
        const uint4 x = *reinterpret_cast<uint4*>(sb + c * 16);
        uint4 h, l;
        split_tf32(x.x, h.x, l.x); split_tf32(x.y, h.y, l.y); split_tf32(x.z, h.z, l.z); split_tf32(x.w, h.w, l.w);
        *reinterpret_cast<uint4*>(sb + c * 16) = h;
        *reinterpret_cast<uint4*>(sb + B_BYTES + c * 16) = l;
      }
      }
      fence_proxy_async();            // generic-proxy writes -> visible to the tensor core (async proxy)
      mbar_arrive(&conv_bar[s]);
      TC_STAMP(3);
    }

    // ---- epilogue: TMEM -> registers -> smem transpose -> coalesced global
    mbar_wait(tmem_full_bar, 0);
    TC_STAMP(4);
    tcgen05_fence_after();
    float* tile = reinterpret_cast<float*>(smem) + warp * (32 * 33);   // stage memory is idle now (all MMAs retired)
    const bool first_split = (blockIdx.z == 0);
    // accumulate without an activation: red.add instead of load->add->store (see gemm_ps.cu: serialised dependent loads)
    const bool atomic = gridDim.z > 1 || (p.accumulate && p.act == ACT_NONE && (int64_t)p.M * p.N <= (int64_t)(4 << 20));
    const int q4 = warp & 3;                       // TMEM lane quarter this warp may read (warp id mod 4)
    if (atomic && p.bulk_red) {
      // Accumulating epilogue without per-element atomics: the CTA's [128 x BN] tile is parked in shared memory (row pitch
      // BN + 4 floats: 16-byte stores of 32 different rows spread over all banks) and every row leaves as ONE bulk
      // reduction (cp.reduce.async.bulk ... .add.f32, performed by the TMA unit at L2): 128 instructions per CTA instead of
      // 128 x BN / 32 warp-wide red.add, whose issue rate is what bounded the epilogue (3.4 us for a 128 x 128 tile).
      constexpr int PITCH = BN + 4;
      float* ctile = reinterpret_cast<float*>(smem);
      for (int cc = (warp >> 2); cc < BN / 32; cc += CONV_WARPS / 4) {
        uint32_t v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(cc * 32);
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
              "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
              "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
              "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        float* row = ctile + (q4 * 32 + lane) * PITCH + cc * 32;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          float4 o;
          float* po = &o.x;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int n = n0 + cc * 32 + j + e;
            float bsum = 0.f;
            if (first_split && n < p.N) {
              if (p.bias) bsum += __ldg(p.bias + n);
              if (p.bias2) bsum += __ldg(p.bias2 + n);
            }
            po[e] = fmaf(p.alpha, __uint_as_float(v[j + e]), bsum);
          }
          *reinterpret_cast<float4*>(row + j) = o;
        }
      }
      tcgen05_fence_before();
      fence_proxy_async();                                    // generic-proxy smem writes -> the bulk copies' async-proxy reads
      asm volatile("bar.sync 1, %0;" ::"n"(CONV_THREADS) : "memory");
      if (t < BM) {
        const int m = m0 + t;
        const int ncols = min(BN, p.N - n0);
        if (m < p.M && ncols > 0 && !p.dbg_noepi) {
          asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                       ::"l"(p.C + (int64_t)m * p.ldc + n0), "r"(smem_u32(ctile + t * PITCH)), "r"((uint32_t)ncols * 4u) : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
      }
    } else
#pragma unroll 1
    for (int cc = (warp >> 2); cc < BN / 32; cc += CONV_WARPS / 4) {
      uint32_t v[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(cc * 32);
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
            "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
            "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
          : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (atomic && p.red_v4) {
        // 16-byte vector atomics: the warp's 32 x 32 chunk goes through a pitch-36 tile (16-byte stores and loads both
        // conflict-free) and leaves as 8 instructions of 4 rows x 128 bytes instead of 32 scalar red.add per thread --
        // the epilogue of these latency-bound GEMMs is bound by the red ISSUE rate (~14 clk per warp-wide red).
        float* t36 = reinterpret_cast<float*>(smem) + warp * (32 * 36);
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(t36 + lane * 36 + j) =
              make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
        __syncwarp();
        const int cg = lane & 7, nv = n0 + cc * 32 + 4 * cg;
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (first_split && nv < p.N) {             // (N % 4 == 0: a column group is inside or outside as a whole)
          if (p.bias) { const float4 t4 = __ldg(reinterpret_cast<const float4*>(p.bias + nv)); b4.x += t4.x; b4.y += t4.y; b4.z += t4.z; b4.w += t4.w; }
          if (p.bias2) { const float4 t4 = __ldg(reinterpret_cast<const float4*>(p.bias2 + nv)); b4.x += t4.x; b4.y += t4.y; b4.z += t4.z; b4.w += t4.w; }
        }
        if (nv < p.N) {
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) {
            const int r = rr * 4 + (lane >> 3), m = m0 + q4 * 32 + r;
            if (m < p.M && !p.dbg_noepi) {
              const float4 x = *reinterpret_cast<const float4*>(t36 + r * 36 + 4 * cg);
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p.C + (int64_t)m * p.ldc + nv),
                           "f"(fmaf(p.alpha, x.x, b4.x)), "f"(fmaf(p.alpha, x.y, b4.y)), "f"(fmaf(p.alpha, x.z, b4.z)),
                           "f"(fmaf(p.alpha, x.w, b4.w)) : "memory");
            }
          }
        }
        __syncwarp();
        continue;
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) tile[lane * 33 + j] = __uint_as_float(v[j]);   // thread = row, j = column
      __syncwarp();
      const int n = n0 + cc * 32 + lane;
      float bsum = 0.f;
      if (n < p.N && first_split) {
        if (p.bias) bsum += __ldg(p.bias + n);
        if (p.bias2) bsum += __ldg(p.bias2 + n);
      }
      if (n < p.N) {
#pragma unroll 4
        for (int r = 0; r < 32; ++r) {
          const int m = m0 + q4 * 32 + r;
          if (m >= p.M) break;
          float x = fmaf(p.alpha, tile[r * 33 + lane], bsum);
          float* cp = p.C + (int64_t)m * p.ldc + n;
          if (atomic) {
            if (!p.dbg_noepi) atomicAdd(cp, x);      // C was zeroed (or holds the value to accumulate onto) by the launcher
          } else {
            if (p.accumulate) x += *cp;
            x = act_apply(x, p.act);
            *cp = x;
          }
        }
      }
      __syncwarp();
    }
    tcgen05_fence_before();
    TC_STAMP(5);
  }

  __syncthreads();
  if (warp == MMA_WARP) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(BN));
  }
}

// ------------------------------------------------------------------------------------------------ host side
// 2-D fp32 tensor map: inner (contiguous) extent `inner`, outer extent `outer`, row pitch ld floats, box {32, box_outer}.
static bool make_map(CUtensorMap* map, const float* base, int64_t inner, int64_t outer, int64_t ld, int box_outer,
                     bool mn_major, bool plain = false) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  cuuint32_t box[2] = {32u, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   plain ? CU_TENSOR_MAP_SWIZZLE_NONE : mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// the bulk-reduction epilogue (ALEXCAP_TC_BULK_RED=1) needs 16-byte aligned row segments.  OFF by default: measured slower
// than red.add on the cfg2 step (2.106 vs 2.078 ms) -- the CTA has to wait for its bulk group before it may exit, which puts
// the L2 round trip that fire-and-forget red.add hides back on the kernel's critical path.
static bool bulk_red_ok(const float* C, int64_t ldc, int N) {
  static const bool on = [] { const char* v = getenv("ALEXCAP_TC_BULK_RED"); return v && v[0] == '1'; }();
  return on && aligned16(C) && (ldc & 3) == 0 && (N & 3) == 0;
}

template <int BN, bool A_MN, bool B_MN, int MODE, bool LA = false>
static cudaError_t launch(cudaStream_t st, const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, dim3 grid) {
  constexpr int STAGES = (BN == 128) ? 3 : 4;
  constexpr size_t SMEM = (size_t)STAGES * (2 * BM * BK * 4 + 2 * BN * BK * 4) + 1024 /*align*/ + 256 /*barriers*/;
  static bool configured = false;
  auto kern = gemm_tc_kernel<BN, A_MN, B_MN, STAGES, MODE, LA>;
  if (!configured) {
    ACK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    configured = true;
  }
  ALEXCAP_NOTE_LAUNCH();
  return launch_pdl(kern, grid, dim3(NUM_THREADS), SMEM, st, ta, tb, p);
}

}  // namespace tc

bool gemm_tc_supported(int M, int N, int K, const float* A, int64_t lda, const float* B, int64_t ldb, const float* C) {
  (void)M; (void)N; (void)K; (void)C;
  return (lda % 4 == 0) && (ldb % 4 == 0) && aligned16(A) && aligned16(B) && tc::get_encode() != nullptr;
}

cudaError_t gemm_tc_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                        const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                        const float* bias, const float* bias2, bool accumulate, int act, int force_splits, int mode) {
  using namespace tc;
  if (M <= 0 || N <= 0 || K <= 0) return cudaSuccess;
  if (!gemm_tc_supported(M, N, K, A, lda, B, ldb, C)) return cudaErrorNotSupported;
  if (mode >= 1 && force_splits == 0 && split_arena() && gemm_ps_wanted(M, N, K)) {
    // big GEMM inside a training step: split the operands once in HBM, feed the tensor pipe by TMA alone (gemm_ps.cu)
    const cudaError_t e = gemm_ps_f32(st, ta, tb, M, N, K, alpha, A, lda, B, ldb, C, ldc, bias, bias2, accumulate, act, 0,
                                      mode == 2 ? 3 : 1, split_arena());
    if (e != cudaErrorMemoryAllocation) return e;       // arena full: convert in the kernel instead
  }
  // an operand whose bf16 copies were written directly by its producer has no fp32 data to convert: fail loudly
  if (split_arena_is_produced(split_arena(), A) || split_arena_is_produced(split_arena(), B)) return cudaErrorInvalidValue;
  const bool a_mn = ta;        // transa: A stored [K,M] -> M contiguous
  const bool b_mn = !tb;       // transb=0: B stored [K,N] -> N contiguous
  const int num_kb = ceil_div(K, BK);
  const int tiles128 = ceil_div(M, BM) * ceil_div(N, 128);
  static const int bn_env = [] { const char* v = getenv("ALEXCAP_TC_BN"); return v ? atoi(v) : 0; }();
  static const int minkb_env = [] { const char* v = getenv("ALEXCAP_TC_MINKB"); return v ? atoi(v) : 2; }();
  const int BN = (bn_env == 64 || bn_env == 128) ? bn_env : (tiles128 >= 120 || N > 64 * 1024 || (int64_t)N * K >= (1 << 20)) ? 128 : 64;
  // (latency-bound M = 128 GEMMs, measured in a PDL chain: the wide tile halves the A-tile re-reads from L2 once the weight
  //  matrix is >= 4 MB; two k-blocks per split are enough -- scripts/bench_gemm.py --small)
  const int tiles = ceil_div(M, BM) * ceil_div(N, BN);
  // split-K (fp32 red.add into C) when the output grid alone cannot fill 148 SMs; no split with an activation
  int splits = 1;
  if (force_splits > 0) splits = force_splits;
  else if (act == ACT_NONE && tiles < 100) {
    splits = 148 / tiles;
    const int max_by_k = num_kb / minkb_env > 0 ? num_kb / minkb_env : 1;     // keep >= 2 k-blocks per CTA
    if (splits > max_by_k) splits = max_by_k;
    if (splits < 1) splits = 1;
  }
  if (act != ACT_NONE) splits = 1;
  int kb_per = ceil_div(num_kb, splits);
  splits = ceil_div(num_kb, kb_per);       // no empty split
  if (splits > 1 && !accumulate) ACK(cudaMemset2DAsync(C, ldc * sizeof(float), 0, (size_t)N * sizeof(float), M, st));

  CUtensorMap mapA, mapB;
  const bool plain = (mode >= 1);      // the bf16 converter transposes MN-major boxes itself: un-swizzled boxes
  bool ok = a_mn ? make_map(&mapA, A, M, K, lda, BK, true, plain) : make_map(&mapA, A, K, M, lda, BM, false);
  ok = ok && (b_mn ? make_map(&mapB, B, N, K, ldb, BK, true, plain) : make_map(&mapB, B, K, N, ldb, BN, false));
  if (!ok) return cudaErrorInvalidValue;

  Params p;
  p.C = C; p.bias = bias; p.bias2 = bias2; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.alpha = alpha;
  p.accumulate = accumulate ? 1 : 0; p.act = act; p.kb_per_split = kb_per; p.num_kb = num_kb;
  p.b_const = (gemm_b_is_weight() && pdl_enabled()) ? 1 : 0;
  p.prof = nullptr;
  p.bulk_red = bulk_red_ok(C, ldc, N) ? 1 : 0;
  { static const int v4 = [] { const char* v = getenv("ALEXCAP_TC_RED_V4"); return (v && v[0] == '0') ? 0 : 1; }();
    p.red_v4 = (v4 && aligned16(C) && (ldc & 3) == 0 && (N & 3) == 0) ? 1 : 0; }
  { static const int noepi = [] { const char* v = getenv("ALEXCAP_DBG_NOEPI"); return (v && v[0] == '1') ? 1 : 0; }(); p.dbg_noepi = noepi; }
  dim3 grid(ceil_div(N, BN), ceil_div(M, BM), splits);
#define TC_DISPATCH(BN_, MODE_)                                                                 \
  do {                                                                                          \
    if (!a_mn && !b_mn) return launch<BN_, false, false, MODE_>(st, mapA, mapB, p, grid);       \
    if (!a_mn && b_mn) return launch<BN_, false, true, MODE_>(st, mapA, mapB, p, grid);         \
    if (a_mn && b_mn) return launch<BN_, true, true, MODE_>(st, mapA, mapB, p, grid);           \
    return launch<BN_, true, false, MODE_>(st, mapA, mapB, p, grid);                            \
  } while (0)
  if (mode == 1) {
    if (BN == 128) TC_DISPATCH(128, 1);
    TC_DISPATCH(64, 1);
  }
  if (mode == 2) {
    if (BN == 128) TC_DISPATCH(128, 2);
    TC_DISPATCH(64, 2);
  }
  if (BN == 128) TC_DISPATCH(128, 0);
  TC_DISPATCH(64, 0);
#undef TC_DISPATCH
}

// Fused LSTM cell -> GEMM (kernels.h: LstmA).  128-column tiles, split-K down to ONE k-block (32 hidden units) per CTA when the
// output grid is small: every column tile repeats the cell math of its k-blocks, so the fusion is only taken when that
// repetition is spread thin (<= 2 k-blocks per CTA).
bool gemm_tc_lstm_supported(int M, int N, int H) {
  using namespace tc;
  if (M <= 0 || N <= 0 || H <= 0 || (H & 7) || get_encode() == nullptr) return false;
  const int num_kb = ceil_div(H, BK);
  const int tiles = ceil_div(M, BM) * ceil_div(N, 128);
  int splits = 148 / tiles;
  if (splits > num_kb) splits = num_kb;
  if (splits < 1) splits = 1;
  return ceil_div(num_kb, splits) <= 2;
}

cudaError_t gemm_tc_lstm_f32(cudaStream_t st, int M, int N, int H, const LstmA& la, const float* B, int64_t ldb, float* C,
                             int64_t ldc, const float* bias, int mode) {
  using namespace tc;
  if (M <= 0 || N <= 0 || H <= 0) return cudaSuccess;
  if ((H & 7) || (ldb & 3) || (la.g_ld & 3) || (la.h_ld & 3) || (la.acts_ld & 3) || !aligned16(B) || !aligned16(la.g) ||
      !aligned16(la.c_prev) || !aligned16(la.c_out) || !aligned16(la.h_out) || !aligned16(la.acts) || get_encode() == nullptr ||
      (mode != 1 && mode != 2))
    return cudaErrorNotSupported;
  constexpr int BN = 128;
  const int num_kb = ceil_div(H, BK);
  const int tiles = ceil_div(M, BM) * ceil_div(N, BN);
  int splits = 148 / tiles;
  if (splits > num_kb) splits = num_kb;
  if (splits < 1) splits = 1;
  int kb_per = ceil_div(num_kb, splits);
  if (kb_per > 2) return cudaErrorNotSupported;
  splits = ceil_div(num_kb, kb_per);
  CUtensorMap mapB;
  if (!make_map(&mapB, B, H, N, ldb, BN, false)) return cudaErrorInvalidValue;
  Params p;
  p.C = C; p.bias = bias; p.bias2 = nullptr; p.M = M; p.N = N; p.K = H; p.ldc = ldc; p.alpha = 1.f;
  p.accumulate = 1; p.act = ACT_NONE; p.kb_per_split = kb_per; p.num_kb = num_kb;
  p.b_const = pdl_enabled() ? 1 : 0;
  p.dbg_noepi = 0;
  p.la = la;
  p.bulk_red = bulk_red_ok(C, ldc, N) ? 1 : 0;
  { static const int v4 = [] { const char* v = getenv("ALEXCAP_TC_RED_V4"); return (v && v[0] == '0') ? 0 : 1; }();
    p.red_v4 = (v4 && aligned16(C) && (ldc & 3) == 0 && (N & 3) == 0) ? 1 : 0; }
  {
    static unsigned long long* prof = [] {
      const char* v = getenv("ALEXCAP_TC_PROF_PTR");
      return v ? reinterpret_cast<unsigned long long*>(strtoull(v, nullptr, 0)) : nullptr;
    }();
    p.prof = prof;
  }
  dim3 grid(ceil_div(N, BN), ceil_div(M, BM), splits);
  if (mode == 2) return launch<BN, false, false, 2, true>(st, mapB, mapB, p, grid);
  return launch<BN, false, false, 1, true>(st, mapB, mapB, p, grid);
}

}  // namespace alexcap
