// gemm_ps.cu -- persistent tcgen05 GEMM over PRE-SPLIT bf16 operands: the tensor pipe fed by TMA alone.
//
//   C[M,N] = act(alpha * op(A) op(B) + bias + bias2) (+ C)        fp32 in HBM on both sides, same contract as gemm_f32
//
// gemm_tc.cu converts fp32 tiles to bf16 inside the kernel; that pass (TMA write + converter read + converter write + MMA
// reads of every shared-memory byte) is what bounds it (~155-215 TFLOP/s).  Here each fp32 operand is split ONCE in HBM
// by split_bf16_kernel into hi = bf16(x) and (fp32-grade mode) lo = bf16(x - hi), cached in a bump arena for every GEMM
// of the training step that reuses the operand, and the GEMM kernel does nothing but TMA -> shared memory -> tcgen05.mma:
//   NPROD = 1 : D += A_hi B_hi                                       (bf16 configs)
//   NPROD = 3 : D += A_lo B_hi ; D += A_hi B_lo ; D += A_hi B_hi     (fp32 configs, ~2^-16 relative operand error)
// Persistent: one CTA per SM loops over (tile, k-split) work items; two TMEM accumulators so the epilogue of item i
// overlaps the main loop of item i+1.
//   warp 0     : TMA producer (cp.async.bulk.tensor.2d, 128B swizzle, STAGES-deep mbarrier ring across items)
//   warp 1     : TMEM allocator + single-thread tcgen05.mma issuer (tcgen05.commit frees smem stages / publishes accumulators)
//   warps 2..9 : epilogue (tcgen05.ld -> registers -> smem transpose -> coalesced st / red.add), two warps per TMEM lane quarter
// Operand layouts in shared memory (bf16, 64-element = 128-byte rows, SWIZZLE_128B):
//   K-major  : one TMA box {64 k, rows};   descriptor SBO = 1024 (8 rows), one K=16 MMA advances 32 bytes in the row
//   MN-major : rows/64 boxes {64 mn, 64 k}; descriptor LBO = 8192 (next 64-wide mn block), SBO = 1024 (next 8 k rows),
//              one K=16 MMA advances 16 k-rows = 2048 bytes
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "kernels.h"
#include "tc_ptx.cuh"

namespace alexcap {

namespace ps {

using namespace tc;

constexpr int BM = 128;
constexpr int BK = 64;             // bf16 elements per k-block = one 128-byte swizzle row
constexpr int UMMA_K = 16;         // kind::f16
constexpr int EPI_WARPS = 8;
constexpr int TMA_WARP = 0, MMA_WARP = 1, EPI_WARP0 = 2;
constexpr int NUM_THREADS = (2 + EPI_WARPS) * 32;
constexpr uint32_t RING_BYTES = 196608;                  // operand ring (all variants use exactly this much)
constexpr uint32_t EPI_BYTES = EPI_WARPS * 32 * 33 * 4;  // per-warp 32x32 transpose tiles
constexpr uint32_t MN_BOX_BYTES = 64 * 128;              // one {64 mn, 64 k} box

struct Params {
  float* C; const float* bias; const float* bias2;
  int M, N, K; int64_t ldc; float alpha; int accumulate, act;
  int tiles_m, tiles_n, splits, kb_per_split, num_kb, n_items;
  // fused vocabulary projection + cross-entropy epilogues (EPI = 1: statistics, EPI = 2: gradient); C is not written
  const int32_t* tgt;           // [M] target class per row
  float4* stats;                // EPI 1 out: [M][tiles_n * 2] partial (max, sum exp(x - max), sum x, x[target])
  const float* row_lse;         // EPI 2 in: [M] log-sum-exp of the row
  const float* denom;           // EPI 2 in: CE divisor (device scalar)
  uint16_t* dl_hi; uint16_t* dl_lo; int64_t dl_ld;   // EPI 2 out: d logits as bf16 hi (+ lo) copies [M, dl_ld]
  float eps; int ignore_index;
  int vec4;                     // C rows allow 16-byte accesses (C 16-byte aligned, ldc % 4 == 0, N % 4 == 0): vector red.add / stores
};

// ------------------------------------------------------------------------------------------------ split pass
// X [rows, cols] fp32 (row pitch ldx) -> hi (and lo) [rows, ldo] bf16, ldo = cols rounded up to 8; pad columns are zero.
__global__ void __launch_bounds__(256) split_bf16_kernel(const float* __restrict__ X, int64_t ldx, int rows, int cols,
                                                         uint16_t* __restrict__ hi, uint16_t* __restrict__ lo, int64_t ldo,
                                                         int groups) {      // groups: 8-column groups written per row (ldo / 8, or less for a sub-view)
  const int64_t total = (int64_t)rows * groups;
  const bool vec = ((ldx & 3) == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / groups), c = (int)(i % groups) * 8;
    const float* src = X + (int64_t)r * ldx + c;
    float v[8];
    if (vec && c + 8 <= cols) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src) + 1);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = (c + j < cols) ? __ldg(src + j) : 0.f;
    }
    uint4 o;
    o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]); o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
    *reinterpret_cast<uint4*>(hi + (int64_t)r * ldo + c) = o;
    if (lo) {
      uint4 l;
#define LO_PAIR(W, A, B) l.W = pack_bf16(A - __uint_as_float(o.W << 16), B - __uint_as_float(o.W & 0xffff0000u))
      LO_PAIR(x, v[0], v[1]); LO_PAIR(y, v[2], v[3]); LO_PAIR(z, v[4], v[5]); LO_PAIR(w, v[6], v[7]);
#undef LO_PAIR
      *reinterpret_cast<uint4*>(lo + (int64_t)r * ldo + c) = l;
    }
  }
}

// ------------------------------------------------------------------------------------------------ kernel helpers
struct Item { int m0, n0, kb0, nkb, split; };

template <int BN>
__device__ __forceinline__ Item decode_item(const Params& p, int w) {
  Item it;
  it.split = w % p.splits;
  const int t = w / p.splits;
  it.m0 = (t % p.tiles_m) * BM;            // m fastest: concurrent CTAs share the B tile through L2
  it.n0 = (t / p.tiles_m) * BN;
  it.kb0 = it.split * p.kb_per_split;
  it.nkb = min(p.num_kb, it.kb0 + p.kb_per_split) - it.kb0;
  return it;
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
}

// one 32x32 sub-tile (thread = accumulator row, v[j] = column j) -> smem transpose -> 128-byte coalesced rows
__device__ __forceinline__ void store_chunk(const uint32_t* v, float* tile, const Params& p, int mrow0, int ncol0, int lane,
                                            bool atomic, bool first_split) {
#pragma unroll
  for (int j = 0; j < 32; ++j) tile[lane * 33 + j] = __uint_as_float(v[j]);
  __syncwarp();
  if (p.vec4) {
    // 16-byte accesses: lane = (row lane / 8 of a 4-row group, 4 columns lane % 8) -> 8 instructions of 4 rows x 128 bytes
    // instead of 32 of 1 row x 128 bytes.  Split-K / accumulate: red.global.add.v4.f32 (the epilogue of the split-K GEMMs is
    // bound by the red issue rate); plain outputs: st.global.v4.  The pitch-33 tile is read conflict-free
    // (bank = row + 4 * column group + e).
    const int cg = lane & 7, nv = ncol0 + 4 * cg;
    if (nv < p.N) {                                   // (N % 4 == 0: a column group is inside or outside as a whole)
      float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (first_split) {
        if (p.bias) { const float4 t4 = __ldg(reinterpret_cast<const float4*>(p.bias + nv)); b4.x += t4.x; b4.y += t4.y; b4.z += t4.z; b4.w += t4.w; }
        if (p.bias2) { const float4 t4 = __ldg(reinterpret_cast<const float4*>(p.bias2 + nv)); b4.x += t4.x; b4.y += t4.y; b4.z += t4.z; b4.w += t4.w; }
      }
      const bool plain = !atomic && !p.accumulate && p.act == ACT_NONE;
#pragma unroll
      for (int rr = 0; rr < 8; ++rr) {
        const int r = rr * 4 + (lane >> 3), m = mrow0 + r;
        if (m < p.M) {
          const float* tr = tile + r * 33 + 4 * cg;
          float4 x = make_float4(fmaf(p.alpha, tr[0], b4.x), fmaf(p.alpha, tr[1], b4.y), fmaf(p.alpha, tr[2], b4.z), fmaf(p.alpha, tr[3], b4.w));
          float4* cp = reinterpret_cast<float4*>(p.C + (int64_t)m * p.ldc + nv);
          if (atomic) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(cp), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w) : "memory");
          } else if (plain) {
            *cp = x;
          } else {
            if (p.accumulate) { const float4 c = *cp; x.x += c.x; x.y += c.y; x.z += c.z; x.w += c.w; }
            x.x = act_apply(x.x, p.act); x.y = act_apply(x.y, p.act); x.z = act_apply(x.z, p.act); x.w = act_apply(x.w, p.act);
            *cp = x;
          }
        }
      }
    }
    __syncwarp();
    return;
  }
  const int n = ncol0 + lane;
  if (n < p.N) {
    float bsum = 0.f;
    if (first_split) {
      if (p.bias) bsum += __ldg(p.bias + n);
      if (p.bias2) bsum += __ldg(p.bias2 + n);
    }
#pragma unroll 4
    for (int r = 0; r < 32; ++r) {
      const int m = mrow0 + r;
      if (m >= p.M) break;
      float x = fmaf(p.alpha, tile[r * 33 + lane], bsum);
      float* cp = p.C + (int64_t)m * p.ldc + n;
      if (atomic) {
        atomicAdd(cp, x);            // C was zeroed (or holds the value to accumulate onto) by the launcher
      } else {
        if (p.accumulate) x += *cp;
        x = act_apply(x, p.act);
        *cp = x;
      }
    }
  }
  __syncwarp();
}

// ---- fused vocabulary projection + CustomCrossEntropyLoss (AlexCap/AttentionLanguageModule.py:79 + AlexCap/CustomLoss.py:7-14).
// The logits tile never leaves the SM: thread = one row of the 128-row tile, v[j] = logit of column ncol0 + j (before bias).
// EPI 1 folds the chunk into the row's running (max, sum exp, sum x, x[target]); EPI 2 turns it into the gradient
//   d logit = (exp(x - lse) - eps / C - (1 - eps) [n == y]) * [y != ignore] / denom
// and writes it as the bf16 hi (+ lo = bf16(g - hi)) operand copies the two backward GEMMs read (dW_o = dlogits^T h,
// dH = dlogits W_o): the B x T x V' logits are never materialised, in fp32 or otherwise.
struct RowStat { float m, s, sx, xt; };
__device__ __forceinline__ void ce_stats_chunk(const uint32_t* v, const Params& p, int ncol0, int y, RowStat& st) {
  float x[32];
  float mc = -INFINITY;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const int n = ncol0 + j;
    const bool ok = n < p.N;
    x[j] = ok ? fmaf(p.alpha, __uint_as_float(v[j]), __ldg(p.bias + (ok ? n : 0))) : -INFINITY;
    mc = fmaxf(mc, x[j]);
  }
  if (mc == -INFINITY) return;                         // chunk entirely past the last class
  const float mn = fmaxf(st.m, mc);
  float acc = 0.f, sx = 0.f, xt = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const int n = ncol0 + j;
    if (n < p.N) {
      acc += __expf(x[j] - mn);
      sx += x[j];
      if (n == y) xt = x[j];
    }
  }
  st.s = st.s * __expf(st.m - mn) + acc;               // (exp(-inf) = 0 on the first chunk)
  st.m = mn; st.sx += sx; st.xt += xt;
}
// `tile`: the warp's 32 x 33 float staging tile.  The thread that owns row `lane` of the chunk computes the 32 gradients of
// its row and parks them in the tile; the warp then walks the tile 8 rows at a time, 4 lanes per row x 8 columns per lane, so
// that every store instruction writes 8 contiguous 64-byte row segments (whole 32-byte sectors) of the bf16 copies -- a
// thread-per-row store would touch 32 different rows with 16 bytes each.
__device__ __forceinline__ void ce_grad_chunk(const uint32_t* v, const Params& p, float* tile, int mrow0, int lane, int ncol0,
                                              int y, float lse, float scale) {
  // scale = [y != ignore] / denom (0 for rows past M); columns >= N get 0 (pad columns of the copies up to dl_ld)
  const float u = p.eps / (float)p.N;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const int n = ncol0 + j;
    const bool ok = n < p.N;
    const float x = fmaf(p.alpha, __uint_as_float(v[j]), __ldg(p.bias + (ok ? n : 0)));
    float g = __expf(x - lse) - u;
    if (n == y) g -= (1.f - p.eps);
    tile[lane * 33 + j] = ok ? g * scale : 0.f;
  }
  __syncwarp();
  const int sub = lane >> 2, cg = (lane & 3) * 8;
  if (ncol0 + cg < p.dl_ld) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int rr = sub + 8 * i, m = mrow0 + rr;
      if (m < p.M) {
        float g[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = tile[rr * 33 + cg + k];
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hi[k]) : "f"(g[2 * k + 1]), "f"(g[2 * k]));
          const float r0 = g[2 * k] - __uint_as_float(hi[k] << 16), r1 = g[2 * k + 1] - __uint_as_float(hi[k] & 0xffff0000u);
          asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo[k]) : "f"(r1), "f"(r0));
        }
        const int64_t off = (int64_t)m * p.dl_ld + ncol0 + cg;
        *reinterpret_cast<uint4*>(p.dl_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        if (p.dl_lo) *reinterpret_cast<uint4*>(p.dl_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      }
    }
  }
  __syncwarp();
}

// ------------------------------------------------------------------------------------------------ kernel
template <int BN, bool A_MN, bool B_MN, int NPROD, int EPI = 0>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_ps_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
               const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl, const Params p) {
  constexpr uint32_t A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
  constexpr uint32_t NOP = (NPROD == 3) ? 2 : 1;                     // tiles per operand: hi (+ lo)
  constexpr uint32_t STAGE_BYTES = NOP * (A_BYTES + B_BYTES);        // [A_hi][A_lo][B_hi][B_lo]
  constexpr int STAGES = RING_BYTES / STAGE_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN;                             // two accumulators
  static_assert(STAGES >= 2 && TMEM_COLS <= 512, "tile configuration");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* epi = reinterpret_cast<float*>(smem + RING_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + RING_BYTES + EPI_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;      // [2] accumulator complete
  uint64_t* tempty_bar = tfull_bar + 2;          // [2] accumulator drained by the epilogue
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == MMA_WARP) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  if (warp == TMA_WARP && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmAh) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmBh) : "memory");
    if (NPROD == 3) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmAl) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmBl) : "memory");
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  if (warp == TMA_WARP) {
    // ================================================================ TMA producer
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int w = blockIdx.x; w < p.n_items; w += gridDim.x) {
        const Item it = decode_item<BN>(p, w);
        for (int i = 0; i < it.nkb; ++i) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* st = smem + s * STAGE_BYTES;
          mbar_expect_tx(&full_bar[s], STAGE_BYTES);
          const int k0 = (it.kb0 + i) * BK;
#pragma unroll
          for (uint32_t o = 0; o < NOP; ++o) {
            const CUtensorMap* ma = o ? &tmAl : &tmAh;
            uint8_t* sa = st + o * A_BYTES;
            if (A_MN) {
#pragma unroll
              for (int j = 0; j < BM / 64; ++j) tma_load_2d(sa + j * MN_BOX_BYTES, ma, it.m0 + 64 * j, k0, &full_bar[s]);
            } else {
              tma_load_2d(sa, ma, k0, it.m0, &full_bar[s]);
            }
            const CUtensorMap* mb = o ? &tmBl : &tmBh;
            uint8_t* sb = st + NOP * A_BYTES + o * B_BYTES;
            if (B_MN) {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * MN_BOX_BYTES, mb, it.n0 + 64 * j, k0, &full_bar[s]);
            } else {
              tma_load_2d(sb, mb, k0, it.n0, &full_bar[s]);
            }
          }
          if (++s == STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ================================================================ MMA issuer (one elected thread)
    // instruction descriptor (cute::UMMA::InstrDescriptor): D = f32, A = B = bf16, majors, N >> 3, M >> 4
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    int s = 0; uint32_t ph = 0;
    int li = 0;
    for (int w = blockIdx.x; w < p.n_items; w += gridDim.x, ++li) {
      const Item it = decode_item<BN>(p, w);
      const int acc = li & 1;
      mbar_wait(&tempty_bar[acc], ((li >> 1) & 1) ^ 1);        // the epilogue has drained this accumulator
      tcgen05_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      for (int i = 0; i < it.nkb; ++i) {
        mbar_wait(&full_bar[s], ph);
        tcgen05_fence_after();
        if (lane == 0) {
          const uint32_t a_hi = smem_u32(smem + s * STAGE_BYTES), a_lo = a_hi + A_BYTES;
          const uint32_t b_hi = a_hi + NOP * A_BYTES, b_lo = b_hi + B_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint32_t a_off = A_MN ? k * 2048 : k * 32, b_off = B_MN ? k * 2048 : k * 32;
            const uint64_t dah = A_MN ? make_desc(a_hi + a_off, MN_BOX_BYTES, 1024) : make_desc(a_hi + a_off, 0, 1024);
            const uint64_t dbh = B_MN ? make_desc(b_hi + b_off, MN_BOX_BYTES, 1024) : make_desc(b_hi + b_off, 0, 1024);
            if (NPROD == 3) {
              const uint64_t dal = A_MN ? make_desc(a_lo + a_off, MN_BOX_BYTES, 1024) : make_desc(a_lo + a_off, 0, 1024);
              const uint64_t dbl = B_MN ? make_desc(b_lo + b_off, MN_BOX_BYTES, 1024) : make_desc(b_lo + b_off, 0, 1024);
              umma_bf16(d_tmem, dal, dbh, idesc, (i | k) != 0);
              umma_bf16(d_tmem, dah, dbl, idesc, 1);
              umma_bf16(d_tmem, dah, dbh, idesc, 1);
            } else {
              umma_bf16(d_tmem, dah, dbh, idesc, (i | k) != 0);
            }
          }
          umma_commit(&empty_bar[s]);                         // smem stage reusable once these MMAs retire
          if (i == it.nkb - 1) umma_commit(&tfull_bar[acc]);  // accumulator complete
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
  } else {
    // ================================================================ epilogue warps
    const int ew = warp - EPI_WARP0;
    const int q4 = warp & 3;                      // TMEM lane quarter this warp may read (warp id mod 4)
    const int half = ew >> 2;                     // which half of the tile's columns
    constexpr int CPW = BN / 64;                  // 32-column chunks per warp
    float* tile = epi + ew * (32 * 33);
    // accumulate without an activation onto a SMALL output (the per-step GEMMs: one wave of CTAs) also goes through
    // red.add (fire-and-forget, one writer per element: the same single fp32 add as `x += *cp`): the load->add->store
    // form serialises 64 dependent global loads per warp, ~40 us per launch on the M = 512 per-step GEMMs of cfg3
    // (profiles/r1k_gemm_step.txt).  Large outputs keep load->add->store: many tiles per CTA hide that latency.
    const bool atomic = p.splits > 1 || (p.accumulate && p.act == ACT_NONE && (int64_t)p.M * p.N <= (int64_t)(4 << 20));
    int li = 0;
    for (int w = blockIdx.x; w < p.n_items; w += gridDim.x, ++li) {
      const Item it = decode_item<BN>(p, w);
      const int acc = li & 1;
      mbar_wait(&tfull_bar[acc], (li >> 1) & 1);
      tcgen05_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(acc * BN);
      // fused cross-entropy epilogues: this thread's row and its target
      const int row = it.m0 + q4 * 32 + lane;
      const bool row_ok = row < p.M;
      int y = -1; float lse = 0.f, scale = 0.f;
      RowStat rs{-INFINITY, 0.f, 0.f, 0.f};
      if (EPI != 0 && row_ok) {
        y = __ldg(p.tgt + row);
        if (EPI == 2) {
          lse = __ldg(p.row_lse + row);
          const float dn = __ldg(p.denom);
          scale = (y != p.ignore_index && dn > 0.f) ? 1.f / dn : 0.f;
        }
      }
#pragma unroll 1
      for (int g = 0; g < CPW; g += 2) {
        uint32_t v0[32], v1[32];
        const int c0 = half * CPW + g, c1 = c0 + 1;
        tmem_ld32(tbase + (uint32_t)(c0 * 32), v0);
        tmem_ld32(tbase + (uint32_t)(c1 * 32), v1);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (g + 2 >= CPW) {                       // last read of this accumulator: hand it back to the MMA warp
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        }
        if (EPI == 0) {
          store_chunk(v0, tile, p, it.m0 + q4 * 32, it.n0 + c0 * 32, lane, atomic, it.split == 0);
          store_chunk(v1, tile, p, it.m0 + q4 * 32, it.n0 + c1 * 32, lane, atomic, it.split == 0);
        } else if (EPI == 1) {
          if (row_ok) { ce_stats_chunk(v0, p, it.n0 + c0 * 32, y, rs); ce_stats_chunk(v1, p, it.n0 + c1 * 32, y, rs); }
        } else {
          ce_grad_chunk(v0, p, tile, it.m0 + q4 * 32, lane, it.n0 + c0 * 32, y, lse, scale);     // (whole warp: rows past M have scale 0)
          ce_grad_chunk(v1, p, tile, it.m0 + q4 * 32, lane, it.n0 + c1 * 32, y, lse, scale);
        }
      }
      if (EPI == 1 && row_ok)
        p.stats[(int64_t)row * (p.tiles_n * 2) + (it.n0 / BN) * 2 + half] = make_float4(rs.m, rs.s, rs.sx, rs.xt);
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS));
  }
}

// ------------------------------------------------------------------------------------------------ CTA-pair kernel
// The same persistent pipeline with tcgen05 cta_group::2: the two CTAs of a cluster (an SM pair) work on ONE 256 x 256 output
// tile.  Each CTA loads its own 128 rows of A and only ITS HALF (128 columns) of the B tile; the leader CTA issues one
// M = 256, N = 256 MMA per k16 that reads A and B from both SMs' shared memory and writes each CTA's 128 accumulator rows
// into that CTA's TMEM.  Per CTA and 64-deep k-block that is 16 KB (A) + 16 KB (B half) per operand copy for 128 x 256 x 64
// MACs -- against 16 + 32 KB for the single-CTA 128 x 256 tile: a third less L2 -> shared-memory traffic per flop, which is
// what bounds this kernel, and 3 ring stages instead of 2 in the bf16x3 mode.
//   full_bar  (leader's): count 1 = the leader's arrive.expect_tx for the bytes of BOTH CTAs; both CTAs' TMA loads complete_tx
//                         on it (.cta_group::2 form, peer bit of the barrier address cleared).  The peer's bytes may land before
//                         the leader has armed the phase (the transaction count goes negative, the pending arrival keeps the
//                         phase open); the peer cannot be a whole ring ahead because its stage is only freed by the leader's
//                         MMA commit.  (A per-k-block remote arrive by the peer producer was measured: it stalls that thread
//                         ~1 us per k-block and halved the bf16 rate.)
//   empty_bar / tfull_bar (own, both CTAs): tcgen05.commit ... multicast::cluster with mask 0b11
//   tempty_bar (leader's): count 2 * EPI_WARPS: the epilogue warps of both CTAs (the peer's arrive remotely)
constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;       // Sm100MmaPeerBitMask: shared::cluster address of the pair's even CTA
constexpr int BM2 = 256, BN2 = 256;

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(leader_bar)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

__device__ __forceinline__ Item decode_item2(const Params& p, int w) {
  Item it;
  it.split = w % p.splits;
  const int t = w / p.splits;
  it.m0 = (t % p.tiles_m) * BM2;
  it.n0 = (t / p.tiles_m) * BN2;
  it.kb0 = it.split * p.kb_per_split;
  it.nkb = min(p.num_kb, it.kb0 + p.kb_per_split) - it.kb0;
  return it;
}

template <bool A_MN, bool B_MN, int NPROD, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_ps2_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
                const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl, const Params p) {
  constexpr uint32_t A_BYTES = 128 * BK * 2, B_BYTES = 128 * BK * 2;          // per CTA: its A rows, its half of the B tile
  constexpr uint32_t NOP = (NPROD == 3) ? 2 : 1;
  constexpr uint32_t STAGE_BYTES = NOP * (A_BYTES + B_BYTES);
  constexpr int STAGES = RING_BYTES / STAGE_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* epi = reinterpret_cast<float*>(smem + RING_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + RING_BYTES + EPI_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tfull_bar[a], 1); mbar_init(&tempty_bar[a], 2 * EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == MMA_WARP) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  if (warp == TMA_WARP && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmAh) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmBh) : "memory");
    if (NPROD == 3) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmAl) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmBl) : "memory");
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();                              // both CTAs' barriers are initialised before anyone signals them
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  if (warp == TMA_WARP) {
    // ================================================================ TMA producer (one per CTA)
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (int w = cluster_id; w < p.n_items; w += n_clusters) {
        const Item it = decode_item2(p, w);
        const int m0 = it.m0 + 128 * (int)rank, n0 = it.n0 + 128 * (int)rank;
        for (int i = 0; i < it.nkb; ++i) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* st = smem + s * STAGE_BYTES;
          const uint32_t lbar = smem_u32(&full_bar[s]) & PEER_MASK;          // the LEADER's barrier of this stage
          if (leader) mbar_expect_tx(&full_bar[s], 2 * STAGE_BYTES);
          const int k0 = (it.kb0 + i) * BK;
#pragma unroll
          for (uint32_t o = 0; o < NOP; ++o) {
            const CUtensorMap* ma = o ? &tmAl : &tmAh;
            uint8_t* sa = st + o * A_BYTES;
            if (A_MN) {
#pragma unroll
              for (int j = 0; j < 2; ++j) tma_load_2d_pair(sa + j * MN_BOX_BYTES, ma, m0 + 64 * j, k0, lbar);
            } else {
              tma_load_2d_pair(sa, ma, k0, m0, lbar);
            }
            const CUtensorMap* mb = o ? &tmBl : &tmBh;
            uint8_t* sb = st + NOP * A_BYTES + o * B_BYTES;
            if (B_MN) {
#pragma unroll
              for (int j = 0; j < 2; ++j) tma_load_2d_pair(sb + j * MN_BOX_BYTES, mb, n0 + 64 * j, k0, lbar);
            } else {
              tma_load_2d_pair(sb, mb, k0, n0, lbar);
            }
          }
          if (++s == STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ================================================================ MMA issuer: the leader CTA's elected thread
    if (leader) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             ((uint32_t)(BN2 >> 3) << 17) | ((uint32_t)(BM2 >> 4) << 24);
      int s = 0; uint32_t ph = 0;
      int li = 0;
      for (int w = cluster_id; w < p.n_items; w += n_clusters, ++li) {
        const Item it = decode_item2(p, w);
        const int acc = li & 1;
        mbar_wait(&tempty_bar[acc], ((li >> 1) & 1) ^ 1);        // both CTAs' epilogues have drained this accumulator
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN2);
        for (int i = 0; i < it.nkb; ++i) {
          mbar_wait(&full_bar[s], ph);
          tcgen05_fence_after();
          if (lane == 0) {
            const uint32_t a_hi = smem_u32(smem + s * STAGE_BYTES), a_lo = a_hi + A_BYTES;
            const uint32_t b_hi = a_hi + NOP * A_BYTES, b_lo = b_hi + B_BYTES;
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint32_t a_off = A_MN ? k * 2048 : k * 32, b_off = B_MN ? k * 2048 : k * 32;
              const uint64_t dah = A_MN ? make_desc(a_hi + a_off, MN_BOX_BYTES, 1024) : make_desc(a_hi + a_off, 0, 1024);
              const uint64_t dbh = B_MN ? make_desc(b_hi + b_off, MN_BOX_BYTES, 1024) : make_desc(b_hi + b_off, 0, 1024);
              if (NPROD == 3) {
                const uint64_t dal = A_MN ? make_desc(a_lo + a_off, MN_BOX_BYTES, 1024) : make_desc(a_lo + a_off, 0, 1024);
                const uint64_t dbl = B_MN ? make_desc(b_lo + b_off, MN_BOX_BYTES, 1024) : make_desc(b_lo + b_off, 0, 1024);
                umma_bf16_pair(d_tmem, dal, dbh, idesc, (i | k) != 0);
                umma_bf16_pair(d_tmem, dah, dbl, idesc, 1);
                umma_bf16_pair(d_tmem, dah, dbh, idesc, 1);
              } else {
                umma_bf16_pair(d_tmem, dah, dbh, idesc, (i | k) != 0);
              }
            }
            umma_commit_pair(&empty_bar[s]);                         // frees the stage in BOTH CTAs
            if (i == it.nkb - 1) umma_commit_pair(&tfull_bar[acc]);  // accumulator complete, both CTAs
          }
          __syncwarp();
          if (++s == STAGES) { s = 0; ph ^= 1; }
        }
      }
    }
  } else {
    // ================================================================ epilogue warps (both CTAs: own 128 rows x 256 columns)
    const int ew = warp - EPI_WARP0;
    const int q4 = warp & 3;
    const int half = ew >> 2;
    constexpr int CPW = BN2 / 64;
    float* tile = epi + ew * (32 * 33);
    const bool atomic = p.splits > 1 || (p.accumulate && p.act == ACT_NONE && (int64_t)p.M * p.N <= (int64_t)(4 << 20));
    const uint32_t ltempty0 = smem_u32(&tempty_bar[0]) & PEER_MASK, ltempty1 = smem_u32(&tempty_bar[1]) & PEER_MASK;
    int li = 0;
    for (int w = cluster_id; w < p.n_items; w += n_clusters, ++li) {
      const Item it = decode_item2(p, w);
      const int m0 = it.m0 + 128 * (int)rank;
      const int acc = li & 1;
      mbar_wait(&tfull_bar[acc], (li >> 1) & 1);
      tcgen05_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(acc * BN2);
      const int row = m0 + q4 * 32 + lane;
      const bool row_ok = row < p.M;
      int y = -1; float lse = 0.f, scale = 0.f;
      RowStat rs{-INFINITY, 0.f, 0.f, 0.f};
      if (EPI != 0 && row_ok) {
        y = __ldg(p.tgt + row);
        if (EPI == 2) {
          lse = __ldg(p.row_lse + row);
          const float dn = __ldg(p.denom);
          scale = (y != p.ignore_index && dn > 0.f) ? 1.f / dn : 0.f;
        }
      }
#pragma unroll 1
      for (int g = 0; g < CPW; g += 2) {
        uint32_t v0[32], v1[32];
        const int c0 = half * CPW + g, c1 = c0 + 1;
        tmem_ld32(tbase + (uint32_t)(c0 * 32), v0);
        tmem_ld32(tbase + (uint32_t)(c1 * 32), v1);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (g + 2 >= CPW) {                       // last read of this accumulator: hand it back to the LEADER's MMA warp
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(acc ? ltempty1 : ltempty0);
        }
        if (EPI == 0) {
          store_chunk(v0, tile, p, m0 + q4 * 32, it.n0 + c0 * 32, lane, atomic, it.split == 0);
          store_chunk(v1, tile, p, m0 + q4 * 32, it.n0 + c1 * 32, lane, atomic, it.split == 0);
        } else if (EPI == 1) {
          if (row_ok) { ce_stats_chunk(v0, p, it.n0 + c0 * 32, y, rs); ce_stats_chunk(v1, p, it.n0 + c1 * 32, y, rs); }
        } else {
          ce_grad_chunk(v0, p, tile, m0 + q4 * 32, lane, it.n0 + c0 * 32, y, lse, scale);
          ce_grad_chunk(v1, p, tile, m0 + q4 * 32, lane, it.n0 + c1 * 32, y, lse, scale);
        }
      }
      if (EPI == 1 && row_ok)
        p.stats[(int64_t)row * (p.tiles_n * 2) + (it.n0 / BN2) * 2 + half] = make_float4(rs.m, rs.s, rs.sx, rs.xt);
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();                              // the peer may still be reading / signalling this CTA's shared memory
  if (warp == MMA_WARP) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS));
  }
}

// ------------------------------------------------------------------------------------------------ host side
// 2-D bf16 tensor map: inner (contiguous) extent, outer extent, row pitch ld elements, box {64, box_outer}, 128B swizzle.
static bool make_map_bf16(CUtensorMap* map, const uint16_t* base, int64_t inner, int64_t outer, int64_t ld, int box_outer) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<uint16_t*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

template <int BN, bool A_MN, bool B_MN, int NPROD, int EPI = 0>
static cudaError_t launch(cudaStream_t st, const CUtensorMap* maps, const Params& p, int grid) {
  constexpr size_t SMEM = (size_t)RING_BYTES + EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static bool configured = false;
  auto kern = gemm_ps_kernel<BN, A_MN, B_MN, NPROD, EPI>;
  if (!configured) {
    ACK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    configured = true;
  }
  ALEXCAP_NOTE_LAUNCH();
  return launch_pdl(kern, dim3(grid), dim3(NUM_THREADS), SMEM, st, maps[0], maps[1], maps[2], maps[3], p);
}

template <bool A_MN, bool B_MN, int NPROD, int EPI = 0>
static cudaError_t launch2(cudaStream_t st, const CUtensorMap* maps, const Params& p, int grid) {
  constexpr size_t SMEM = (size_t)RING_BYTES + EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static bool configured = false;
  auto kern = gemm_ps2_kernel<A_MN, B_MN, NPROD, EPI>;
  if (!configured) {
    ACK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    configured = true;
  }
  ALEXCAP_NOTE_LAUNCH();
  return launch_pdl(kern, dim3(grid), dim3(NUM_THREADS), SMEM, st, maps[0], maps[1], maps[2], maps[3], p);
}
// CTA-pair (cta_group::2) tiles.  ALEXCAP_PS_PAIR=0 never, =1 whenever the shape allows, default = where they measured
// faster on B200 with the 16-byte epilogue (profiles/r2m_gemm_epilogue.txt; with the scalar epilogue of
// profiles/r2_gemm_pair.txt the epilogue bounded both kernels at K = 512 and the pair tiles only paid off from K = 1024):
// bf16x3 mode from K >= 512 (cfg2 vocabulary projection 62.5 -> 54.0 us, Ws projection 43.6 -> 35.0 us), bf16
// single-product mode from K >= 1024 (cfg3 vocabulary projection 178 -> 166 us = 1077 TFLOP/s, dH 37 -> 26 us; at K = 512
// the single-CTA kernel is still ahead, 35.5 vs 38.6 us).
static bool ps_vec4_on() {       // ALEXCAP_PS_RED_V4=0: scalar red.add (A/B timing)
  static const bool on = [] { const char* v = getenv("ALEXCAP_PS_RED_V4"); return !(v && v[0] == '0'); }();
  return on;
}
static bool pair_wanted(int M, int N, int K, int nprod) {
  static const int mode = [] { const char* v = getenv("ALEXCAP_PS_PAIR"); return v ? (v[0] == '1' ? 1 : 0) : 2; }();
  if (mode == 0 || M < 256 || N < 256) return false;
  const int64_t tiles = (int64_t)ceil_div(M, BM2) * ceil_div(N, BN2);
  if (tiles < 8) return false;
  if (mode == 1) return true;
  static const int k3 = [] { const char* v = getenv("ALEXCAP_PS_PAIR_K3"); return v ? atoi(v) : 512; }();
  static const int k1 = [] { const char* v = getenv("ALEXCAP_PS_PAIR_K1"); return v ? atoi(v) : 1024; }();
  if (nprod == 3) return K >= k3;
  return k1 > 0 ? K >= k1 : (K >= 2048 && tiles < 148);      // (ALEXCAP_PS_PAIR_K1=0: the round-2 rule, for A/B timing)
}

struct Split { uint16_t* hi; uint16_t* lo; int64_t ld; };

// hi/lo bf16 copy of the fp32 matrix X [rows, cols] (pitch ld) out of the arena: cached per source for the arena's lifetime
// X [rows, cols] (pitch ld) as a sub-block of a mirrored matrix: same pitch, inside its extents, column offset % 8 == 0
static bool find_view(const SplitArena* a, const float* X, int rows, int cols, int64_t ld, bool with_lo, Split* out) {
  for (int i = 0; i < a->n; ++i) {
    const SplitArena::Entry& e = a->e[i];
    if (!e.mirror || e.ld != ld || X < e.src || (with_lo && !e.lo)) continue;
    const int64_t off = X - e.src, r0 = off / ld, c0 = off % ld;
    if (r0 + rows > e.rows || c0 + cols > e.cols || (c0 & 7)) continue;
    out->hi = e.hi + r0 * e.ldo + c0;
    out->lo = e.lo ? e.lo + r0 * e.ldo + c0 : nullptr;
    out->ld = e.ldo;
    return true;
  }
  return false;
}

static cudaError_t get_split(cudaStream_t st, SplitArena* a, const float* X, int rows, int cols, int64_t ld, bool with_lo,
                             Split* out) {
  for (int i = 0; i < a->n; ++i) {
    const SplitArena::Entry& e = a->e[i];
    if (e.src == X && e.rows == rows && e.cols == cols && e.ld == ld && (e.lo != nullptr || !with_lo)) {
      out->hi = e.hi; out->lo = e.lo; out->ld = e.ldo;
      return cudaSuccess;
    }
  }
  if (find_view(a, X, rows, cols, ld, with_lo, out)) return cudaSuccess;      // a sub-block of a producer-written mirror
  const int64_t ldo = ((int64_t)cols + 7) & ~int64_t(7);
  const size_t one = (((size_t)rows * ldo * 2) + 1023) & ~size_t(1023);
  const size_t need = one * (with_lo ? 2 : 1);
  if (!a->base || a->used + need > a->bytes || a->n >= SplitArena::MAX_ENTRIES) return cudaErrorMemoryAllocation;
  SplitArena::Entry& e = a->e[a->n];
  e.src = X; e.rows = rows; e.cols = cols; e.ld = ld; e.ldo = ldo; e.produced = false; e.mirror = false;
  e.hi = reinterpret_cast<uint16_t*>(a->base + a->used);
  e.lo = with_lo ? reinterpret_cast<uint16_t*>(a->base + a->used + one) : nullptr;
  a->used += need;
  e.end = a->used;
  a->n++;
  const int64_t total = (int64_t)rows * (ldo >> 3);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  ALEXCAP_NOTE_LAUNCH();
  split_bf16_kernel<<<blocks, 256, 0, st>>>(X, ld, rows, cols, e.hi, e.lo, ldo, (int)(ldo >> 3));
  ACK(cudaGetLastError());
  out->hi = e.hi; out->lo = e.lo; out->ld = ldo;
  return cudaSuccess;
}

}  // namespace ps

SplitArena*& split_arena() {
  static thread_local SplitArena* a = nullptr;
  return a;
}

void split_arena_release(SplitArena* a, size_t mark) {
  while (a->n > 0 && a->e[a->n - 1].end > mark) a->n--;
  a->used = mark;
}

// Entry for a matrix whose PRODUCER writes the bf16 copies itself (CE kernel -> dlogits, attn_dws -> dWs): same key as the
// fp32 matrix would have had, no split pass.  False when the arena cannot hold it (the producer then writes fp32 as before).
bool split_arena_reserve(SplitArena* a, const float* key, int rows, int cols, int64_t ld, bool with_lo, uint16_t** hi,
                         uint16_t** lo, int64_t* ldo_out) {
  if (!a || !a->base || a->n >= SplitArena::MAX_ENTRIES) return false;
  const int64_t ldo = ((int64_t)cols + 7) & ~int64_t(7);
  const size_t one = (((size_t)rows * ldo * 2) + 1023) & ~size_t(1023);
  const size_t need = one * (with_lo ? 2 : 1);
  if (a->used + need > a->bytes) return false;
  SplitArena::Entry& e = a->e[a->n];
  e.src = key; e.rows = rows; e.cols = cols; e.ld = ld; e.ldo = ldo; e.produced = true; e.mirror = false;
  e.hi = reinterpret_cast<uint16_t*>(a->base + a->used);
  e.lo = with_lo ? reinterpret_cast<uint16_t*>(a->base + a->used + one) : nullptr;
  a->used += need;
  e.end = a->used;
  a->n++;
  *hi = e.hi; *lo = e.lo; *ldo_out = ldo;
  return true;
}

bool split_arena_reserve_mirror(SplitArena* a, const float* src, int rows, int cols, int64_t ld, bool with_lo, uint16_t** hi,
                                uint16_t** lo, int64_t* ldo_out) {
  if (!split_arena_reserve(a, src, rows, cols, ld, with_lo, hi, lo, ldo_out)) return false;
  a->e[a->n - 1].mirror = true;
  return true;
}

cudaError_t split_arena_fill_view(cudaStream_t st, SplitArena* a, const float* X, int rows, int cols, int64_t ld) {
  ps::Split sp;
  if (!a || !ps::find_view(a, X, rows, cols, ld, false, &sp)) return cudaErrorInvalidValue;
  if (cols & 7) return cudaErrorInvalidValue;           // (the split kernel zero-pads to 8 columns: must not spill into the neighbours)
  const int64_t total = (int64_t)rows * (cols >> 3);
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  ALEXCAP_NOTE_LAUNCH();
  ps::split_bf16_kernel<<<blocks, 256, 0, st>>>(X, ld, rows, cols, sp.hi, sp.lo, sp.ld, cols >> 3);
  return cudaGetLastError();
}

// The copies a producer registered under `key` (split_arena_reserve) earlier in the step; false if there are none.
bool split_arena_find(const SplitArena* a, const float* key, uint16_t** hi, uint16_t** lo, int64_t* ldo_out) {
  if (!a) return false;
  for (int i = a->n - 1; i >= 0; --i)
    if (a->e[i].src == key && a->e[i].produced) {
      *hi = a->e[i].hi; *lo = a->e[i].lo; *ldo_out = a->e[i].ldo;
      return true;
    }
  return false;
}

// bf16 (hi) copy of X out of the arena -- made now if no GEMM has split X yet.  Used to stream Ws / feat as bf16 rows through
// the attention kernels in ALEXCAP_BF16 mode.
cudaError_t split_arena_get_hi(cudaStream_t st, SplitArena* a, const float* X, int rows, int cols, int64_t ld, const uint16_t** hi,
                               int64_t* ldo) {
  if (!a) return cudaErrorMemoryAllocation;
  ps::Split sp;
  cudaError_t e = ps::get_split(st, a, X, rows, cols, ld, false, &sp);
  if (e != cudaSuccess) return e;
  *hi = sp.hi; *ldo = sp.ld;
  return cudaSuccess;
}

bool split_arena_is_produced(const SplitArena* a, const float* p) {
  if (!a) return false;
  for (int i = 0; i < a->n; ++i)
    if (a->e[i].src == p && a->e[i].produced && !a->e[i].mirror) return true;      // (a mirror has its fp32 matrix)
  return false;
}

bool gemm_ps_wanted(int M, int N, int K) {
  static const bool off = [] { const char* v = getenv("ALEXCAP_GEMM_PRESPLIT"); return v && v[0] == '0'; }();
  return !off && M >= 256 && N >= 64 && K >= 128 && 2.0 * M * N * K >= 1e9;
}

cudaError_t gemm_ps_f32(cudaStream_t st, bool ta, bool tb, int M, int N, int K, float alpha,
                        const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                        const float* bias, const float* bias2, bool accumulate, int act, int force_splits, int nprod,
                        SplitArena* arena) {
  using namespace ps;
  if (M <= 0 || N <= 0 || K <= 0) return cudaSuccess;
  if (!arena || !get_encode()) return cudaErrorMemoryAllocation;
  const bool a_mn = ta, b_mn = !tb;
  const bool with_lo = (nprod == 3);
  const size_t mark = arena->used;
  Split sa, sb;
  cudaError_t e = get_split(st, arena, A, ta ? K : M, ta ? M : K, lda, with_lo, &sa);
  if (e == cudaSuccess) e = get_split(st, arena, B, tb ? N : K, tb ? K : N, ldb, with_lo, &sb);
  if (e != cudaSuccess) { split_arena_release(arena, mark); return e; }

  int nsm = 148;
  { int dev = 0; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev); }
  if (gemm_cta_cap() > 0 && gemm_cta_cap() < nsm) nsm = gemm_cta_cap();
  // ---- CTA-pair 256 x 256 tiles (cta_group::2) when the output is large enough to give every SM pair work
  if (pair_wanted(M, N, K, nprod)) {
    const int ncl = nsm / 2;
    const int num_kb = ceil_div(K, BK);
    const int tiles_m = ceil_div(M, BM2), tiles_n = ceil_div(N, BN2), tiles = tiles_m * tiles_n;
    int splits = 1;
    if (force_splits > 0) splits = force_splits;
    else if (act == ACT_NONE && tiles < (2 * ncl) / 3) {
      splits = ncl / tiles;
      const int max_by_k = num_kb / 2 > 0 ? num_kb / 2 : 1;
      if (splits > max_by_k) splits = max_by_k;
      if (splits < 1) splits = 1;
    }
    if (act != ACT_NONE) splits = 1;
    const int kb_per = ceil_div(num_kb, splits);
    splits = ceil_div(num_kb, kb_per);
    if (splits > 1 && !accumulate) ACK(cudaMemset2DAsync(C, ldc * sizeof(float), 0, (size_t)N * sizeof(float), M, st));
    CUtensorMap maps[4];
    bool ok = true;
    for (int o = 0; o < (with_lo ? 2 : 1); ++o) {
      const uint16_t* pa = o ? sa.lo : sa.hi;
      const uint16_t* pb = o ? sb.lo : sb.hi;
      ok = ok && (a_mn ? make_map_bf16(&maps[o], pa, M, K, sa.ld, 64) : make_map_bf16(&maps[o], pa, K, M, sa.ld, 128));
      ok = ok && (b_mn ? make_map_bf16(&maps[2 + o], pb, N, K, sb.ld, 64) : make_map_bf16(&maps[2 + o], pb, K, N, sb.ld, 128));
    }
    if (!with_lo) { maps[1] = maps[0]; maps[3] = maps[2]; }
    if (!ok) return cudaErrorInvalidValue;
    Params p = {};
    p.C = C; p.bias = bias; p.bias2 = bias2; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.alpha = alpha;
    p.accumulate = accumulate ? 1 : 0; p.act = act;
    p.vec4 = (ps_vec4_on() && aligned16(C) && (ldc & 3) == 0 && (N & 3) == 0) ? 1 : 0;
    p.tiles_m = tiles_m; p.tiles_n = tiles_n; p.splits = splits; p.kb_per_split = kb_per; p.num_kb = num_kb;
    p.n_items = tiles * splits;
    const int grid = 2 * (p.n_items < ncl ? p.n_items : ncl);
#define PS2_DISPATCH(NP_)                                                                  \
  do {                                                                                     \
    if (!a_mn && !b_mn) return launch2<false, false, NP_>(st, maps, p, grid);              \
    if (!a_mn && b_mn) return launch2<false, true, NP_>(st, maps, p, grid);                \
    if (a_mn && b_mn) return launch2<true, true, NP_>(st, maps, p, grid);                  \
    return launch2<true, false, NP_>(st, maps, p, grid);                                   \
  } while (0)
    if (with_lo) PS2_DISPATCH(3);
    PS2_DISPATCH(1);
#undef PS2_DISPATCH
  }

  static const int bn_env = [] { const char* v = getenv("ALEXCAP_PS_BN"); return v ? atoi(v) : 0; }();
  const int BN = (bn_env == 128 || bn_env == 256) ? bn_env : ((N >= 1024 && (int64_t)ceil_div(M, BM) * ceil_div(N, 256) >= 296) ? 256 : 128);
  const int num_kb = ceil_div(K, BK);
  const int tiles_m = ceil_div(M, BM), tiles_n = ceil_div(N, BN);
  const int tiles = tiles_m * tiles_n;
  // split-K (fp32 red.add into C) when the output tiles alone cannot fill 148 SMs; never with an activation
  int splits = 1;
  if (force_splits > 0) splits = force_splits;
  else if (act == ACT_NONE && tiles < 100) {
    splits = 148 / tiles;
    const int max_by_k = num_kb / 2 > 0 ? num_kb / 2 : 1;
    if (splits > max_by_k) splits = max_by_k;
    if (splits < 1) splits = 1;
  }
  if (act != ACT_NONE) splits = 1;
  int kb_per = ceil_div(num_kb, splits);
  splits = ceil_div(num_kb, kb_per);       // no empty split
  if (splits > 1 && !accumulate) ACK(cudaMemset2DAsync(C, ldc * sizeof(float), 0, (size_t)N * sizeof(float), M, st));

  CUtensorMap maps[4];
  bool ok = true;
  for (int o = 0; o < (with_lo ? 2 : 1); ++o) {
    const uint16_t* pa = o ? sa.lo : sa.hi;
    const uint16_t* pb = o ? sb.lo : sb.hi;
    ok = ok && (a_mn ? make_map_bf16(&maps[o], pa, M, K, sa.ld, 64) : make_map_bf16(&maps[o], pa, K, M, sa.ld, BM));
    ok = ok && (b_mn ? make_map_bf16(&maps[2 + o], pb, N, K, sb.ld, 64) : make_map_bf16(&maps[2 + o], pb, K, N, sb.ld, BN));
  }
  if (!with_lo) { maps[1] = maps[0]; maps[3] = maps[2]; }
  if (!ok) return cudaErrorInvalidValue;

  Params p;
  p.C = C; p.bias = bias; p.bias2 = bias2; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.alpha = alpha;
  p.accumulate = accumulate ? 1 : 0; p.act = act;
  p.vec4 = (ps_vec4_on() && aligned16(C) && (ldc & 3) == 0 && (N & 3) == 0) ? 1 : 0;
  p.tiles_m = tiles_m; p.tiles_n = tiles_n; p.splits = splits; p.kb_per_split = kb_per; p.num_kb = num_kb;
  p.n_items = tiles * splits;
  const int grid = p.n_items < nsm ? p.n_items : nsm;
#define PS_DISPATCH(BN_, NP_)                                                              \
  do {                                                                                     \
    if (!a_mn && !b_mn) return launch<BN_, false, false, NP_>(st, maps, p, grid);          \
    if (!a_mn && b_mn) return launch<BN_, false, true, NP_>(st, maps, p, grid);            \
    if (a_mn && b_mn) return launch<BN_, true, true, NP_>(st, maps, p, grid);              \
    return launch<BN_, true, false, NP_>(st, maps, p, grid);                               \
  } while (0)
  if (BN == 256) {
    if (with_lo) PS_DISPATCH(256, 3);
    PS_DISPATCH(256, 1);
  }
  if (with_lo) PS_DISPATCH(128, 3);
  PS_DISPATCH(128, 1);
#undef PS_DISPATCH
}

// ------------------------------------------------------------------------------------------------ fused vocabulary CE
namespace ps {

// one warp per row: fold the row's partial statistics into lse and the masked label-smoothed loss (CustomLoss.py:7)
//   loss_row = ((1 - eps) (lse - x[y]) + eps (lse - mean_c x[c])) / denom      (0 for ignored rows)
__global__ void __launch_bounds__(256) ce_combine_kernel(const float4* __restrict__ stats, int np, const int32_t* __restrict__ tgt,
                                                         int M, int C, float eps, int ignore_index, const float* __restrict__ denom,
                                                         float* __restrict__ row_lse, float* __restrict__ row_loss) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= M) return;
  const float4* sp = stats + (int64_t)row * np;
  float m = -INFINITY;
  for (int i = lane; i < np; i += 32) m = fmaxf(m, sp[i].x);
  m = warp_max(m);
  float s = 0.f, sx = 0.f, xt = 0.f;
  for (int i = lane; i < np; i += 32) {
    const float4 q = sp[i];
    if (q.x > -INFINITY) s = fmaf(q.y, expf(q.x - m), s);
    sx += q.z; xt += q.w;
  }
  s = warp_sum(s); sx = warp_sum(sx); xt = warp_sum(xt);
  if (lane == 0) {
    const float lse = m + logf(s);
    row_lse[row] = lse;
    const float dn = denom[0];
    const float inv_d = dn > 0.f ? 1.f / dn : 0.f;
    row_loss[row] = (tgt[row] == ignore_index) ? 0.f : ((1.f - eps) * (lse - xt) + eps * (lse - sx / (float)C)) * inv_d;
  }
}

}  // namespace ps

bool gemm_ps_ce_wanted(int M, int N, int K) { return gemm_ps_wanted(M, N, K) && N >= 256; }
int gemm_ps_ce_partials(int N) { return 2 * ceil_div(N, (N >= 1024) ? 256 : 128); }

// pass 1 (want_grad = false): stats[M][gemm_ps_ce_partials(N)] from logits = A B^T + bias, then ce_combine -> row_lse, row_loss;
// pass 2 (want_grad = true): the same product again, gradient written as bf16 copies.  A [M,K] (lda), B [N,K] (ldb), both
// K-major (the vocabulary projection h W_o^T); operand copies come from / go to `arena` like gemm_ps_f32.
cudaError_t gemm_ps_ce(cudaStream_t st, bool want_grad, int M, int N, int K, const float* A, int64_t lda, const float* B, int64_t ldb,
                       const float* bias, const int32_t* tgt, float eps, int ignore_index, const float* denom, float4* stats,
                       float* row_lse, float* row_loss, uint16_t* dl_hi, uint16_t* dl_lo, int64_t dl_ld, int nprod,
                       SplitArena* arena) {
  using namespace ps;
  if (M <= 0 || N <= 0 || K <= 0) return cudaSuccess;
  if (!arena || !get_encode() || !bias || !tgt) return cudaErrorInvalidValue;
  const bool with_lo = (nprod == 3);
  Split sa, sb;
  ACK(get_split(st, arena, A, M, K, lda, with_lo, &sa));
  ACK(get_split(st, arena, B, N, K, ldb, with_lo, &sb));
  const int BN = (N >= 1024) ? 256 : 128;
  Params p = {};
  p.bias = bias; p.M = M; p.N = N; p.K = K; p.alpha = 1.f;
  p.tiles_m = ceil_div(M, BM); p.tiles_n = ceil_div(N, BN); p.splits = 1; p.num_kb = ceil_div(K, BK); p.kb_per_split = p.num_kb;
  p.n_items = p.tiles_m * p.tiles_n;
  p.tgt = tgt; p.stats = stats; p.row_lse = row_lse; p.denom = denom; p.dl_hi = dl_hi; p.dl_lo = dl_lo; p.dl_ld = dl_ld;
  p.eps = eps; p.ignore_index = ignore_index;
  if (want_grad && (!dl_hi || (dl_ld & 7) || (with_lo && !dl_lo))) return cudaErrorInvalidValue;
  CUtensorMap maps[4];
  bool ok = true;
  for (int o = 0; o < (with_lo ? 2 : 1); ++o) {
    ok = ok && make_map_bf16(&maps[o], o ? sa.lo : sa.hi, K, M, sa.ld, BM);
    ok = ok && make_map_bf16(&maps[2 + o], o ? sb.lo : sb.hi, K, N, sb.ld, BN);
  }
  if (!with_lo) { maps[1] = maps[0]; maps[3] = maps[2]; }
  if (!ok) return cudaErrorInvalidValue;
  int nsm = 148;
  { int dev = 0; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev); }
  if (gemm_cta_cap() > 0 && gemm_cta_cap() < nsm) nsm = gemm_cta_cap();
  const int grid = p.n_items < nsm ? p.n_items : nsm;
  cudaError_t e;
  if (BN == 256 && pair_wanted(M, N, K, nprod)) {
    // CTA-pair tiles: same 256-column partition of the classes (the statistics layout does not change), 256-row tiles
    p.tiles_m = ceil_div(M, BM2); p.n_items = p.tiles_m * p.tiles_n;
    bool ok2 = true;
    for (int o = 0; o < (with_lo ? 2 : 1); ++o) ok2 = ok2 && make_map_bf16(&maps[2 + o], o ? sb.lo : sb.hi, K, N, sb.ld, 128);
    if (!with_lo) maps[3] = maps[2];
    if (!ok2) return cudaErrorInvalidValue;
    const int ncl = nsm / 2;
    const int grid2 = 2 * (p.n_items < ncl ? p.n_items : ncl);
    if (!want_grad) {
      e = with_lo ? launch2<false, false, 3, 1>(st, maps, p, grid2) : launch2<false, false, 1, 1>(st, maps, p, grid2);
      ACK(e);
      ALEXCAP_NOTE_LAUNCH();
      ce_combine_kernel<<<ceil_div(M, 8), 256, 0, st>>>(stats, p.tiles_n * 2, tgt, M, N, eps, ignore_index, denom, row_lse, row_loss);
      return cudaGetLastError();
    }
    return with_lo ? launch2<false, false, 3, 2>(st, maps, p, grid2) : launch2<false, false, 1, 2>(st, maps, p, grid2);
  }
  if (!want_grad) {
    if (BN == 256) e = with_lo ? launch<256, false, false, 3, 1>(st, maps, p, grid) : launch<256, false, false, 1, 1>(st, maps, p, grid);
    else e = with_lo ? launch<128, false, false, 3, 1>(st, maps, p, grid) : launch<128, false, false, 1, 1>(st, maps, p, grid);
    ACK(e);
    ALEXCAP_NOTE_LAUNCH();
    ce_combine_kernel<<<ceil_div(M, 8), 256, 0, st>>>(stats, p.tiles_n * 2, tgt, M, N, eps, ignore_index, denom, row_lse, row_loss);
    return cudaGetLastError();
  }
  if (BN == 256) e = with_lo ? launch<256, false, false, 3, 2>(st, maps, p, grid) : launch<256, false, false, 1, 2>(st, maps, p, grid);
  else e = with_lo ? launch<128, false, false, 3, 2>(st, maps, p, grid) : launch<128, false, false, 1, 2>(st, maps, p, grid);
  return e;
}

}  // namespace alexcap
