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
};

// ------------------------------------------------------------------------------------------------ split pass
// X [rows, cols] fp32 (row pitch ldx) -> hi (and lo) [rows, ldo] bf16, ldo = cols rounded up to 8; pad columns are zero.
__global__ void __launch_bounds__(256) split_bf16_kernel(const float* __restrict__ X, int64_t ldx, int rows, int cols,
                                                         uint16_t* __restrict__ hi, uint16_t* __restrict__ lo, int64_t ldo) {
  const int groups = (int)(ldo >> 3);
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
        if (p.act == ACT_TANH) x = tanhf(x);
        else if (p.act == ACT_RELU) x = fmaxf(x, 0.f);
        *cp = x;
      }
    }
  }
  __syncwarp();
}

// ------------------------------------------------------------------------------------------------ kernel
template <int BN, bool A_MN, bool B_MN, int NPROD>
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
        store_chunk(v0, tile, p, it.m0 + q4 * 32, it.n0 + c0 * 32, lane, atomic, it.split == 0);
        store_chunk(v1, tile, p, it.m0 + q4 * 32, it.n0 + c1 * 32, lane, atomic, it.split == 0);
      }
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS));
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

template <int BN, bool A_MN, bool B_MN, int NPROD>
static cudaError_t launch(cudaStream_t st, const CUtensorMap* maps, const Params& p, int grid) {
  constexpr size_t SMEM = (size_t)RING_BYTES + EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
  static bool configured = false;
  auto kern = gemm_ps_kernel<BN, A_MN, B_MN, NPROD>;
  if (!configured) {
    ACK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM));
    configured = true;
  }
  ALEXCAP_NOTE_LAUNCH();
  return launch_pdl(kern, dim3(grid), dim3(NUM_THREADS), SMEM, st, maps[0], maps[1], maps[2], maps[3], p);
}

struct Split { uint16_t* hi; uint16_t* lo; int64_t ld; };

// hi/lo bf16 copy of the fp32 matrix X [rows, cols] (pitch ld) out of the arena: cached per source for the arena's lifetime
static cudaError_t get_split(cudaStream_t st, SplitArena* a, const float* X, int rows, int cols, int64_t ld, bool with_lo,
                             Split* out) {
  for (int i = 0; i < a->n; ++i) {
    const SplitArena::Entry& e = a->e[i];
    if (e.src == X && e.rows == rows && e.cols == cols && e.ld == ld && (e.lo != nullptr || !with_lo)) {
      out->hi = e.hi; out->lo = e.lo; out->ld = e.ldo;
      return cudaSuccess;
    }
  }
  const int64_t ldo = ((int64_t)cols + 7) & ~int64_t(7);
  const size_t one = (((size_t)rows * ldo * 2) + 1023) & ~size_t(1023);
  const size_t need = one * (with_lo ? 2 : 1);
  if (!a->base || a->used + need > a->bytes || a->n >= SplitArena::MAX_ENTRIES) return cudaErrorMemoryAllocation;
  SplitArena::Entry& e = a->e[a->n];
  e.src = X; e.rows = rows; e.cols = cols; e.ld = ld; e.ldo = ldo; e.produced = false;
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
  split_bf16_kernel<<<blocks, 256, 0, st>>>(X, ld, rows, cols, e.hi, e.lo, ldo);
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
  e.src = key; e.rows = rows; e.cols = cols; e.ld = ld; e.ldo = ldo; e.produced = true;
  e.hi = reinterpret_cast<uint16_t*>(a->base + a->used);
  e.lo = with_lo ? reinterpret_cast<uint16_t*>(a->base + a->used + one) : nullptr;
  a->used += need;
  e.end = a->used;
  a->n++;
  *hi = e.hi; *lo = e.lo; *ldo_out = ldo;
  return true;
}

bool split_arena_is_produced(const SplitArena* a, const float* p) {
  if (!a) return false;
  for (int i = 0; i < a->n; ++i)
    if (a->e[i].src == p && a->e[i].produced) return true;
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
  p.tiles_m = tiles_m; p.tiles_n = tiles_n; p.splits = splits; p.kb_per_split = kb_per; p.num_kb = num_kb;
  p.n_items = tiles * splits;
  int nsm = 148;
  { int dev = 0; if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev); }
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

}  // namespace alexcap
