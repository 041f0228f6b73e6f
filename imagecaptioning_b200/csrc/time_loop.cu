// time_loop.cu -- the teacher-forced recurrence of the attention decoder (AlexCap/AttentionLanguageModule.py:72-82) as ONE
// persistent, weight-stationary kernel: every time step of the forward pass inside a single cooperative launch.
//
// What it replaces: the 4-launch-per-step chain of api.cu (query GEMM -> attention -> gates GEMM -> LSTM pointwise; 68
// launches at S = 17) whose GEMMs are latency-bound at M = B = 128 (launch + prologue + first weight fetch per launch).
//
// Structure (grid = max(H/4, B) CTAs <= #SMs, one per SM, 19 warps):
//   * CTA c < H/4 OWNS hidden units [4c, 4c+4): the 16 gate rows (i,f,g,o x 4 units) of W_hh and of W_ih[:, E:], plus a
//     slice of [W_U ; W_beta] rows (the next step's query and beta pre-activations).  They are loaded ONCE, as bf16 hi/lo
//     pairs in the UMMA K-major SWIZZLE_128B layout, and stay in shared memory for all S steps (cfg2: 80 KB per SM).
//   * per step, three phases separated by grid barriers (a release/acquire counter in global memory):
//       P1  [gates_hh | q | bpre] = h_{t-1} [W_hh ; W_U ; W_beta]_c^T : h_{t-1} (bf16 hi/lo, written by P3 of the previous
//           step) streams through a TMA ring as the A operand (M = batch), the resident weights are the B operand (N = 16 +
//           QN), tcgen05.mma x3 (lo*hi + hi*lo + hi*hi), accumulators in TMEM.  The epilogue stores only q / bpre; the 16
//           gate columns STAY in TMEM.
//       P2  the fused attention step (attention_stream.cuh body, one image per CTA); its finaliser also writes gctx as
//           bf16 hi/lo pairs.
//       P3  gates += gctx_t W_ih[:, E:]_c^T accumulated ON TOP of the TMEM columns P1 left; the epilogue reads them once,
//           adds the hoisted embedding projection, and does the LSTM pointwise math in registers (a thread owns one batch
//           row x the CTA's 4 hidden units; c_t never leaves its registers between steps): h_t, c_t, the gate activations
//           for the backward pass, and h_t as the bf16 hi/lo operand of the next step's P1.
//     -> north_star (3): the LSTM pointwise is the tcgen05 epilogue of the gates GEMM and feeds the next step's query GEMM
//        without a kernel boundary; north_star (5) forward half: persistent, weights on chip across steps.
//   * the attention ring and the GEMM A-operand ring are the SAME 128 KB of shared memory (the phases alternate).
// Cost model: each owning CTA must see the whole [B, K] activation (N-slicing is what allows the pointwise epilogue), i.e.
// 2 x 256 KB of L2 -> SM traffic per CTA per step at cfg2; that, not the MMAs (< 0.2 us), is what the GEMM phases cost.
#include <cuda.h>
#include <stdlib.h>

#include "attention_stream.cuh"
#include "kernels.h"
#include "tc_ptx.cuh"

namespace alexcap {

bool& time_loop_enabled();

namespace tloop {

using namespace sattn;

constexpr int LOAD_WARP = SA_CWARPS + 1, MMA_WARP = SA_CWARPS + 2;
constexpr int TL_THREADS = SA_THREADS + 64;          // attention roles + GEMM loader warp + MMA warp
constexpr int GST = 4;                               // A-operand stages (alias the attention ring)
constexpr uint32_t A_HALF = 128 * 128;               // [128 rows][64 bf16] = 16 KB
constexpr uint32_t A_STAGE = 2 * A_HALF;             // hi | lo
constexpr int RING_FLOATS = GST * (int)A_STAGE / 4;  // 128 KB
constexpr int STAGE_FLOATS = (int)A_STAGE / 4;       // attention ring stage = GEMM stage
constexpr int UPC = 4;                               // hidden units per owning CTA -> 16 gate columns
constexpr uint32_t SPIN_LIMIT = 1u << 26;

struct LoopFwdP {
  SAttnFwdP att;                       // geometry + time-invariant pointers; the per-step ones are set in the kernel
  float* QB; float* XH; float* DGQ; float* Call; float* ALPHA; float* CTX; float* BETA;
  uint16_t *Hb_hi, *Hb_lo, *Gb_hi, *Gb_lo;
  const float* bU; const float* bBeta;
  unsigned* bar;
  int S, B, H, D, L, NG, QN, NA, KBA, KBB;   // NA = 16 + QN rows of the P1 weight slice; k-blocks of P1 / P3
  int64_t HD, DH, W5;
  uint32_t wa_bytes, wb_bytes;               // bytes of ONE (hi or lo) copy of the P1 / P3 weight slice
  uint32_t fixed_off, gbar_off, w_off;       // shared-memory map (bytes from the 1024-aligned base)
  uint32_t tmem_cols;
  int extra_syncs;                           // debug (ALEXCAP_LOOP_EXTRA_SYNC=n): n more grid barriers per step (measures their cost)
  unsigned long long* prof;                  // debug: per-step phase time stamps of CTA 0 (ALEXCAP_LOOP_PROF=1), else NULL
};

__device__ __forceinline__ void grid_sync(unsigned* ctr, unsigned target) {
  asm volatile("fence.proxy.async;" ::: "memory");       // generic-proxy global writes -> TMA reads issued by other CTAs
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
    unsigned v, spins = 0;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
      if (++spins > SPIN_LIMIT) __trap();                // a protocol bug must trap, not hang the GPU box
    } while (v < target);
  }
  __syncthreads();
  asm volatile("fence.proxy.async;" ::: "memory");
}

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define TL_STAMP(slot) do { if (p.prof && blockIdx.x == 0 && threadIdx.x == 0) p.prof[(int64_t)t * 8 + (slot)] = gtime(); } while (0)

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}

// ------------------------------------------------------------------------------------------------ weight packing
// Per owning CTA c, NA rows for P1 -- 16 gate rows of W_hh (gate g, unit 4c+j -> row 4g+j), then QN rows of [W_U ; W_beta]
// (columns QN*c .. of [q | bpre]; rows past H+D are zero) -- and 16 rows for P3 (the same gate rows of W_ih[:, E:]), each
// as bf16 hi and lo = bf16(x - hi), row pitch K (a multiple of 8).
__global__ void __launch_bounds__(256) pack_loop_weights_kernel(const float* __restrict__ w_hh, const float* __restrict__ w_ih,
                                                                const float* __restrict__ w_u, const float* __restrict__ w_beta,
                                                                int H, int D, int E, int NG, int QN, uint16_t* __restrict__ a_hi,
                                                                uint16_t* __restrict__ a_lo, uint16_t* __restrict__ b_hi,
                                                                uint16_t* __restrict__ b_lo) {
  const int NA = 16 + QN, H8 = H >> 3, D8 = D >> 3;
  const int64_t na = (int64_t)NG * NA * H8, nb = (int64_t)NG * 16 * D8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < na + nb; i += (int64_t)gridDim.x * blockDim.x) {
    const float* src = nullptr;
    uint16_t *hi, *lo;
    if (i < na) {
      const int k8 = (int)(i % H8);
      const int64_t row = i / H8;
      const int c = (int)(row / NA), j = (int)(row % NA);
      if (j < 16) src = w_hh + ((int64_t)(j >> 2) * H + UPC * c + (j & 3)) * H;
      else {
        const int n = QN * c + (j - 16);
        if (n < H) src = w_u + (int64_t)n * H;
        else if (n < H + D) src = w_beta + (int64_t)(n - H) * H;
      }
      if (src) src += k8 * 8;
      hi = a_hi + row * H + k8 * 8; lo = a_lo + row * H + k8 * 8;
    } else {
      const int64_t i2 = i - na;
      const int k8 = (int)(i2 % D8);
      const int64_t row = i2 / D8;
      const int c = (int)(row / 16), j = (int)(row % 16);
      src = w_ih + ((int64_t)(j >> 2) * H + UPC * c + (j & 3)) * (E + D) + E + k8 * 8;
      hi = b_hi + row * D + k8 * 8; lo = b_lo + row * D + k8 * 8;
    }
    float4 x0 = f4_zero(), x1 = f4_zero();
    if (src) { x0 = __ldg(reinterpret_cast<const float4*>(src)); x1 = __ldg(reinterpret_cast<const float4*>(src) + 1); }
    uint2 h0, l0, h1, l1;
    split_bf16x4(x0, h0, l0); split_bf16x4(x1, h1, l1);
    *reinterpret_cast<uint4*>(hi) = make_uint4(h0.x, h0.y, h1.x, h1.y);
    *reinterpret_cast<uint4*>(lo) = make_uint4(l0.x, l0.y, l1.x, l1.y);
  }
}

// rows x cols fp32 (pitch ld) -> bf16 hi / lo (pitch cols, cols % 8 == 0): h_0 for the first step's P1
__global__ void __launch_bounds__(256) split_rows_kernel(const float* __restrict__ X, int64_t ld, int rows, int cols,
                                                         uint16_t* __restrict__ hi, uint16_t* __restrict__ lo) {
  const int c8 = cols >> 3;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (int64_t)rows * c8; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / c8), k = (int)(i % c8) * 8;
    const float4 x0 = __ldg(reinterpret_cast<const float4*>(X + (int64_t)r * ld + k));
    const float4 x1 = __ldg(reinterpret_cast<const float4*>(X + (int64_t)r * ld + k) + 1);
    uint2 h0, l0, h1, l1;
    split_bf16x4(x0, h0, l0); split_bf16x4(x1, h1, l1);
    *reinterpret_cast<uint4*>(hi + (int64_t)r * cols + k) = make_uint4(h0.x, h0.y, h1.x, h1.y);
    *reinterpret_cast<uint4*>(lo + (int64_t)r * cols + k) = make_uint4(l0.x, l0.y, l1.x, l1.y);
  }
}

// ------------------------------------------------------------------------------------------------ the forward loop
template <int NJ, bool FULL>
__global__ void __launch_bounds__(TL_THREADS, 1)
attn_loop_fwd_kernel(const __grid_constant__ CUtensorMap tmHh, const __grid_constant__ CUtensorMap tmHl,
                     const __grid_constant__ CUtensorMap tmGh, const __grid_constant__ CUtensorMap tmGl,
                     const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
                     const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl, const LoopFwdP p) {
  using namespace tc;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* sm = reinterpret_cast<float*>(smem);                         // attention ring (= GEMM A stages) + attention fixed part
  uint64_t* gfull = reinterpret_cast<uint64_t*>(smem + p.gbar_off);   // [GST]
  uint64_t* gempty = gfull + GST;                                     // [GST]
  uint64_t* tfull = gempty + GST;                                     // accumulator complete (one per GEMM)
  uint64_t* wfull = tfull + 1;                                        // resident weights landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);
  uint8_t* wsm = smem + p.w_off;                                      // WA_hi | WA_lo | WB_hi | WB_lo

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cta = blockIdx.x;
  const bool gemm_cta = cta < p.NG, att_cta = cta < p.B;
  const int S = p.S, B = p.B, H = p.H, D = p.D;

  if (tid == 0) {
    attn_ring_init(sm, p.att.NST, p.att.stage_floats);
    for (int s = 0; s < GST; ++s) { mbar_init(&gfull[s], 1); mbar_init(&gempty[s], 1); }
    mbar_init(tfull, 1); mbar_init(wfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == MMA_WARP) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(p.tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // resident weight slices: one TMA box per 64-wide k-block
  if (gemm_cta && warp == LOAD_WARP && lane == 0) {
    mbar_expect_tx(wfull, 2 * (p.wa_bytes + p.wb_bytes));
    const uint32_t a_kb = (uint32_t)p.NA * 128u;
    for (int kb = 0; kb < p.KBA; ++kb) {
      tma_load_2d(wsm + kb * a_kb, &tmAh, kb * 64, cta * p.NA, wfull);
      tma_load_2d(wsm + p.wa_bytes + kb * a_kb, &tmAl, kb * 64, cta * p.NA, wfull);
    }
    uint8_t* wb = wsm + 2 * p.wa_bytes;
    for (int kb = 0; kb < p.KBB; ++kb) {
      tma_load_2d(wb + kb * 2048, &tmBh, kb * 64, cta * 16, wfull);
      tma_load_2d(wb + p.wb_bytes + kb * 2048, &tmBl, kb * 64, cta * 16, wfull);
    }
  }

  // epilogue threads: batch row r, the CTA's 4 hidden units; c_{t-1} lives in registers
  const int r = warp * 32 + lane;                       // (warps 0..3)
  const bool epi = gemm_cta && warp < 4;
  const bool epi_row = epi && r < B;
  float4 c_prev = f4_zero();
  if (epi_row) c_prev = *reinterpret_cast<const float4*>(p.Call + (int64_t)r * H + UPC * cta);

  Cursor ak{0, 0u};                                     // attention ring cursor (consumers and producer keep their own copy)
  int akg = 0;
  int gs = 0; uint32_t gph = 0;                         // GEMM A-stage cursor (loader lane and MMA lane)
  uint32_t n_gemm = 0;                                  // GEMMs completed so far (tfull parity)
  unsigned bar_target = 0;
  const uint32_t idescA = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(((p.NA + 15) & ~15) >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const uint32_t idescB = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(16 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

  for (int t = 0; t < S; ++t) {
    // ======================================================================== P1: [gates_hh | q | bpre] = h_{t-1} W_A^T
    // ======================================================================== P3: gates += gctx_t W_B^T, LSTM pointwise
#pragma unroll 1
    for (int phase = 0; phase < 2; ++phase) {
      if (phase == 0) TL_STAMP(0);
      if (phase == 1) {
        // ---------------------------------------------------------------------- P2: fused attention step
        TL_STAMP(2);
        if (att_cta && warp <= SA_CWARPS) {
          SAttnFwdP a = p.att;
          a.q = p.QB + (int64_t)t * B * p.HD; a.bpre = a.q + H;
          a.alpha = p.ALPHA + (int64_t)t * p.L;
          a.ctx = p.CTX + (int64_t)t * B * D; a.beta = p.BETA + (int64_t)t * B * D;
          a.gctx = p.XH + (int64_t)t * B * p.DH;
          a.gctx_hi = p.Gb_hi + (int64_t)t * B * D; a.gctx_lo = p.Gb_lo + (int64_t)t * B * D;
          attn_fwd_body<NJ, FULL, 1, false, true>(a, sm, cta, ak, akg);      // Ws in exponential form (ACT_EXP2C)
        }
        if (p.prof) __syncthreads();
        TL_STAMP(3);
        bar_target += gridDim.x;
        grid_sync(p.bar, bar_target);
        TL_STAMP(4);
      }
      if (gemm_cta) {
        const int KB = phase ? p.KBB : p.KBA;
        if (warp == LOAD_WARP) {
          if (lane == 0) {
            const CUtensorMap* mh = phase ? &tmGh : &tmHh;
            const CUtensorMap* ml = phase ? &tmGl : &tmHl;
            const int row0 = t * B;                       // h_{t-1} is slot t of Hb; gctx_t is slot t of Gb
            for (int kb = 0; kb < KB; ++kb) {
              mbar_wait(&gempty[gs], gph ^ 1);
              uint8_t* st = smem + gs * A_STAGE;
              mbar_expect_tx(&gfull[gs], A_STAGE);
              tma_load_2d(st, mh, kb * 64, row0, &gfull[gs]);
              tma_load_2d(st + A_HALF, ml, kb * 64, row0, &gfull[gs]);
              if (++gs == GST) { gs = 0; gph ^= 1; }
            }
          }
        } else if (warp == MMA_WARP) {
          if (t == 0 && phase == 0) mbar_wait(wfull, 0);
          tcgen05_fence_after();
          const uint32_t wbase = smem_u32(wsm) + (phase ? 2 * p.wa_bytes : 0u);
          const uint32_t wlo = phase ? p.wb_bytes : p.wa_bytes;
          const uint32_t w_kb = phase ? 2048u : (uint32_t)p.NA * 128u;
          const uint32_t idesc = phase ? idescB : idescA;
          for (int kb = 0; kb < KB; ++kb) {
            mbar_wait(&gfull[gs], gph);
            tcgen05_fence_after();
            if (lane == 0) {
              const uint32_t a_hi = smem_u32(smem + gs * A_STAGE), a_lo = a_hi + A_HALF;
              const uint32_t b_hi = wbase + kb * w_kb, b_lo = b_hi + wlo;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint64_t dah = make_desc(a_hi + k * 32, 0, 1024), dal = make_desc(a_lo + k * 32, 0, 1024);
                const uint64_t dbh = make_desc(b_hi + k * 32, 0, 1024), dbl = make_desc(b_lo + k * 32, 0, 1024);
                umma_bf16(tmem_base, dal, dbh, idesc, (phase | kb | k) != 0);   // P3 accumulates onto P1's gate columns
                umma_bf16(tmem_base, dah, dbl, idesc, 1);
                umma_bf16(tmem_base, dah, dbh, idesc, 1);
              }
              umma_commit(&gempty[gs]);
              if (kb == KB - 1) umma_commit(tfull);
            }
            __syncwarp();
            if (++gs == GST) { gs = 0; gph ^= 1; }
          }
        } else if (epi) {
          // everything the epilogue needs from global memory is fetched BEFORE waiting for the accumulator: the loads travel
          // under the A-operand stream and the MMAs
          const uint32_t tlane = tmem_base + ((uint32_t)(warp * 32) << 16);
          if (phase == 0) {
            // q / bpre columns [QN*cta, QN*cta + QN) of step t, + bias
            const int nq0 = p.QN * cta;
            float4 b0 = f4_zero(), b1 = f4_zero();
            if (p.QN == 8 && epi_row && nq0 < H + D) {
              const float* bias = (nq0 < H) ? p.bU + nq0 : p.bBeta + (nq0 - H);
              b0 = __ldg(reinterpret_cast<const float4*>(bias)); b1 = __ldg(reinterpret_cast<const float4*>(bias) + 1);
            }
            mbar_wait(tfull, n_gemm & 1);
            tcgen05_fence_after();
            for (int j0 = 0; j0 < p.QN; j0 += 8) {
              uint32_t v[8];
              tmem_ld8(tlane + 16 + j0, v);
              asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
              const int n0 = nq0 + j0;
              if (epi_row && n0 < H + D) {
                if (p.QN != 8) {
                  const float* bias = (n0 < H) ? p.bU + n0 : p.bBeta + (n0 - H);
                  b0 = __ldg(reinterpret_cast<const float4*>(bias)); b1 = __ldg(reinterpret_cast<const float4*>(bias) + 1);
                }
                float4* out = reinterpret_cast<float4*>(p.QB + ((int64_t)t * B + r) * p.HD + n0);
                out[0] = make_float4(__uint_as_float(v[0]) + b0.x, __uint_as_float(v[1]) + b0.y, __uint_as_float(v[2]) + b0.z,
                                     __uint_as_float(v[3]) + b0.w);
                out[1] = make_float4(__uint_as_float(v[4]) + b1.x, __uint_as_float(v[5]) + b1.y, __uint_as_float(v[6]) + b1.z,
                                     __uint_as_float(v[7]) + b1.w);
              }
            }
          } else {
            // LSTM cell (AttentionLanguageModule.py:78, gate order i,f,g,o) straight out of TMEM
            float* g = p.DGQ + ((int64_t)t * B + (epi_row ? r : 0)) * p.W5 + UPC * cta;      // hoisted Emb W_ih[:, :E]^T + b_ih + b_hh
            float4 gi = f4_zero(), gf = f4_zero(), gg = f4_zero(), go = f4_zero();
            if (epi_row) {
              gi = *reinterpret_cast<const float4*>(g); gf = *reinterpret_cast<const float4*>(g + H);
              gg = *reinterpret_cast<const float4*>(g + 2 * H); go = *reinterpret_cast<const float4*>(g + 3 * H);
            }
            mbar_wait(tfull, n_gemm & 1);
            tcgen05_fence_after();
            uint32_t v[16];
            tmem_ld16(tlane, v);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (epi_row) {
              float4 cn, hn;
#define CELL(x, j)                                                                                              \
  gi.x = sigmoid_f(gi.x + __uint_as_float(v[j])); gf.x = sigmoid_f(gf.x + __uint_as_float(v[4 + j]));           \
  gg.x = tanhf(gg.x + __uint_as_float(v[8 + j])); go.x = sigmoid_f(go.x + __uint_as_float(v[12 + j]));           \
  cn.x = fmaf(gf.x, c_prev.x, gi.x * gg.x); hn.x = go.x * tanhf(cn.x);
              CELL(x, 0) CELL(y, 1) CELL(z, 2) CELL(w, 3)
#undef CELL
              // h_t first: it is what the next phase of every other CTA waits for
              uint2 h2, l2;
              split_bf16x4(hn, h2, l2);
              *reinterpret_cast<uint2*>(p.Hb_hi + ((int64_t)(t + 1) * B + r) * H + UPC * cta) = h2;
              *reinterpret_cast<uint2*>(p.Hb_lo + ((int64_t)(t + 1) * B + r) * H + UPC * cta) = l2;
              *reinterpret_cast<float4*>(p.XH + ((int64_t)(t + 1) * B + r) * p.DH + D + UPC * cta) = hn;
              *reinterpret_cast<float4*>(p.Call + ((int64_t)(t + 1) * B + r) * H + UPC * cta) = cn;
              *reinterpret_cast<float4*>(g) = gi; *reinterpret_cast<float4*>(g + H) = gf;
              *reinterpret_cast<float4*>(g + 2 * H) = gg; *reinterpret_cast<float4*>(g + 3 * H) = go;
              c_prev = cn;
            }
          }
          tcgen05_fence_before();
        }
      }
      ++n_gemm;
      if (p.prof) __syncthreads();
      TL_STAMP(phase ? 5 : 1);
      bar_target += gridDim.x;
      grid_sync(p.bar, bar_target);
      if (phase) TL_STAMP(6);
    }
    for (int x = 0; x < p.extra_syncs; ++x) { bar_target += gridDim.x; grid_sync(p.bar, bar_target); }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols));
  }
}

static bool make_map_bf16(CUtensorMap* map, const uint16_t* base, int64_t inner, int64_t outer, int64_t ld, int box_outer) {
  tc::EncodeTiledFn enc = tc::get_encode();
  if (!enc) return false;
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1u, 1u};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<uint16_t*>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct Plan {
  bool ok; int NG, QN, NA, KBA, KBB, grid, RH, RD, seg_pad;
  uint32_t wa_bytes, wb_bytes, fixed_off, gbar_off, w_off, tmem_cols; size_t smem;
};

static int sm_count() {
  static int n = 0;
  if (!n) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev); if (n <= 0) n = 148; }
  return n;
}

static Plan make_plan(int B, int L, int D, int H) {
  Plan pl = {};
  if (!time_loop_enabled() || B > 128 || (H % 8) || (D % 8) || H % UPC || H > 512 || D > 1024) return pl;
  pl.NG = H / UPC;
  const int nsm = sm_count();
  if (pl.NG > nsm || B > nsm) return pl;
  const int chunks = (H + D) / 8;
  pl.QN = 8 * ((chunks + pl.NG - 1) / pl.NG);
  pl.NA = 16 + pl.QN;
  const int na_pad = (pl.NA + 15) & ~15;
  if (na_pad > 256) return pl;
  pl.tmem_cols = 32;
  while ((int)pl.tmem_cols < na_pad) pl.tmem_cols *= 2;
  pl.KBA = (H + 63) / 64; pl.KBB = (D + 63) / 64;
  pl.wa_bytes = (uint32_t)pl.KBA * pl.NA * 128u; pl.wb_bytes = (uint32_t)pl.KBB * 2048u;
  pl.grid = pl.NG > B ? pl.NG : B;
  auto rows_for = [&](int n) { int r = 2 * SA_GWARPS; while (r > 1 && r * n > STAGE_FLOATS) --r; return r > L ? L : r; };
  pl.RH = rows_for(H); pl.RD = rows_for(D);
  if (H + D > STAGE_FLOATS) return pl;                          // header piece
  pl.seg_pad = (L + 3) & ~3;
  // attention fixed part (attn_fwd_body with NC = 1): barriers, red, es, scratch, bp_s, segb/segflag
  const size_t att_fixed = 2 * SA_MAX_NST * 8 + (size_t)SA_CTHREADS * 16 + (size_t)(pl.seg_pad + 40 + D) * 4 + 2 * 2 * 4 + 16;
  pl.fixed_off = RING_FLOATS * 4;
  pl.gbar_off = (uint32_t)((pl.fixed_off + att_fixed + 15) & ~size_t(15));
  // the P1 MMA reads 16-row groups: the last k-block of a NA = 24-row slice reads 8 rows past its end -> keep 1 KB of slack
  pl.w_off = (uint32_t)((pl.gbar_off + (2 * GST + 2) * 8 + 16 + 1023) & ~size_t(1023));
  pl.smem = (size_t)pl.w_off + 2 * (size_t)(pl.wa_bytes + pl.wb_bytes) + 1024 /*tail slack*/ + 1024 /*alignment*/;
  if (pl.smem > 227 * 1024) return pl;
  pl.ok = true;
  return pl;
}

static unsigned long long* prof_buffer() {
  static unsigned long long* buf = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    const char* v = getenv("ALEXCAP_LOOP_PROF");
    if (v && v[0] == '1' && cudaMalloc(&buf, 2 * 64 * 8 * sizeof(unsigned long long)) != cudaSuccess) buf = nullptr;
  }
  return buf;
}

}  // namespace tloop

// Measured on B200 at cfg2 (profiles/r2_time_loop_phases.txt): 42.1 us per step inside the persistent kernel against ~35 us for the
// PDL-chained per-step launches, so the chain stays the default; ALEXCAP_TIME_LOOP=1 or alexcap_set_time_loop(1) selects the loop.
bool& time_loop_enabled() {
  static bool on = [] { const char* v = getenv("ALEXCAP_TIME_LOOP"); return v && v[0] == '1'; }();
  return on;
}

// debug hook (ALEXCAP_LOOP_PROF=1): copies the phase time stamps (ns, %globaltimer) of the last forward / backward loop launch
// -- [dir][step][8] -- to `host_out`; returns the number of uint64 written (0 when profiling is off).  Synchronises the device.
int time_loop_profile_read(unsigned long long* host_out, int max_entries) {
  unsigned long long* b = tloop::prof_buffer();
  if (!b) return 0;
  const int n = max_entries < 2 * 64 * 8 ? max_entries : 2 * 64 * 8;
  if (cudaDeviceSynchronize() != cudaSuccess) return 0;
  if (cudaMemcpy(host_out, b, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost) != cudaSuccess) return 0;
  return n;
}

bool attn_loop_supported(int B, int L, int D, int H) { return tloop::make_plan(B, L, D, H).ok; }

size_t attn_loop_workspace_elems_bf16(int B, int S, int D, int H) {
  const tloop::Plan pl = tloop::make_plan(B, 1, D, H);
  if (!pl.ok) return 0;
  return 2 * ((size_t)(S + 1) * B * H + (size_t)S * B * D + (size_t)pl.NG * pl.NA * H + (size_t)pl.NG * 16 * D) + 4096;
}

// Forward recurrence of one training step (replaces the per-step loop of attn_fwd_core).  bf16ws: attn_loop_workspace_elems_bf16
// elements; bar: 4 unsigned (zeroed here).
cudaError_t attn_loop_fwd(cudaStream_t st, int B, int L, int D, int H, int E, int S, const float* feat, const float* Ws,
                          const float* w_hh, const float* w_ih, const float* w_u, const float* b_u, const float* w_beta,
                          const float* b_beta, const float* wv, float* QB, float* XH, float* DGQ, float* Call, float* ALPHA,
                          float* CTX, float* BETA, uint16_t* bf16ws, unsigned* bar) {
  using namespace tloop;
  const Plan pl = make_plan(B, L, D, H);
  if (!pl.ok || !tc::get_encode()) return cudaErrorInvalidValue;
  uint16_t* Hb_hi = bf16ws; uint16_t* Hb_lo = Hb_hi + (size_t)(S + 1) * B * H;
  uint16_t* Gb_hi = Hb_lo + (size_t)(S + 1) * B * H; uint16_t* Gb_lo = Gb_hi + (size_t)S * B * D;
  uint16_t* WA_hi = Gb_lo + (size_t)S * B * D; uint16_t* WA_lo = WA_hi + (size_t)pl.NG * pl.NA * H;
  uint16_t* WB_hi = WA_lo + (size_t)pl.NG * pl.NA * H; uint16_t* WB_lo = WB_hi + (size_t)pl.NG * 16 * D;
  const int64_t DH = D + H, HD = H + D, W5 = 5 * (int64_t)H + D;

  ALEXCAP_NOTE_LAUNCH();
  pack_loop_weights_kernel<<<296, 256, 0, st>>>(w_hh, w_ih, w_u, w_beta, H, D, E, pl.NG, pl.QN, WA_hi, WA_lo, WB_hi, WB_lo);
  ACK(cudaGetLastError());
  ALEXCAP_NOTE_LAUNCH();
  split_rows_kernel<<<ceil_div(B * (H / 8), 256), 256, 0, st>>>(XH + D, DH, B, H, Hb_hi, Hb_lo);
  ACK(cudaGetLastError());
  ACK(cudaMemsetAsync(bar, 0, 4 * sizeof(unsigned), st));

  CUtensorMap maps[8];
  bool ok = make_map_bf16(&maps[0], Hb_hi, H, (int64_t)(S + 1) * B, H, 128) && make_map_bf16(&maps[1], Hb_lo, H, (int64_t)(S + 1) * B, H, 128) &&
            make_map_bf16(&maps[2], Gb_hi, D, (int64_t)S * B, D, 128) && make_map_bf16(&maps[3], Gb_lo, D, (int64_t)S * B, D, 128) &&
            make_map_bf16(&maps[4], WA_hi, H, (int64_t)pl.NG * pl.NA, H, pl.NA) && make_map_bf16(&maps[5], WA_lo, H, (int64_t)pl.NG * pl.NA, H, pl.NA) &&
            make_map_bf16(&maps[6], WB_hi, D, (int64_t)pl.NG * 16, D, 16) && make_map_bf16(&maps[7], WB_lo, D, (int64_t)pl.NG * 16, D, 16);
  if (!ok) return cudaErrorInvalidValue;

  LoopFwdP p = {};
  SAttnFwdP& a = p.att;
  a.Ws = Ws; a.feat = feat; a.q_ld = HD; a.bpre_ld = HD; a.wv = wv; a.alpha_ld = (int64_t)S * L; a.gctx_ld = DH;
  a.scratch = reinterpret_cast<float*>(bar); a.counters = bar;       // never dereferenced: every image lives in one CTA
  a.B = B; a.L = L; a.D = D; a.H = H; a.rpi = 1; a.rp = L; a.RH = pl.RH; a.RD = pl.RD; a.NST = GST; a.stage_floats = STAGE_FLOATS;
  a.maxseg = 2; a.seg_pad = pl.seg_pad; a.slot_floats = 4 + D; a.dry = 0; a.l2_frac = 0.f; a.gb_ld = D;
  p.QB = QB; p.XH = XH; p.DGQ = DGQ; p.Call = Call; p.ALPHA = ALPHA; p.CTX = CTX; p.BETA = BETA;
  p.Hb_hi = Hb_hi; p.Hb_lo = Hb_lo; p.Gb_hi = Gb_hi; p.Gb_lo = Gb_lo; p.bU = b_u; p.bBeta = b_beta; p.bar = bar;
  p.S = S; p.B = B; p.H = H; p.D = D; p.L = L; p.NG = pl.NG; p.QN = pl.QN; p.NA = pl.NA; p.KBA = pl.KBA; p.KBB = pl.KBB;
  p.HD = HD; p.DH = DH; p.W5 = W5; p.wa_bytes = pl.wa_bytes; p.wb_bytes = pl.wb_bytes;
  p.fixed_off = pl.fixed_off; p.gbar_off = pl.gbar_off; p.w_off = pl.w_off; p.tmem_cols = pl.tmem_cols;
  p.prof = (S <= 64) ? prof_buffer() : nullptr;
  { static const int xs = [] { const char* v = getenv("ALEXCAP_LOOP_EXTRA_SYNC"); return v ? atoi(v) : 0; }(); p.extra_syncs = xs; }

  const int nj = ceil_div(H, 128);
  void* args[] = {&maps[0], &maps[1], &maps[2], &maps[3], &maps[4], &maps[5], &maps[6], &maps[7], &p};
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pl.grid); cfg.blockDim = dim3(TL_THREADS); cfg.dynamicSmemBytes = pl.smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;      // all CTAs co-resident: they wait for each other
  cfg.attrs = at; cfg.numAttrs = 1;
#define TL_LAUNCH(NJ, FULL)                                                                                        \
  do {                                                                                                             \
    auto kern = attn_loop_fwd_kernel<NJ, FULL>;                                                                    \
    static size_t granted = 0;                                                                                     \
    if (pl.smem > granted) {                                                                                       \
      ACK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));                  \
      granted = pl.smem;                                                                                           \
    }                                                                                                              \
    ALEXCAP_NOTE_LAUNCH();                                                                                         \
    return cudaLaunchKernelExC(&cfg, reinterpret_cast<const void*>(kern), args);                                   \
  } while (0)
  const bool full = (H == 128 * nj);
  if (nj <= 1) { if (full) TL_LAUNCH(1, true); TL_LAUNCH(1, false); }
  if (nj <= 2) { if (full) TL_LAUNCH(2, true); TL_LAUNCH(2, false); }
  if (full) TL_LAUNCH(4, true);
  TL_LAUNCH(4, false);
#undef TL_LAUNCH
}

}  // namespace alexcap
