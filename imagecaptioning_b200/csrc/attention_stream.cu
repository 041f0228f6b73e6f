// attention_stream.cu -- the fused additive soft-attention step (forward and backward) as PERSISTENT, TMA-fed
// streaming kernels (the default attention engine; attention.cu keeps the direct-load cluster kernels as engine 0).
//
// Math: AlexCap/AttentionLanguageModule.py:15-22 (Attention.forward) + :73-75 (beta gate), per timestep.
//
// The step is HBM-bound: Ws[b] (L x H) and feat[b] (L x D) are read exactly once per launch.  What the first kernels
// lost was not bytes but time: one cluster per image meant 512 CTAs for 444 resident slots (a second, mostly empty
// wave), and every CTA drained its loads at each phase boundary.  Here:
//   * the B*L rows are cut into EQUAL contiguous ranges, one per persistent CTA (grid = SMs x ctas_per_sm), so every SM
//     moves the same number of bytes; a range is walked image segment by image segment;
//   * a producer warp streams the segment's Ws rows, then its feat rows, in pieces of R rows through an NST-deep
//     shared-memory ring with 1-D TMA bulk copies (cp.async.bulk + mbarrier complete_tx); it runs ahead across phase
//     and segment boundaries, so the bytes in flight per SM are set by the ring (~100-200 KB), not by the compiler;
//   * consumer warps compute out of shared memory: scores warp-per-row, context thread-per-column;
//   * an image cut across CTAs is combined flash-decoding style through a small global scratch: every segment
//     publishes (max, sum, unnormalised context) [backward: partial sums that are LINEAR in the softmax-backward
//     scalar], bumps the image's counter, and the CTA that arrives last finalises the image.  No CTA ever waits for
//     another one.
#include <stdio.h>
#include <stdlib.h>

#include <mutex>
#include <vector>

#include "common.cuh"
#include "kernels.h"
#include "pipe.cuh"
#include "attention_stream.cuh"

namespace alexcap {

using namespace sattn;

namespace {

// Fraction of the streamed Ws / feat lines loaded with an L2 evict_last policy (ALEXCAP_ATT_L2_FRAC, 0 = plain loads).
float l2_fraction() {
  static const float f = [] {
    const char* v = getenv("ALEXCAP_ATT_L2_FRAC");
    const float x = (v && *v) ? (float)atof(v) : 0.f;
    if (x > 0.f) {        // evict_last lines live in the persisting set-aside of L2: size it (ALEXCAP_ATT_L2_MB, default = max)
      int dev = 0, maxb = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&maxb, cudaDevAttrMaxPersistingL2CacheSize, dev);
      const char* mb = getenv("ALEXCAP_ATT_L2_MB");
      size_t want = (mb && *mb) ? (size_t)atoll(mb) << 20 : (size_t)maxb;
      if (want > (size_t)maxb) want = (size_t)maxb;
      cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
      fprintf(stderr, "[alexcap] persisting L2 set-aside %zu MB (max %d MB)\n", want >> 20, maxb >> 20);
    }
    return x;
  }();
  return f < 0.f ? 0.f : (f > 1.f ? 1.f : f);
}
int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}
struct Tune { int cps, rows, nst, piece_kb; };
// Debug/tuning knobs (read once): CTAs per SM, rows per piece, ring depth.
const Tune& tune_override() {
  static Tune t = {env_int("ALEXCAP_ATT_CPS", 0), env_int("ALEXCAP_ATT_ROWS", 0), env_int("ALEXCAP_ATT_NST", 0),
                   env_int("ALEXCAP_ATT_PIECE_KB", 0)};
  return t;
}

// debug: device buffer (>= 528 u64) receiving CTA 0's phase time stamps, address given in ALEXCAP_ATT_PROF_PTR (scripts/attn_prof.py)
unsigned long long* prof_ptr() {
  static unsigned long long* p = [] {
    const char* v = getenv("ALEXCAP_ATT_PROF_PTR");
    return v ? reinterpret_cast<unsigned long long*>(strtoull(v, nullptr, 0)) : nullptr;
  }();
  return p;
}

// ALEXCAP_ATT_PREFETCH=1: fetch row pieces before griddepcontrol.wait where the caller says the rows are complete.  Measured:
// no gain on the step (cfg2 2.0537 vs 2.0554 ms, cfg3 5.9095 vs 5.8980) -- the passes are bound by the ~72 GB/s per-SM
// ingest while 128 SMs stream at once, not by the first pieces' latency -- so it stays off.
bool prefetch_before_wait() {
  static const bool on = [] { const char* v = getenv("ALEXCAP_ATT_PREFETCH"); return v && v[0] == '1'; }();
  return on;
}

// Static header members (backward: q, beta, ctx, the reference row of Ws) copied before griddepcontrol.wait, only d gctx after
// it (ALEXCAP_ATT_HDR_SPLIT=0 disables).
bool header_split() {
  static const bool on = [] { const char* v = getenv("ALEXCAP_ATT_HDR_SPLIT"); return !(v && v[0] == '0'); }();
  return on;
}

int sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

struct Geometry {
  int nCTA, rp, RH, RD, NST, stage_floats, maxseg, seg_rows;   // rp: rows per CTA; RH / RD: rows per piece of Ws / feat rows
};

Geometry pick_geometry(int B, int L, int D, int H, size_t fixed_smem_bytes, int header_floats, int esz = 4) {
  Geometry g;
  const int64_t total = (int64_t)B * L;
  const int cps = tune_override().cps > 0 ? tune_override().cps : 1;
  int64_t n = (int64_t)sm_count() * cps;
  const int64_t by_rows = (total + 7) / 8;                       // at least 8 rows per CTA
  if (n > by_rows) n = by_rows;
  if (n < 1) n = 1;
  g.rp = (int)((total + n - 1) / n);
  // An image that lives in ONE CTA needs no cross-CTA combine (no scratch, fence or atomic): when the batch fits the
  // grid, give every CTA exactly one image rather than cutting the rows into perfectly equal ranges.
  static const bool split_ok = env_int("ALEXCAP_ATT_SPLIT_IMAGES", 0) != 0;     // experiment: equal row ranges over ALL SMs even when B <= #SMs
  if (!split_ok && B <= n && g.rp <= L) g.rp = L;
  // More images than CTAs: whole images per CTA (ceil(B / n) each) instead of equal row ranges when that costs at most
  // `whole_pct` % more rows on the fullest CTA -- no image is cut, so no partials, fence, atomic or finaliser pass.
  // Measured: cfg3 (B = 512: 4 images = 196 rows per CTA on 128 CTAs instead of 170 rows on 148) 5.586 -> 5.210 ms per step,
  // cfg5 (B = 256) 8.236 -> 7.458 ms: the cut images cost far more than the 15 % of extra rows.
  static const int whole_pct = env_int("ALEXCAP_ATT_WHOLE_PCT", 35);      // (0: always equal row ranges, as before)
  if (whole_pct > 0 && B > n) {
    const int64_t per = (B + n - 1) / n;
    if (per * L * 100 <= (int64_t)g.rp * (100 + whole_pct)) g.rp = (int)(per * L);
  }
  g.nCTA = (int)((total + g.rp - 1) / g.rp);
  g.seg_rows = g.rp < L ? g.rp : L;
  g.maxseg = (g.rp + L - 1) / L + 1;
  // Rows per piece, chosen per row width so that BOTH passes fill a stage: with H != D (cfg3: H = 1024, D = 2048) one
  // common row count left the narrow pass with half-empty stages and half of its warp-per-row consumers idle.
  const size_t piece_cap = (size_t)(tune_override().piece_kb > 0 ? tune_override().piece_kb : 40) * 1024;
  auto rows_for = [&](int n) {
    int r = tune_override().rows > 0 ? tune_override().rows : 2 * SA_GWARPS;
    while (r > 1 && (size_t)r * n * esz > piece_cap) --r;            // pieces of at most 40 KB (16 rows of <= 2.5 KB: 32 KB at cfg2)
    return r > g.seg_rows ? g.seg_rows : r;
  };
  g.RH = rows_for(H); g.RD = rows_for(D);
  g.stage_floats = (g.RH * H > g.RD * D ? g.RH * H : g.RD * D) * esz / 4;     // (rows of esz-byte elements; H, D % 8 == 0 when esz == 2)
  if (g.stage_floats < header_floats) g.stage_floats = header_floats;
  const size_t budget = (cps == 1 ? 216 : 100) * 1024;
  int nst = tune_override().nst > 0 ? tune_override().nst
                                    : (int)((budget - fixed_smem_bytes) / ((size_t)g.stage_floats * 4));
  if (nst > SA_MAX_NST) nst = SA_MAX_NST;
  // The ring depth MUST be a multiple of the number of consumer groups: then every stage is always consumed by the
  // same group, which therefore observes every phase of that stage's `full` barrier in order.  (Parity waits alias:
  // a group arriving a whole ring ahead at a stage whose previous use -- owned by the other group -- is still
  // incomplete would see "parity matches the preceding phase" and read the stage before it is filled.)
  nst = nst / SA_NG * SA_NG;
  if (nst < 2 * SA_NG) nst = 2 * SA_NG;
  g.NST = nst;
  return g;
}

// Process-lifetime device scratch for the cross-CTA combine (grown on demand, never freed or moved so that captured
// CUDA graphs stay valid).  One attention launch in flight per device (the library's single-stream contract).
struct Scratch { float* buf = nullptr; size_t floats = 0; unsigned* counters = nullptr; size_t ncount = 0; };
cudaError_t get_scratch(size_t floats, size_t ncount, float** buf, unsigned** counters) {
  static Scratch per_dev[64];
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  int dev = 0;
  ACK(cudaGetDevice(&dev));
  Scratch& s = per_dev[dev & 63];
  if (floats > s.floats) {
    size_t want = floats < ((size_t)4 << 20) ? ((size_t)4 << 20) : floats * 2;
    float* p = nullptr;
    ACK(cudaMalloc(&p, want * sizeof(float)));     // the previous block (if any) is intentionally kept alive
    s.buf = p; s.floats = want;
  }
  if (ncount > s.ncount) {
    size_t want = ncount < 8192 ? 8192 : ncount * 2;
    unsigned* p = nullptr;
    ACK(cudaMalloc(&p, want * sizeof(unsigned)));
    ACK(cudaMemset(p, 0, want * sizeof(unsigned)));
    s.counters = p; s.ncount = want;
  }
  *buf = s.buf; *counters = s.counters;
  return cudaSuccess;
}

template <typename K>
cudaError_t set_smem(K kernel, size_t bytes) {
  static size_t granted = 48 * 1024;
  static K granted_for = nullptr;
  if (bytes <= 48 * 1024) return cudaSuccess;
  if (granted_for == kernel && bytes <= granted) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) { granted = bytes; granted_for = kernel; }
  return e;
}

}  // namespace

bool attn_stream_supported(int D, int H) {
  return !(H & 3) && !(D & 3) && H <= 1024 && D <= 4 * SA_GTHREADS * SA_NC;
}

// ------------------------------------------------------------------------------------------ forward step
// Stand-alone kernels: one launch per time step around the shared bodies (attention_stream.cuh).
template <int NJ, bool FULL, bool BF, bool EW>
__global__ void __launch_bounds__(SA_THREADS, 1) attn_fwd_stream_kernel(const SAttnFwdP a) {
  extern __shared__ __align__(128) float sm[];
  if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) a.prof[0] = sa_gtime();
  if (threadIdx.x == 0) attn_ring_init(sm, a.NST, a.stage_floats);
  __syncthreads();
  pdl_trigger();                      // (griddepcontrol.wait is executed inside the body: a.pdl)
  Cursor k{0, 0u};
  int kg = 0;
  attn_fwd_body<NJ, FULL, SA_NC, BF, EW>(a, sm, (int)blockIdx.x, k, kg);
  if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) {        // debug timeline: (entry, body entry, end) of every launch
    const unsigned long long n = a.prof[7]++ & 63ull;
    unsigned long long* tl = a.prof + 16 + 4 * n;
    tl[0] = a.prof[0]; tl[1] = a.prof[1]; tl[2] = sa_gtime();
  }
}

template <bool ALLWHOLE, bool BF, bool EW>
__global__ void __launch_bounds__(SA_THREADS, 1) attn_bwd_stream_kernel(const SAttnBwdP a) {
  extern __shared__ __align__(128) float sm[];
  if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) a.prof[0] = sa_gtime();
  if (threadIdx.x == 0) attn_ring_init(sm, a.NST, a.stage_floats);
  __syncthreads();
  pdl_trigger();                      // (griddepcontrol.wait is executed inside the body: a.pdl)
  Cursor k{0, 0u};
  int kg = 0;
  attn_bwd_body<ALLWHOLE, BF, EW>(a, sm, (int)blockIdx.x, k, kg);
  if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) {
    const unsigned long long n = a.prof[7]++ & 63ull;
    unsigned long long* tl = a.prof + 8 + 256 + 4 * n;          // (a.prof = buffer + 8 for the backward kernel)
    tl[0] = a.prof[0]; tl[1] = a.prof[1]; tl[2] = sa_gtime();
  }
}

static size_t fwd_fixed_smem(int D, int seg_pad, int maxseg) {
  return 2 * SA_MAX_NST * 8 + (size_t)SA_NC * SA_CTHREADS * 16 + (size_t)(seg_pad + 40 + D) * 4 + (size_t)2 * maxseg * 4 + 16;
}

cudaError_t attn_step_fwd_stream(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                                 const float* q, int64_t q_ld, const float* bpre, int64_t bpre_ld, const float* wv,
                                 float* alpha, int64_t alpha_ld, float* ctx, float* beta, float* gctx, int64_t gctx_ld,
                                 int rows_per_image, bool rows_bf16, bool ws_exp, bool static_ready, uint16_t* gctx_b16, int64_t gb_ld) {
  if (!attn_stream_supported(D, H) || (q_ld & 3) || (bpre_ld & 3) || (gctx_ld & 3)) return cudaErrorInvalidValue;
  if (rows_bf16 && ((D & 7) || (H & 7))) return cudaErrorInvalidValue;      // 16-byte bulk copies of bf16 rows
  if (B <= 0 || L <= 0) return cudaSuccess;
  const int esz = rows_bf16 ? 2 : 4;
  const Geometry g0 = pick_geometry(B, L, D, H, 0, H + D, esz);      // maxseg first (it sizes the fixed part)
  const Geometry g = pick_geometry(B, L, D, H, fwd_fixed_smem(D, (L + 3) & ~3, g0.maxseg), H + D, esz);
  const int seg_pad = (g.seg_rows + 3) & ~3;
  const int slot_floats = 4 + D;
  float* scratch = nullptr;
  unsigned* counters = nullptr;
  ACK(get_scratch((size_t)g.nCTA * g.maxseg * slot_floats, (size_t)B, &scratch, &counters));
  SAttnFwdP a{Ws, feat, q, q_ld, bpre, bpre_ld, wv, alpha, alpha_ld, ctx, beta, gctx, gctx_ld, scratch, counters,
              B, L, D, H, rows_per_image < 1 ? 1 : rows_per_image, g.rp, g.RH, g.RD, g.NST, g.stage_floats, g.maxseg, seg_pad,
              slot_floats, env_int("ALEXCAP_ATT_DRY", 0), l2_fraction()};
  a.bf16 = rows_bf16 ? 1 : 0;
  a.ws_exp = ws_exp ? 1 : 0;
  a.pdl = !static_ready ? 1 : (prefetch_before_wait() ? 2 : (header_split() ? 3 : 1));
  a.prof = prof_ptr();
  if (gctx_b16) {
    if ((gb_ld & 3) || (reinterpret_cast<uintptr_t>(gctx_b16) & 7)) return cudaErrorInvalidValue;
    a.gctx_hi = gctx_b16; a.gctx_lo = nullptr; a.gb_ld = gb_ld;
  }
  const size_t smem = (size_t)g.NST * g.stage_floats * 4 + fwd_fixed_smem(D, seg_pad, g.maxseg);
  const int nj = ceil_div(H, 128);
#define LAUNCH_FWD_K(NJ, FULL, BF, EW)                                                      \
  do {                                                                                      \
    ACK(set_smem(attn_fwd_stream_kernel<NJ, FULL, BF, EW>, smem));                          \
    ALEXCAP_NOTE_LAUNCH();                                                                  \
    ACK(launch_pdl(attn_fwd_stream_kernel<NJ, FULL, BF, EW>, dim3(g.nCTA), dim3(SA_THREADS), smem, st, a));   \
    return cudaGetLastError();                                                              \
  } while (0)
#define LAUNCH_FWD_E(NJ, FULL, BF)                                                          \
  do { if (ws_exp) LAUNCH_FWD_K(NJ, FULL, BF, true); LAUNCH_FWD_K(NJ, FULL, BF, false); } while (0)
#define LAUNCH_FWD(NJ)                                                                      \
  do {                                                                                      \
    if (H == 128 * NJ) { if (rows_bf16) LAUNCH_FWD_E(NJ, true, true); LAUNCH_FWD_E(NJ, true, false); }   \
    if (rows_bf16) LAUNCH_FWD_E(NJ, false, true);                                           \
    LAUNCH_FWD_E(NJ, false, false);                                                         \
  } while (0)
  if (nj <= 1) LAUNCH_FWD(1);
  else if (nj <= 2) LAUNCH_FWD(2);
  else if (nj <= 4) LAUNCH_FWD(4);
  else LAUNCH_FWD(8);
#undef LAUNCH_FWD_K
#undef LAUNCH_FWD_E
#undef LAUNCH_FWD
}


// ------------------------------------------------------------------------------------------ backward step
static size_t bwd_fixed_smem(int D, int H, int seg_pad, int rp_pad, int maxseg) {
  return 2 * SA_MAX_NST * 8 + (size_t)4 * SA_CTHREADS * 16 + (size_t)(2 * D + 2 * H + seg_pad + rp_pad + 40) * 4 +
         (size_t)2 * maxseg * 4 + 16;
}

cudaError_t attn_step_bwd_stream(cudaStream_t st, int B, int L, int D, int H, const float* Ws, const float* feat,
                                 const float* q, int64_t q_ld, const float* wv, const float* alpha, int64_t alpha_ld,
                                 const float* ctx, const float* beta, const float* dgctx, int64_t dgctx_ld,
                                 const float* dalpha_ext, int64_t ext_ld, const float* reg_r,
                                 float* dq, int64_t dq_ld, float* dbpre, int64_t dbpre_ld, float* de, int64_t de_ld,
                                 float* dwv_part, float* dfeat, bool rows_bf16, bool ws_exp, bool static_ready,
                                 uint16_t* dq_b16, uint16_t* dbpre_b16, int64_t db_ld) {
  if (!attn_stream_supported(D, H) || (q_ld & 3) || (dgctx_ld & 3) || (dq_ld & 3) || (dbpre_ld & 3))
    return cudaErrorInvalidValue;
  if (rows_bf16 && ((D & 7) || (H & 7))) return cudaErrorInvalidValue;
  if (B <= 0 || L <= 0) return cudaSuccess;
  const int esz = rows_bf16 ? 2 : 4;
  const int header = 2 * H + 3 * D;
  const Geometry g0 = pick_geometry(B, L, D, H, 0, header, esz);
  const Geometry g = pick_geometry(B, L, D, H, bwd_fixed_smem(D, H, (L + 3) & ~3, (g0.rp + 3) & ~3, g0.maxseg), header, esz);
  const int seg_pad = (g.seg_rows + 3) & ~3, rp_pad = (g.rp + 3) & ~3;
  const int slot_floats = 4 + 4 * H;
  float* scratch = nullptr;
  unsigned* counters = nullptr;
  ACK(get_scratch((size_t)g.nCTA * g.maxseg * slot_floats, (size_t)B, &scratch, &counters));
  SAttnBwdP a{Ws, feat, q, q_ld, wv, alpha, alpha_ld, ctx, beta, dgctx, dgctx_ld, dalpha_ext, ext_ld, reg_r,
              dq, dq_ld, dbpre, dbpre_ld, de, de_ld, dwv_part, dfeat, scratch, counters,
              B, L, D, H, g.rp, g.RH, g.RD, g.NST, g.stage_floats, g.maxseg, seg_pad, rp_pad, slot_floats, l2_fraction()};
  const size_t smem = (size_t)g.NST * g.stage_floats * 4 + bwd_fixed_smem(D, H, seg_pad, rp_pad, g.maxseg);
  a.bf16 = rows_bf16 ? 1 : 0;
  a.ws_exp = ws_exp ? 1 : 0;
  a.pdl = !static_ready ? 1 : (prefetch_before_wait() ? 2 : (header_split() ? 3 : 1));
  a.prof = prof_ptr() ? prof_ptr() + 8 : nullptr;
  if (dq_b16 || dbpre_b16) {
    if (!dq_b16 || !dbpre_b16 || (db_ld & 3) || (reinterpret_cast<uintptr_t>(dq_b16) & 7) || (reinterpret_cast<uintptr_t>(dbpre_b16) & 7))
      return cudaErrorInvalidValue;
    a.dq_b16 = dq_b16; a.dbpre_b16 = dbpre_b16; a.db_ld = db_ld;
  }
#define LAUNCH_BWD_K(AW, BF, EW)                                                            \
  do {                                                                                      \
    ACK(set_smem(attn_bwd_stream_kernel<AW, BF, EW>, smem));                                \
    ALEXCAP_NOTE_LAUNCH();                                                                  \
    return launch_pdl(attn_bwd_stream_kernel<AW, BF, EW>, dim3(g.nCTA), dim3(SA_THREADS), smem, st, a);   \
  } while (0)
#define LAUNCH_BWD(AW, BF)                                                                  \
  do { if (ws_exp) LAUNCH_BWD_K(AW, BF, true); LAUNCH_BWD_K(AW, BF, false); } while (0)
  if (g.rp % L == 0) { if (rows_bf16) LAUNCH_BWD(true, true); LAUNCH_BWD(true, false); }
  if (rows_bf16) LAUNCH_BWD(false, true);
  LAUNCH_BWD(false, false);
#undef LAUNCH_BWD_K
#undef LAUNCH_BWD
}


}  // namespace alexcap
