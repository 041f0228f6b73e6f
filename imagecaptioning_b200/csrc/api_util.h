// api_util.h -- error reporting and workspace carving shared by the C-ABI translation units.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/alexcap.h"
#include "common.cuh"
#include "kernels.h"

namespace alexcap {

char* err_buf();   // thread-local, 512 bytes (api.cu)

inline int fail(int code, const char* msg) {
  snprintf(err_buf(), 512, "%s", msg);
  return code;
}

#define API_CK(expr)                                                                                          \
  do {                                                                                                        \
    cudaError_t e__ = (expr);                                                                                 \
    if (e__ != cudaSuccess) {                                                                                 \
      snprintf(alexcap::err_buf(), 512, "%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
      return ALEXCAP_E_CUDA;                                                                                  \
    }                                                                                                         \
  } while (0)

// bump allocator over the caller's workspace; with base == nullptr it only measures.
struct Bump {
  char* base; size_t off;
  explicit Bump(void* b) : base(static_cast<char*>(b)), off(0) {}
  template <typename T>
  T* take(size_t n) {
    off = (off + 255) & ~size_t(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};

inline int check_dims_common(const alexcap_dims* d) {
  if (!d) return fail(ALEXCAP_E_BADARG, "dims is NULL");
  if (d->B <= 0 || d->L <= 0 || d->D <= 0 || d->H <= 0 || d->E <= 0 || d->Vp <= 3 || d->S <= 1)
    return fail(ALEXCAP_E_BADARG, "non-positive dimension in alexcap_dims");
  if ((d->D & 3) || (d->H & 3) || (d->E & 3)) return fail(ALEXCAP_E_BADARG, "D, H and E must be multiples of 4");
  if (d->H > 1024) return fail(ALEXCAP_E_UNSUPPORTED, "H > 1024 not supported by the attention kernels");
  if (d->dtype != ALEXCAP_F32 && d->dtype != ALEXCAP_BF16) return fail(ALEXCAP_E_UNSUPPORTED, "dtype must be ALEXCAP_F32 or ALEXCAP_BF16");
  if (d->dropout_p < 0.f || d->dropout_p >= 1.f) return fail(ALEXCAP_E_BADARG, "dropout_p must be in [0,1)");
  return 0;
}

// dims.dtype == ALEXCAP_BF16: every GEMM of the call runs on the tcgen05 bf16 engine (operands rounded to bf16 in shared
// memory, fp32 accumulation in TMEM); parameters, activations, gradients and the attention/softmax/CE arithmetic stay
// fp32 ("bf16 with fp32 master/accumulate" of BASELINE.json configs 3 and 5).  Scoped: the process-wide engine is restored.
struct DtypeScope {
  int saved;
  explicit DtypeScope(const alexcap_dims* d) : saved(gemm_engine()) {
    if (d && d->dtype == ALEXCAP_BF16) gemm_engine() = 2;
  }
  ~DtypeScope() { gemm_engine() = saved; }
};

// Big GEMMs of the call may keep bf16 hi/lo copies of their operands in this slice of the caller's workspace (gemm_ps.cu);
// a copy is reused by every later GEMM of the call that reads the same matrix (feat, W_o, dlogits, dgates, ...).
struct ArenaScope {
  SplitArena arena;
  SplitArena* saved;
  ArenaScope(void* base, size_t bytes) : saved(split_arena()) {
    arena.base = static_cast<uint8_t*>(base); arena.bytes = bytes;
    split_arena() = base ? &arena : nullptr;
  }
  // A training step run as several phase calls keeps its copies from one call to the next: `keep` is the bookkeeping of
  // the previous call on this thread; it is reused when it describes the same arena memory and `fresh` is false.
  ArenaScope(void* base, size_t bytes, SplitArena* keep, bool fresh) : saved(split_arena()) {
    if (fresh || keep->base != static_cast<uint8_t*>(base) || keep->bytes != bytes) {
      keep->base = static_cast<uint8_t*>(base); keep->bytes = bytes; keep->used = 0; keep->n = 0;
    }
    split_arena() = base ? keep : nullptr;
  }
  ~ArenaScope() { split_arena() = saved; }
};
// bytes that hold every operand copy one attention / LSTM training step can ask for
inline size_t split_arena_bytes(int64_t B, int64_t L, int64_t D, int64_t H, int64_t E, int64_t Vp, int64_t S, int64_t chunk_rows) {
  const int64_t W5 = 5 * H + D, DH = D + H;
  const int64_t act = B * L * (D + H) + 2 * S * B * (W5 + DH + E + 5 * H) + chunk_rows * (Vp + 8) + S * B * H;
  const int64_t wts = Vp * H + 2 * (4 * H * (E + D) + 4 * H * DH + (H + D) * H + W5 * H + H * D + 4 * H * H);
  return (size_t)(act + wts) * 4 + (size_t)SplitArena::MAX_ENTRIES * 2048 + (1 << 20);
}

// ---- vocabulary projection + CustomCrossEntropyLoss + vocabulary backward, chunked over rows (north_star item 4).
// The logits exist only as a workspace chunk of `vocab_chunk_rows` rows: the projection GEMM produces it, the one-pass CE
// kernel turns it into its gradient in place, the two backward GEMMs consume it, and the next chunk overwrites it -- the
// caller never sees logits and the workspace is bounded (ALEXCAP_VOCAB_CHUNK_MB, default 4096 MB).  Measured on B200 at
// cfg2: chunks small enough to stay L2-resident (24 MB: 5 chunks) make the step 9 % SLOWER -- the K = chunk-rows dW_o
// GEMMs and their read-modify-write of dW_o cost more than the ~0.4 GB of HBM traffic they save -- so the default keeps
// one chunk whenever it fits the budget.
inline int vocab_chunk_rows(int SB, int64_t ldV) {
  static const int64_t mb = [] { const char* v = getenv("ALEXCAP_VOCAB_CHUNK_MB"); return (v && *v) ? atoll(v) : 4096ll; }();
  int64_t r = (mb << 20) / (ldV * 4);
  r = r / 128 * 128;
  if (r < 128) r = 128;
  return r > SB ? SB : (int)r;
}
inline int vocab_ce_chunked(cudaStream_t st, int SB, int Vp, int H, const float* hout, int64_t hld, const float* Wo,
                            const float* bo, const int32_t* tgt_tm, int64_t ldV, float* logits_chunk, float eps,
                            int ignore_index, const float* denom, float* row_loss, float* dWo, float* dbo, float* dHout,
                            float* cs_scratch) {
  const int RC = vocab_chunk_rows(SB, ldV);
  for (int r0 = 0; r0 < SB; r0 += RC) {
    const int rows = SB - r0 < RC ? SB - r0 : RC;
    const float* h = hout + (int64_t)r0 * hld;
    // logits = h W_o^T + b_o  (AttentionLanguageModule.py:79 / LanguageModule.py:106)
    API_CK(gemm_f32(st, false, true, rows, Vp, H, 1.f, h, hld, Wo, H, logits_chunk, ldV, bo, nullptr, false, ACT_NONE));
    SplitArena* arena = split_arena();
    const size_t mark = arena ? arena->used : 0;      // the chunk buffer is rewritten per chunk: drop its bf16 copy afterwards
    // When the two backward GEMMs will read bf16 copies anyway, the CE kernel writes them itself (no fp32 gradient, no
    // split pass: -174 MB of HBM traffic per cfg2 step) and db_o is summed from the copies.
    uint16_t *dl_hi = nullptr, *dl_lo = nullptr;
    int64_t dl_ld = 0;
    const bool direct = arena && gemm_engine() >= 2 && gemm_ps_wanted(Vp, H, rows) && gemm_ps_wanted(rows, H, Vp) &&
                        split_arena_reserve(arena, logits_chunk, rows, Vp, ldV, gemm_engine() == 3, &dl_hi, &dl_lo, &dl_ld);
    API_CK(ce_rows(st, logits_chunk, tgt_tm + r0, rows, Vp, ldV, eps, ignore_index, denom, row_loss + r0,
                   direct ? nullptr : logits_chunk, dl_hi, dl_lo, dl_ld));
    // dW_o += dlogits^T h ; db_o += colsum(dlogits) ; dH = dlogits W_o
    API_CK(gemm_f32(st, true, false, Vp, H, rows, 1.f, logits_chunk, ldV, h, hld, dWo, H, nullptr, nullptr, r0 > 0, ACT_NONE));
    if (direct) API_CK(col_sum_bf16(st, dl_hi, dl_lo, rows, Vp, dl_ld, dbo, r0 > 0, cs_scratch));
    else API_CK(col_sum(st, logits_chunk, rows, Vp, ldV, dbo, r0 > 0, cs_scratch));
    API_CK(gemm_f32(st, false, false, rows, H, Vp, 1.f, logits_chunk, ldV, Wo, H, dHout + (int64_t)r0 * H, H, nullptr, nullptr,
                    false, ACT_NONE));
    if (arena && RC < SB) split_arena_release(arena, mark);
  }
  return 0;
}

// ---- the same block with the logits NEVER materialised (north_star item 4): the projection runs on the pre-split tcgen05
// engine with the cross-entropy in its epilogue (gemm_ps.cu: EPI 1 = per-row online-softmax statistics straight from TMEM,
// EPI 2 = the product recomputed and turned into d logits, written only as the bf16 operand copies the two backward GEMMs
// read).  Used whenever the three GEMMs qualify for the pre-split engine; the chunked path above remains for the exact-fp32
// engines (0 / 1) and for tiny shapes.
inline bool vocab_ce_fused_ok(int SB, int Vp, int H) {
  static const bool off = [] { const char* v = getenv("ALEXCAP_VOCAB_FUSED"); return v && v[0] == '0'; }();
  return !off && gemm_engine() >= 2 && gemm_ps_ce_wanted(SB, Vp, H) && gemm_ps_wanted(Vp, H, SB) && gemm_ps_wanted(SB, H, Vp);
}
inline int vocab_ce_fused(cudaStream_t st, int SB, int Vp, int H, const float* hout, int64_t hld, const float* Wo, const float* bo,
                          const int32_t* tgt_tm, int64_t ldV, float4* stats, float* row_lse, float eps, int ignore_index,
                          const float* denom, float* row_loss, float* dWo, float* dbo, float* dHout, float* cs_scratch,
                          bool defer_dwo = false) {
  SplitArena* arena = split_arena();
  if (!arena) return fail(ALEXCAP_E_BADARG, "vocab_ce_fused: no operand arena");
  const int nprod = gemm_engine() == 3 ? 3 : 1;
  // the gradient exists only as operand copies; their arena key is the statistics buffer (never read as an fp32 matrix)
  const float* key = reinterpret_cast<const float*>(stats);
  uint16_t *dl_hi = nullptr, *dl_lo = nullptr;
  int64_t dl_ld = 0;
  if (!split_arena_reserve(arena, key, SB, Vp, ldV, nprod == 3, &dl_hi, &dl_lo, &dl_ld))
    return fail(ALEXCAP_E_WORKSPACE, "vocab_ce_fused: operand arena too small for the d logits copies");
  API_CK(gemm_ps_ce(st, false, SB, Vp, H, hout, hld, Wo, H, bo, tgt_tm, eps, ignore_index, denom, stats, row_lse, row_loss,
                    nullptr, nullptr, 0, nprod, arena));
  API_CK(gemm_ps_ce(st, true, SB, Vp, H, hout, hld, Wo, H, bo, tgt_tm, eps, ignore_index, denom, stats, row_lse, row_loss,
                    dl_hi, dl_lo, dl_ld, nprod, arena));
  // dW_o = dlogits^T h ; db_o = colsum(dlogits) ; dH = dlogits W_o      (all three read the copies)
  // (defer_dwo: the reverse-time loop only needs dH; the caller runs vocab_ce_dwo() on a side stream underneath that loop)
  if (!defer_dwo) {
    API_CK(gemm_f32(st, true, false, Vp, H, SB, 1.f, key, ldV, hout, hld, dWo, H, nullptr, nullptr, false, ACT_NONE));
    API_CK(col_sum_bf16(st, dl_hi, dl_lo, SB, Vp, dl_ld, dbo, false, cs_scratch));
  }
  API_CK(gemm_f32(st, false, false, SB, H, Vp, 1.f, key, ldV, Wo, H, dHout, H, nullptr, nullptr, false, ACT_NONE));
  return 0;
}
// The two vocabulary gradients that nothing downstream in the step reads: dW_o = dlogits^T h, db_o = colsum(dlogits), from the
// operand copies vocab_ce_fused() left in the arena (same key).  `cs_scratch` must not be shared with work on another stream.
inline int vocab_ce_dwo(cudaStream_t st, int SB, int Vp, int H, const float* hout, int64_t hld, int64_t ldV, const float4* stats,
                        float* dWo, float* dbo, float* cs_scratch) {
  SplitArena* arena = split_arena();
  const float* key = reinterpret_cast<const float*>(stats);
  uint16_t *dl_hi = nullptr, *dl_lo = nullptr;
  int64_t dl_ld = 0;
  if (!split_arena_find(arena, key, &dl_hi, &dl_lo, &dl_ld))
    return fail(ALEXCAP_E_WORKSPACE, "vocab_ce_dwo: the d logits copies are not in the operand arena");
  // (the HBM-streaming column sum first: it then runs beside the loss finalisation, not beside the loop's attention kernels --
  //  measured 1.873 vs 1.902 ms for the other order)
  API_CK(col_sum_bf16(st, dl_hi, dl_lo, SB, Vp, dl_ld, dbo, false, cs_scratch));
  API_CK(gemm_f32(st, true, false, Vp, H, SB, 1.f, key, ldV, hout, hld, dWo, H, nullptr, nullptr, false, ACT_NONE));
  return 0;
}

// Per-step GEMMs of the time loop: B = weights (or their concatenations built at the start of the call, many kernels and at
// least one fully-ordered node earlier), so the GEMM kernel may prefetch its B tiles before griddepcontrol.wait.
struct WeightBScope {
  bool saved;
  WeightBScope() : saved(gemm_b_is_weight()) { gemm_b_is_weight() = true; }
  ~WeightBScope() { gemm_b_is_weight() = saved; }
};

// Decoding must reproduce the reference's token ids exactly (fp32): never run it on a reduced-precision engine.
struct ExactScope {
  int saved;
  ExactScope() : saved(gemm_engine()) { if (saved >= 2) gemm_engine() = 1; }
  ~ExactScope() { gemm_engine() = saved; }
};

}  // namespace alexcap
