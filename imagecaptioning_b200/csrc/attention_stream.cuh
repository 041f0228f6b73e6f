// attention_stream.cuh -- the fused additive soft-attention step (forward and backward) as device-function BODIES shared by
// the stand-alone streaming kernels (attention_stream.cu: one launch per time step) and the persistent time-loop kernels
// (time_loop.cu: every time step inside one launch).  See attention_stream.cu for the design notes.
#pragma once
#include "common.cuh"
#include "pipe.cuh"

#ifndef ALEXCAP_ATT_NG
#define ALEXCAP_ATT_NG 2
#endif

namespace alexcap {
namespace sattn {

// Consumers: SA_NG groups of SA_GWARPS warps.  Inside a pass, piece k belongs to group k % SA_NG, so SA_NG pieces are
// being computed at any time (one row per warp is a long dependent chain: tanh, 5-step shuffle reduction).
constexpr int SA_NG = ALEXCAP_ATT_NG;
constexpr int SA_GWARPS = 8;
constexpr int SA_GTHREADS = SA_GWARPS * 32;
constexpr int SA_CWARPS = SA_NG * SA_GWARPS;        // consumer warps
constexpr int SA_CTHREADS = SA_CWARPS * 32;
constexpr int SA_THREADS = SA_CTHREADS + 32;        // + one producer warp
constexpr int SA_MAX_NST = 16;
constexpr int SA_NC = 2;                            // float4 column chunks per consumer thread (fwd): D <= 4*256*SA_NC


// 4 consecutive row elements (index i4 counts groups of 4) as float4, from fp32 rows or from bf16 rows (ALEXCAP_BF16 storage of
// the streamed Ws / feat rows: half the bytes per attention launch)
template <bool BF>
__device__ __forceinline__ float4 ld_row4(const void* base, int i4) {
  if (BF) {
    const uint2 u = reinterpret_cast<const uint2*>(base)[i4];
    return make_float4(__uint_as_float(u.x << 16), __uint_as_float(u.x & 0xffff0000u), __uint_as_float(u.y << 16),
                       __uint_as_float(u.y & 0xffff0000u));
  }
  return reinterpret_cast<const float4*>(base)[i4];
}

// fp32 x4 -> bf16 hi (round to nearest) and lo = bf16(x - hi), packed pairs
__device__ __forceinline__ void split_bf16x4(const float4& v, uint2& hi, uint2& lo) {
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hi.x) : "f"(v.y), "f"(v.x));
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hi.y) : "f"(v.w), "f"(v.z));
  const float r0 = v.x - __uint_as_float(hi.x << 16), r1 = v.y - __uint_as_float(hi.x & 0xffff0000u);
  const float r2 = v.z - __uint_as_float(hi.y << 16), r3 = v.w - __uint_as_float(hi.y & 0xffff0000u);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo.x) : "f"(r1), "f"(r0));
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lo.y) : "f"(r3), "f"(r2));
}

// fp32 x4 -> bf16 (round to nearest), packed pairs: what ps::split_bf16_kernel would write as the `hi` copy
__device__ __forceinline__ uint2 to_bf16x4(const float4& v) {
  uint2 o;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o.x) : "f"(v.y), "f"(v.x));
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(o.y) : "f"(v.w), "f"(v.z));
  return o;
}

// The ring protocol, shared by producer and consumers.  A cursor (stage index, use parity) walks the pieces in
// order; it is advanced incrementally (no integer division per piece).
struct Cursor {
  int s; uint32_t ph;
  __device__ __forceinline__ void next(int NST) { if (++s == NST) { s = 0; ph ^= 1u; } }
};
struct Ring {
  float* base; uint64_t* full; uint64_t* empty; int NST, stage_floats;
  uint64_t policy = 0; bool hinted = false;      // L2 eviction policy of the streamed rows (producer thread only)
  __device__ __forceinline__ float* stage(const Cursor& c) const { return base + (int64_t)c.s * stage_floats; }
  __device__ __forceinline__ void wait_full(const Cursor& c) const { pipe::mbar_wait(&full[c.s], c.ph); }
  __device__ __forceinline__ void release(const Cursor& c) const { pipe::mbar_arrive(&empty[c.s]); }
  // producer side: begin(cursor, total bytes of the piece), then one or more copy(cursor, ...)
  __device__ __forceinline__ void begin(const Cursor& c, uint32_t bytes) const {
    pipe::mbar_wait(&empty[c.s], c.ph ^ 1u);       // passes immediately on the first use of a stage
    pipe::mbar_expect_tx(&full[c.s], bytes);
  }
  __device__ __forceinline__ void copy(const Cursor& c, int dst_float_off, const float* src, uint32_t bytes) const {
    pipe::bulk_g2s(base + (int64_t)c.s * stage_floats + dst_float_off, src, bytes, &full[c.s]);
  }
  __device__ __forceinline__ void copy_rows(const Cursor& c, const float* src, uint32_t bytes) const {
    if (hinted) pipe::bulk_g2s_hint(base + (int64_t)c.s * stage_floats, src, bytes, &full[c.s], policy);
    else pipe::bulk_g2s(base + (int64_t)c.s * stage_floats, src, bytes, &full[c.s]);
  }
};

// Small per-image vectors that open every segment (the "header piece"): up to 5 rows copied back to back.  Byte-addressed:
// a member may be fp32 or bf16 (the reference row of Ws in bf16-storage mode).
struct Header { const char* p[5]; int64_t ld[5]; int n[5]; int count; unsigned dyn = 0x1fu; };   // p: base, ld: bytes per image, n: bytes to copy
// dyn: bit i set = member i is written by the kernel that precedes this launch (the others were final long before it)
__device__ __forceinline__ void hdr_set(Header& hd, int i, const void* p, int64_t ld_elems, int n_elems, int esz) {
  hd.p[i] = static_cast<const char*>(p); hd.ld[i] = ld_elems * esz; hd.n[i] = n_elems * esz;
}

// Producer: for every image segment of this CTA's row range stream [header] [X0 rows] [X1 rows] (R0- / R1-row pieces).
// esz: bytes per element of the X0 / X1 rows (4, or 2 in bf16-storage mode).
// `k` is the producer's ring cursor: a persistent caller (time_loop.cu) keeps it across calls.
// `defer` (stand-alone kernels launched with programmatic dependent launch; the caller guarantees that everything but the
// header members flagged in `hd.dyn` was complete before the PRECEDING kernel started):
//   1: the first header's stage is armed and its static members are copied at once; griddepcontrol.wait follows and only then
//      the dynamic members (backward: d gctx) and the row pieces -- the latency of the static copies hides under the
//      predecessor's tail;
//   2: as 1, and the row pieces that fill the rest of the ring are also fetched before the wait (experiment, see
//      attention_stream.cu: prefetch_before_wait).
__device__ __forceinline__ void produce(const Ring& ring, Cursor& k, int64_t row0, int64_t row1, int L, int R0, int R1, int rpi,
                                        const Header& hd, const void* X0, int n0, const void* X1, int n1, int esz,
                                        int defer = 0) {
  bool pending = false;                              // the first header's dynamic copies have not been issued yet
  Cursor kh{0, 0u};
  int hb = 0, since = 0;
  // which: 0 = the static members, 1 = the dynamic members, 2 = all
  auto issue_header = [&](const Cursor& kc, int b, int which) {
    int off = 0;                                     // bytes (every member is a multiple of 16 bytes)
    for (int i = 0; i < hd.count; ++i) {
      const int dyn = (int)((hd.dyn >> i) & 1u);
      if (which == 2 || which == dyn)
        pipe::bulk_g2s(reinterpret_cast<char*>(ring.stage(kc)) + off, hd.p[i] + (int64_t)b * hd.ld[i], (uint32_t)hd.n[i], &ring.full[kc.s]);
      off += hd.n[i];
    }
  };
  for (int64_t lo = row0; lo < row1;) {
    const int b = (int)(lo / L);
    int64_t hi = (int64_t)(b + 1) * L;
    if (hi > row1) hi = row1;
    const int l0 = (int)(lo - (int64_t)b * L), nl = (int)(hi - lo);
    const int64_t src_row = (int64_t)(b / rpi) * L + l0;
    {
      uint32_t bytes = 0;
      for (int i = 0; i < hd.count; ++i) bytes += (uint32_t)hd.n[i];
      ring.begin(k, bytes);
      if (defer && lo == row0) {
        issue_header(k, b, 0);
        if (defer == 1) { pdl_wait(); issue_header(k, b, 1); }
        else { pending = true; kh = k; hb = b; since = 0; }
      } else {
        issue_header(k, b, 2);
      }
      k.next(ring.NST);
    }
    for (int pass = 0; pass < 2; ++pass) {
      const char* X = static_cast<const char*>(pass ? X1 : X0);
      const int n = pass ? n1 : n0;
      const int R = pass ? R1 : R0;
      for (int r0 = 0; r0 < nl; r0 += R, k.next(ring.NST)) {
        if (pending && ++since >= ring.NST) {        // the ring is full: the next stage is the header's own
          pdl_wait();
          issue_header(kh, hb, 1);
          pending = false;
        }
        const int rows = min(R, nl - r0);
        const uint32_t bytes = (uint32_t)rows * (uint32_t)n * (uint32_t)esz;
        ring.begin(k, bytes);
        ring.copy_rows(k, reinterpret_cast<const float*>(X + (src_row + r0) * n * esz), bytes);
      }
    }
    if (pending) { pdl_wait(); issue_header(kh, hb, 1); pending = false; }
    lo = hi;
  }
}

// mbarriers of a ring whose storage starts at `sm` (layout of the bodies below): one thread, before any use
__device__ __forceinline__ void attn_ring_init(float* sm, int NST, int stage_floats) {
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + (int64_t)NST * stage_floats);
  uint64_t* empty = full + SA_MAX_NST;
  for (int s = 0; s < NST; ++s) { pipe::mbar_init(&full[s], 1); pipe::mbar_init(&empty[s], SA_GWARPS); }
  pipe::mbar_fence_init();
}

// debug (ALEXCAP_ATT_PROF_PTR=<device address of >= 16 u64>): %globaltimer stamps of CTA 0 at the phase boundaries
__device__ __forceinline__ unsigned long long sa_gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define SA_STAMP(slot) do { if (a.prof && cta == 0 && threadIdx.x == 0) a.prof[slot] = sa_gtime(); } while (0)

// r = 1 / (1 + e^{2(w + q)}) (tanh = 1 - 2 r).  EW = false: w is the raw pre-activation, qv = C q (ex2 + rcp per element).
// EW = true: w is EXP2C(Ws) as stored by the hoisted GEMM, qv = EXP2C(q): one FMA + ONE MUFU (common.cuh, exp2c).
template <bool EW>
__device__ __forceinline__ float r_of(float w, float qv) {
  return EW ? rcp_ftz(fmaf(w, qv, 1.f)) : tanh_r(fmaf(w, TANH_C, qv));
}
template <bool EW>
__device__ __forceinline__ float4 q_stage(const float4 t) {
  return EW ? make_float4(exp2c(t.x), exp2c(t.y), exp2c(t.z), exp2c(t.w))
            : make_float4(TANH_C * t.x, TANH_C * t.y, TANH_C * t.z, TANH_C * t.w);
}

// ------------------------------------------------------------------------------------------ forward step
struct SAttnFwdP {
  const float* Ws; const float* feat; const float* q; int64_t q_ld; const float* bpre; int64_t bpre_ld;
  const float* wv; float* alpha; int64_t alpha_ld; float* ctx; float* beta; float* gctx; int64_t gctx_ld;
  float* scratch; unsigned* counters;
  int B, L, D, H, rpi, rp, RH, RD, NST, stage_floats, maxseg, seg_pad, slot_floats, dry;
  float l2_frac;
  // optional bf16 hi / lo = bf16(x - hi) copies of gctx (row stride gb_ld elements), written beside the fp32 value by a CTA
  // that finalises a whole image: the A operand of the persistent time loop's gates GEMM (time_loop.cu)
  uint16_t* gctx_hi = nullptr; uint16_t* gctx_lo = nullptr; int64_t gb_ld = 0;
  int bf16 = 0;       // Ws / feat point at bf16 rows (same shapes): selects the BF instantiation
  int ws_exp = 0;     // Ws holds EXP2C(W(features)) instead of the pre-activation: selects the EW instantiation
  int pdl = 0;        // 0: no griddepcontrol.wait in the body (persistent caller); 1: every thread waits at body entry;
                      // 2 / 3: the consumers wait at body entry, the producer inside produce() (`defer` = 2 / 1)
  unsigned long long* prof = nullptr;
};

// Scratch slot of (CTA c, its j-th segment): [0]=local max, [1]=local sum, [4 .. 4+D) unnormalised partial context.
// NJ: number of float4 per lane along H (H <= 128*NJ).
// The kernel body as a device function: `sm` = this CTA's shared memory for the step (ring first, fixed part after it),
// `cta` = index of the CTA's row range, (k, kg) = ring cursor and piece-owner counter of the calling thread -- a persistent
// caller keeps them across calls; the ring's mbarriers must have been initialised by the caller (attn_ring_init).
// Only the SA_THREADS threads of the attention roles may enter (warps 0..SA_CWARPS-1 consumers, warp SA_CWARPS producer).
// NC: float4 column chunks per consumer thread in the context pass (D <= 4 * SA_GTHREADS * NC).
template <int NJ, bool FULL, int NC, bool BF = false, bool EW = false>
__device__ __forceinline__ void attn_fwd_body(const SAttnFwdP& a, float* sm, const int cta, Cursor& k, int& kg) {
  const int L = a.L, D = a.D, H4 = a.H >> 2, D4 = D >> 2, RH = a.RH, RD = a.RD;
  Ring ring;
  ring.base = sm; ring.NST = a.NST; ring.stage_floats = a.stage_floats;
  ring.full = reinterpret_cast<uint64_t*>(sm + (int64_t)a.NST * a.stage_floats);
  ring.empty = ring.full + SA_MAX_NST;
  float4* red = reinterpret_cast<float4*>(ring.empty + SA_MAX_NST);             // [NC][SA_CTHREADS]
  float* es = reinterpret_cast<float*>(red + NC * SA_CTHREADS);              // [seg_pad] scores, then exp(e - m)
  float* scratch = es + a.seg_pad;                                              // [40]
  float* bp_s = scratch + 40;                                                   // [D] bpre[b,:] of the current segment
  int* segb = reinterpret_cast<int*>(bp_s + D);                                 // [maxseg] image of segment j (-1: done)
  int* segflag = segb + a.maxseg;                                               // [maxseg] this CTA finalises it

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t total = (int64_t)a.B * L;
  const int64_t row0 = (int64_t)cta * a.rp;
  const int64_t row1 = (row0 + a.rp < total) ? row0 + a.rp : total;

  if (warp == SA_CWARPS) {
    if (lane == 0) {
      Header hd;
      hdr_set(hd, 0, a.q, a.q_ld, a.H, 4);
      hdr_set(hd, 1, a.bpre, a.bpre_ld, D, 4);
      hd.count = 2;
      if (a.l2_frac > 0.f) { ring.policy = pipe::l2_policy_evict_last(a.l2_frac); ring.hinted = true; }
      if (a.pdl == 1) pdl_wait();
      produce(ring, k, row0, row1, L, RH, RD, a.rpi, hd, a.Ws, a.H, a.feat, D, BF ? 2 : 4, a.pdl == 2 ? 2 : (a.pdl == 3 ? 1 : 0));
    }
    return;
  }
  if (a.pdl) pdl_wait();

  // ================================================================ consumers
  SA_STAMP(1);
  const int grp = warp / SA_GWARPS, gw = warp % SA_GWARPS, gtid = tid % SA_GTHREADS;
  // e[l] = sum_h v_h tanh(z_h) = sum_h v_h - 2 sum_h v_h r_h with r = 1/(2^(C z)+1): the constant drops out of the
  // softmax, so the score kept is  sum_h (-2 v_h) r_h  (5 instructions and 2 MUFU per element).
  float4 v4[NJ];
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    const int f = lane + 32 * j;
    const float4 v = (FULL || f < H4) ? __ldg(reinterpret_cast<const float4*>(a.wv) + f) : f4_zero();
    v4[j] = make_float4(-2.f * v.x, -2.f * v.y, -2.f * v.z, -2.f * v.w);
  }
  const bool wide = D4 >= SA_GTHREADS;
  const int G = wide ? 1 : SA_GTHREADS / D4;
  const int c = wide ? gtid : gtid % D4;
  const int g = wide ? 0 : gtid / D4;

  int j_seg = 0;                  // kg: piece number modulo SA_NG (which consumer group owns the piece)
  for (int64_t lo = row0; lo < row1; ++j_seg) {
    const int b = (int)(lo / L);
    int64_t hi = (int64_t)(b + 1) * L;
    if (hi > row1) hi = row1;
    const int l0 = (int)(lo - (int64_t)b * L), nl = (int)(hi - lo);
    lo = hi;
    float* slot = a.scratch + ((int64_t)cta * a.maxseg + j_seg) * a.slot_floats;
    const bool whole = (l0 == 0 && nl == L);      // the image lives in this CTA alone: finalise locally
    if (tid == 0) segb[j_seg] = whole ? -1 : b;

    // ---- header piece: q[b,:] | bpre[b,:] (streamed through the ring like everything else: its latency is hidden)
    float4 q4[NJ];
    ring.wait_full(k);
    {
      const float4* qs = reinterpret_cast<const float4*>(ring.stage(k));
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int f = lane + 32 * j;
        const float4 t = (FULL || f < H4) ? qs[f] : f4_zero();
        q4[j] = q_stage<EW>(t);
      }
      for (int cc = tid; cc < D4; cc += SA_CTHREADS) reinterpret_cast<float4*>(bp_s)[cc] = qs[H4 + cc];
    }
    pipe::named_sync(1, SA_CTHREADS);
    if (kg == grp && lane == 0) ring.release(k);
    k.next(ring.NST); kg = (kg + 1 == SA_NG) ? 0 : kg + 1;

    SA_STAMP(2);
    // ---- pass 1: e[l] = w_v . tanh(Ws[b,l,:] + q[b,:])     two rows per warp in flight   (AttentionLanguageModule.py:18-19)
    for (int r0 = 0; r0 < nl; r0 += RH, k.next(ring.NST), kg = (kg + 1 == SA_NG) ? 0 : kg + 1) {
      if (kg != grp) continue;
      ring.wait_full(k);
      if (a.dry) { __syncwarp(); if (lane == 0) ring.release(k); continue; }   // debug: pure streaming rate
      const void* st = ring.stage(k);
      const int rows = min(RH, nl - r0);
      for (int rr = gw; rr < rows; rr += 2 * SA_GWARPS) {
        const int rr2 = rr + SA_GWARPS;
        const bool has2 = rr2 < rows;
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int f = lane + 32 * j;
          if (FULL || f < H4) {
            const float4 w0 = ld_row4<BF>(st, rr * H4 + f);
            s0 = fmaf(v4[j].x, r_of<EW>(w0.x, q4[j].x), s0); s0 = fmaf(v4[j].y, r_of<EW>(w0.y, q4[j].y), s0);
            s0 = fmaf(v4[j].z, r_of<EW>(w0.z, q4[j].z), s0); s0 = fmaf(v4[j].w, r_of<EW>(w0.w, q4[j].w), s0);
            if (has2) {
              const float4 w1 = ld_row4<BF>(st, rr2 * H4 + f);
              s1 = fmaf(v4[j].x, r_of<EW>(w1.x, q4[j].x), s1); s1 = fmaf(v4[j].y, r_of<EW>(w1.y, q4[j].y), s1);
              s1 = fmaf(v4[j].z, r_of<EW>(w1.z, q4[j].z), s1); s1 = fmaf(v4[j].w, r_of<EW>(w1.w, q4[j].w), s1);
            }
          }
        }
        s0 = warp_sum(s0); s1 = warp_sum(s1);
        if (lane == 0) {                          // + b_v omitted: softmax is shift invariant (alpha identical)
          es[r0 + rr] = s0;
          if (has2) es[r0 + rr2] = s1;
        }
      }
      __syncwarp();
      if (lane == 0) ring.release(k);
    }
    pipe::named_sync(1, SA_CTHREADS);

    SA_STAMP(3);
    // ---- local softmax statistics                                              (:20, split over the segments)
    float mx = -INFINITY;
    for (int i = tid; i < nl; i += SA_CTHREADS) mx = fmaxf(mx, es[i]);
    mx = pipe::group_max(mx, scratch, SA_CWARPS, 1);
    float sum = 0.f;
    for (int i = tid; i < nl; i += SA_CTHREADS) {
      const float e = __expf(es[i] - mx);
      es[i] = e;
      sum += e;
    }
    sum = pipe::group_sum(sum, scratch, SA_CWARPS, 1);        // (its barriers also publish es[])
    const float inv_local = 1.f / sum;
    // split image: unnormalised alpha now, the finaliser rescales it
    for (int i = tid; i < nl; i += SA_CTHREADS) a.alpha[(int64_t)b * a.alpha_ld + l0 + i] = whole ? es[i] * inv_local : es[i];
    if (!whole && tid == 0) { slot[0] = mx; slot[1] = sum; }

    SA_STAMP(4);
    // ---- pass 2: unnormalised partial context  sum_{own l} p[l] * feat[b,l,:]                           (:21)
    float4 acc[NC];
#pragma unroll
    for (int ci = 0; ci < NC; ++ci) acc[ci] = f4_zero();
    for (int r0 = 0; r0 < nl; r0 += RD, k.next(ring.NST), kg = (kg + 1 == SA_NG) ? 0 : kg + 1) {
      if (kg != grp) continue;
      ring.wait_full(k);
      if (a.dry) { __syncwarp(); if (lane == 0) ring.release(k); continue; }   // debug: pure streaming rate
      const void* st = ring.stage(k);
      const int rows = min(RD, nl - r0);
      if (g < G) {
#pragma unroll 4
        for (int rr = g; rr < rows; rr += G) {
          const float p = es[r0 + rr];
#pragma unroll
          for (int ci = 0; ci < NC; ++ci) {
            const int c2 = c + ci * SA_GTHREADS;
            if (c2 < D4) f4_fma(acc[ci], p, ld_row4<BF>(st, rr * D4 + c2));
          }
        }
      }
      __syncwarp();
      if (lane == 0) ring.release(k);
    }
    SA_STAMP(5);
    // reduce over the consumer groups and the row sub-groups inside each
#pragma unroll
    for (int ci = 0; ci < NC; ++ci) red[ci * SA_CTHREADS + tid] = acc[ci];
    pipe::named_sync(1, SA_CTHREADS);
    if (grp == 0 && g == 0) {
      float4* pctx = reinterpret_cast<float4*>(slot + 4);
#pragma unroll
      for (int ci = 0; ci < NC; ++ci) {
        const int c2 = c + ci * SA_GTHREADS;
        if (c2 < D4) {
          float4 t = f4_zero();
          for (int gi = 0; gi < SA_NG; ++gi)
            for (int gg = 0; gg < G; ++gg) f4_add(t, red[ci * SA_CTHREADS + gi * SA_GTHREADS + gg * D4 + c]);
          if (whole) {                                                           // ctx / beta / gctx  (:21, :74-75)
            t.x *= inv_local; t.y *= inv_local; t.z *= inv_local; t.w *= inv_local;
            const float4 bp = reinterpret_cast<const float4*>(bp_s)[c2];
            float4 be;
            be.x = sigmoid_f(bp.x); be.y = sigmoid_f(bp.y); be.z = sigmoid_f(bp.z); be.w = sigmoid_f(bp.w);
            reinterpret_cast<float4*>(a.ctx + (int64_t)b * D)[c2] = t;
            reinterpret_cast<float4*>(a.beta + (int64_t)b * D)[c2] = be;
            const float4 gc = make_float4(be.x * t.x, be.y * t.y, be.z * t.z, be.w * t.w);
            reinterpret_cast<float4*>(a.gctx + (int64_t)b * a.gctx_ld)[c2] = gc;
            if (a.gctx_hi) {
              uint2 h2, l2;
              split_bf16x4(gc, h2, l2);
              reinterpret_cast<uint2*>(a.gctx_hi + (int64_t)b * a.gb_ld)[c2] = h2;
              if (a.gctx_lo) reinterpret_cast<uint2*>(a.gctx_lo + (int64_t)b * a.gb_ld)[c2] = l2;
            }
          } else {
            pctx[c2] = t;
          }
        }
      }
    }
    pipe::named_sync(1, SA_CTHREADS);             // es / red are reused by the next segment
    SA_STAMP(6);
  }
  const int nsegs = j_seg;

  // ---- publish the split segments with ONE fence; an image is finalised by the CTA whose segment arrives last
  if ((int64_t)a.rp % L == 0) return;               // every image lives in one CTA: nothing to combine
  __threadfence();
  pipe::named_sync(1, SA_CTHREADS);
  if (tid < nsegs) {
    const int b = segb[tid];
    int fl = 0;
    if (b >= 0) {
      const int c_first = (int)(((int64_t)b * L) / a.rp), c_last = (int)(((int64_t)(b + 1) * L - 1) / a.rp);
      fl = (atomicAdd(a.counters + b, 1u) == (unsigned)(c_last - c_first)) ? 1 : 0;
    }
    segflag[tid] = fl;
  }
  pipe::named_sync(1, SA_CTHREADS);
  for (int j = 0; j < nsegs; ++j) {
    if (!segflag[j]) continue;
    const int b = segb[j];
    const int c_first = (int)(((int64_t)b * L) / a.rp), c_last = (int)(((int64_t)(b + 1) * L - 1) / a.rp);
    const int nseg = c_last - c_first + 1;
    __threadfence();
    // M = max m_s, S = sum s_s e^{m_s-M};  alpha = p e^{m_s-M}/S;  ctx = sum_s (e^{m_s-M}/S) pctx_s
    float* fin = reinterpret_cast<float*>(red);                 // [2*nseg] (m_s, s_s), one global round trip
    for (int i = tid; i < nseg; i += SA_CTHREADS) {
      const int cc = c_first + i;
      const int jj = b - (int)(((int64_t)cc * a.rp) / L);
      const float* sl = a.scratch + ((int64_t)cc * a.maxseg + jj) * a.slot_floats;
      fin[2 * i] = __ldcg(sl); fin[2 * i + 1] = __ldcg(sl + 1);
    }
    pipe::named_sync(1, SA_CTHREADS);
    float M = -INFINITY, S = 0.f;
    for (int i = 0; i < nseg; ++i) M = fmaxf(M, fin[2 * i]);
    for (int i = 0; i < nseg; ++i) S = fmaf(fin[2 * i + 1], __expf(fin[2 * i] - M), S);
    const float inv = 1.f / S;
    float4 o[NC];
#pragma unroll
    for (int ci = 0; ci < NC; ++ci) o[ci] = f4_zero();
    for (int i = 0; i < nseg; ++i) {
      const int cc = c_first + i;
      const int jj = b - (int)(((int64_t)cc * a.rp) / L);
      const float* sl = a.scratch + ((int64_t)cc * a.maxseg + jj) * a.slot_floats;
      const float scale = __expf(fin[2 * i] - M) * inv;
      // rows of image b owned by CTA cc
      const int64_t s_lo = max((int64_t)cc * a.rp, (int64_t)b * L), s_hi = min((int64_t)(cc + 1) * a.rp, (int64_t)(b + 1) * L);
      float* al = a.alpha + (int64_t)b * a.alpha_ld + (s_lo - (int64_t)b * L);
      for (int t = tid; t < (int)(s_hi - s_lo); t += SA_CTHREADS) al[t] = __ldcg(al + t) * scale;
#pragma unroll
      for (int ci = 0; ci < NC; ++ci) {
        const int c2 = tid + ci * SA_CTHREADS;
        if (c2 < D4) f4_fma(o[ci], scale, __ldcg(reinterpret_cast<const float4*>(sl + 4) + c2));
      }
    }
#pragma unroll
    for (int ci = 0; ci < NC; ++ci) {                                         // ctx / beta / gctx  (:21, :74-75)
      const int c2 = tid + ci * SA_CTHREADS;
      if (c2 < D4) {
        const float4 bp = __ldg(reinterpret_cast<const float4*>(a.bpre + (int64_t)b * a.bpre_ld) + c2);
        float4 be;
        be.x = sigmoid_f(bp.x); be.y = sigmoid_f(bp.y); be.z = sigmoid_f(bp.z); be.w = sigmoid_f(bp.w);
        reinterpret_cast<float4*>(a.ctx + (int64_t)b * D)[c2] = o[ci];
        reinterpret_cast<float4*>(a.beta + (int64_t)b * D)[c2] = be;
        const float4 gc = make_float4(be.x * o[ci].x, be.y * o[ci].y, be.z * o[ci].z, be.w * o[ci].w);
        reinterpret_cast<float4*>(a.gctx + (int64_t)b * a.gctx_ld)[c2] = gc;
        if (a.gctx_hi) {
          uint2 h2, l2;
          split_bf16x4(gc, h2, l2);
          reinterpret_cast<uint2*>(a.gctx_hi + (int64_t)b * a.gb_ld)[c2] = h2;
          if (a.gctx_lo) reinterpret_cast<uint2*>(a.gctx_lo + (int64_t)b * a.gb_ld)[c2] = l2;
        }
      }
    }
    if (tid == 0) a.counters[b] = 0;              // ready for the next launch
    pipe::named_sync(1, SA_CTHREADS);             // fin (red) is reused by the next image
  }
}


// ------------------------------------------------------------------------------------------ backward step
// With dalpha[l] = dctx . feat[l] (+ upstream terms), dot = sum_L alpha*dalpha over the WHOLE image and
// de[l] = alpha[l] (dalpha[l] - dot), the score backward
//     dq   = w_v * sum_l de[l] (1 - t_l^2),      dw_v += sum_l de[l] t_l,      t_l = tanh(Ws[l] + q)
// is linear in dot.  A segment s therefore works against its LOCAL reference ref_s = sum_s alpha dalpha / sum_s alpha
// (known after its feat pass; equal to dot when the segment is the whole image, and close to it otherwise, which keeps the
// correction below free of cancellation): X1 = sum alpha (dalpha - ref_s)(1-t^2), X2 = sum alpha (1-t^2),
// Y1 = sum alpha (dalpha - ref_s) t, Y2 = sum alpha t.  The finaliser knows dot and forms, with d_s = dot - ref_s,
// dq = w_v sum_s (X1 - d_s X2), dw_v += sum_s (Y1 - d_s Y2), de = alpha (dalpha - dot).
struct SAttnBwdP {
  const float* Ws; const float* feat; const float* q; int64_t q_ld; const float* wv;
  const float* alpha; int64_t alpha_ld; const float* ctx; const float* beta; const float* dgctx; int64_t dgctx_ld;
  const float* ext; int64_t ext_ld; const float* reg_r;
  float* dq; int64_t dq_ld; float* dbpre; int64_t dbpre_ld; float* de; int64_t de_ld; float* dwv_part; float* dfeat;
  float* scratch; unsigned* counters;
  int B, L, D, H, rp, RH, RD, NST, stage_floats, maxseg, seg_pad, rp_pad, slot_floats;
  float l2_frac;
  int bf16 = 0;       // Ws / feat point at bf16 rows (same shapes): selects the BF instantiation
  uint16_t* dq_b16 = nullptr; uint16_t* dbpre_b16 = nullptr; int64_t db_ld = 0;   // optional bf16 mirrors of dq / dbpre
  int ws_exp = 0;     // Ws holds EXP2C(W(features)) (see r_of): selects the EW instantiation
  int pdl = 0;        // as SAttnFwdP::pdl (2: feat pieces are prefetched before the producer's griddepcontrol.wait)
  unsigned long long* prof = nullptr;
};

// Scratch slot: [0] = partial dot, [1] = ref_s, [4 + i*H .. ) i = 0..3 : X1, X2, Y1, Y2.
// ALLWHOLE: every image lives in one CTA (rp % L == 0): the split-path accumulators x2 / y2 are never needed.
template <bool ALLWHOLE, bool BF = false, bool EW = false>
__device__ __forceinline__ void attn_bwd_body(const SAttnBwdP& a, float* sm, const int cta, Cursor& k, int& kg) {
  const int L = a.L, D = a.D, H = a.H, H4 = H >> 2, D4 = D >> 2, RH = a.RH, RD = a.RD;
  Ring ring;
  ring.base = sm; ring.NST = a.NST; ring.stage_floats = a.stage_floats;
  ring.full = reinterpret_cast<uint64_t*>(sm + (int64_t)a.NST * a.stage_floats);
  ring.empty = ring.full + SA_MAX_NST;
  float4* red = reinterpret_cast<float4*>(ring.empty + SA_MAX_NST);   // [4][SA_CTHREADS]
  float* dctx_s = reinterpret_cast<float*>(red + 4 * SA_CTHREADS);    // [D]
  float* cx_s = dctx_s + D;                                           // [D] ctx[b,:]
  float* q_s = cx_s + D;                                              // [H]
  float* w0_s = q_s + H;                                              // [H] Ws[b,0,:]
  float* al_all = w0_s + H;                                            // [rp_pad] alpha of this CTA's whole row range
  float* da_s = al_all + a.rp_pad;                                    // [seg_pad]
  float* scratch = da_s + a.seg_pad;                                  // [40]
  int* segb = reinterpret_cast<int*>(scratch + 40);                   // [maxseg]
  int* segflag = segb + a.maxseg;                                     // [maxseg]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t total = (int64_t)a.B * L;
  const int64_t row0 = (int64_t)cta * a.rp;
  const int64_t row1 = (row0 + a.rp < total) ? row0 + a.rp : total;

  if (warp == SA_CWARPS) {
    if (lane == 0) {
      Header hd;                                  // [q | dgctx | beta | ctx] of the segment's image
      hdr_set(hd, 0, a.q, a.q_ld, H, 4);
      hdr_set(hd, 1, a.dgctx, a.dgctx_ld, D, 4);
      hdr_set(hd, 2, a.beta, D, D, 4);
      hdr_set(hd, 3, a.ctx, D, D, 4);
      hdr_set(hd, 4, a.Ws, (int64_t)L * H, H, BF ? 2 : 4);            // Ws[b,0,:]: the reference row of pass D
      hd.count = 5;
      if (a.l2_frac > 0.f) { ring.policy = pipe::l2_policy_evict_last(a.l2_frac); ring.hinted = true; }
      if (a.pdl == 1) pdl_wait();
      hd.dyn = 1u << 1;                                                // only d gctx comes from the preceding kernel
      produce(ring, k, row0, row1, L, RD, RH, 1, hd, a.feat, D, a.Ws, H, BF ? 2 : 4, a.pdl == 2 ? 2 : (a.pdl == 3 ? 1 : 0));     // feat pieces, then Ws pieces
    }
    return;
  }
  if (a.pdl) pdl_wait();

  SA_STAMP(1);
  const int grp = warp / SA_GWARPS, gw = warp % SA_GWARPS, gtid = tid % SA_GTHREADS;
  const bool wide = H4 >= SA_GTHREADS;        // H4 <= 256 (H <= 1024)
  const int G = wide ? 1 : SA_GTHREADS / H4;
  const int f = wide ? gtid : gtid % H4;
  const int g = wide ? 0 : gtid / H4;

  for (int i = tid; i < (int)(row1 - row0); i += SA_CTHREADS) {
    const int64_t row = row0 + i;
    const int bb = (int)(row / L);
    al_all[i] = __ldg(a.alpha + (int64_t)bb * a.alpha_ld + (row - (int64_t)bb * L));
  }

  int j_seg = 0;                  // kg: piece number modulo SA_NG (which consumer group owns the piece)
  for (int64_t lo = row0; lo < row1; ++j_seg) {
    const int b = (int)(lo / L);
    int64_t hi = (int64_t)(b + 1) * L;
    if (hi > row1) hi = row1;
    const int l0 = (int)(lo - (int64_t)b * L), nl = (int)(hi - lo);
    const float* al_s = al_all + (lo - row0);
    lo = hi;
    float* slot = a.scratch + ((int64_t)cta * a.maxseg + j_seg) * a.slot_floats;
    const bool whole = (l0 == 0 && nl == L);      // the image lives in this CTA alone: finalise locally
    if (tid == 0) segb[j_seg] = whole ? -1 : b;

    // upstream d alpha terms of this thread's rows (added after pass B): loaded now, so their latency is off the per-piece path
    float up0 = 0.f;
    if (tid < nl) {
      if (a.ext) up0 += __ldg(a.ext + (int64_t)b * a.ext_ld + l0 + tid);
      if (a.reg_r) up0 += __ldg(a.reg_r + (int64_t)b * L + l0 + tid);
    }
    // ---- A: header piece -> q_s, dctx_s;  gate backward  gctx = beta*ctx, beta = sigmoid(bpre)   (AttentionLanguageModule.py:74-75)
    ring.wait_full(k);
    {
      const float4* hq = reinterpret_cast<const float4*>(ring.stage(k));
      const float4* hdg = hq + H4; const float4* hbe = hdg + D4; const float4* hcx = hbe + D4;
      for (int ff = tid; ff < H4; ff += SA_CTHREADS) {
        reinterpret_cast<float4*>(q_s)[ff] = hq[ff];
        reinterpret_cast<float4*>(w0_s)[ff] = ld_row4<BF>(hcx + D4, ff);
      }
      for (int c = tid; c < D4; c += SA_CTHREADS) {
        const float4 dg = hdg[c], be = hbe[c];
        reinterpret_cast<float4*>(dctx_s)[c] = make_float4(dg.x * be.x, dg.y * be.y, dg.z * be.z, dg.w * be.w);
        const float4 cx = hcx[c];
        reinterpret_cast<float4*>(cx_s)[c] = cx;
        if (l0 == 0) {                             // the segment holding the image's first row writes d bpre
          float4 db;
          db.x = dg.x * cx.x * be.x * (1.f - be.x); db.y = dg.y * cx.y * be.y * (1.f - be.y);
          db.z = dg.z * cx.z * be.z * (1.f - be.z); db.w = dg.w * cx.w * be.w * (1.f - be.w);
          reinterpret_cast<float4*>(a.dbpre + (int64_t)b * a.dbpre_ld)[c] = db;
          if (a.dbpre_b16) reinterpret_cast<uint2*>(a.dbpre_b16 + (int64_t)b * a.db_ld)[c] = to_bf16x4(db);
        }
      }
    }
    pipe::named_sync(1, SA_CTHREADS);              // (also publishes al_all on the first segment)
    if (kg == grp && lane == 0) ring.release(k);
    k.next(ring.NST); kg = (kg + 1 == SA_NG) ? 0 : kg + 1;

    SA_STAMP(2);
    // ---- B: d alpha[l] = dctx . feat[b,l,:] (+ upstream terms);  dfeat += alpha[l] * dctx        (:21)
    // (Centring the dot product -- dctx . (feat_l - ctx) -- was measured: no accuracy gain, 40 % more pass-B instructions.)
    float4* dfb = a.dfeat ? reinterpret_cast<float4*>(a.dfeat) + ((int64_t)b * L + l0) * D4 : nullptr;
    for (int r0 = 0; r0 < nl; r0 += RD, k.next(ring.NST), kg = (kg + 1 == SA_NG) ? 0 : kg + 1) {
      if (kg != grp) continue;
      ring.wait_full(k);
      const void* st = ring.stage(k);
      const int rows = min(RD, nl - r0);
      for (int rr = gw; rr < rows; rr += 2 * SA_GWARPS) {
        const int rr2 = rr + SA_GWARPS;
        const bool has2 = rr2 < rows;
        const int i0 = r0 + rr, i1 = r0 + rr2;
        const float al0 = al_s[i0], al1 = has2 ? al_s[i1] : 0.f;
        float d0 = 0.f, d1 = 0.f;
        pipe::f32x2 p0 = 0ull, p1 = 0ull;            // packed partial dot products (FFMA2)
#pragma unroll 4
        for (int c = lane; c < D4; c += 32) {
          const float4 dc = reinterpret_cast<const float4*>(dctx_s)[c];
          const float4 x0 = ld_row4<BF>(st, rr * D4 + c);
          const float4 x1 = has2 ? ld_row4<BF>(st, rr2 * D4 + c) : x0;
          const pipe::f32x2 dca = pipe::pk2(dc.x, dc.y), dcb = pipe::pk2(dc.z, dc.w);
          p0 = pipe::fma2(pipe::pk2(x0.x, x0.y), dca, p0); p0 = pipe::fma2(pipe::pk2(x0.z, x0.w), dcb, p0);
          p1 = pipe::fma2(pipe::pk2(x1.x, x1.y), dca, p1); p1 = pipe::fma2(pipe::pk2(x1.z, x1.w), dcb, p1);
          if (dfb) {
            float4 o = dfb[(int64_t)i0 * D4 + c];
            f4_fma(o, al0, dc);
            dfb[(int64_t)i0 * D4 + c] = o;
            if (has2) {
              float4 o2 = dfb[(int64_t)i1 * D4 + c];
              f4_fma(o2, al1, dc);
              dfb[(int64_t)i1 * D4 + c] = o2;
            }
          }
        }
        { float u0, u1; pipe::upk2(p0, u0, u1); d0 = u0 + u1; pipe::upk2(p1, u0, u1); d1 = u0 + u1; }
        d0 = warp_sum(d0); d1 = warp_sum(d1);
        if (lane == 0) {
          da_s[i0] = d0;
          if (has2) da_s[i1] = d1;
        }
      }
      __syncwarp();
      if (lane == 0) ring.release(k);
    }
    pipe::named_sync(1, SA_CTHREADS);
    SA_STAMP(3);
    float part = 0.f, asum = 0.f;
    for (int i = tid; i < nl; i += SA_CTHREADS) {
      float dv = da_s[i];
      if (i == tid) dv += up0;
      else {
        if (a.ext) dv += __ldg(a.ext + (int64_t)b * a.ext_ld + l0 + i);
        if (a.reg_r) dv += __ldg(a.reg_r + (int64_t)b * L + l0 + i);
      }
      da_s[i] = dv;                                        // (each element is owned by the one thread that reads it below)
      if (!whole) a.de[(int64_t)b * a.de_ld + l0 + i] = dv;   // raw d alpha; the finaliser turns it into d e
      part = fmaf(al_s[i], dv, part); asum += al_s[i];
    }
    part = pipe::group_sum(part, scratch, SA_CWARPS, 1);
    if (!whole) asum = pipe::group_sum(asum, scratch, SA_CWARPS, 1);     // (uniform per segment; a whole image has sum alpha = 1)
    const float ref = whole ? part : (asum > 0.f ? part / asum : 0.f);    // whole image: ref == dot (sum alpha = 1)
    if (!whole && tid == 0) { slot[0] = part; slot[1] = ref; }
    if (whole)                                            // de = alpha (dalpha - dot)                            (:20)
      for (int i = tid; i < nl; i += SA_CTHREADS) a.de[(int64_t)b * a.de_ld + i] = al_s[i] * (da_s[i] - part);

    SA_STAMP(4);
    // ---- D: score backward partial sums, thread-per-column over the segment's Ws rows               (:18-19)
    float4 x1 = f4_zero(), x2 = f4_zero(), y1 = f4_zero(), y2 = f4_zero();
    // sum_l de_l = 0 over the image, so any per-(image, h) constant may be subtracted from t_l and from t_l^2:
    // with tc = tanh(Ws[b,0,h] + q_h) the sums run over (t_l - tc) and (tc^2 - t_l^2), which are SMALL whenever t varies
    // little along l -- the case in which the textbook form sum_l de_l (1 - t_l^2) cancels catastrophically in fp32.
    float4 qq = (g < G && f < H4) ? reinterpret_cast<const float4*>(q_s)[f] : f4_zero();
    qq = q_stage<EW>(qq);
    // In terms of r = 1 / (2^z + 1) (t = 1 - 2r), rc = r of the reference row, dr = r - rc:
    //   t - tc = -2 dr ;  tc^2 - t^2 = 4 dr (tc - dr)      -> 9 instructions per element, the factors applied after the loop
    float4 rc = f4_zero(), tc = f4_zero();
    if (g < G && f < H4) {
      const float4 w0 = reinterpret_cast<const float4*>(w0_s)[f];
      rc.x = r_of<EW>(w0.x, qq.x); rc.y = r_of<EW>(w0.y, qq.y);
      rc.z = r_of<EW>(w0.z, qq.z); rc.w = r_of<EW>(w0.w, qq.w);
      tc = make_float4(fmaf(-2.f, rc.x, 1.f), fmaf(-2.f, rc.y, 1.f), fmaf(-2.f, rc.z, 1.f), fmaf(-2.f, rc.w, 1.f));
    }
    // Packed fp32x2 arithmetic (FFMA2 / FADD2 / FMUL2): per element pair 6 packed instructions + 4 MUFU.  Accumulated per
    // column: Y = sum ad*dr and Z = sum ad*dr^2 (and the alpha-weighted pair for split images); X = tc*Y - Z afterwards.
    using pipe::f32x2; using pipe::pk2; using pipe::upk2; using pipe::fma2; using pipe::add2; using pipe::mul2;
    const f32x2 c2 = pk2(TANH_C, TANH_C), one2 = pk2(1.f, 1.f);
    const f32x2 qa = pk2(qq.x, qq.y), qb = pk2(qq.z, qq.w), nra = pk2(-rc.x, -rc.y), nrb = pk2(-rc.z, -rc.w);
    f32x2 ya = 0ull, yb = 0ull, za = 0ull, zb = 0ull, y2a = 0ull, y2b = 0ull, z2a = 0ull, z2b = 0ull;   // +0.f pairs
    for (int r0 = 0; r0 < nl; r0 += RH, k.next(ring.NST), kg = (kg + 1 == SA_NG) ? 0 : kg + 1) {
      if (kg != grp) continue;
      ring.wait_full(k);
      const void* st = ring.stage(k);
      const int rows = min(RH, nl - r0);
      if (g < G && f < H4) {
#pragma unroll 2
        for (int rr = g; rr < rows; rr += G) {
          const float al = al_s[r0 + rr], ad = al * (da_s[r0 + rr] - ref);
          const f32x2 ad2 = pk2(ad, ad);
          const float4 w = ld_row4<BF>(st, rr * H4 + f);
          float d0, d1, d2, d3;
          if (EW) {                                  // w = EXP2C(Ws), q = EXP2C(q): 1 + e^{2(w+q)} is one packed FMA
            upk2(fma2(pk2(w.x, w.y), qa, one2), d0, d1);
            upk2(fma2(pk2(w.z, w.w), qb, one2), d2, d3);
          } else {
            float e0, e1, e2, e3;
            upk2(fma2(pk2(w.x, w.y), c2, qa), e0, e1);
            upk2(fma2(pk2(w.z, w.w), c2, qb), e2, e3);
            upk2(add2(pk2(ex2_ftz(e0), ex2_ftz(e1)), one2), d0, d1);
            upk2(add2(pk2(ex2_ftz(e2), ex2_ftz(e3)), one2), d2, d3);
          }
          const f32x2 dra = add2(pk2(rcp_ftz(d0), rcp_ftz(d1)), nra), drb = add2(pk2(rcp_ftz(d2), rcp_ftz(d3)), nrb);
          ya = fma2(ad2, dra, ya); yb = fma2(ad2, drb, yb);
          za = fma2(mul2(ad2, dra), dra, za); zb = fma2(mul2(ad2, drb), drb, zb);
          if (!ALLWHOLE) {
            const f32x2 al2 = pk2(al, al);
            y2a = fma2(al2, dra, y2a); y2b = fma2(al2, drb, y2b);
            z2a = fma2(mul2(al2, dra), dra, z2a); z2b = fma2(mul2(al2, drb), drb, z2b);
          }
        }
      }
      __syncwarp();
      if (lane == 0) ring.release(k);
    }
    {
      float4 Y, Z;
      upk2(ya, Y.x, Y.y); upk2(yb, Y.z, Y.w); upk2(za, Z.x, Z.y); upk2(zb, Z.z, Z.w);
      x1 = make_float4(4.f * fmaf(tc.x, Y.x, -Z.x), 4.f * fmaf(tc.y, Y.y, -Z.y), 4.f * fmaf(tc.z, Y.z, -Z.z), 4.f * fmaf(tc.w, Y.w, -Z.w));
      y1 = make_float4(-2.f * Y.x, -2.f * Y.y, -2.f * Y.z, -2.f * Y.w);
      if (!ALLWHOLE) {
        upk2(y2a, Y.x, Y.y); upk2(y2b, Y.z, Y.w); upk2(z2a, Z.x, Z.y); upk2(z2b, Z.z, Z.w);
        x2 = make_float4(4.f * fmaf(tc.x, Y.x, -Z.x), 4.f * fmaf(tc.y, Y.y, -Z.y), 4.f * fmaf(tc.z, Y.z, -Z.z), 4.f * fmaf(tc.w, Y.w, -Z.w));
        y2 = make_float4(-2.f * Y.x, -2.f * Y.y, -2.f * Y.z, -2.f * Y.w);
      }
    }
    SA_STAMP(5);
    // reduce over the consumer groups and the row sub-groups inside each
    red[tid] = x1; red[2 * SA_CTHREADS + tid] = y1;
    if (!ALLWHOLE) { red[SA_CTHREADS + tid] = x2; red[3 * SA_CTHREADS + tid] = y2; }
    pipe::named_sync(1, SA_CTHREADS);
    if (grp == 0 && g == 0 && f < H4) {
      x1 = x2 = y1 = y2 = f4_zero();
      for (int gi = 0; gi < SA_NG; ++gi)
        for (int gg = 0; gg < G; ++gg) {
          const int o = gi * SA_GTHREADS + gg * H4 + f;
          f4_add(x1, red[o]); f4_add(y1, red[2 * SA_CTHREADS + o]);
          if (!ALLWHOLE) { f4_add(x2, red[SA_CTHREADS + o]); f4_add(y2, red[3 * SA_CTHREADS + o]); }
        }
      if (whole) {                                        // ref == dot: dq = w_v X1, dw_v += Y1
        const float4 v = __ldg(reinterpret_cast<const float4*>(a.wv) + f);
        const float4 dqv = make_float4(v.x * x1.x, v.y * x1.y, v.z * x1.z, v.w * x1.w);
        reinterpret_cast<float4*>(a.dq + (int64_t)b * a.dq_ld)[f] = dqv;
        if (a.dq_b16) reinterpret_cast<uint2*>(a.dq_b16 + (int64_t)b * a.db_ld)[f] = to_bf16x4(dqv);
        float4* out = reinterpret_cast<float4*>(a.dwv_part + (int64_t)b * H) + f;
        float4 o = *out;
        f4_add(o, y1);
        *out = o;
      } else {
        float4* sl4 = reinterpret_cast<float4*>(slot + 4);
        sl4[f] = x1; sl4[H4 + f] = x2; sl4[2 * H4 + f] = y1; sl4[3 * H4 + f] = y2;
      }
    }
    pipe::named_sync(1, SA_CTHREADS);             // smem vectors are reused by the next segment
    SA_STAMP(6);
  }
  const int nsegs = j_seg;

  // ---- publish the split segments with ONE fence; an image is finalised by the CTA whose segment arrives last
  if ((int64_t)a.rp % L == 0) return;               // every image lives in one CTA: nothing to combine
  __threadfence();
  pipe::named_sync(1, SA_CTHREADS);
  if (tid < nsegs) {
    const int b = segb[tid];
    int fl = 0;
    if (b >= 0) {
      const int c_first = (int)(((int64_t)b * L) / a.rp), c_last = (int)(((int64_t)(b + 1) * L - 1) / a.rp);
      fl = (atomicAdd(a.counters + b, 1u) == (unsigned)(c_last - c_first)) ? 1 : 0;
    }
    segflag[tid] = fl;
  }
  pipe::named_sync(1, SA_CTHREADS);
  for (int j = 0; j < nsegs; ++j) {
    if (!segflag[j]) continue;
    const int b = segb[j];
    const int c_first = (int)(((int64_t)b * L) / a.rp), c_last = (int)(((int64_t)(b + 1) * L - 1) / a.rp);
    const int nseg = c_last - c_first + 1;
    __threadfence();
    float* fin = reinterpret_cast<float*>(red);                 // [2*nseg] (partial dot, ref_s)
    for (int i = tid; i < nseg; i += SA_CTHREADS) {
      const int cc = c_first + i;
      const int jj = b - (int)(((int64_t)cc * a.rp) / L);
      const float* sl = a.scratch + ((int64_t)cc * a.maxseg + jj) * a.slot_floats;
      fin[2 * i] = __ldcg(sl); fin[2 * i + 1] = __ldcg(sl + 1);
    }
    pipe::named_sync(1, SA_CTHREADS);
    float dot = 0.f;
    for (int i = 0; i < nseg; ++i) dot += fin[2 * i];
    if (tid < H4) {
      float4 s1 = f4_zero(), t1 = f4_zero();
      for (int i = 0; i < nseg; ++i) {
        const int cc = c_first + i;
        const int jj = b - (int)(((int64_t)cc * a.rp) / L);
        const float4* sl4 = reinterpret_cast<const float4*>(a.scratch + ((int64_t)cc * a.maxseg + jj) * a.slot_floats + 4);
        const float ds = dot - fin[2 * i + 1];
        f4_add(s1, __ldcg(sl4 + tid)); f4_fma(s1, -ds, __ldcg(sl4 + H4 + tid));
        f4_add(t1, __ldcg(sl4 + 2 * H4 + tid)); f4_fma(t1, -ds, __ldcg(sl4 + 3 * H4 + tid));
      }
      const float4 v = __ldg(reinterpret_cast<const float4*>(a.wv) + tid);
      const float4 dqv = make_float4(v.x * s1.x, v.y * s1.y, v.z * s1.z, v.w * s1.w);
      reinterpret_cast<float4*>(a.dq + (int64_t)b * a.dq_ld)[tid] = dqv;
      if (a.dq_b16) reinterpret_cast<uint2*>(a.dq_b16 + (int64_t)b * a.db_ld)[tid] = to_bf16x4(dqv);
      float4* out = reinterpret_cast<float4*>(a.dwv_part + (int64_t)b * H) + tid;
      float4 o = *out;
      f4_add(o, t1);
      *out = o;
    }
    float* deb = a.de + (int64_t)b * a.de_ld;
    const float* alb = a.alpha + (int64_t)b * a.alpha_ld;
    for (int i = tid; i < L; i += SA_CTHREADS) deb[i] = __ldg(alb + i) * (__ldcg(deb + i) - dot);
    if (tid == 0) a.counters[b] = 0;
    pipe::named_sync(1, SA_CTHREADS);             // fin (red) is reused by the next image
  }
}


}  // namespace sattn
}  // namespace alexcap
