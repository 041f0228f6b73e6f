// api_util.h -- error reporting and workspace carving shared by the C-ABI translation units.
#pragma once
#include <stdio.h>
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

// Decoding must reproduce the reference's token ids exactly (fp32): never run it on a reduced-precision engine.
struct ExactScope {
  int saved;
  ExactScope() : saved(gemm_engine()) { if (saved >= 2) gemm_engine() = 1; }
  ~ExactScope() { gemm_engine() = saved; }
};

}  // namespace alexcap
