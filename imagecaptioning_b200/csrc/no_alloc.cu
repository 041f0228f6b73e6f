// no_alloc.cu -- alexcap_cross_entropy_static: alexcap_cross_entropy (api.cu) for callers that may not allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The scratch the existing entry point takes from the caller (row losses [N] + n_valid [1]) lives here in ONE file-scope
// __device__ array of a fixed, modest size (ALEXCAP_CE_STATIC_MAX_ROWS rows -> 512 KB per GPU that loads the library, held
// until the process exits).  In host code the array's NAME is a host-side stand-in, not a device address: the device address
// is looked up once per device (with that device current) with cudaGetSymbolAddress and that pointer is what the kernels get.
// Every call shares the array and it keeps the last call's contents; the region a call uses is cleared at its start, on the
// caller's stream.  Calls that overlap on the GPU (different streams, no ordering between them) RACE on it.
#include <mutex>

#include "api_util.h"

namespace alexcap {
namespace {

constexpr int kMaxRows = ALEXCAP_CE_STATIC_MAX_ROWS;
constexpr int kMaxDevices = 64;

// [0, N) row losses, [N] n_valid (the layout alexcap_cross_entropy uses inside its caller-provided scratch)
static __device__ __align__(256) float g_ce_static_scratch[2 * kMaxRows];

std::mutex g_addr_mu;
float* g_addr[kMaxDevices] = {};

// device address of g_ce_static_scratch on the CURRENT device (each device that loaded the module has its own instance)
int scratch_address(float** out) {
  int dev = -1;
  API_CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(ALEXCAP_E_UNSUPPORTED, "alexcap_cross_entropy_static: device ordinal >= 64");
  std::lock_guard<std::mutex> lk(g_addr_mu);
  if (!g_addr[dev]) {
    float* p = nullptr;
    API_CK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_ce_static_scratch));
    if (!p) return fail(ALEXCAP_E_CUDA, "alexcap_cross_entropy_static: cudaGetSymbolAddress returned NULL");
    g_addr[dev] = p;
  }
  *out = g_addr[dev];
  return 0;
}

}  // namespace
}  // namespace alexcap

using namespace alexcap;

extern "C" {

int32_t alexcap_cross_entropy_static_max_rows(void) { return kMaxRows; }

int alexcap_cross_entropy_static(const float* logits, const int32_t* tgt, int32_t N, int32_t C, int64_t ld,
                                 float label_smoothing, int32_t ignore_index, float* loss_out, float* dlogits, void* stream) {
  if (!logits || !tgt || !loss_out || N <= 0 || C <= 0 || ld < C)
    return fail(ALEXCAP_E_BADARG, "alexcap_cross_entropy_static: bad arguments");
  if (N > kMaxRows) {
    snprintf(err_buf(), 512, "alexcap_cross_entropy_static: N = %d rows exceeds the static scratch (max %d rows); use "
             "alexcap_cross_entropy with caller-provided scratch", (int)N, kMaxRows);
    return ALEXCAP_E_UNSUPPORTED;
  }
  float* scratch = nullptr;
  if (int rc = scratch_address(&scratch)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the array holds whatever the previous call left: start this call's region from zero
  API_CK(cudaMemsetAsync(scratch, 0, (size_t)(N + 1) * sizeof(float), st));
  API_CK(count_valid(st, tgt, N, ignore_index, scratch + N));
  API_CK(ce_rows(st, logits, tgt, N, C, ld, label_smoothing, ignore_index, scratch + N, scratch, dlogits));
  API_CK(reduce_sum(st, scratch, N, loss_out));
  API_CK(cudaMemcpyAsync(loss_out + 1, scratch + N, sizeof(float), cudaMemcpyDeviceToDevice, st));
  return 0;
}

}  // extern "C"
