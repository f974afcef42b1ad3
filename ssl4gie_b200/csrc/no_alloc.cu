// Allocation-free variant of one entry point that needs scratch memory: s4g_layernorm_bwd_static.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  s4g_layernorm_bwd writes one partial row
// [3 C] per block for the deterministic dgamma / dbeta reduction (reduce.cu) into a scratch that s4g_init allocates with
// cudaMalloc; this variant keeps that scratch in a file-scope __device__ array instead, so nothing is allocated by the call or
// by any set-up it needs (s4g_init is not required).
//
// Limits and sharing rules (also in include/ssl4gie_b200.h):
//   * the array holds 148 blocks x 3 x 768 floats = 1 363 968 bytes: C <= 768 (the kernels exist for C = 512 and 768), any row
//     count; larger shapes are rejected with S4G_ERR_WORKSPACE.  Every GPU that loads the library holds the array until exit;
//   * every call on a device shares the one array, which keeps the previous call's partial rows.  Nothing has to start at zero
//     (every row that is read is fully overwritten by the same call first), so no clearing is done;
//   * calls that overlap on the GPU (different streams, or graph branches without an edge between them) race on the array:
//     serialise them.
#include <mutex>

#include "kernels.h"

namespace {

constexpr int STATIC_LN_MAX_C = 768;
constexpr size_t STATIC_LN_BYTES = (size_t)148 * 3 * STATIC_LN_MAX_C * sizeof(float);   // == ln_bwd_scratch_bytes(768)

static __device__ __align__(256) unsigned char g_static_ln_scratch[STATIC_LN_BYTES];

// In host code `g_static_ln_scratch` is only a handle for the symbol, not a device address: the address is looked up once per
// device, with that device current, and cached.
constexpr int MAX_DEVICES = 64;
void* g_static_ln_ptr[MAX_DEVICES] = {};
std::mutex g_static_ln_mutex;

int static_ln_scratch(void** out) {
  int dev = 0;
  S4G_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) { s4g_set_error("layernorm_bwd_static: device index out of range"); return S4G_ERR_STATE; }
  std::lock_guard<std::mutex> lock(g_static_ln_mutex);
  if (!g_static_ln_ptr[dev]) {
    void* p = nullptr;
    S4G_CUDA(cudaGetSymbolAddress(&p, g_static_ln_scratch));
    if (!p) { s4g_set_error("layernorm_bwd_static: cudaGetSymbolAddress returned a null address"); return S4G_ERR_CUDA; }
    g_static_ln_ptr[dev] = p;
  }
  *out = g_static_ln_ptr[dev];
  return 0;
}

}  // namespace

extern "C" int s4g_layernorm_bwd_static(const void* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                                        const float* dres, float* dx, float* dgamma, float* dbeta, int rows, int C, int dtype,
                                        void* stream) {
  if (!dy || !x || !gamma || !mean || !rstd || !dx || !dgamma || !dbeta || rows < 1 || C < 1) {
    s4g_set_error("layernorm_bwd_static: bad argument");
    return S4G_ERR_STATE;
  }
  if (C > STATIC_LN_MAX_C || ln_bwd_scratch_bytes(C) > STATIC_LN_BYTES) {
    s4g_set_error("layernorm_bwd_static: C = " + std::to_string(C) + " needs " + std::to_string(ln_bwd_scratch_bytes(C)) +
                  " bytes of scratch, the static array holds " + std::to_string(STATIC_LN_BYTES) + " (C <= 768); use s4g_layernorm_bwd");
    return S4G_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  S4G_TRY(static_ln_scratch(&scratch));
  Reducer r;
  r.scratch = scratch; r.scratch_bytes = STATIC_LN_BYTES; r.defer = nullptr;   // combine launched by this call, nothing queued
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == S4G_DTYPE_F32)
    return k_layernorm_bwd<float>(reinterpret_cast<const float*>(dy), x, gamma, mean, rstd, dres, dx, dgamma, dbeta, rows, C, st, nullptr, nullptr, &r);
  if (dtype == S4G_DTYPE_BF16)
    return k_layernorm_bwd<bf16>(reinterpret_cast<const bf16*>(dy), x, gamma, mean, rstd, dres, dx, dgamma, dbeta, rows, C, st, nullptr, nullptr, &r);
  s4g_set_error("bad dtype");
  return S4G_ERR_SHAPE;
}
