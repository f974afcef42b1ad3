// Shared device/host helpers for libssl4gie_b200 (sm_100a only).
#pragma once
#include <cstdlib>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------------------------
// error plumbing: kernels never throw / exit; C-ABI entry points return negative S4G_ERR_* codes
// ---------------------------------------------------------------------------------------------
void s4g_set_error(const std::string& msg);
int s4g_cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define S4G_CUDA(call)                                                        \
  do {                                                                        \
    cudaError_t _e = (call);                                                  \
    if (_e != cudaSuccess) return s4g_cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)
#define S4G_KERNEL_CHECK() S4G_CUDA(cudaGetLastError())
#define S4G_TRY(expr)          \
  do {                         \
    int _r = (expr);           \
    if (_r != 0) return _r;    \
  } while (0)

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---------------------------------------------------------------------------------------------
// scalar conversion
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// 8-wide vector access (16 B for bf16, 32 B for fp32): the unit every NHWC elementwise kernel works in
template <typename T> struct Vec8;
template <> struct Vec8<float> {
  float v[8];
  __device__ __forceinline__ void load(const float* p) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  __device__ __forceinline__ void store(float* p) const {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <> struct Vec8<bf16> {
  float v[8];
  __device__ __forceinline__ void load(const bf16* p) {
    uint4 r = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 f = __bfloat1622float2(h[i]);
      v[2 * i] = f.x; v[2 * i + 1] = f.y;
    }
  }
  __device__ __forceinline__ void store(bf16* p) const {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = r;
  }
};

// raw 8-element load kept in its memory format (4 registers for bf16) until get(): lets a kernel keep several
// loads in flight without paying fp32 registers for them
template <typename T> struct Raw8;
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) { a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 4); }
  __device__ __forceinline__ void get(float (&f)[8]) const { f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w; }
};
template <> struct Raw8<bf16> {
  uint4 r;
  __device__ __forceinline__ void load(const bf16* p) { r = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void get(float (&f)[8]) const {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { f[2 * i] = __uint_as_float(w[i] << 16); f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Programmatic dependent launch: a kernel launched with programmaticStreamSerializationAllowed may start (barrier init, TMEM
// allocation, descriptor prefetch) while the previous kernel of the stream drains; pdl_wait() returns once that kernel has
// completed and its writes are visible.  pdl_trigger() lets the NEXT kernel of the stream do the same with this one.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// Host side of the same: launch `kernel` so that it may begin before the previous kernel of `st` has drained.  Only for
// kernels that execute pdl_wait() before their first global-memory access.  S4G_PDL=0 launches them the ordinary way.
inline int pdl_enabled() {
  static const int v = [] { const char* e = getenv("S4G_PDL"); return e ? (atoi(e) != 0) : 1; }();
  return v;
}
template <typename... KA, typename... A>
inline cudaError_t launch_dep(void (*kernel)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KA>(args)...);
}
__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float dgelu_f(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}

// fast GELU pieces for the bf16 tensor-core epilogues: erf by Abramowitz-Stegun 7.1.26 (|err| < 1.5e-7, far below bf16
// resolution); cdf = Phi(x), pdf = phi(x) share one exponential
__device__ __forceinline__ void gelu_parts_fast(float x, float& cdf, float& pdf) {
  const float z = fabsf(x) * 0.70710678118654752440f;
  const float t = __frcp_rn(fmaf(0.3275911f, z, 1.0f));
  const float e = __expf(-z * z);
  const float poly = t * (0.254829592f + t * (-0.284496736f + t * (1.421413741f + t * (-1.453152027f + t * 1.061405429f))));
  const float erfa = fmaf(-poly, e, 1.0f);
  cdf = 0.5f * (1.0f + copysignf(erfa, x));
  pdf = 0.39894228040143267794f * e;
}
__device__ __forceinline__ float gelu_fast(float x) { float c, p; gelu_parts_fast(x, c, p); return x * c; }
__device__ __forceinline__ float dgelu_fast(float x) { float c, p; gelu_parts_fast(x, c, p); return fmaf(x, p, c); }

// counter-based RNG for dropout (recomputed in backward, never stored)
__device__ __forceinline__ uint32_t hash_u32(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return (uint32_t)x;
}
__device__ __forceinline__ bool dropout_keep(uint64_t seed, uint64_t idx, float p) {
  return (hash_u32(seed * 0x9E3779B97F4A7C15ULL + idx) >> 8) * (1.0f / 16777216.0f) >= p;
}

// ---------------------------------------------------------------------------------------------
// Epilogue shared by the SIMT (validation) and tcgen05 contraction kernels.
//   v = acc (+ bias[col]) ; mode-specific transform ; (+ residual) ; store as T or fp32
// ---------------------------------------------------------------------------------------------
enum { EPI_NONE = 0, EPI_GELU = 1, EPI_DGELU = 2, EPI_RELU = 3, EPI_MUL = 4 };   // EPI_MUL: v *= aux (a stored derivative)

struct EpiParams {
  const float* bias;   // [N] fp32 or null
  const void* res;     // residual [M, ldr] (fp32 if res_f32 else T) or null
  const void* aux;     // EPI_DGELU: pre-activation u (T) [M, ldo]; EPI_MUL: elementwise factor (T) [M, ldo]
  void* out;           // [M, ldo]
  void* out2;          // EPI_GELU: pre-activation copy (T) or null
  int ldo, ldr;
  int mode;
  int out_f32, res_f32;
  float alpha;         // scales acc before everything else (1.0 default)
  int accumulate;      // out += (fp32 outputs only)
  // optional fused row statistic of the tcgen05 bf16 (TMA-store) epilogue: for every 64-column group g of the output (one
  // attention head) dot_out[(row / dot_rows) * dot_groups * dot_rows + g * dot_rows + row % dot_rows] = sum_c out[row, 64 g + c] *
  // dot_aux[row, 64 g + c]  -- the attention backward's delta = rowsum(dO o O), produced where dO is produced
  const void* dot_aux;   // bf16 [M, ldo] or null
  float* dot_out;
  int dot_rows, dot_groups;
  // optional fused COLUMN statistics of the same epilogue (BatchNorm statistics of a conv output, N <= 256 in one column tile),
  // taken from the bf16-rounded values it stores: every CTA writes ONE row [2 N] = (sum | sum of squares) over all the rows it
  // produced: stat_part[cta * 2 N + ...], umma_fprop_stat_rows(geometry) rows in total, then summed in a fixed order by
  // finalize_kernel (reduce.cu): deterministic, no atomics, and the separate statistics pass over the tensor disappears
  float* stat_part;
};

template <typename T>
__device__ __forceinline__ void epi_store1(const EpiParams& e, long long row, int col, float v) {
  v *= e.alpha;
  if (e.bias) v += e.bias[col];
  const long long o = row * e.ldo + col;
  if (e.mode == EPI_GELU) {
    if (e.out2) reinterpret_cast<T*>(e.out2)[o] = from_f<T>(v);
    v = gelu_f(v);
  } else if (e.mode == EPI_DGELU) {
    v *= dgelu_f(to_f(reinterpret_cast<const T*>(e.aux)[o]));
  } else if (e.mode == EPI_RELU) {
    v = fmaxf(v, 0.f);
  } else if (e.mode == EPI_MUL) {
    v *= to_f(reinterpret_cast<const T*>(e.aux)[o]);
  }
  if (e.res) {
    const long long r = row * e.ldr + col;
    v += e.res_f32 ? reinterpret_cast<const float*>(e.res)[r] : to_f(reinterpret_cast<const T*>(e.res)[r]);
  }
  if (e.out_f32) {
    float* p = reinterpret_cast<float*>(e.out) + o;
    *p = e.accumulate ? (*p + v) : v;
  } else {
    reinterpret_cast<T*>(e.out)[o] = from_f<T>(v);
  }
}
