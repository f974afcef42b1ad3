// Multi-head self-attention of the timm 0.6.12 Block (reference call site Models/models.py:333-334):
// softmax(q k^T * d^-0.5) v for 12 heads x 64 dims over 197 tokens, forward and backward.
//   *_simt : fp32 FFMA kernels for the 1e-4 validation mode (and the cross-check of the tensor-core kernels)
//   *_mma  : bf16 tensor-core kernels (attention_mma.cu)
// qkv is the [B*N, 3*H*64] output of attn.qkv (q | k | v, head-major inside each third).
#include "kernels.h"

namespace {
constexpr int D = 64;

template <typename T>
__global__ void __launch_bounds__(256) attn_fwd_simt_kernel(const T* __restrict__ qkv, T* __restrict__ o,
                                                            float* __restrict__ lse, int N, int H) {
  extern __shared__ float sm[];
  float* Ks = sm;                    // [N][65]
  float* Vs = Ks + N * 65;           // [N][64]
  float* qs = Vs + N * 64;           // [8][64]
  float* ps = qs + 8 * 64;           // [8][NP]
  const int NP = (N + 31) / 32 * 32;
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int C3 = 3 * H * D, C = H * D;
  const T* base = qkv + (long long)b * N * C3 + h * D;
  for (int i = threadIdx.x; i < N * D; i += 256) {
    const int t = i / D, d = i % D;
    Ks[t * 65 + d] = to_f(base[(long long)t * C3 + C + d]);
    Vs[t * 64 + d] = to_f(base[(long long)t * C3 + 2 * C + d]);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* q = qs + warp * 64;
  float* p = ps + warp * NP;
  const float scale = 0.125f;
  for (int i = warp; i < N; i += 8) {
    q[lane] = to_f(base[(long long)i * C3 + lane]);
    q[lane + 32] = to_f(base[(long long)i * C3 + lane + 32]);
    __syncwarp();
    float mx = -INFINITY;
    for (int j = lane; j < N; j += 32) {
      float s = 0.f;
#pragma unroll 16
      for (int d = 0; d < D; ++d) s = fmaf(q[d], Ks[j * 65 + d], s);
      s *= scale;
      p[j] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < N; j += 32) { const float e = expf(p[j] - mx); p[j] = e; sum += e; }
    sum = warp_sum(sum);
    __syncwarp();
    float a0 = 0.f, a1 = 0.f;
    for (int j = 0; j < N; ++j) { const float pj = p[j]; a0 = fmaf(pj, Vs[j * 64 + lane], a0); a1 = fmaf(pj, Vs[j * 64 + lane + 32], a1); }
    const float inv = 1.f / sum;
    T* orow = o + ((long long)b * N + i) * C + h * D;
    orow[lane] = from_f<T>(a0 * inv);
    orow[lane + 32] = from_f<T>(a1 * inv);
    if (lane == 0) lse[((long long)b * H + h) * N + i] = mx + logf(sum);
    __syncwarp();
  }
}

template <typename T>
__global__ void __launch_bounds__(256) attn_bwd_simt_kernel(const T* __restrict__ qkv, const T* __restrict__ o,
                                                            const T* __restrict__ dout, const float* __restrict__ lse,
                                                            T* __restrict__ dqkv, int N, int H) {
  extern __shared__ float sm[];
  const int NP = (N + 31) / 32 * 32;
  float* Ks = sm;                 // [N][65]
  float* Vs = Ks + N * 65;        // [N][65]
  float* dK = Vs + N * 65;        // [N][64]
  float* dV = dK + N * 64;        // [N][64]
  float* qs = dV + N * 64;        // [8][64]
  float* dos = qs + 8 * 64;       // [8][64]
  float* ps = dos + 8 * 64;       // [8][NP]
  float* dss = ps + 8 * NP;       // [8][NP]
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int C3 = 3 * H * D, C = H * D;
  const T* base = qkv + (long long)b * N * C3 + h * D;
  T* dbase = dqkv + (long long)b * N * C3 + h * D;
  for (int i = threadIdx.x; i < N * D; i += 256) {
    const int t = i / D, d = i % D;
    Ks[t * 65 + d] = to_f(base[(long long)t * C3 + C + d]);
    Vs[t * 65 + d] = to_f(base[(long long)t * C3 + 2 * C + d]);
    dK[i] = 0.f; dV[i] = 0.f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* q = qs + warp * 64;
  float* dO = dos + warp * 64;
  float* p = ps + warp * NP;
  float* ds = dss + warp * NP;
  const float scale = 0.125f;
  for (int i0 = 0; i0 < N; i0 += 8) {
    const int i = i0 + warp;
    const bool ok = i < N;
    if (ok) {
      const long long r = ((long long)b * N + i);
      q[lane] = to_f(base[(long long)i * C3 + lane]);
      q[lane + 32] = to_f(base[(long long)i * C3 + lane + 32]);
      const float d0 = to_f(dout[r * C + h * D + lane]), d1 = to_f(dout[r * C + h * D + lane + 32]);
      dO[lane] = d0; dO[lane + 32] = d1;
      float dl = d0 * to_f(o[r * C + h * D + lane]) + d1 * to_f(o[r * C + h * D + lane + 32]);
      dl = warp_sum(dl);
      const float l = lse[((long long)b * H + h) * N + i];
      __syncwarp();
      for (int j = lane; j < N; j += 32) {
        float s = 0.f, dp = 0.f;
#pragma unroll 16
        for (int d = 0; d < D; ++d) { s = fmaf(q[d], Ks[j * 65 + d], s); dp = fmaf(dO[d], Vs[j * 65 + d], dp); }
        const float pj = expf(s * scale - l);
        p[j] = pj;
        ds[j] = pj * (dp - dl) * scale;
      }
      __syncwarp();
      float a0 = 0.f, a1 = 0.f;
      for (int j = 0; j < N; ++j) { const float v = ds[j]; a0 = fmaf(v, Ks[j * 65 + lane], a0); a1 = fmaf(v, Ks[j * 65 + lane + 32], a1); }
      dbase[(long long)i * C3 + lane] = from_f<T>(a0);
      dbase[(long long)i * C3 + lane + 32] = from_f<T>(a1);
    } else {
      for (int j = lane; j < NP; j += 32) { p[j] = 0.f; ds[j] = 0.f; }
      q[lane] = q[lane + 32] = 0.f; dO[lane] = dO[lane + 32] = 0.f;
    }
    __syncthreads();
    for (int e = threadIdx.x; e < N * D; e += 256) {
      const int j = e / D, d = e % D;
      float ak = 0.f, av = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) { ak = fmaf(dss[w * NP + j], qs[w * 64 + d], ak); av = fmaf(ps[w * NP + j], dos[w * 64 + d], av); }
      dK[e] += ak; dV[e] += av;
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < N * D; i += 256) {
    const int t = i / D, d = i % D;
    dbase[(long long)t * C3 + C + d] = from_f<T>(dK[i]);
    dbase[(long long)t * C3 + 2 * C + d] = from_f<T>(dV[i]);
  }
}
}  // namespace

template <typename T> int attn_fwd_simt(const T* qkv, T* o, float* lse, int B, int N, int H, cudaStream_t st) {
  const int NP = (N + 31) / 32 * 32;
  const size_t smem = sizeof(float) * ((size_t)N * 65 + (size_t)N * 64 + 8 * 64 + 8 * NP);
  S4G_CUDA(cudaFuncSetAttribute(attn_fwd_simt_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  attn_fwd_simt_kernel<T><<<B * H, 256, smem, st>>>(qkv, o, lse, N, H);
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T>
int attn_bwd_simt(const T* qkv, const T* o, const T* dout, const float* lse, T* dqkv, int B, int N, int H, cudaStream_t st) {
  const int NP = (N + 31) / 32 * 32;
  const size_t smem = sizeof(float) * ((size_t)N * 65 * 2 + (size_t)N * 64 * 2 + 16 * 64 + 16 * NP);
  if (smem > 227 * 1024) { s4g_set_error("attn_bwd_simt: sequence too long for shared memory"); return S4G_ERR_SHAPE; }
  S4G_CUDA(cudaFuncSetAttribute(attn_bwd_simt_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  attn_bwd_simt_kernel<T><<<B * H, 256, smem, st>>>(qkv, o, dout, lse, dqkv, N, H);
  S4G_KERNEL_CHECK();
  return 0;
}
template int attn_fwd_simt<float>(const float*, float*, float*, int, int, int, cudaStream_t);
template int attn_fwd_simt<bf16>(const bf16*, bf16*, float*, int, int, int, cudaStream_t);
template int attn_bwd_simt<float>(const float*, const float*, const float*, const float*, float*, int, int, int, cudaStream_t);
template int attn_bwd_simt<bf16>(const bf16*, const bf16*, const bf16*, const float*, bf16*, int, int, int, cudaStream_t);
