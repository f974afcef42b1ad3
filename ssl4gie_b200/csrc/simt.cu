// Validation-mode contraction kernels: plain FFMA implicit GEMM (fprop / dgrad-as-fprop) and wgrad.
// Used for the fp32 1e-4 parity mode (tcgen05 kind::tf32 has a 10-bit mantissa, not enough) and as
// the on-device cross-check of the tcgen05 kernels in tests.  NOT the bf16 production path.
#include "kernels.h"

namespace {

constexpr int BM = 64, BN = 64, BK = 16;

// out[m][n] = sum_{tap,ci} X[pix(m) shifted by tap][ci] * W[n][tap*Cin+ci]
template <typename T>
__global__ void __launch_bounds__(256) conv_fprop_simt_kernel(ConvGeom g, const T* __restrict__ X,
                                                              const T* __restrict__ W, EpiParams e) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const long long M = (long long)g.Nimg * g.Ho * g.Wo;
  const int N = g.Cout, K = g.R * g.S * g.Cin;
  const long long m0 = (long long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int lr = tid >> 2, lk = (tid & 3) * 4;  // loader: row within tile, k offset
  // decode this thread's A row once
  const long long am = m0 + lr;
  int an = 0, aho = 0, awo = 0;
  const bool arow_ok = am < M;
  if (arow_ok) {
    an = (int)(am / (g.Ho * g.Wo));
    int rem = (int)(am % (g.Ho * g.Wo));
    aho = rem / g.Wo; awo = rem % g.Wo;
  }
  const int bn = n0 + lr;
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < K; k0 += BK) {
    const int tap = k0 / g.Cin, ci0 = k0 % g.Cin;
    const int r = tap / g.S, s = tap % g.S;
    float av[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
    if (arow_ok) {
      const int hi = aho * g.stride + r - g.pad, wi = awo * g.stride + s - g.pad;
      if (hi >= 0 && hi < g.Hi && wi >= 0 && wi < g.Wi) {
        const T* p = X + (((long long)an * g.Hi + hi) * g.Wi + wi) * g.ldx + ci0 + lk;
#pragma unroll
        for (int i = 0; i < 4; ++i) av[i] = to_f(p[i]);
      }
    }
    if (bn < N) {
      const T* p = W + (long long)bn * K + k0 + lk;
#pragma unroll
      for (int i = 0; i < 4; ++i) bv[i] = to_f(p[i]);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) { As[lk + i][lr] = av[i]; Bs[lk + i][lr] = bv[i]; }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; b[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < N) epi_store1<T>(e, m, n, acc[i][j]);
    }
  }
}

// dWp[co][tap*Cin+ci] (+)= sum_m dY[m][co] * X[pix(m) shifted by tap][ci];  fp32 out, split over m
template <typename T>
__global__ void __launch_bounds__(256) conv_wgrad_simt_kernel(ConvGeom g, const T* __restrict__ dY,
                                                              const T* __restrict__ X, float* __restrict__ dW,
                                                              long long m_per_split, int use_atomic) {
  __shared__ float As[BK][BM + 4];  // [pixel][co]
  __shared__ float Bs[BK][BN + 4];  // [pixel][col]
  const int tid = threadIdx.x;
  const long long M = (long long)g.Nimg * g.Ho * g.Wo;
  const int K = g.R * g.S * g.Cin;
  const int co0 = blockIdx.x * BM, c0 = blockIdx.y * BN;
  const long long mb = (long long)blockIdx.z * m_per_split;
  const long long me = min(M, mb + m_per_split);
  const int lp = tid >> 4, lc = (tid & 15) * 4;  // loader: pixel within chunk, 4 consecutive cols
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // column -> (tap, ci) for the 4 columns this thread loads (4-aligned and Cin%4==0 => same tap)
  const int col = c0 + lc;
  const bool col_ok = col < K;
  const int tap = col_ok ? col / g.Cin : 0, ci = col_ok ? col % g.Cin : 0;
  const int r = tap / g.S, s = tap % g.S;

  for (long long p0 = mb; p0 < me; p0 += BK) {
    const long long m = p0 + lp;
    float av[4] = {0.f, 0.f, 0.f, 0.f}, bv[4] = {0.f, 0.f, 0.f, 0.f};
    if (m < me) {
      if (co0 + lc < g.Cout) {
        const T* p = dY + m * g.ldy + co0 + lc;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (co0 + lc + i < g.Cout) av[i] = to_f(p[i]);
      }
      if (col_ok) {
        const int n = (int)(m / (g.Ho * g.Wo));
        const int rem = (int)(m % (g.Ho * g.Wo));
        const int hi = (rem / g.Wo) * g.stride + r - g.pad, wi = (rem % g.Wo) * g.stride + s - g.pad;
        if (hi >= 0 && hi < g.Hi && wi >= 0 && wi < g.Wi) {
          const T* p = X + (((long long)n * g.Hi + hi) * g.Wi + wi) * g.ldx + ci;
#pragma unroll
          for (int i = 0; i < 4; ++i) bv[i] = to_f(p[i]);
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) { As[lp][lc + i] = av[i]; Bs[lp][lc + i] = bv[i]; }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[kk][ty * 4 + i]; b[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int co = co0 + ty * 4 + i;
    if (co >= g.Cout) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c0 + tx * 4 + j;
      if (c >= K) continue;
      float* o = dW + (long long)co * K + c;
      if (use_atomic) atomicAdd(o, acc[i][j]); else *o = acc[i][j];
    }
  }
}

}  // namespace

template <typename T>
int simt_conv_fprop(const ConvGeom& g, const T* X, const T* W, const EpiParams& e, cudaStream_t st) {
  const long long M = (long long)g.Nimg * g.Ho * g.Wo;
  if (g.Cin % 16 != 0) { s4g_set_error("simt_conv_fprop: Cin must be a multiple of 16"); return S4G_ERR_SHAPE; }
  dim3 grid(cdiv(M, BM), cdiv(g.Cout, BN));
  conv_fprop_simt_kernel<T><<<grid, 256, 0, st>>>(g, X, W, e);
  S4G_KERNEL_CHECK();
  return 0;
}

template <typename T>
int simt_conv_wgrad(const ConvGeom& g, const T* dY, const T* X, float* dW, cudaStream_t st) {
  const long long M = (long long)g.Nimg * g.Ho * g.Wo;
  const int K = g.R * g.S * g.Cin;
  if (g.Cin % 4 != 0) { s4g_set_error("simt_conv_wgrad: Cin must be a multiple of 4"); return S4G_ERR_SHAPE; }
  const int tiles = cdiv(g.Cout, BM) * cdiv(K, BN);
  int splits = (int)min((long long)cdiv(M, 512), (long long)max(1, 592 / tiles));
  long long mps = ((M + splits - 1) / splits + BK - 1) / BK * BK;
  splits = cdiv(M, mps);
  if (splits > 1) S4G_CUDA(cudaMemsetAsync(dW, 0, sizeof(float) * (size_t)g.Cout * K, st));
  dim3 grid(cdiv(g.Cout, BM), cdiv(K, BN), splits);
  conv_wgrad_simt_kernel<T><<<grid, 256, 0, st>>>(g, dY, X, dW, mps, splits > 1);
  S4G_KERNEL_CHECK();
  return 0;
}

template int simt_conv_fprop<float>(const ConvGeom&, const float*, const float*, const EpiParams&, cudaStream_t);
template int simt_conv_fprop<bf16>(const ConvGeom&, const bf16*, const bf16*, const EpiParams&, cudaStream_t);
template int simt_conv_wgrad<float>(const ConvGeom&, const float*, const float*, float*, cudaStream_t);
template int simt_conv_wgrad<bf16>(const ConvGeom&, const bf16*, const bf16*, float*, cudaStream_t);
