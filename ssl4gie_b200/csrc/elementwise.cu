// HBM-bound kernels of the ViT-B/16 + DPT path: patchify, token assembly, LayerNorm, GELU', ReLU, residual
// adds, pixel (un)shuffle for k==s transposed convs, stride-2 im2col, bilinear x2 (align_corners=True),
// BatchNorm (train) statistics / apply / backward, the two head tails, weight packing.
// All NHWC kernels move 8 channels per thread (16 B bf16 / 32 B fp32), coalesced along C.
#include <algorithm>

#include "kernels.h"

#define GRID1D(n, bs) dim3((unsigned)min((long long)cdiv((n), (bs)), (long long)148 * 32)), dim3(bs)
#define FOR_GRID(i, n) \
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (long long)gridDim.x * blockDim.x)

// ------------------------------------------------------------------------------------------ patch embed
// timm PatchEmbed (reference Models/models.py:326): Conv2d(3,768,k16,s16) == GEMM over k = c*256 + i*16 + j
template <typename T>
__global__ void patchify_kernel(const float* __restrict__ img, T* __restrict__ out, int B) {
  pdl_trigger();  // the next kernel of the stream, if launched programmatically dependent, may start its prologue now
  const long long n = (long long)B * 196 * 96;  // 768/8 groups per patch
  FOR_GRID(t, n) {
    const int g = (int)(t % 96);
    const long long pidx = t / 96;
    const int p = (int)(pidx % 196), b = (int)(pidx / 196);
    const int c = g / 32, i = (g % 32) / 2, j0 = (g % 2) * 8;
    const int ph = p / 14, pw = p % 14;
    const float* src = img + (((long long)b * 3 + c) * 224 + ph * 16 + i) * 224 + pw * 16 + j0;
    Vec8<T> v;
    float4 a = *reinterpret_cast<const float4*>(src), bq = *reinterpret_cast<const float4*>(src + 4);
    v.v[0] = a.x; v.v[1] = a.y; v.v[2] = a.z; v.v[3] = a.w; v.v[4] = bq.x; v.v[5] = bq.y; v.v[6] = bq.z; v.v[7] = bq.w;
    v.store(out + pidx * 768 + g * 8);
  }
}
template <typename T> int k_patchify(const float* img, T* out, int B, cudaStream_t st) {
  const long long n = (long long)B * 196 * 96;
  patchify_kernel<T><<<GRID1D(n, 256), 0, st>>>(img, out, B);
  S4G_KERNEL_CHECK();
  return 0;
}

// _pos_embed (Models/models.py:330 / :445-448): x[b,0] = cls + pos[0]; x[b,1+p] = pe[b,p] + pos[1+p]
__global__ void assemble_tokens_kernel(const float* __restrict__ pe, const float* __restrict__ cls,
                                       const float* __restrict__ pos, float* __restrict__ x, int B) {
  pdl_trigger();  // the next kernel of the stream, if launched programmatically dependent, may start its prologue now
  const long long n = (long long)B * 197 * 192;
  FOR_GRID(t, n) {
    const int c4 = (int)(t % 192);
    const long long row = t / 192;
    const int tok = (int)(row % 197), b = (int)(row / 197);
    const float4 pz = reinterpret_cast<const float4*>(pos)[tok * 192 + c4];
    float4 v = tok == 0 ? reinterpret_cast<const float4*>(cls)[c4]
                        : reinterpret_cast<const float4*>(pe)[((long long)b * 196 + tok - 1) * 192 + c4];
    v.x += pz.x; v.y += pz.y; v.z += pz.z; v.w += pz.w;
    reinterpret_cast<float4*>(x)[t] = v;
  }
}
int k_assemble_tokens(const float* pe, const float* cls, const float* pos, float* x, int B, cudaStream_t st) {
  assemble_tokens_kernel<<<GRID1D((long long)B * 197 * 192, 256), 0, st>>>(pe, cls, pos, x, B);
  S4G_KERNEL_CHECK();
  return 0;
}
// dpos[tok] = sum_b dx[b,tok]; dcls = sum_b dx[b,0]  (either may be null)
__global__ void assemble_tokens_bwd_kernel(const float* __restrict__ dx, float* __restrict__ dcls,
                                           float* __restrict__ dpos, int B) {
  const int n = 197 * 768;
  FOR_GRID(t, n) {
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dx[(long long)b * n + t];
    if (dpos) dpos[t] = s;
    if (dcls && t < 768) dcls[t] = s;
  }
}
int k_assemble_tokens_bwd(const float* dx, float* dpe, float* dcls, float* dpos, int B, cudaStream_t st) {
  (void)dpe;
  assemble_tokens_bwd_kernel<<<GRID1D(197 * 768, 256), 0, st>>>(dx, dcls, dpos, B);
  S4G_KERNEL_CHECK();
  return 0;
}

// 4 consecutive elements of T <-> float4
__device__ __forceinline__ float4 load4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 load4(const bf16* p) {
  const uint2 r = *reinterpret_cast<const uint2*>(p);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&r.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void store4(bf16* p, float4 v) {
  uint2 r;
  *reinterpret_cast<__nv_bfloat162*>(&r.x) = __floats2bfloat162_rn(v.x, v.y);
  *reinterpret_cast<__nv_bfloat162*>(&r.y) = __floats2bfloat162_rn(v.z, v.w);
  *reinterpret_cast<uint2*>(p) = r;
}

// ------------------------------------------------------------------------------------------ LayerNorm
// one warp per token row, fp32 statistics (two-pass in registers), T output
template <typename T, int C>
__global__ void __launch_bounds__(256) layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, T* __restrict__ y,
                                                            float* __restrict__ mean, float* __restrict__ rstd, int rows,
                                                            float eps) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  constexpr int V = C / 128;  // float4 per lane
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (long long)row * C);
  float4 v[V];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < V; ++i) { v[i] = xr[lane + 32 * i]; s += v[i].x + v[i].y + v[i].z + v[i].w; }
  const float mu = warp_sum(s) * (1.0f / C);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const float a = v[i].x - mu, b = v[i].y - mu, c = v[i].z - mu, d = v[i].w - mu;
    q += a * a + b * b + c * c + d * d;
  }
  const float rs = rsqrtf(warp_sum(q) * (1.0f / C) + eps);
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c4 = lane + 32 * i;
    const float4 g = reinterpret_cast<const float4*>(gamma)[c4], b = reinterpret_cast<const float4*>(beta)[c4];
    store4(y + (long long)row * C + c4 * 4, make_float4((v[i].x - mu) * rs * g.x + b.x, (v[i].y - mu) * rs * g.y + b.y,
                                                        (v[i].z - mu) * rs * g.z + b.z, (v[i].w - mu) * rs * g.w + b.w));
  }
}
template <typename T>
int k_layernorm_fwd(const float* x, const float* gamma, const float* beta, T* y, float* mean, float* rstd, int rows,
                    int C, float eps, cudaStream_t st) {
  if (C == 768) S4G_CUDA(launch_dep(layernorm_fwd_kernel<T, 768>, dim3(cdiv(rows, 8)), dim3(256), 0, st, x, gamma, beta, y, mean, rstd, rows, eps));
  else if (C == 512) S4G_CUDA(launch_dep(layernorm_fwd_kernel<T, 512>, dim3(cdiv(rows, 8)), dim3(256), 0, st, x, gamma, beta, y, mean, rstd, rows, eps));
  else { s4g_set_error("layernorm: C must be 768 or 512"); return S4G_ERR_SHAPE; }
  S4G_KERNEL_CHECK();
  return 0;
}

// LayerNorm backward, one warp per token row: lane l owns columns 4l + 128i (i < C/128) of every row its warp visits, so
// the three column accumulators (dgamma, dbeta, and the column sum of the produced dx = bias gradient of the Linear that
// precedes this residual branch) stay in registers across rows, the two per-row reductions are warp shuffles, and there
// is no block barrier inside the row loop.  At the end the warps of a block combine in shared memory in warp order and
// the block writes ONE partial row [3][C]; finalize (reduce.cu) sums the rows in a fixed order -- no atomics.
template <typename T, int C>
__global__ void __launch_bounds__(256, 1) layernorm_bwd_kernel(const T* __restrict__ dy, const float* __restrict__ x,
                                                               const float* __restrict__ gamma,
                                                               const float* __restrict__ mean, const float* __restrict__ rstd,
                                                               const float* __restrict__ dres, float* __restrict__ dx,
                                                               float* __restrict__ part, int rows, T* __restrict__ dx_cast) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  constexpr int V = C / 128;
  __shared__ float red[3][C];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4 gm[V], ag[V], ab[V], ac[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    gm[i] = reinterpret_cast<const float4*>(gamma)[lane + 32 * i];
    ag[i] = ab[i] = ac[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int wstride = gridDim.x * 8;
  for (int row = blockIdx.x * 8 + warp; row < rows; row += wstride) {
    const long long o = (long long)row * C;
    float4 xv[V], dv[V];
#pragma unroll
    for (int i = 0; i < V; ++i) {
      xv[i] = reinterpret_cast<const float4*>(x + o)[lane + 32 * i];
      dv[i] = load4(dy + o + (lane + 32 * i) * 4);
    }
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      xv[i] = make_float4((xv[i].x - mu) * rs, (xv[i].y - mu) * rs, (xv[i].z - mu) * rs, (xv[i].w - mu) * rs);   // xhat
      ab[i].x += dv[i].x; ab[i].y += dv[i].y; ab[i].z += dv[i].z; ab[i].w += dv[i].w;
      ag[i].x += dv[i].x * xv[i].x; ag[i].y += dv[i].y * xv[i].y; ag[i].z += dv[i].z * xv[i].z; ag[i].w += dv[i].w * xv[i].w;
      dv[i] = make_float4(dv[i].x * gm[i].x, dv[i].y * gm[i].y, dv[i].z * gm[i].z, dv[i].w * gm[i].w);           // g = dy * gamma
      s1 += dv[i].x + dv[i].y + dv[i].z + dv[i].w;
      s2 += dv[i].x * xv[i].x + dv[i].y * xv[i].y + dv[i].z * xv[i].z + dv[i].w * xv[i].w;
    }
    s1 = warp_sum(s1) * (1.0f / C);
    s2 = warp_sum(s2) * (1.0f / C);
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float4 rv = dres ? reinterpret_cast<const float4*>(dres + o)[lane + 32 * i] : make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 r = make_float4(rs * (dv[i].x - s1 - xv[i].x * s2) + rv.x, rs * (dv[i].y - s1 - xv[i].y * s2) + rv.y,
                                   rs * (dv[i].z - s1 - xv[i].z * s2) + rv.z, rs * (dv[i].w - s1 - xv[i].w * s2) + rv.w);
      reinterpret_cast<float4*>(dx + o)[lane + 32 * i] = r;
      if (dx_cast) store4(dx_cast + o + (lane + 32 * i) * 4, r);
      ac[i].x += r.x; ac[i].y += r.y; ac[i].z += r.z; ac[i].w += r.w;
    }
  }
  // warps add their register sums into the block's row one after the other (warp 0 stores, 1..7 accumulate)
  for (int w = 0; w < 8; ++w) {
    if (warp == w) {
#pragma unroll
      for (int i = 0; i < V; ++i) {
        float4* r0 = reinterpret_cast<float4*>(&red[0][0]) + lane + 32 * i;
        float4* r1 = reinterpret_cast<float4*>(&red[1][0]) + lane + 32 * i;
        float4* r2 = reinterpret_cast<float4*>(&red[2][0]) + lane + 32 * i;
        if (w == 0) { *r0 = ag[i]; *r1 = ab[i]; *r2 = ac[i]; }
        else {
          float4 t = *r0; t.x += ag[i].x; t.y += ag[i].y; t.z += ag[i].z; t.w += ag[i].w; *r0 = t;
          t = *r1; t.x += ab[i].x; t.y += ab[i].y; t.z += ab[i].z; t.w += ab[i].w; *r1 = t;
          t = *r2; t.x += ac[i].x; t.y += ac[i].y; t.z += ac[i].z; t.w += ac[i].w; *r2 = t;
        }
      }
    }
    __syncthreads();
  }
  float* prow = part + (size_t)blockIdx.x * 3 * C;
  for (int c = threadIdx.x; c < 3 * C; c += 256) prow[c] = (&red[0][0])[c];
}
template <typename T>
int k_layernorm_bwd(const T* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                    const float* dres, float* dx, float* dgamma, float* dbeta, int rows, int C, cudaStream_t st, T* dx_cast,
                    float* dxsum, const Reducer* red) {
  const int grid = std::min(cdiv(rows, 8), 148);
  float* part = reinterpret_cast<float*>(reduce_scratch(red, ln_bwd_scratch_bytes(C)));
  if (!part) return S4G_ERR_WORKSPACE;
  if (C == 768)
    S4G_CUDA(launch_dep(layernorm_bwd_kernel<T, 768>, dim3(grid), dim3(256), 0, st, dy, x, gamma, mean, rstd, dres, dx, part, rows, dx_cast));
  else if (C == 512)
    S4G_CUDA(launch_dep(layernorm_bwd_kernel<T, 512>, dim3(grid), dim3(256), 0, st, dy, x, gamma, mean, rstd, dres, dx, part, rows, dx_cast));
  else { s4g_set_error("layernorm: C must be 768 or 512"); return S4G_ERR_SHAPE; }
  S4G_KERNEL_CHECK();
  // three outputs of one partial matrix [grid][3*C]; queued jobs of one producer stay together (library scratch: the
  // three combines are launched one after the other before anything else touches the scratch)
  if (red && red->scratch && red->defer) {
    S4G_TRY(k_finalize(red, part, dgamma, grid, C, 3 * C, 0, 0, st));
    S4G_TRY(k_finalize(red, part + C, dbeta, grid, C, 3 * C, 0, 0, st));
    if (dxsum) S4G_TRY(k_finalize(red, part + 2 * C, dxsum, grid, C, 3 * C, 0, 0, st));
    return 0;
  }
  FinalizeJobs q; q.n = 0; q.total_blocks = 0; q.launches = 0;
  Reducer rq; rq.scratch = part; rq.scratch_bytes = ln_bwd_scratch_bytes(C); rq.defer = &q;
  S4G_TRY(k_finalize(&rq, part, dgamma, grid, C, 3 * C, 0, 0, st));
  S4G_TRY(k_finalize(&rq, part + C, dbeta, grid, C, 3 * C, 0, 0, st));
  if (dxsum) S4G_TRY(k_finalize(&rq, part + 2 * C, dxsum, grid, C, 3 * C, 0, 0, st));
  return k_finalize_flush(&q, st);
}

// ------------------------------------------------------------------------------------------ small elementwise
// out = dy * gelu'(u)   (bf16: A&S erf, error far below bf16 resolution; fp32: erff)
template <typename T> __global__ void __launch_bounds__(256) dgelu_mul_kernel(const T* __restrict__ dy, const T* __restrict__ u, T* __restrict__ out, long long n8) {
  pdl_trigger();  // the next kernel of the stream, if launched programmatically dependent, may start its prologue now
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < n8; i0 += 2 * stride) {
    Vec8<T> a[2], b[2];
#pragma unroll
    for (int q = 0; q < 2; ++q)
      if (i0 + q * stride < n8) { a[q].load(dy + (i0 + q * stride) * 8); b[q].load(u + (i0 + q * stride) * 8); }
#pragma unroll
    for (int q = 0; q < 2; ++q)
      if (i0 + q * stride < n8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) a[q].v[k] *= sizeof(T) == 2 ? dgelu_fast(b[q].v[k]) : dgelu_f(b[q].v[k]);
        a[q].store(out + (i0 + q * stride) * 8);
      }
  }
}
// out = dy * gelu'(u) on an [M, N] matrix, plus colsum[c] += sum_rows out[:, c] (the bias gradient of the Linear that
// produced u): same block shape as colsum_kernel -- 32 column groups of 8 x 8 row lanes -- so the second pass over the
// 77 MB tensor that a separate column-sum launch would make disappears.
template <typename T>
__global__ void __launch_bounds__(256, 3) dgelu_mul_colsum_kernel(const T* __restrict__ dy, const T* __restrict__ u, T* __restrict__ out,
                                                               float* __restrict__ part, long long M, int N, long long rows_per_block) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  __shared__ float red[8][256 + 8];
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + cg * 8;
  const long long r0 = (long long)blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float s[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = 0.f;
  if (col < N) {
    for (long long rb = r0 + rl; rb < r1; rb += 16) {
      Vec8<T> a[2], b[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const long long r = min(rb + q * 8, r1 - 1);
        a[q].load(dy + r * N + col); b[q].load(u + r * N + col);
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const long long r = rb + q * 8;
        if (r < r1) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            a[q].v[k] *= sizeof(T) == 2 ? dgelu_fast(b[q].v[k]) : dgelu_f(b[q].v[k]);
            s[k] += to_f(from_f<T>(a[q].v[k]));   // the sum of what is stored (what a separate pass would read)
          }
          a[q].store(out + r * N + col);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[rl][cg * 8 + i] = s[i];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < N) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
    part[(size_t)blockIdx.y * N + c] = t;   // one partial row per row block; finalize (reduce.cu) sums them in order
  }
}
// colsum (optional, overwritten): [N] column sums of the result for an [n / N, N] matrix
template <typename T> int k_dgelu_mul(const T* dy, const T* u, T* out, long long n, cudaStream_t st, float* colsum, int N, const Reducer* red) {
  if (colsum && N % 8 == 0 && n % N == 0) {
    const long long M = n / N;
    // 3 resident blocks per SM (launch bound above): one wave of 148 * 3 blocks
    const int rblocks = (int)std::max<long long>(1, std::min<long long>(cdiv(M, 64), (148 * 3) / cdiv(N, 256)));
    float* part = reinterpret_cast<float*>(reduce_scratch(red, colsum_scratch_bytes(N)));
    if (!part) return S4G_ERR_WORKSPACE;
    S4G_CUDA(launch_dep(dgelu_mul_colsum_kernel<T>, dim3(dim3(cdiv(N, 256), rblocks)), dim3(256), 0, st, dy, u, out, part, M, N, cdiv(M, rblocks)));
    S4G_KERNEL_CHECK();
    return k_finalize(red, part, colsum, rblocks, N, N, 0, 0, st);
  } else {
    if (colsum) { s4g_set_error("dgelu_mul: bad column-sum shape"); return S4G_ERR_SHAPE; }
    dgelu_mul_kernel<T><<<GRID1D(n / 16, 256), 0, st>>>(dy, u, out, n / 8);
  }
  S4G_KERNEL_CHECK();
  return 0;
}
// a = gelu(u); gprime (optional, may alias u): gelu'(u) -- the backward then needs one multiply per element (EPI_MUL in the
// fc2 dgrad epilogue) instead of a pass that re-evaluates erf / exp over the 77 MB tensor.  Both come from one (cdf, pdf).
template <typename T> __global__ void __launch_bounds__(256) gelu_kernel(const T* u, T* __restrict__ out, T* gprime, long long n8) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < n8; i0 += 2 * stride) {
    Vec8<T> b[2], g[2];
#pragma unroll
    for (int q = 0; q < 2; ++q)
      if (i0 + q * stride < n8) b[q].load(u + (i0 + q * stride) * 8);
#pragma unroll
    for (int q = 0; q < 2; ++q)
      if (i0 + q * stride < n8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float x = b[q].v[k];
          if (sizeof(T) == 2) {
            float c, pdf;
            gelu_parts_fast(x, c, pdf);
            b[q].v[k] = x * c;
            g[q].v[k] = fmaf(x, pdf, c);
          } else {
            b[q].v[k] = gelu_f(x);
            g[q].v[k] = dgelu_f(x);
          }
        }
        b[q].store(out + (i0 + q * stride) * 8);
        if (gprime) g[q].store(gprime + (i0 + q * stride) * 8);
      }
  }
}
template <typename T> int k_gelu(const T* u, T* out, long long n, cudaStream_t st, T* gprime) {
  S4G_CUDA(launch_dep(gelu_kernel<T>, GRID1D(n / 16, 256), 0, st, u, out, gprime, n / 8));
  S4G_KERNEL_CHECK();
  return 0;
}

// column sums of a [M, N] matrix (bias gradients).  Block = 32 column groups (8 columns, one 16 B load each) x 8 row
// lanes, rows walked with a 4-deep unroll; one partial row per row block, summed in order by finalize (reduce.cu).
template <typename T> __global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ dy, float* __restrict__ part, long long M, int N, int ld, long long rows_per_block) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  __shared__ float red[8][256 + 8];
  const int cg = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + cg * 8;
  const long long r0 = (long long)blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float s[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s[i] = 0.f;
  if (col < N) {
    for (long long rb = r0 + rl; rb < r1; rb += 32) {
      Vec8<T> v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const long long r = rb + u * 8;
        if (r < r1) v[u].load(dy + r * ld + col);
        else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[u].v[i] = 0.f;
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] += v[u].v[i];
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red[rl][cg * 8 + i] = s[i];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c < N) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
    part[(size_t)blockIdx.y * N + c] = t;
  }
}
template <typename T> int k_colsum(const T* dy, float* out, long long M, int N, int ld, cudaStream_t st, const Reducer* red) {
  if (N % 8 || ld % 8) { s4g_set_error("colsum: N and ld must be multiples of 8"); return S4G_ERR_SHAPE; }
  const int gx = cdiv(N, 256);
  const long long want = std::max<long long>(1, (148 * 4) / gx);
  const long long rpb = std::max<long long>(64, (M + want - 1) / want);
  dim3 grid(gx, cdiv(M, rpb));
  float* part = reinterpret_cast<float*>(reduce_scratch(red, colsum_scratch_bytes(N)));
  if (!part) return S4G_ERR_WORKSPACE;
  S4G_CUDA(launch_dep(colsum_kernel<T>, dim3(grid), dim3(256), 0, st, dy, part, M, N, ld, rpb));
  S4G_KERNEL_CHECK();
  return k_finalize(red, part, out, (int)grid.y, N, N, 0, 0, st);
}

// Slice(1)+Transpose+Unflatten (Models/DPT_decoder.py:503-515): token rows 1..196 ARE the NHWC map
template <typename T> __global__ void tokens_to_nhwc_kernel(const float* __restrict__ x, T* __restrict__ out, int B) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const long long n = (long long)B * 196 * 96;
  FOR_GRID(t, n) {
    const int g = (int)(t % 96);
    const long long pidx = t / 96;
    const int p = (int)(pidx % 196), b = (int)(pidx / 196);
    Vec8<float> v;
    v.load(x + ((long long)b * 197 + 1 + p) * 768 + g * 8);
    Vec8<T> o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.v[k] = v.v[k];
    o.store(out + pidx * 768 + g * 8);
  }
}
template <typename T> int k_tokens_to_nhwc(const float* x, T* out, int B, cudaStream_t st) {
  S4G_CUDA(launch_dep(tokens_to_nhwc_kernel<T>, GRID1D((long long)B * 196 * 96, 256), 0, st, x, out, B));
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> __global__ void tap_grad_kernel(const T* __restrict__ dtap, float* __restrict__ dx, int B, int accumulate, T* __restrict__ cast_out) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const long long n = (long long)B * 197 * 96;
  FOR_GRID(t, n) {
    const int g = (int)(t % 96);
    const long long row = t / 96;
    const int tok = (int)(row % 197), b = (int)(row / 197);
    float* d = dx + row * 768 + g * 8;
    Vec8<float> o;
    if (accumulate) o.load(d);
    else {
#pragma unroll
      for (int k = 0; k < 8; ++k) o.v[k] = 0.f;
    }
    if (tok > 0) {
      Vec8<T> v;
      v.load(dtap + ((long long)b * 196 + tok - 1) * 768 + g * 8);
#pragma unroll
      for (int k = 0; k < 8; ++k) o.v[k] += v.v[k];
    } else if (accumulate && !cast_out) {
      continue;
    }
    o.store(d);
    if (cast_out) {
      Vec8<T> c;
#pragma unroll
      for (int k = 0; k < 8; ++k) c.v[k] = o.v[k];
      c.store(cast_out + row * 768 + g * 8);
    }
  }
}
template <typename T> int k_tap_grad(const T* dtap, float* dx, int B, int accumulate, cudaStream_t st, T* cast_out) {
  S4G_CUDA(launch_dep(tap_grad_kernel<T>, GRID1D((long long)B * 197 * 96, 256), 0, st, dtap, dx, B, accumulate, cast_out));
  S4G_KERNEL_CHECK();
  return 0;
}

template <typename T> __global__ void cast_f32_kernel(const float* __restrict__ in, T* __restrict__ out, long long n) {
  FOR_GRID(i, n) out[i] = from_f<T>(in[i]);
}
template <typename T> int k_cast_f32(const float* in, T* out, long long n, cudaStream_t st) {
  cast_f32_kernel<T><<<GRID1D(n, 256), 0, st>>>(in, out, n);
  S4G_KERNEL_CHECK();
  return 0;
}

template <typename T> __global__ void relu_kernel(const T* __restrict__ x, T* __restrict__ y, long long n8) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  FOR_GRID(i, n8) {
    Vec8<T> a;
    a.load(x + i * 8);
#pragma unroll
    for (int k = 0; k < 8; ++k) a.v[k] = fmaxf(a.v[k], 0.f);
    a.store(y + i * 8);
  }
}
template <typename T> int k_relu(const T* x, T* y, long long n, cudaStream_t st) {
  S4G_CUDA(launch_dep(relu_kernel<T>, GRID1D(n / 8, 256), 0, st, x, y, n / 8));
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> __global__ void relu_bwd_add_kernel(const T* __restrict__ dr, const T* __restrict__ x, const T* __restrict__ dskip, T* __restrict__ dx, long long n8) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  FOR_GRID(i, n8) {
    Vec8<T> a, b, c;
    a.load(dr + i * 8); b.load(x + i * 8);
    if (dskip) c.load(dskip + i * 8);
#pragma unroll
    for (int k = 0; k < 8; ++k) a.v[k] = (b.v[k] > 0.f ? a.v[k] : 0.f) + (dskip ? c.v[k] : 0.f);
    a.store(dx + i * 8);
  }
}
template <typename T> int k_relu_bwd_add(const T* dr, const T* x, const T* dskip, T* dx, long long n, cudaStream_t st) {
  S4G_CUDA(launch_dep(relu_bwd_add_kernel<T>, GRID1D(n / 8, 256), 0, st, dr, x, dskip, dx, n / 8));
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> __global__ void add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y, long long n8) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  FOR_GRID(i, n8) {
    Vec8<T> u, v;
    u.load(a + i * 8); v.load(b + i * 8);
#pragma unroll
    for (int k = 0; k < 8; ++k) u.v[k] += v.v[k];
    u.store(y + i * 8);
  }
}
template <typename T> int k_add(const T* a, const T* b, T* y, long long n, cudaStream_t st) {
  S4G_CUDA(launch_dep(add_kernel<T>, GRID1D(n / 8, 256), 0, st, a, b, y, n / 8));
  S4G_KERNEL_CHECK();
  return 0;
}

// ConvTranspose2d with kernel == stride (Models/DPT_decoder.py:341-350,365-374) = GEMM to
// ycol[p][(kh,kw,co)] then this scatter: out[b, h*k+kh, w*k+kw, co] = ycol + bias[co]
template <typename T> __global__ void pixel_shuffle_kernel(const T* __restrict__ ycol, const float* __restrict__ bias, T* __restrict__ out, int B, int H, int W, int C, int k, int unshuffle) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int c8 = C / 8;
  const long long n = (long long)B * H * k * W * k * c8;
  FOR_GRID(t, n) {
    const int g = (int)(t % c8);
    long long r = t / c8;
    const int ow = (int)(r % (W * k)); r /= (W * k);
    const int oh = (int)(r % (H * k));
    const int b = (int)(r / (H * k));
    const int h = oh / k, kh = oh % k, w = ow / k, kw = ow % k;
    const long long src = ((((long long)b * H + h) * W + w) * (k * k) + kh * k + kw) * C + g * 8;
    const long long dst = (((long long)b * H * k + oh) * (W * k) + ow) * C + g * 8;
    Vec8<T> v;
    if (unshuffle) { v.load(out + dst); v.store(const_cast<T*>(ycol) + src); }
    else {
      v.load(ycol + src);
#pragma unroll
      for (int i = 0; i < 8; ++i) v.v[i] += bias[g * 8 + i];
      v.store(out + dst);
    }
  }
}
template <typename T> int k_pixel_shuffle(const T* ycol, const float* bias, T* out, int B, int H, int W, int C, int k, cudaStream_t st) {
  const long long n = (long long)B * H * k * W * k * (C / 8);
  S4G_CUDA(launch_dep(pixel_shuffle_kernel<T>, GRID1D(n, 256), 0, st, ycol, bias, out, B, H, W, C, k, 0));
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> int k_pixel_unshuffle(const T* dout, T* dycol, int B, int H, int W, int C, int k, cudaStream_t st) {
  const long long n = (long long)B * H * k * W * k * (C / 8);
  S4G_CUDA(launch_dep(pixel_shuffle_kernel<T>, GRID1D(n, 256), 0, st, dycol, nullptr, const_cast<T*>(dout), B, H, W, C, k, 1));
  S4G_KERNEL_CHECK();
  return 0;
}

// Conv2d 3x3 stride 2 pad 1 (Models/DPT_decoder.py:403-409) as explicit im2col + GEMM (only one small layer)
template <typename T> __global__ void im2col_s2_kernel(const T* __restrict__ x, T* __restrict__ col, int B, int H, int W, int C) {
  pdl_trigger();  // the next kernel of the stream, if launched programmatically dependent, may start its prologue now
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2, c8 = C / 8;
  const long long n = (long long)B * Ho * Wo * 9 * c8;
  FOR_GRID(t, n) {
    const int g = (int)(t % c8);
    long long r = t / c8;
    const int tap = (int)(r % 9); r /= 9;
    const int wo = (int)(r % Wo); r /= Wo;
    const int ho = (int)(r % Ho);
    const int b = (int)(r / Ho);
    const int hi = ho * 2 + tap / 3 - 1, wi = wo * 2 + tap % 3 - 1;
    Vec8<T> v;
    if (hi >= 0 && hi < H && wi >= 0 && wi < W) v.load(x + (((long long)b * H + hi) * W + wi) * C + g * 8);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v.v[i] = 0.f;
    }
    v.store(col + t * 8);
  }
}
template <typename T> int k_im2col_s2(const T* x, T* col, int B, int H, int W, int C, cudaStream_t st) {
  const long long n = (long long)B * ((H + 1) / 2) * ((W + 1) / 2) * 9 * (C / 8);
  im2col_s2_kernel<T><<<GRID1D(n, 256), 0, st>>>(x, col, B, H, W, C);
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> __global__ void col2im_s2_kernel(const T* __restrict__ dcol, T* __restrict__ dx, int B, int H, int W, int C) {
  pdl_trigger();  // the next kernel of the stream, if launched programmatically dependent, may start its prologue now
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2, c8 = C / 8;
  const long long n = (long long)B * H * W * c8;
  FOR_GRID(t, n) {
    const int g = (int)(t % c8);
    long long r = t / c8;
    const int wi = (int)(r % W); r /= W;
    const int hi = (int)(r % H);
    const int b = (int)(r / H);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    for (int kr = 0; kr < 3; ++kr) {
      const int hn = hi + 1 - kr;
      if (hn < 0 || (hn & 1) || hn / 2 >= Ho) continue;
      for (int ks = 0; ks < 3; ++ks) {
        const int wn = wi + 1 - ks;
        if (wn < 0 || (wn & 1) || wn / 2 >= Wo) continue;
        Vec8<T> v;
        v.load(dcol + ((((long long)b * Ho + hn / 2) * Wo + wn / 2) * 9 + kr * 3 + ks) * C + g * 8);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += v.v[i];
      }
    }
    Vec8<T> o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = acc[i];
    o.store(dx + t * 8);
  }
}
template <typename T> int k_col2im_s2(const T* dcol, T* dx, int B, int H, int W, int C, cudaStream_t st) {
  col2im_s2_kernel<T><<<GRID1D((long long)B * H * W * (C / 8), 256), 0, st>>>(dcol, dx, B, H, W, C);
  S4G_KERNEL_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------ bilinear x2
// F.interpolate(scale_factor=2, mode="bilinear", align_corners=True) (Models/DPT_decoder.py:295-297,93-98):
// src = dst * (H-1)/(2H-1); ATen computes src, floor index and lambda in fp32 — mirrored here.
__device__ __forceinline__ void bil_src(int o, int in_size, float scale, int& i0, int& i1, float& l0, float& l1) {
  const float s = scale * o;
  i0 = (int)s;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = s - i0;
  l0 = 1.f - l1;
}
// A thread owns (b, output row oh, 8 channels) and walks the output columns: the two source rows are blended once per SOURCE column
// (c(w) = lh0 x[h0, w] + lh1 x[h1, w]) and every output column is lw0 c(w0) + lw1 c(w0 + 1); w0 never decreases, so each x element
// is loaded once per output row (2 loads per 2 output columns instead of 8) and the kernel is bound by its 16 B stores.
template <typename T> __global__ void __launch_bounds__(256) bilinear2x_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int B, int H, int W, int C) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int c8 = C / 8, Ho = 2 * H, Wo = 2 * W;
  const float sh = (float)(H - 1) / (float)(Ho - 1), sw = (float)(W - 1) / (float)(Wo - 1);
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)B * Ho * c8) return;
  const int g = (int)(idx % c8), oh = (int)((idx / c8) % Ho), b = (int)(idx / ((long long)c8 * Ho));
  int h0, h1; float lh0, lh1;
  bil_src(oh, H, sh, h0, h1, lh0, lh1);
  const T* r0 = x + (((size_t)b * H + h0) * W) * C + g * 8;
  const T* r1 = x + (((size_t)b * H + h1) * W) * C + g * 8;
  T* out = y + (((size_t)b * Ho + oh) * Wo) * C + g * 8;
  auto column = [&](int w, float (&c)[8]) {
    Raw8<T> a, bq;
    a.load(r0 + (size_t)w * C); bq.load(r1 + (size_t)w * C);
    float fa[8], fb[8];
    a.get(fa); bq.get(fb);
#pragma unroll
    for (int i = 0; i < 8; ++i) c[i] = lh0 * fa[i] + lh1 * fb[i];
  };
  float c0[8], c1[8];
  column(0, c0);
  column(min(1, W - 1), c1);
  int cur = 0;
  for (int ow = 0; ow < Wo; ++ow) {
    int w0, w1; float lw0, lw1;
    bil_src(ow, W, sw, w0, w1, lw0, lw1);
    if (w0 > cur) {
      ++cur;
#pragma unroll
      for (int i = 0; i < 8; ++i) c0[i] = c1[i];
      column(min(cur + 1, W - 1), c1);
    }
    Vec8<T> o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = lw0 * c0[i] + lw1 * (w1 == w0 ? c0[i] : c1[i]);
    o.store(out + (size_t)ow * C);
  }
}
template <typename T> int k_bilinear2x_fwd(const T* x, T* y, int B, int H, int W, int C, cudaStream_t st) {
  S4G_CUDA(launch_dep(bilinear2x_fwd_kernel<T>, dim3(cdiv((long long)B * 2 * H * (C / 8), 256)), dim3(256), 0, st, x, y, B, H, W, C));
  S4G_KERNEL_CHECK();
  return 0;
}
// gather form of the adjoint (deterministic, no atomics): dx[i] = sum_o w(o,i) dy[o]
__device__ __forceinline__ float bil_w(int o, int i, int in_size, float scale) {
  int i0, i1; float l0, l1;
  bil_src(o, in_size, scale, i0, i1, l0, l1);
  return (i0 == i ? l0 : 0.f) + (i1 == i ? l1 : 0.f);
}
__device__ __forceinline__ void bil_range(int i, int in_size, int& lo, int& hi) {
  // outputs o with floor(o*s) in {i-1, i}: o in ((i-1)/s, (i+1)/s); s = (in-1)/(2in-1) ~ 0.5 => |o - 2i| <= 3
  lo = max(0, 2 * i - 3);
  hi = min(2 * in_size - 1, 2 * i + 3);
}
// the 7 candidate outputs 2i-3 .. 2i+3 of input i with their weights (0 for non-contributing / out of range)
__device__ __forceinline__ void bil_taps(int i, int in_size, float scale, float (&wt)[7]) {
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    const int o = 2 * i - 3 + k;
    wt[k] = (o >= 0 && o < 2 * in_size) ? bil_w(o, i, in_size, scale) : 0.f;
  }
}
// A thread owns (b, input row h, 8 channels) and walks the output columns ow = 0 .. 2W-1 once: the (up to 5 of 7 candidate) output
// rows that touch h are combined vertically per column, and the column's value is added to the two input columns w0(ow), w0 + 1 it
// interpolates between -- w0 never decreases, so two running accumulators suffice and each dx element is written exactly once.
// Every dy element is loaded by the two input rows it belongs to (the second time from cache); the earlier gather form loaded it
// ~16 times (7 x 7 candidate window per input pixel) and ran at L1 / L2 speed: 220 us for the 112 x 112 map against ~100 us of
// HBM time.  Deterministic (fixed order, no atomics).
template <typename T> __global__ void __launch_bounds__(256) bilinear2x_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int B, int H, int W, int C) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int c8 = C / 8, Ho = 2 * H, Wo = 2 * W;
  const float sh = (float)(H - 1) / (float)(Ho - 1), sw = (float)(W - 1) / (float)(Wo - 1);
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)B * H * c8) return;
  const int g = (int)(idx % c8), h = (int)((idx / c8) % H), b = (int)(idx / ((long long)c8 * H));
  float wh[7];
  bil_taps(h, H, sh, wh);
  const T* base = dy + ((size_t)b * Ho * Wo) * C + g * 8;
  T* out = dx + (((size_t)b * H + h) * W) * C + g * 8;
  float a0[8], a1[8];   // accumulators of input columns cur and cur + 1
#pragma unroll
  for (int i = 0; i < 8; ++i) a0[i] = a1[i] = 0.f;
  int cur = 0;
  for (int ow0 = 0; ow0 < Wo; ow0 += 2) {   // Wo is even
    Raw8<T> v[2][7];
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int kh = 0; kh < 7; ++kh)
        if (wh[kh] != 0.f) v[u][kh].load(base + ((size_t)(2 * h - 3 + kh) * Wo + ow0 + u) * C);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      float col[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) col[i] = 0.f;
#pragma unroll
      for (int kh = 0; kh < 7; ++kh) {
        if (wh[kh] != 0.f) {
          float f[8];
          v[u][kh].get(f);
#pragma unroll
          for (int i = 0; i < 8; ++i) col[i] = fmaf(wh[kh], f[i], col[i]);
        }
      }
      int w0, w1; float l0, l1;
      bil_src(ow0 + u, W, sw, w0, w1, l0, l1);
      if (w0 > cur) {   // column `cur` is complete
        Vec8<T> o;
#pragma unroll
        for (int i = 0; i < 8; ++i) { o.v[i] = a0[i]; a0[i] = a1[i]; a1[i] = 0.f; }
        o.store(out + (size_t)cur * C);
        ++cur;
      }
      if (w1 == w0) l0 += l1;   // last input column: both taps are the same element
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a0[i] = fmaf(l0, col[i], a0[i]);
        if (w1 != w0) a1[i] = fmaf(l1, col[i], a1[i]);
      }
    }
  }
  Vec8<T> o;
#pragma unroll
  for (int i = 0; i < 8; ++i) o.v[i] = a0[i];
  o.store(out + (size_t)cur * C);
  if (cur + 1 < W) {
#pragma unroll
    for (int i = 0; i < 8; ++i) o.v[i] = a1[i];
    o.store(out + (size_t)(cur + 1) * C);
  }
}
template <typename T> int k_bilinear2x_bwd(const T* dy, T* dx, int B, int H, int W, int C, cudaStream_t st) {
  S4G_CUDA(launch_dep(bilinear2x_bwd_kernel<T>, dim3(cdiv((long long)B * H * (C / 8), 256)), dim3(256), 0, st, dy, dx, B, H, W, C));
  S4G_KERNEL_CHECK();
  return 0;
}
__global__ void bilinear2x_f32_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int B, int H, int W) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int Ho = 2 * H, Wo = 2 * W;
  const float sh = (float)(H - 1) / (float)(Ho - 1), sw = (float)(W - 1) / (float)(Wo - 1);
  const long long n = (long long)B * Ho * Wo;
  FOR_GRID(t, n) {
    const int ow = (int)(t % Wo), oh = (int)((t / Wo) % Ho), b = (int)(t / ((long long)Wo * Ho));
    int h0, h1, w0, w1; float lh0, lh1, lw0, lw1;
    bil_src(oh, H, sh, h0, h1, lh0, lh1);
    bil_src(ow, W, sw, w0, w1, lw0, lw1);
    const float* p = x + (long long)b * H * W;
    y[t] = lh0 * (lw0 * p[h0 * W + w0] + lw1 * p[h0 * W + w1]) + lh1 * (lw0 * p[h1 * W + w0] + lw1 * p[h1 * W + w1]);
  }
}
int k_bilinear2x_f32_fwd(const float* x, float* y, int B, int H, int W, cudaStream_t st) {
  S4G_CUDA(launch_dep(bilinear2x_f32_fwd_kernel, GRID1D((long long)B * 4 * H * W, 256), 0, st, x, y, B, H, W));
  S4G_KERNEL_CHECK();
  return 0;
}
__global__ void bilinear2x_f32_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int B, int H, int W) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int Ho = 2 * H, Wo = 2 * W;
  const float sh = (float)(H - 1) / (float)(Ho - 1), sw = (float)(W - 1) / (float)(Wo - 1);
  const long long n = (long long)B * H * W;
  FOR_GRID(t, n) {
    const int w = (int)(t % W), h = (int)((t / W) % H), b = (int)(t / ((long long)W * H));
    int hlo, hhi, wlo, whi;
    bil_range(h, H, hlo, hhi);
    bil_range(w, W, wlo, whi);
    float acc = 0.f;
    const float* p = dy + (long long)b * Ho * Wo;
    for (int oh = hlo; oh <= hhi; ++oh) {
      const float wh = bil_w(oh, h, H, sh);
      if (wh == 0.f) continue;
      for (int ow = wlo; ow <= whi; ++ow) acc += wh * bil_w(ow, w, W, sw) * p[oh * Wo + ow];
    }
    dx[t] = acc;
  }
}
int k_bilinear2x_f32_bwd(const float* dy, float* dx, int B, int H, int W, cudaStream_t st) {
  S4G_CUDA(launch_dep(bilinear2x_f32_bwd_kernel, GRID1D((long long)B * H * W, 256), 0, st, dy, dx, B, H, W));
  S4G_KERNEL_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------ BatchNorm2d (train)
// nn.BatchNorm2d(256) eps 1e-5 momentum 0.1 (Models/DPT_decoder.py:204-206,492). x is [M, C] NHWC, C <= 256.
// Block = 256 threads = (C/8) channel groups x (2048/C) row lanes.
template <typename T, bool BWD>
__global__ void __launch_bounds__(256) bn_reduce_kernel(const T* __restrict__ a, const T* __restrict__ x,
                                                        const float* __restrict__ mean, const float* __restrict__ rstd,
                                                        const float* __restrict__ gamma, const float* __restrict__ beta,
                                                        double* __restrict__ sums, int relu, long long M, int C,
                                                        long long rows_per_block) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  // FWD: a = x, sums = {sum x, sum x^2}.  BWD: a = dy, sums = {sum dy', sum dy' * xhat}
  extern __shared__ float sm[];  // [2][lanes][C]
  const int c8 = C / 8, lanes = 256 / c8;
  const int g = threadIdx.x % c8, rl = threadIdx.x / c8;
  const long long r0 = (long long)blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float s1[8], s2[8], mu[8], rs[8], gm[8], bt[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    s1[i] = 0.f; s2[i] = 0.f;
    if (BWD) { mu[i] = mean[g * 8 + i]; rs[i] = rstd[g * 8 + i]; gm[i] = gamma[g * 8 + i]; bt[i] = beta[g * 8 + i]; }
  }
  for (long long rb = r0 + rl; rb < r1; rb += 4 * lanes) {
    Vec8<T> v[4], xv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long r = rb + (long long)u * lanes;
      if (r < r1) {
        v[u].load(a + r * C + g * 8);
        if (BWD) xv[u].load(x + r * C + g * 8);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) { v[u].v[i] = 0.f; xv[u].v[i] = 0.f; }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (BWD) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float xh = (xv[u].v[i] - mu[i]) * rs[i];
          float d = v[u].v[i];
          if (relu && xh * gm[i] + bt[i] <= 0.f) d = 0.f;
          s1[i] += d; s2[i] += d * xh;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) { s1[i] += v[u].v[i]; s2[i] += v[u].v[i] * v[u].v[i]; }
      }
    }
  }
  float* p1 = sm + (size_t)rl * C + g * 8;
  float* p2 = sm + (size_t)(lanes + rl) * C + g * 8;
#pragma unroll
  for (int i = 0; i < 8; ++i) { p1[i] = s1[i]; p2[i] = s2[i]; }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * C; c += 256) {
    const int which = c / C, ch = c % C;
    double s = 0.0;
    for (int l = 0; l < lanes; ++l) s += (double)sm[(size_t)(which * lanes + l) * C + ch];
    sums[(size_t)blockIdx.x * 2 * C + which * C + ch] = s;   // partial row of this block (finalize sums the rows in order)
  }
}
// backward sums {sum dy', sum dy' * xhat}: a thread accumulates sum dy' and sum dy' * x for its 8 channels (only the
// ReLU-mask constants live in registers, loads stay in their raw 16 B form, 4 in flight per operand);
// xhat = (x - mean) * rstd is applied once per block in double.
template <typename T>
__global__ void __launch_bounds__(256, 2) bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                               const float* __restrict__ mean, const float* __restrict__ rstd,
                                                               const float* __restrict__ gamma, const float* __restrict__ beta,
                                                               double* __restrict__ sums, int relu, long long M, int C,
                                                               long long rows_per_block) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  extern __shared__ float sm[];  // [2][lanes][C]
  const int c8 = C / 8, lanes = 256 / c8;
  const int g = threadIdx.x % c8, rl = threadIdx.x / c8;
  const long long r0 = (long long)blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float s1[8], s2[8], sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    s1[i] = 0.f; s2[i] = 0.f;
    const int c = g * 8 + i;
    sc[i] = relu ? rstd[c] * gamma[c] : 0.f;
    sh[i] = relu ? beta[c] - mean[c] * sc[i] : 1.f;   // relu == 0: the mask test below is always false
  }
  const T* pd = dy + g * 8;
  const T* px = x + g * 8;
  for (long long rb = r0 + rl; rb < r1; rb += 4 * lanes) {
    Raw8<T> d[4], xv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long r = min(rb + (long long)u * lanes, r1 - 1);   // clamped; masked out below
      d[u].load(pd + r * C);
      xv[u].load(px + r * C);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (rb + (long long)u * lanes < r1) {
        float a[8], b[8];
        d[u].get(a); xv[u].get(b);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float dd = fmaf(b[i], sc[i], sh[i]) <= 0.f ? 0.f : a[i];
          s1[i] += dd; s2[i] = fmaf(dd, b[i], s2[i]);
        }
      }
    }
  }
  float* p1 = sm + (size_t)rl * C + g * 8;
  float* p2 = sm + (size_t)(lanes + rl) * C + g * 8;
#pragma unroll
  for (int i = 0; i < 8; ++i) { p1[i] = s1[i]; p2[i] = s2[i]; }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += 256) {
    double a = 0.0, b = 0.0;
    for (int l = 0; l < lanes; ++l) { a += (double)sm[(size_t)l * C + c]; b += (double)sm[(size_t)(lanes + l) * C + c]; }
    double* prow = sums + (size_t)blockIdx.x * 2 * C;   // partial row of this block
    prow[c] = a;
    prow[C + c] = (b - (double)mean[c] * a) * (double)rstd[c];
  }
}
static inline long long bn_rows_per_block(long long M) { return max((long long)32, (M + 148 * 4 - 1) / (148 * 4)); }
template <typename T> int k_bn_stats(const T* x, double* stats, long long M, int C, cudaStream_t st, const Reducer* red) {
  if (C % 8 || 256 % (C / 8)) { s4g_set_error("bn: unsupported C"); return S4G_ERR_SHAPE; }
  const long long rpb = bn_rows_per_block(M);
  const int nb = cdiv(M, rpb);
  double* part = reinterpret_cast<double*>(reduce_scratch(red, bn_scratch_bytes(C)));
  if (!part) return S4G_ERR_WORKSPACE;
  S4G_CUDA(launch_dep(bn_reduce_kernel<T, false>, dim3(nb), dim3(256), 2 * 2048 * sizeof(float), st, x, nullptr, nullptr, nullptr, nullptr, nullptr, part, 0, M, C, rpb));
  S4G_KERNEL_CHECK();
  return k_finalize(red, part, stats, nb, 2 * C, 2 * C, 1, 1, st);
}
template <typename T>
int k_bn_bwd_reduce(const T* dy, const T* x, const float* mean, const float* rstd, const float* gamma,
                    const float* beta, double* sums, int relu, long long M, int C, cudaStream_t st, const Reducer* red) {
  if (C % 8 || 256 % (C / 8)) { s4g_set_error("bn: unsupported C"); return S4G_ERR_SHAPE; }
  const long long rpb = bn_rows_per_block(M);
  const int nb = cdiv(M, rpb);
  double* part = reinterpret_cast<double*>(reduce_scratch(red, bn_scratch_bytes(C)));
  if (!part) return S4G_ERR_WORKSPACE;
  S4G_CUDA(launch_dep(bn_bwd_reduce_kernel<T>, dim3(nb), dim3(256), 2 * 2048 * sizeof(float), st, dy, x, mean, rstd, gamma, beta, part, relu, M, C, rpb));
  S4G_KERNEL_CHECK();
  return k_finalize(red, part, sums, nb, 2 * C, 2 * C, 1, 1, st);
}
__global__ void bn_finalize_kernel(const double* __restrict__ stats, double count, float* mean, float* rstd,
                                   float* running_mean, float* running_var, long long* nbt, float momentum, float eps, int C) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double mu = stats[c] / count;
  double var = stats[C + c] / count - mu * mu;
  var = var < 0.0 ? 0.0 : var;
  mean[c] = (float)mu;
  rstd[c] = (float)(1.0 / sqrt(var + (double)eps));
  if (running_mean) {
    const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mu;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    if (c == 0 && nbt) *nbt += 1;
  }
}
int k_bn_finalize(const double* stats, double count, float* mean, float* rstd, float* running_mean, float* running_var,
                  long long* nbt, float momentum, float eps, int C, cudaStream_t st) {
  S4G_CUDA(launch_dep(bn_finalize_kernel, dim3(cdiv(C, 256)), dim3(256), 0, st, stats, count, mean, rstd, running_mean, running_var, nbt, momentum, eps, C));
  S4G_KERNEL_CHECK();
  return 0;
}
__global__ void bn_eval_prepare_kernel(const float* rm, const float* rv, float* mean, float* rstd, float eps, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) { mean[c] = rm[c]; rstd[c] = rsqrtf(rv[c] + eps); }
}
int k_bn_eval_prepare(const float* rm, const float* rv, float* mean, float* rstd, float eps, int C, cudaStream_t st) {
  bn_eval_prepare_kernel<<<cdiv(C, 256), 256, 0, st>>>(rm, rv, mean, rstd, eps, C);
  S4G_KERNEL_CHECK();
  return 0;
}
// Block = 256 threads = (C/8) channel groups x (2048/C) row lanes; a thread keeps its 8 channels' constants in
// registers and walks rows with a 4-deep unroll (4 x 16 B loads in flight per operand).
template <typename T>
__global__ void __launch_bounds__(256, 3) bn_apply_kernel(const T* __restrict__ x, const float* __restrict__ mean,
                                                          const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, const T* __restrict__ skip,
                                                          T* __restrict__ y, int relu, long long M, int C) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int c8 = C / 8, lanes = 256 / c8;
  const int g = threadIdx.x % c8, rl = threadIdx.x / c8;
  float sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = g * 8 + i;
    sc[i] = rstd[c] * gamma[c];
    sh[i] = beta[c] - mean[c] * sc[i];
  }
  // loads stay packed (Raw8: 4 registers per 16 B for bf16) until they are used, so that three blocks fit on an SM with 4 (+ 4
  // with a skip operand) loads in flight per thread: the fp32-unpacked version held 2 blocks per SM and ran at 2.3-3.2 TB/s
  const long long stride = (long long)gridDim.x * lanes;
  const T* px = x + g * 8;
  const T* ps = skip ? skip + g * 8 : nullptr;
  T* py = y + g * 8;
  for (long long rb = (long long)blockIdx.x * lanes + rl; rb < M; rb += 4 * stride) {
    Raw8<T> v[4], sk[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long r = min(rb + u * stride, M - 1);
      v[u].load(px + r * C);
      if (ps) sk[u].load(ps + r * C);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long r = rb + u * stride;
      if (r < M) {
        float a[8], b[8];
        v[u].get(a);
        if (ps) sk[u].get(b);
        Vec8<T> o;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float t = fmaf(a[i], sc[i], sh[i]);
          if (relu) t = fmaxf(t, 0.f);
          if (ps) t += b[i];
          o.v[i] = t;
        }
        o.store(py + r * C);
      }
    }
  }
}
static inline unsigned bn_grid(long long M, int C, int per_sm = 8) {
  const int lanes = 256 / (C / 8);
  return (unsigned)std::min<long long>(cdiv(M, (long long)lanes * 4), 148 * per_sm);   // grid-stride beyond
}
template <typename T>
int k_bn_apply(const T* x, const float* mean, const float* rstd, const float* gamma, const float* beta, const T* skip,
               T* y, int relu, long long M, int C, cudaStream_t st) {
  if (C % 8 || 256 % (C / 8)) { s4g_set_error("bn: unsupported C"); return S4G_ERR_SHAPE; }
  S4G_CUDA(launch_dep(bn_apply_kernel<T>, dim3(bn_grid(M, C, 6)), dim3(256), 0, st, x, mean, rstd, gamma, beta, skip, y, relu, M, C));
  S4G_KERNEL_CHECK();
  return 0;
}
// dx = gamma*rstd*(dy' - sum1/count - xhat*sum2/count) = sc*dy' - (c0 + x*c1) with per-channel sc = gamma*rstd,
// c1 = sc*rstd*sum2/count, c0 = sc*sum1/count - mean*c1; block 0 also writes dgamma = sum2, dbeta = sum1.
// The ReLU mask uses the same fmaf(x, sc, sh) as the forward bn_apply_kernel.
template <typename T>
__global__ void __launch_bounds__(256, 2) bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                              const float* __restrict__ mean, const float* __restrict__ rstd,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              const double* __restrict__ sums, double count,
                                                              T* __restrict__ dx, float* dgamma, float* dbeta, int relu,
                                                              long long M, int C) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  const int c8 = C / 8, lanes = 256 / c8;
  const int g = threadIdx.x % c8, rl = threadIdx.x / c8;
  // SyncBatchNorm (count = M * world): torch keeps grad_weight / grad_bias LOCAL and DDP averages them over ranks; the
  // sums here are already global, so hand the gradient all-reduce (sum, then 1/world) the global sum / world.
  if (blockIdx.x == 0 && dgamma) {
    const double gs = (double)M / count;
    for (int c = threadIdx.x; c < C; c += blockDim.x) { dbeta[c] = (float)(sums[c] * gs); dgamma[c] = (float)(sums[C + c] * gs); }
  }
  const float inv = (float)(1.0 / count);
  float sc[8], sh[8], c0[8], c1[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = g * 8 + i;
    const float mu = mean[c], rs = rstd[c];
    sc[i] = gamma[c] * rs;
    sh[i] = relu ? beta[c] - mu * sc[i] : 1.f;
    c1[i] = sc[i] * (float)sums[C + c] * inv * rs;
    c0[i] = sc[i] * (float)sums[c] * inv - mu * c1[i];
  }
  const float msk = relu ? 1.f : 0.f;   // relu == 0: fmaf(x, 0, 1) > 0 keeps every element
  const long long stride = (long long)gridDim.x * lanes;
  const T* pd = dy + g * 8;
  const T* px = x + g * 8;
  T* po = dx + g * 8;
  for (long long rb = (long long)blockIdx.x * lanes + rl; rb < M; rb += 4 * stride) {
    Raw8<T> d[4], xv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long r = min(rb + u * stride, M - 1);
      d[u].load(pd + r * C);
      xv[u].load(px + r * C);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long r = rb + u * stride;
      if (r < M) {
        float a[8], b[8];
        d[u].get(a); xv[u].get(b);
        Vec8<T> o;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float dd = fmaf(b[i], sc[i] * msk, sh[i]) <= 0.f ? 0.f : a[i];
          o.v[i] = fmaf(sc[i], dd, -fmaf(b[i], c1[i], c0[i]));
        }
        o.store(po + r * C);
      }
    }
  }
}
template <typename T>
int k_bn_bwd_apply(const T* dy, const T* x, const float* mean, const float* rstd, const float* gamma, const float* beta,
                   const double* sums, double count, T* dx, float* dgamma, float* dbeta, int relu, long long M, int C,
                   cudaStream_t st) {
  if (C % 8 || 256 % (C / 8)) { s4g_set_error("bn: unsupported C"); return S4G_ERR_SHAPE; }
  S4G_CUDA(launch_dep(bn_bwd_apply_kernel<T>, dim3(bn_grid(M, C)), dim3(256), 0, st, dy, x, mean, rstd, gamma, beta, sums, count, dx, dgamma, dbeta, relu, M, C));
  S4G_KERNEL_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------ seg head tail
// output_conv[1..4] (Models/DPT_decoder.py:492-495): BN -> ReLU -> Dropout(0.1) -> Conv1x1(256->1).
// One warp per pixel row (C = 256 = 32 lanes x 8 channels).
// keep-scale (0 or 1/(1-p)) of 8 consecutive elements starting at idx (idx % 8 == 0): from the explicit byte mask
// (parity tests) or from two 64-bit counter hashes -> eight 16-bit uniform values (recomputed in backward, never stored)
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}
__device__ __forceinline__ void drop_scale8(const uint8_t* mask, float p, uint64_t seed, long long idx, float (&k)[8]) {
  const float s = 1.f / (1.f - p);
  if (mask) {
    const uint2 m = *reinterpret_cast<const uint2*>(mask + idx);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      k[i] = ((m.x >> (8 * i)) & 0xff) ? s : 0.f;
      k[4 + i] = ((m.y >> (8 * i)) & 0xff) ? s : 0.f;
    }
  } else if (p <= 0.f) {
#pragma unroll
    for (int i = 0; i < 8; ++i) k[i] = 1.f;
  } else {
    const uint32_t thr = (uint32_t)(p * 65536.f);
    const uint64_t h0 = mix64(seed * 0x9E3779B97F4A7C15ULL + (uint64_t)idx);
    const uint64_t h1 = mix64(h0 ^ 0xD6E8FEB86659FD93ULL);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      k[i] = ((uint32_t)(h0 >> (16 * i)) & 0xffffu) >= thr ? s : 0.f;
      k[4 + i] = ((uint32_t)(h1 >> (16 * i)) & 0xffffu) >= thr ? s : 0.f;
    }
  }
}
template <typename T>
__global__ void __launch_bounds__(256, 3) seghead_fwd_kernel(const T* __restrict__ x, const float* __restrict__ mean,
                                                             const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, const float* __restrict__ w,
                                                             const float* __restrict__ b, const uint8_t* __restrict__ mask,
                                                             float p, uint64_t seed, float* __restrict__ small, long long M,
                                                             const unsigned long long* __restrict__ seed_dev) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  if (seed_dev) seed = *seed_dev;   // device-resident dropout seed (CUDA-graph replays: a by-value seed would be baked in)
  const int lane = threadIdx.x & 31;
  float sc[8], sh[8], wv[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane * 8 + i;
    sc[i] = rstd[c] * gamma[c]; sh[i] = beta[c] - mean[c] * sc[i]; wv[i] = w[c];
  }
  const float b0 = b[0];
  const long long stride = (long long)gridDim.x * 8;
  const T* px = x + lane * 8;
  for (long long rb = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); rb < M; rb += 4 * stride) {
    Raw8<T> v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u].load(px + min(rb + u * stride, M - 1) * 256);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long row = rb + u * stride;
      if (row < M) {
        float k[8], f[8];
        v[u].get(f);
        drop_scale8(mask, p, seed, row * 256 + lane * 8, k);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) s += fmaxf(fmaf(f[i], sc[i], sh[i]), 0.f) * k[i] * wv[i];
        s = warp_sum(s);
        if (lane == 0) small[row] = s + b0;
      }
    }
  }
}
template <typename T>
int k_seghead_fwd(const T* x, const float* mean, const float* rstd, const float* gamma, const float* beta,
                  const float* w, const float* b, const uint8_t* mask, float p, uint64_t seed, float* small,
                  long long M, int C, cudaStream_t st, const unsigned long long* seed_dev) {
  if (C != 256) { s4g_set_error("seghead: C must be 256"); return S4G_ERR_SHAPE; }
  S4G_CUDA(launch_dep(seghead_fwd_kernel<T>, dim3((unsigned)min((long long)cdiv(M, 8), (long long)148 * 16)), dim3(256), 0, st, x, mean, rstd, gamma, beta, w, b, mask, p, seed, small, M, seed_dev));
  S4G_KERNEL_CHECK();
  return 0;
}
// pass 1: per-channel sums {sum dy', sum dy'*xhat} (double), dw[c] = sum dsmall * z, db = sum dsmall.
// A thread accumulates sum dy' and sum dy'*x; xhat = (x - mean) * rstd is applied once per block in double.
template <typename T>
__global__ void __launch_bounds__(256, 2) seghead_bwd_reduce_kernel(const float* __restrict__ dsmall, const T* __restrict__ x,
                                                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                    const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                    const float* __restrict__ w, const uint8_t* __restrict__ mask,
                                                                    float p, uint64_t seed, double* __restrict__ sums, long long M,
                                                                    const unsigned long long* __restrict__ seed_dev) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  if (seed_dev) seed = *seed_dev;
  __shared__ float sm[3][8][256];
  __shared__ float sdb[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float sc[8], sh[8], wv[8], s1[8], s2[8], s3[8];
  float sd = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane * 8 + i;
    sc[i] = rstd[c] * gamma[c]; sh[i] = beta[c] - mean[c] * sc[i]; wv[i] = w[c];
    s1[i] = s2[i] = s3[i] = 0.f;
  }
  const long long stride = (long long)gridDim.x * 8;
  const T* px = x + lane * 8;
  for (long long rb = (long long)blockIdx.x * 8 + warp; rb < M; rb += 4 * stride) {
    Raw8<T> v[4];
    float dsv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long row = min(rb + u * stride, M - 1);
      v[u].load(px + row * 256);
      dsv[u] = dsmall[row];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long row = rb + u * stride;
      if (row < M) {
        const float ds = dsv[u];
        sd += ds;
        float k[8], f[8];
        v[u].get(f);
        drop_scale8(mask, p, seed, row * 256 + lane * 8, k);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float y = fmaf(f[i], sc[i], sh[i]);
          const float dk = ds * k[i];
          const float dy = y > 0.f ? dk * wv[i] : 0.f;
          s1[i] += dy; s2[i] = fmaf(dy, f[i], s2[i]); s3[i] = fmaf(dk, fmaxf(y, 0.f), s3[i]);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) { sm[0][warp][lane * 8 + i] = s1[i]; sm[1][warp][lane * 8 + i] = s2[i]; sm[2][warp][lane * 8 + i] = s3[i]; }
  if (lane == 0) sdb[warp] = sd;
  __syncthreads();
  const int c = threadIdx.x;
  double a = 0.0, bq = 0.0; float cw = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) { a += sm[0][k][c]; bq += sm[1][k][c]; cw += sm[2][k][c]; }
  // partial row of this block: [0, 512) BN sums, [512, 768) dw, [768] db (doubles; finalize sums the rows in order)
  double* prow = sums + (size_t)blockIdx.x * 776;
  prow[c] = a;
  prow[256 + c] = (bq - (double)mean[c] * a) * (double)rstd[c];
  prow[512 + c] = (double)cw;
  if (c == 0) { float t = 0.f; for (int k = 0; k < 8; ++k) t += sdb[k]; prow[768] = (double)t; }
}
template <typename T>
int k_seghead_bwd_reduce(const float* dsmall, const T* x, const float* mean, const float* rstd, const float* gamma,
                         const float* beta, const float* w, const uint8_t* mask, float p, uint64_t seed, double* sums,
                         float* dw, float* db, long long M, int C, cudaStream_t st, const Reducer* red,
                         const unsigned long long* seed_dev) {
  if (C != 256) { s4g_set_error("seghead: C must be 256"); return S4G_ERR_SHAPE; }
  const int nb = (int)min((long long)cdiv(M, 8), (long long)148 * 4);
  double* part = reinterpret_cast<double*>(reduce_scratch(red, seghead_scratch_bytes()));
  if (!part) return S4G_ERR_WORKSPACE;
  S4G_CUDA(launch_dep(seghead_bwd_reduce_kernel<T>, dim3(nb), dim3(256), 0, st, dsmall, x, mean, rstd, gamma, beta, w, mask, p, seed, part, M, seed_dev));
  S4G_KERNEL_CHECK();
  // one grouped combine for the three outputs (the BN sums are consumed by the very next kernel: never left queued)
  FinalizeJobs q; q.n = 0; q.total_blocks = 0; q.launches = 0;
  Reducer rq; rq.scratch = part; rq.scratch_bytes = seghead_scratch_bytes(); rq.defer = &q;
  S4G_TRY(k_finalize(&rq, part, sums, nb, 512, 776, 1, 1, st));
  S4G_TRY(k_finalize(&rq, part + 512, dw, nb, 256, 776, 1, 0, st));
  S4G_TRY(k_finalize(&rq, part + 768, db, nb, 1, 776, 1, 0, st));
  return k_finalize_flush(&q, st);
}
template <typename T>
__global__ void __launch_bounds__(256, 3) seghead_bwd_apply_kernel(const float* __restrict__ dsmall, const T* __restrict__ x,
                                                                   const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                   const float* __restrict__ w, const uint8_t* __restrict__ mask,
                                                                   float p, uint64_t seed, const double* __restrict__ sums,
                                                                   double count, T* __restrict__ dx, float* dgamma, float* dbeta,
                                                                   long long M, const unsigned long long* __restrict__ seed_dev) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  if (seed_dev) seed = *seed_dev;
  const int lane = threadIdx.x & 31;
  if (blockIdx.x == 0) {  // global sums / world under SyncBatchNorm, see bn_bwd_apply_kernel
    const double gs = (double)M / count;
    dbeta[threadIdx.x] = (float)(sums[threadIdx.x] * gs); dgamma[threadIdx.x] = (float)(sums[256 + threadIdx.x] * gs);
  }
  const float inv = (float)(1.0 / count);
  // dx = sc*dy - (c0 + x*c1), see bn_bwd_apply_kernel; scw = sc * w folds the 1x1 conv weight into the dy term
  float sc[8], sh[8], scw[8], c0[8], c1[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane * 8 + i;
    const float mu = mean[c], rs = rstd[c];
    sc[i] = gamma[c] * rs; sh[i] = beta[c] - mu * sc[i]; scw[i] = sc[i] * w[c];
    c1[i] = sc[i] * (float)sums[256 + c] * inv * rs;
    c0[i] = sc[i] * (float)sums[c] * inv - mu * c1[i];
  }
  const long long stride = (long long)gridDim.x * 8;
  const T* px = x + lane * 8;
  T* po = dx + lane * 8;
  for (long long rb = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); rb < M; rb += 4 * stride) {
    Raw8<T> v[4];
    float dsv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long row = min(rb + u * stride, M - 1);
      v[u].load(px + row * 256);
      dsv[u] = dsmall[row];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long row = rb + u * stride;
      if (row < M) {
        float k[8], f[8];
        v[u].get(f);
        drop_scale8(mask, p, seed, row * 256 + lane * 8, k);
        Vec8<T> o;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float y = fmaf(f[i], sc[i], sh[i]);
          const float dyw = y > 0.f ? dsv[u] * k[i] : 0.f;
          o.v[i] = fmaf(scw[i], dyw, -fmaf(f[i], c1[i], c0[i]));
        }
        o.store(po + row * 256);
      }
    }
  }
}
template <typename T>
int k_seghead_bwd_apply(const float* dsmall, const T* x, const float* mean, const float* rstd, const float* gamma,
                        const float* beta, const float* w, const uint8_t* mask, float p, uint64_t seed,
                        const double* sums, double count, T* dx, float* dgamma, float* dbeta, long long M, int C,
                        cudaStream_t st, const unsigned long long* seed_dev) {
  if (C != 256) { s4g_set_error("seghead: C must be 256"); return S4G_ERR_SHAPE; }
  S4G_CUDA(launch_dep(seghead_bwd_apply_kernel<T>, dim3((unsigned)min((long long)cdiv(M, 8), (long long)148 * 16)), dim3(256), 0, st, dsmall, x, mean, rstd, gamma, beta, w, mask, p, seed, sums, count, dx, dgamma, dbeta, M, seed_dev));
  S4G_KERNEL_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------ depth head tail
// output_conv[4..5] (Models/DPT_decoder.py:480-481): Conv1x1(32->1) + Sigmoid on the ReLU'd 32-ch map
template <typename T>
__global__ void depthhead_fwd_kernel(const T* __restrict__ h, const float* __restrict__ w, const float* __restrict__ b,
                                     float* __restrict__ out, long long M, int C) {
  FOR_GRID(row, M) {
    float s = b[0];
    for (int c = 0; c < C; c += 8) {
      Vec8<T> v;
      v.load(h + row * C + c);
#pragma unroll
      for (int i = 0; i < 8; ++i) s += v.v[i] * w[c + i];
    }
    out[row] = 1.f / (1.f + __expf(-s));
  }
}
template <typename T>
int k_depthhead_fwd(const T* h, const float* w, const float* b, float* out, long long M, int C, cudaStream_t st) {
  depthhead_fwd_kernel<T><<<GRID1D(M, 256), 0, st>>>(h, w, b, out, M, C);
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T>
__global__ void __launch_bounds__(256) depthhead_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ out,
                                                            const T* __restrict__ h, const float* __restrict__ w,
                                                            T* __restrict__ dh, float* __restrict__ part, long long M, int C) {
  __shared__ float sdw[8][33];
  float lw[32], ldb = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) lw[i] = 0.f;
  FOR_GRID(row, M) {
    const float o = out[row];
    const float ds = dout[row] * o * (1.f - o);
    ldb += ds;
#pragma unroll
    for (int c = 0; c < 32; c += 8) {
      Vec8<T> v, d;
      v.load(h + row * 32 + c);
#pragma unroll
      for (int i = 0; i < 8; ++i) { lw[c + i] += ds * v.v[i]; d.v[i] = ds * w[c + i]; }
      d.store(dh + row * 32 + c);
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 32; ++i) { const float s = warp_sum(lw[i]); if (lane == 0) sdw[warp][i] = s; }
  ldb = warp_sum(ldb);
  if (lane == 0) sdw[warp][32] = ldb;
  __syncthreads();
  // partial row of this block: [0, 32) dw, [32] db -- warps combined in warp order, rows summed in order by finalize
  if (threadIdx.x < 33) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sdw[k][threadIdx.x];
    part[(size_t)blockIdx.x * 33 + threadIdx.x] = t;
  }
}
template <typename T>
int k_depthhead_bwd(const float* dout, const float* out, const T* h, const float* w, T* dh, float* dw, float* db,
                    long long M, int C, cudaStream_t st, const Reducer* red) {
  if (C != 32) { s4g_set_error("depthhead: C must be 32"); return S4G_ERR_SHAPE; }
  const int nb = 148 * 4;
  float* part = reinterpret_cast<float*>(reduce_scratch(red, depthhead_scratch_bytes()));
  if (!part) return S4G_ERR_WORKSPACE;
  depthhead_bwd_kernel<T><<<nb, 256, 0, st>>>(dout, out, h, w, dh, part, M, C);
  S4G_KERNEL_CHECK();
  FinalizeJobs q; q.n = 0; q.total_blocks = 0; q.launches = 0;
  Reducer rq; rq.scratch = part; rq.scratch_bytes = depthhead_scratch_bytes(); rq.defer = &q;
  S4G_TRY(k_finalize(&rq, part, dw, nb, 32, 33, 0, 0, st));
  S4G_TRY(k_finalize(&rq, part + 32, db, nb, 1, 33, 0, 0, st));
  return k_finalize_flush(&q, st);
}

// ------------------------------------------------------------------------------------------ weight packing
template <typename T> __global__ void pack_linear_kernel(const float* __restrict__ w, T* __restrict__ wp, T* __restrict__ wt, int N, int K) {
  __shared__ float tile[32][33];
  const int k = blockIdx.x * 32 + threadIdx.x, n0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int n = n0 + j;
    float v = 0.f;
    if (n < N && k < K) { v = w[(long long)n * K + k]; wp[(long long)n * K + k] = from_f<T>(v); }
    tile[j][threadIdx.x] = v;
  }
  __syncthreads();
  if (!wt) return;
  const int n = n0 + threadIdx.x;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int kk = blockIdx.x * 32 + j;
    if (n < N && kk < K) wt[(long long)kk * N + n] = from_f<T>(tile[threadIdx.x][j]);
  }
}
template <typename T> int k_pack_linear(const float* w, T* wp, T* wt, int N, int K, cudaStream_t st) {
  pack_linear_kernel<T><<<dim3(cdiv(K, 32), cdiv(N, 32)), dim3(32, 8), 0, st>>>(w, wp, wt, N, K);
  S4G_KERNEL_CHECK();
  return 0;
}
// All nn.Linear / 1x1-conv weights of the model in ONE launch (the per-layer launches are latency-bound: 58 of them
// took 0.44 ms for 0.1 ms worth of HBM traffic).  The job table travels in kernel-parameter space.
__device__ __forceinline__ void store2(bf16* p, float a, float b) { *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b); }
__device__ __forceinline__ void store2(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }

// 64 x 64 tiles (PACK_TILE, kernels.h): 16 B loads of the fp32 master weights (a warp reads two 256 B row segments), 8 B stores
// of the bf16 copy, and the transposed copy written as 128 B row segments -- 680 MB per step move at HBM speed instead of in
// 64 / 128 B pieces (the 32 x 32 scalar version ran at 2.5 TB/s)
template <typename T> __global__ void __launch_bounds__(256) pack_linear_grouped_kernel(const __grid_constant__ PackLinearJobs jobs) {
  pdl_trigger();  // the next kernel of the stream, if launched programmatically dependent, may start its prologue now
  __shared__ float tile[PACK_TILE][PACK_TILE + 1];
  int lo = 0, hi = jobs.n - 1;
  while (lo < hi) {   // last job whose first tile is <= blockIdx.x
    const int mid = (lo + hi + 1) >> 1;
    if (jobs.job[mid].tile0 <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const PackLinearJob& jb = jobs.job[lo];
  const int t = blockIdx.x - jb.tile0;
  const int N = jb.N, K = jb.K;
  const int tiles_k = (K + PACK_TILE - 1) / PACK_TILE;
  const int k0 = (t % tiles_k) * PACK_TILE, n0 = (t / tiles_k) * PACK_TILE;
  const float* w = jb.src;
  T* wp = reinterpret_cast<T*>(jb.wp);
  T* wt = reinterpret_cast<T*>(jb.wt);
  const bool vec_k = (K & 3) == 0;
  const int c4 = (threadIdx.x & 15) * 4, r0 = threadIdx.x >> 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int j = r0 + 16 * i, n = n0 + j, k = k0 + c4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (n < N) {
      const long long o = (long long)n * K + k;
      if (vec_k && k + 3 < K) {
        v = *reinterpret_cast<const float4*>(w + o);
        store4(wp + o, v);
      } else {
        float* f = &v.x;
        for (int e = 0; e < 4; ++e)
          if (k + e < K) { f[e] = w[o + e]; wp[o + e] = from_f<T>(f[e]); }
      }
    }
    tile[j][c4] = v.x; tile[j][c4 + 1] = v.y; tile[j][c4 + 2] = v.z; tile[j][c4 + 3] = v.w;
  }
  if (!wt) return;
  __syncthreads();
  const int nl = (threadIdx.x & 31) * 2, n = n0 + nl;
  const bool vec_n = (N & 1) == 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int j = (threadIdx.x >> 5) + 8 * i, kk = k0 + j;
    if (kk >= K || n >= N) continue;
    const long long o = (long long)kk * N + n;
    if (vec_n) store2(wt + o, tile[nl][j], tile[nl + 1][j]);   // N even and n even: n + 1 < N, 4 B aligned
    else { wt[o] = from_f<T>(tile[nl][j]); if (n + 1 < N) wt[o + 1] = from_f<T>(tile[nl + 1][j]); }
  }
}
template <typename T> int k_pack_linear_grouped(const PackLinearJobs& jobs, int total_tiles, cudaStream_t st) {
  if (jobs.n == 0) return 0;
  pack_linear_grouped_kernel<T><<<total_tiles, 256, 0, st>>>(jobs);
  S4G_KERNEL_CHECK();
  return 0;
}
// conv weight [Co][Ci][R][R] -> fprop wf[Co][tap][Ci]; dgrad wd[Ci][tap'][Co] with tap' = flipped tap
template <typename T> __global__ void pack_conv_kernel(const float* __restrict__ w, T* __restrict__ wf, T* __restrict__ wd, int Co, int Ci, int R) {
  const int taps = R * R;
  const long long n = (long long)Co * Ci * taps;
  FOR_GRID(t, n) {
    const int tap = (int)(t % taps);
    const int ci = (int)((t / taps) % Ci), co = (int)(t / ((long long)taps * Ci));
    const float v = w[t];
    wf[((long long)co * taps + tap) * Ci + ci] = from_f<T>(v);
    if (wd) wd[((long long)ci * taps + (taps - 1 - tap)) * Co + co] = from_f<T>(v);
  }
}
template <typename T> int k_pack_conv(const float* w, T* wf, T* wd, int Co, int Ci, int R, cudaStream_t st) {
  pack_conv_kernel<T><<<GRID1D((long long)Co * Ci * R * R, 256), 0, st>>>(w, wf, wd, Co, Ci, R);
  S4G_KERNEL_CHECK();
  return 0;
}
// All 3x3 conv weights in ONE launch: a block owns one output channel (-> wf row [tap][Ci]) or one input channel
// (-> wd row [tap'][Co]) and transposes its 9*C weights through shared memory, so both writes are contiguous.
template <typename T> __global__ void __launch_bounds__(256) pack_conv3_grouped_kernel(const __grid_constant__ PackLinearJobs jobs) {
  pdl_trigger();  // the next kernel of the stream, if launched programmatically dependent, may start its prologue now
  __shared__ float sm[9 * 768];
  int lo = 0, hi = jobs.n - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (jobs.job[mid].tile0 <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const PackLinearJob& jb = jobs.job[lo];
  const int Co = jb.N, Ci = jb.K;
  const int t = blockIdx.x - jb.tile0;
  const float* w = jb.src;
  if (t < Co) {
    const int co = t, n = 9 * Ci;
    const float* src = w + (size_t)co * n;
    for (int i = threadIdx.x; i < n; i += 256) sm[(i % 9) * Ci + i / 9] = src[i];
    __syncthreads();
    T* dst = reinterpret_cast<T*>(jb.wp) + (size_t)co * n;
    for (int i = threadIdx.x; i < n; i += 256) dst[i] = from_f<T>(sm[i]);
  } else {
    const int ci = t - Co, n = 9 * Co;
    for (int i = threadIdx.x; i < n; i += 256) {
      const int co = i / 9, tap = i % 9;
      sm[(8 - tap) * Co + co] = w[((size_t)co * Ci + ci) * 9 + tap];
    }
    __syncthreads();
    T* dst = reinterpret_cast<T*>(jb.wt) + (size_t)ci * n;
    for (int i = threadIdx.x; i < n; i += 256) dst[i] = from_f<T>(sm[i]);
  }
}
// jobs: N = Co, K = Ci (both <= 768), wp = wf, wt = wd, tile0 = first block
template <typename T> int k_pack_conv3_grouped(const PackLinearJobs& jobs, int total_blocks, cudaStream_t st) {
  if (jobs.n == 0) return 0;
  pack_conv3_grouped_kernel<T><<<total_blocks, 256, 0, st>>>(jobs);
  S4G_KERNEL_CHECK();
  return 0;
}
// ConvTranspose weight [Ci][Co][k][k] -> fprop B operand wf[(kh,kw,co)][ci]; dgrad B operand wd[ci][(kh,kw,co)]
template <typename T> __global__ void pack_convT_kernel(const float* __restrict__ w, T* __restrict__ wf, T* __restrict__ wd, int Ci, int Co, int k) {
  const int kk = k * k;
  const long long n = (long long)Ci * Co * kk;
  FOR_GRID(t, n) {
    const int tap = (int)(t % kk);
    const int co = (int)((t / kk) % Co), ci = (int)(t / ((long long)kk * Co));
    const float v = w[t];
    const long long col = (long long)tap * Co + co;
    wf[col * Ci + ci] = from_f<T>(v);
    wd[(long long)ci * kk * Co + col] = from_f<T>(v);
  }
}
template <typename T> int k_pack_convT(const float* w, T* wf, T* wd, int Ci, int Co, int k, cudaStream_t st) {
  pack_convT_kernel<T><<<GRID1D((long long)Ci * Co * k * k, 256), 0, st>>>(w, wf, wd, Ci, Co, k);
  S4G_KERNEL_CHECK();
  return 0;
}
__global__ void unpack_conv_grad_kernel(const float* __restrict__ dwp, float* __restrict__ dw, int Co, int Ci, int R) {
  const int taps = R * R;
  const long long n = (long long)Co * Ci * taps;
  FOR_GRID(t, n) {
    const int tap = (int)(t % taps);
    const int ci = (int)((t / taps) % Ci), co = (int)(t / ((long long)taps * Ci));
    dw[t] = dwp[((long long)co * taps + tap) * Ci + ci];
  }
}
int k_unpack_conv_grad(const float* dwp, float* dw, int Co, int Ci, int R, cudaStream_t st) {
  unpack_conv_grad_kernel<<<GRID1D((long long)Co * Ci * R * R, 256), 0, st>>>(dwp, dw, Co, Ci, R);
  S4G_KERNEL_CHECK();
  return 0;
}
__global__ void unpack_convT_grad_kernel(const float* __restrict__ dwp, float* __restrict__ dw, int Ci, int Co, int k) {
  const int kk = k * k;
  const long long n = (long long)Ci * Co * kk;
  FOR_GRID(t, n) {
    const int tap = (int)(t % kk);
    const int co = (int)((t / kk) % Co), ci = (int)(t / ((long long)kk * Co));
    dw[t] = dwp[((long long)tap * Co + co) * Ci + ci];
  }
}
int k_unpack_convT_grad(const float* dwp, float* dw, int Ci, int Co, int k, cudaStream_t st) {
  unpack_convT_grad_kernel<<<GRID1D((long long)Ci * Co * k * k, 256), 0, st>>>(dwp, dw, Ci, Co, k);
  S4G_KERNEL_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------ instantiation
#define INST(T)                                                                                                        \
  template int k_patchify<T>(const float*, T*, int, cudaStream_t);                                                     \
  template int k_pack_linear_grouped<T>(const PackLinearJobs&, int, cudaStream_t);                                     \
  template int k_pack_conv3_grouped<T>(const PackLinearJobs&, int, cudaStream_t);                                      \
  template int k_layernorm_fwd<T>(const float*, const float*, const float*, T*, float*, float*, int, int, float, cudaStream_t); \
  template int k_layernorm_bwd<T>(const T*, const float*, const float*, const float*, const float*, const float*, float*, float*, float*, int, int, cudaStream_t, T*, float*, const Reducer*); \
  template int k_dgelu_mul<T>(const T*, const T*, T*, long long, cudaStream_t, float*, int, const Reducer*);                          \
  template int k_gelu<T>(const T*, T*, long long, cudaStream_t, T*);                                                      \
  template int k_colsum<T>(const T*, float*, long long, int, int, cudaStream_t, const Reducer*);                                       \
  template int k_tokens_to_nhwc<T>(const float*, T*, int, cudaStream_t);                                               \
  template int k_tap_grad<T>(const T*, float*, int, int, cudaStream_t, T*);                                                \
  template int k_cast_f32<T>(const float*, T*, long long, cudaStream_t);                                               \
  template int k_relu<T>(const T*, T*, long long, cudaStream_t);                                                       \
  template int k_relu_bwd_add<T>(const T*, const T*, const T*, T*, long long, cudaStream_t);                           \
  template int k_add<T>(const T*, const T*, T*, long long, cudaStream_t);                                              \
  template int k_pixel_shuffle<T>(const T*, const float*, T*, int, int, int, int, int, cudaStream_t);                  \
  template int k_pixel_unshuffle<T>(const T*, T*, int, int, int, int, int, cudaStream_t);                              \
  template int k_im2col_s2<T>(const T*, T*, int, int, int, int, cudaStream_t);                                         \
  template int k_col2im_s2<T>(const T*, T*, int, int, int, int, cudaStream_t);                                         \
  template int k_bilinear2x_fwd<T>(const T*, T*, int, int, int, int, cudaStream_t);                                    \
  template int k_bilinear2x_bwd<T>(const T*, T*, int, int, int, int, cudaStream_t);                                    \
  template int k_bn_stats<T>(const T*, double*, long long, int, cudaStream_t, const Reducer*);                                         \
  template int k_bn_apply<T>(const T*, const float*, const float*, const float*, const float*, const T*, T*, int, long long, int, cudaStream_t); \
  template int k_bn_bwd_reduce<T>(const T*, const T*, const float*, const float*, const float*, const float*, double*, int, long long, int, cudaStream_t, const Reducer*); \
  template int k_bn_bwd_apply<T>(const T*, const T*, const float*, const float*, const float*, const float*, const double*, double, T*, float*, float*, int, long long, int, cudaStream_t); \
  template int k_seghead_fwd<T>(const T*, const float*, const float*, const float*, const float*, const float*, const float*, const uint8_t*, float, uint64_t, float*, long long, int, cudaStream_t, const unsigned long long*); \
  template int k_seghead_bwd_reduce<T>(const float*, const T*, const float*, const float*, const float*, const float*, const float*, const uint8_t*, float, uint64_t, double*, float*, float*, long long, int, cudaStream_t, const Reducer*, const unsigned long long*); \
  template int k_seghead_bwd_apply<T>(const float*, const T*, const float*, const float*, const float*, const float*, const float*, const uint8_t*, float, uint64_t, const double*, double, T*, float*, float*, long long, int, cudaStream_t, const unsigned long long*); \
  template int k_depthhead_fwd<T>(const T*, const float*, const float*, float*, long long, int, cudaStream_t);         \
  template int k_depthhead_bwd<T>(const float*, const float*, const T*, const float*, T*, float*, float*, long long, int, cudaStream_t, const Reducer*); \
  template int k_pack_linear<T>(const float*, T*, T*, int, int, cudaStream_t);                                         \
  template int k_pack_conv<T>(const float*, T*, T*, int, int, int, cudaStream_t);                                      \
  template int k_pack_convT<T>(const float*, T*, T*, int, int, int, cudaStream_t);
INST(float)
INST(bf16)
