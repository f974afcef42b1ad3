// Kernels of the MAE pretraining step that the fine-tuning path does not have (reference Models/mae/models_mae.py):
//   random_masking :123-148 (argsort of per-sample noise -> ids_restore / ids_keep / mask),
//   forward_encoder :150-170 (patch embed of the KEPT patches only, + pos, + cls),
//   forward_decoder :172-196 (mask-token append + un-shuffle gather + decoder pos),
//   forward_loss :198-214 (patchify target, per-patch normalisation, MSE on the removed patches),
// and the head-dimension pad / unpad passes that let the 16-head x 32-dim decoder attention run on the 64-dim attention
// kernels.  All reductions are fixed-order (no atomics).
#include "kernels.h"

namespace {

#define FOR_GRID(i, n) \
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (long long)gridDim.x * blockDim.x)

constexpr int GRID_CAP = 148 * 16;
inline dim3 grid1d(long long n, int block) { return dim3((unsigned)std::min<long long>((n + block - 1) / block, GRID_CAP)); }

// ------------------------------------------------------------------------------------------------ masking
// rank[i] = position of noise[i] in the ascending (stable) order == ids_restore[i]; ids_shuffle[rank[i]] = i
__global__ void __launch_bounds__(256) mae_masking_kernel(const float* __restrict__ noise, int* __restrict__ ids_restore,
                                                          int* __restrict__ ids_keep, float* __restrict__ mask, int L, int keep) {
  extern __shared__ float sn[];
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < L; i += blockDim.x) sn[i] = noise[(long long)b * L + i];
  __syncthreads();
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    const float v = sn[i];
    int r = 0;
    for (int j = 0; j < L; ++j) r += (sn[j] < v) || (sn[j] == v && j < i);
    ids_restore[(long long)b * L + i] = r;
    if (r < keep) ids_keep[(long long)b * keep + r] = i;
    mask[(long long)b * L + i] = r < keep ? 0.f : 1.f;
  }
}

// ------------------------------------------------------------------------------------------------ encoder input
// PatchEmbed rows of the kept patches only: out[b*keep + k, c*256 + i*16 + j] = img[b, c, ph*16+i, pw*16+j], p = ids_keep[b,k]
template <typename T>
__global__ void mae_patchify_kept_kernel(const float* __restrict__ img, const int* __restrict__ ids_keep, T* __restrict__ out,
                                         long long rows, int keep) {
  const long long n = rows * 96;
  FOR_GRID(t, n) {
    const int g = (int)(t % 96);
    const long long row = t / 96;
    const int b = (int)(row / keep);
    const int p = ids_keep[row];
    const int c = g / 32, i = (g % 32) / 2, j0 = (g % 2) * 8;
    const int ph = p / 14, pw = p % 14;
    const float* src = img + (((long long)b * 3 + c) * 224 + ph * 16 + i) * 224 + pw * 16 + j0;
    const float4 a = *reinterpret_cast<const float4*>(src), bq = *reinterpret_cast<const float4*>(src + 4);
    Vec8<T> v;
    v.v[0] = a.x; v.v[1] = a.y; v.v[2] = a.z; v.v[3] = a.w; v.v[4] = bq.x; v.v[5] = bq.y; v.v[6] = bq.z; v.v[7] = bq.w;
    v.store(out + row * 768 + g * 8);
  }
}

// x[b,0] = cls + pos[0]; x[b,1+k] = pe[b*keep+k] + pos[1 + ids_keep[b,k]]   (models_mae.py:152-163)
__global__ void mae_assemble_kernel(const float* __restrict__ pe, const float* __restrict__ cls, const float* __restrict__ pos,
                                    const int* __restrict__ ids_keep, float* __restrict__ x, int B, int keep, int C4) {
  const long long n = (long long)B * (keep + 1) * C4;
  FOR_GRID(t, n) {
    const int c4 = (int)(t % C4);
    const long long row = t / C4;
    const int tok = (int)(row % (keep + 1)), b = (int)(row / (keep + 1));
    float4 v, pz;
    if (tok == 0) {
      v = reinterpret_cast<const float4*>(cls)[c4];
      pz = reinterpret_cast<const float4*>(pos)[c4];
    } else {
      const long long src = (long long)b * keep + tok - 1;
      v = reinterpret_cast<const float4*>(pe)[src * C4 + c4];
      pz = reinterpret_cast<const float4*>(pos)[(long long)(1 + ids_keep[src]) * C4 + c4];
    }
    v.x += pz.x; v.y += pz.y; v.z += pz.z; v.w += pz.w;
    reinterpret_cast<float4*>(x)[t] = v;
  }
}
// dpe[b*keep+k] = dx[b,1+k] (cast to T); dcls[c] = sum_b dx[b,0,c] in batch order
template <typename T>
__global__ void mae_assemble_bwd_kernel(const float* __restrict__ dx, T* __restrict__ dpe, float* __restrict__ dcls, int B,
                                        int keep, int C) {
  const int C8 = C / 8;
  const long long n = (long long)B * keep * C8;
  FOR_GRID(t, n) {
    const int g = (int)(t % C8);
    const long long row = t / C8;
    const int b = (int)(row / keep), k = (int)(row % keep);
    Vec8<float> v;
    v.load(dx + ((long long)b * (keep + 1) + 1 + k) * C + g * 8);
    Vec8<T> o;
#pragma unroll
    for (int q = 0; q < 8; ++q) o.v[q] = v.v[q];
    o.store(dpe + row * C + g * 8);
  }
  if (dcls && blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float s = 0.f;
      for (int b = 0; b < B; ++b) s += dx[(long long)b * (keep + 1) * C + c];
      dcls[c] = s;
    }
  }
}

// ------------------------------------------------------------------------------------------------ decoder input
// y[b,0] = de[b,0] + dpos[0]; y[b,1+i] = (r = ids_restore[b,i]) < keep ? de[b,1+r] : mask_token, + dpos[1+i]
__global__ void mae_unshuffle_kernel(const float* __restrict__ de, const float* __restrict__ mask_token,
                                     const float* __restrict__ dpos, const int* __restrict__ ids_restore, float* __restrict__ y,
                                     int B, int L, int keep, int C4) {
  const long long n = (long long)B * (L + 1) * C4;
  FOR_GRID(t, n) {
    const int c4 = (int)(t % C4);
    const long long row = t / C4;
    const int tok = (int)(row % (L + 1)), b = (int)(row / (L + 1));
    float4 v;
    if (tok == 0) {
      v = reinterpret_cast<const float4*>(de)[(long long)b * (keep + 1) * C4 + c4];
    } else {
      const int r = ids_restore[(long long)b * L + tok - 1];
      v = r < keep ? reinterpret_cast<const float4*>(de)[((long long)b * (keep + 1) + 1 + r) * C4 + c4]
                   : reinterpret_cast<const float4*>(mask_token)[c4];
    }
    const float4 pz = reinterpret_cast<const float4*>(dpos)[(long long)tok * C4 + c4];
    v.x += pz.x; v.y += pz.y; v.z += pz.z; v.w += pz.w;
    reinterpret_cast<float4*>(y)[t] = v;
  }
}
// dde[b,0] = dy[b,0]; dde[b,1+r] = dy[b,1+i] for the kept i; part[b, c] = sum over the removed i (ascending) of dy[b,1+i,c]
template <typename T>
__global__ void __launch_bounds__(256) mae_unshuffle_bwd_kernel(const float* __restrict__ dy, const int* __restrict__ ids_restore,
                                                                T* __restrict__ dde, float* __restrict__ part, int L, int keep,
                                                                int C) {
  const int b = blockIdx.x;
  const float* src = dy + (long long)b * (L + 1) * C;
  T* dst = dde + (long long)b * (keep + 1) * C;
  const int* rs = ids_restore + (long long)b * L;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    dst[c] = from_f<T>(src[c]);
    float s = 0.f;
    for (int i = 0; i < L; ++i) {
      const int r = rs[i];
      const float v = src[(long long)(1 + i) * C + c];
      if (r < keep) dst[(long long)(1 + r) * C + c] = from_f<T>(v);
      else s += v;
    }
    part[(long long)b * C + c] = s;
  }
}

// ------------------------------------------------------------------------------------------------ head pad / unpad
// in [rows, heads*32] -> out [rows, heads*64] (upper 32 dims zero); the first `scaled` heads are multiplied by `scale`
template <typename T>
__global__ void head_pad_kernel(const T* __restrict__ in, T* __restrict__ out, long long rows, int heads, int scaled, float scale) {
  const long long n = rows * heads * 8;
  FOR_GRID(t, n) {
    const int g = (int)(t % 8);
    const long long rh = t / 8;
    const int h = (int)(rh % heads);
    Vec8<T> v;
    if (g < 4) {
      v.load(in + rh * 32 + g * 8);
      if (h < scaled) {
#pragma unroll
        for (int q = 0; q < 8; ++q) v.v[q] *= scale;
      }
    } else {
#pragma unroll
      for (int q = 0; q < 8; ++q) v.v[q] = 0.f;
    }
    v.store(out + rh * 64 + g * 8);
  }
}
template <typename T>
__global__ void head_unpad_kernel(const T* __restrict__ in, T* __restrict__ out, long long rows, int heads, int scaled, float scale) {
  const long long n = rows * heads * 4;
  FOR_GRID(t, n) {
    const int g = (int)(t % 4);
    const long long rh = t / 4;
    const int h = (int)(rh % heads);
    Vec8<T> v;
    v.load(in + rh * 64 + g * 8);
    if (h < scaled) {
#pragma unroll
      for (int q = 0; q < 8; ++q) v.v[q] *= scale;
    }
    v.store(out + rh * 32 + g * 8);
  }
}

// ------------------------------------------------------------------------------------------------ prediction rows
// pred [B, L, C] = full [B, L+1, C] without the cls row (models_mae.py:194)
__global__ void drop_cls_kernel(const float* __restrict__ full, float* __restrict__ out, int B, int L, int C4) {
  const long long n = (long long)B * L * C4;
  FOR_GRID(t, n) {
    const long long row = t / C4;
    const int b = (int)(row / L);
    reinterpret_cast<float4*>(out)[t] = reinterpret_cast<const float4*>(full)[t + (long long)(b + 1) * C4];
  }
}
// gradient of the same: T rows [B, L+1, C] with a zero cls row
template <typename T>
__global__ void add_cls_row_kernel(const float* __restrict__ dout, T* __restrict__ dfull, int B, int L, int C8) {
  const long long n = (long long)B * (L + 1) * C8;
  FOR_GRID(t, n) {
    const int g = (int)(t % C8);
    const long long row = t / C8;
    const int tok = (int)(row % (L + 1)), b = (int)(row / (L + 1));
    Vec8<T> o;
    if (tok == 0) {
#pragma unroll
      for (int q = 0; q < 8; ++q) o.v[q] = 0.f;
    } else {
      Vec8<float> v;
      v.load(dout + (((long long)b * L + tok - 1) * C8 + g) * 8);
#pragma unroll
      for (int q = 0; q < 8; ++q) o.v[q] = v.v[q];
    }
    o.store(dfull + t * 8);
  }
}

// ------------------------------------------------------------------------------------------------ loss
// One warp per patch.  target[e], e = (i*16 + j)*3 + c (patchify, models_mae.py:95-107); with norm_pix the target is
// (t - mean) / sqrt(var + 1e-6), var unbiased over the 768 values (:206-209).  lp = mean_e (pred - target)^2 (:211-212).
// Lane l owns the 16-float image row segments (c, i) = pairs l/16 + 2*k, column j = l % 16.
struct PatchTarget {
  float t[24];
  __device__ __forceinline__ void load(const float* img, int b, int p, int lane) {
    const int ph = p / 14, pw = p % 14, j = lane & 15;
#pragma unroll
    for (int k = 0; k < 24; ++k) {
      const int ci = (lane >> 4) + 2 * k;  // 0..47 = c*16 + i
      const int c = ci >> 4, i = ci & 15;
      t[k] = img[(((long long)b * 3 + c) * 224 + ph * 16 + i) * 224 + pw * 16 + j];
    }
  }
  static __device__ __forceinline__ int elem(int k, int lane) {
    const int ci = (lane >> 4) + 2 * k, c = ci >> 4, i = ci & 15, j = lane & 15;
    return (i * 16 + j) * 3 + c;
  }
  __device__ __forceinline__ void normalise(int norm_pix) {
    if (!norm_pix) return;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 24; ++k) s += t[k];
    const float mean = warp_sum(s) * (1.f / 768.f);
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < 24; ++k) { const float d = t[k] - mean; q = fmaf(d, d, q); }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / 767.f) + 1.0e-6f);
#pragma unroll
    for (int k = 0; k < 24; ++k) t[k] = (t[k] - mean) * rstd;
  }
};

__global__ void __launch_bounds__(256) mae_loss_patch_kernel(const float* __restrict__ pred, const float* __restrict__ img,
                                                             float* __restrict__ lp, int n_patches, int L, int norm_pix) {
  const int lane = threadIdx.x & 31;
  for (int pid = blockIdx.x * 8 + (threadIdx.x >> 5); pid < n_patches; pid += gridDim.x * 8) {
    PatchTarget tg;
    tg.load(img, pid / L, pid % L, lane);
    tg.normalise(norm_pix);
    const float* pr = pred + (long long)pid * 768;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 24; ++k) { const float d = pr[PatchTarget::elem(k, lane)] - tg.t[k]; s = fmaf(d, d, s); }
    s = warp_sum(s);
    if (lane == 0) lp[pid] = s * (1.f / 768.f);
  }
}
// loss = sum(lp * mask) / sum(mask) in one block, fixed order (models_mae.py:213); sums[0] = sum(mask) kept for backward
__global__ void __launch_bounds__(256) mae_loss_finalize_kernel(const float* __restrict__ lp, const float* __restrict__ mask,
                                                                double* __restrict__ sums, float* __restrict__ loss, int n) {
  __shared__ double sa[256], sb[256];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) { a += (double)lp[i] * (double)mask[i]; b += (double)mask[i]; }
  sa[threadIdx.x] = a; sb[threadIdx.x] = b;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) { sa[threadIdx.x] += sa[threadIdx.x + o]; sb[threadIdx.x] += sb[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { sums[0] = sb[0]; *loss = (float)(sa[0] / sb[0]); }
}
// dpred[pid, e] = grad_out * mask[pid] * 2 (pred - target) / (768 * sum(mask))
__global__ void __launch_bounds__(256) mae_loss_bwd_kernel(const float* __restrict__ pred, const float* __restrict__ img,
                                                           const float* __restrict__ mask, const double* __restrict__ sums,
                                                           const float* __restrict__ grad_out, float* __restrict__ dpred,
                                                           int n_patches, int L, int norm_pix) {
  const int lane = threadIdx.x & 31;
  const float base = *grad_out * (float)(2.0 / (768.0 * sums[0]));
  for (int pid = blockIdx.x * 8 + (threadIdx.x >> 5); pid < n_patches; pid += gridDim.x * 8) {
    float* dp = dpred + (long long)pid * 768;
    const float w = base * mask[pid];
    if (w == 0.f) {
#pragma unroll
      for (int k = 0; k < 6; ++k) reinterpret_cast<float4*>(dp)[lane + 32 * k] = make_float4(0.f, 0.f, 0.f, 0.f);
      continue;
    }
    PatchTarget tg;
    tg.load(img, pid / L, pid % L, lane);
    tg.normalise(norm_pix);
    const float* pr = pred + (long long)pid * 768;
#pragma unroll
    for (int k = 0; k < 24; ++k) { const int e = PatchTarget::elem(k, lane); dp[e] = w * (pr[e] - tg.t[k]); }
  }
}

}  // namespace

// ================================================================================================ host entry points
int k_mae_masking(const float* noise, int* ids_restore, int* ids_keep, float* mask, int B, int L, int keep, cudaStream_t st) {
  if (L < 1 || L > 4096 || keep < 1 || keep > L) { s4g_set_error("mae_masking: bad sequence length"); return S4G_ERR_SHAPE; }
  mae_masking_kernel<<<B, 256, sizeof(float) * L, st>>>(noise, ids_restore, ids_keep, mask, L, keep);
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> int k_mae_patchify_kept(const float* img, const int* ids_keep, T* out, int B, int keep, cudaStream_t st) {
  const long long rows = (long long)B * keep;
  mae_patchify_kept_kernel<T><<<grid1d(rows * 96, 256), 256, 0, st>>>(img, ids_keep, out, rows, keep);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_mae_assemble(const float* pe, const float* cls, const float* pos, const int* ids_keep, float* x, int B, int keep, int C,
                   cudaStream_t st) {
  mae_assemble_kernel<<<grid1d((long long)B * (keep + 1) * (C / 4), 256), 256, 0, st>>>(pe, cls, pos, ids_keep, x, B, keep, C / 4);
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> int k_mae_assemble_bwd(const float* dx, T* dpe, float* dcls, int B, int keep, int C, cudaStream_t st) {
  mae_assemble_bwd_kernel<T><<<grid1d((long long)B * keep * (C / 8), 256), 256, 0, st>>>(dx, dpe, dcls, B, keep, C);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_mae_unshuffle(const float* de, const float* mask_token, const float* dpos, const int* ids_restore, float* y, int B, int L,
                    int keep, int C, cudaStream_t st) {
  mae_unshuffle_kernel<<<grid1d((long long)B * (L + 1) * (C / 4), 256), 256, 0, st>>>(de, mask_token, dpos, ids_restore, y, B, L, keep, C / 4);
  S4G_KERNEL_CHECK();
  return 0;
}
size_t mae_unshuffle_scratch_bytes(int B, int C) { return sizeof(float) * (size_t)B * C; }
template <typename T>
int k_mae_unshuffle_bwd(const float* dy, const int* ids_restore, T* dde, float* dmask_token, int B, int L, int keep, int C,
                        cudaStream_t st, const Reducer* r) {
  float* part = reinterpret_cast<float*>(reduce_scratch(r, mae_unshuffle_scratch_bytes(B, C)));
  if (!part) return S4G_ERR_WORKSPACE;
  mae_unshuffle_bwd_kernel<T><<<B, 256, 0, st>>>(dy, ids_restore, dde, part, L, keep, C);
  S4G_KERNEL_CHECK();
  return k_finalize(r, part, dmask_token, B, C, C, 0, 0, st);
}
template <typename T> int k_head_pad(const T* in, T* out, long long rows, int heads, int scaled, float scale, cudaStream_t st) {
  head_pad_kernel<T><<<grid1d(rows * heads * 8, 256), 256, 0, st>>>(in, out, rows, heads, scaled, scale);
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> int k_head_unpad(const T* in, T* out, long long rows, int heads, int scaled, float scale, cudaStream_t st) {
  head_unpad_kernel<T><<<grid1d(rows * heads * 4, 256), 256, 0, st>>>(in, out, rows, heads, scaled, scale);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_drop_cls(const float* full, float* out, int B, int L, int C, cudaStream_t st) {
  drop_cls_kernel<<<grid1d((long long)B * L * (C / 4), 256), 256, 0, st>>>(full, out, B, L, C / 4);
  S4G_KERNEL_CHECK();
  return 0;
}
template <typename T> int k_add_cls_row(const float* dout, T* dfull, int B, int L, int C, cudaStream_t st) {
  add_cls_row_kernel<T><<<grid1d((long long)B * (L + 1) * (C / 8), 256), 256, 0, st>>>(dout, dfull, B, L, C / 8);
  S4G_KERNEL_CHECK();
  return 0;
}

size_t mae_loss_scratch_bytes(int B) { return 256 + sizeof(float) * (size_t)B * 196; }
int k_mae_loss_fwd(const float* pred, const float* img, const float* mask, void* scratch, float* loss, int B, int norm_pix,
                   cudaStream_t st) {
  double* sums = reinterpret_cast<double*>(scratch);
  float* lp = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 256);
  const int n = B * 196;
  mae_loss_patch_kernel<<<std::min(cdiv(n, 8), GRID_CAP), 256, 0, st>>>(pred, img, lp, n, 196, norm_pix);
  S4G_KERNEL_CHECK();
  mae_loss_finalize_kernel<<<1, 256, 0, st>>>(lp, mask, sums, loss, n);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_mae_loss_bwd(const float* pred, const float* img, const float* mask, const void* scratch, const float* grad_out,
                   float* dpred, int B, int norm_pix, cudaStream_t st) {
  const int n = B * 196;
  mae_loss_bwd_kernel<<<std::min(cdiv(n, 8), GRID_CAP), 256, 0, st>>>(pred, img, mask, reinterpret_cast<const double*>(scratch),
                                                                      grad_out, dpred, n, 196, norm_pix);
  S4G_KERNEL_CHECK();
  return 0;
}

#define INST(T)                                                                                                     \
  template int k_mae_patchify_kept<T>(const float*, const int*, T*, int, int, cudaStream_t);                        \
  template int k_mae_assemble_bwd<T>(const float*, T*, float*, int, int, int, cudaStream_t);                        \
  template int k_mae_unshuffle_bwd<T>(const float*, const int*, T*, float*, int, int, int, int, cudaStream_t, const Reducer*); \
  template int k_head_pad<T>(const T*, T*, long long, int, int, float, cudaStream_t);                               \
  template int k_head_unpad<T>(const T*, T*, long long, int, int, float, cudaStream_t);                             \
  template int k_add_cls_row<T>(const float*, T*, int, int, int, cudaStream_t);
INST(float)
INST(bf16)
