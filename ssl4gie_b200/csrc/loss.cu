// Losses of the two dense tasks, forward and analytic backward, as warp-reduction kernels.
//   SoftDiceLoss                 reference Binary_segmentation/Metrics/losses.py:10-24
//   ScaleAndShiftInvariantLoss   reference Depth_estimation/Metrics/losses.py:5-146 (batch-based reduction)
// No host synchronisation (the reference's nonzero()/`if divisor == 0` host syncs become device-side selects).
#include "kernels.h"

namespace {

__device__ __forceinline__ double block_sum_d(double v, double* red) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0) for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += red[i];
  return s;  // valid on thread 0
}

// ---------------------------------------------------------------- Dice
// scratch (doubles): sums[B][3] = per image {sum p*t, sum p*p, sum t*t}, then the per-block partial rows part[B][LOSS_GX][3]
// (block (x, b) writes row (b, x); the finalize kernel adds them in x order: no atomics, bit-reproducible)
constexpr int LOSS_GX = 64;   // upper bound of the blocks per image
__global__ void __launch_bounds__(256) dice_sums_kernel(const float* __restrict__ logits, const float* __restrict__ targets,
                                                        double* __restrict__ part, int HW) {
  __shared__ double red[8];
  const int b = blockIdx.y;
  double s0 = 0, s1 = 0, s2 = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < HW; i += gridDim.x * 256) {
    const float p = 1.f / (1.f + expf(-logits[(long long)b * HW + i]));
    const float t = targets[(long long)b * HW + i];
    s0 += p * t; s1 += p * p; s2 += t * t;
  }
  s0 = block_sum_d(s0, red); s1 = block_sum_d(s1, red); s2 = block_sum_d(s2, red);
  if (threadIdx.x == 0) { double* r = part + ((size_t)b * LOSS_GX + blockIdx.x) * 3; r[0] = s0; r[1] = s1; r[2] = s2; }
}
__global__ void __launch_bounds__(256) dice_finalize_kernel(double* __restrict__ sums, const double* __restrict__ part,
                                                            float* __restrict__ loss, int B, int gx, float smooth) {
  __shared__ double score[256];
  double mine = 0.0;
  for (int b = threadIdx.x; b < B; b += 256) {   // a thread owns whole images; images are added in index order below
    double s0 = 0, s1 = 0, s2 = 0;
    const double* r = part + (size_t)b * LOSS_GX * 3;
    for (int x = 0; x < gx; ++x) { s0 += r[x * 3]; s1 += r[x * 3 + 1]; s2 += r[x * 3 + 2]; }
    sums[b * 3] = s0; sums[b * 3 + 1] = s1; sums[b * 3 + 2] = s2;
    const float inter = (float)s0, pp = (float)s1, tt = (float)s2;
    mine += (double)(2.0f * (inter + smooth) / (pp + tt + smooth));
  }
  score[threadIdx.x] = mine;
  __syncthreads();
  if (threadIdx.x == 0) {
    double acc = 0.0;
    for (int i = 0; i < 256 && i < B; ++i) acc += score[i];
    *loss = 1.f - (float)(acc / B);
  }
}
__global__ void dice_bwd_kernel(const float* __restrict__ logits, const float* __restrict__ targets,
                                const double* __restrict__ sums, const float* __restrict__ gout,
                                float* __restrict__ dlogits, int B, int HW, float smooth) {
  const long long n = (long long)B * HW;
  const float g = gout ? *gout : 1.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / HW);
    const float inter = (float)sums[b * 3] + smooth, den = (float)sums[b * 3 + 1] + (float)sums[b * 3 + 2] + smooth;
    const float p = 1.f / (1.f + expf(-logits[i])), t = targets[i];
    const float dscore = 2.f * t / den - 4.f * inter * p / (den * den);
    dlogits[i] = g * (-1.f / B) * dscore * p * (1.f - p);
  }
}

// ---------------------------------------------------------------- SSI
// scratch (doubles): per image b: S[b*8 + {0:a00,1:a01,2:a11,3:b0,4:b1,5:scale,6:shift,7:valid}],
//   then global G[0]=mse_num, G[1]=Mtot, G[2+k]=grad_num_k, G[10+k]=M_k (k < 8), then per image adjoint sums
//   A[b*2+{0:G1,1:G0}], the per-image backward weights Wt[b][9], then the per-block partial rows P[B][LOSS_GX][SSI_PW] every reducing kernel writes (block (x, b)
//   -> row (b, x)) and the next small kernel adds in index order: no atomics, bit-reproducible
constexpr int SSI_MAX_SCALES = 8;
constexpr int SSI_PW = 1 + 2 * SSI_MAX_SCALES;   // widest partial row: mse + grad_num[8] + M[8]
__device__ __forceinline__ double* ssi_S(double* s, int b) { return s + b * 8; }
__device__ __forceinline__ double* ssi_G(double* s, int B) { return s + B * 8; }
__device__ __forceinline__ double* ssi_A(double* s, int B, int b) { return s + B * 8 + 24 + b * 2; }
// per-image backward weights: Wt[b][0] = d total / d (sum m d^2 / 2) of image b, Wt[b][1+k] = d total / d (grad sum k)
__device__ __forceinline__ double* ssi_Wt(double* s, int B, int b) { return s + B * 10 + 24 + b * (1 + SSI_MAX_SCALES); }
__device__ __forceinline__ double* ssi_P(double* s, int B) { return s + B * (11 + SSI_MAX_SCALES) + 24; }

__global__ void __launch_bounds__(256) ssi_sums_kernel(const float* __restrict__ pred, const float* __restrict__ targ,
                                                       double* __restrict__ scratch, int B, int HW) {
  __shared__ double red[8];
  const int b = blockIdx.y;
  double a00 = 0, a01 = 0, a11 = 0, b0 = 0, b1 = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < HW; i += gridDim.x * 256) {
    const float p = pred[(long long)b * HW + i], t = targ[(long long)b * HW + i];
    if (t > 0.f) { a00 += p * p; a01 += p; a11 += 1.0; b0 += p * t; b1 += t; }
  }
  a00 = block_sum_d(a00, red); a01 = block_sum_d(a01, red); a11 = block_sum_d(a11, red);
  b0 = block_sum_d(b0, red); b1 = block_sum_d(b1, red);
  if (threadIdx.x == 0) {
    double* r = ssi_P(scratch, B) + ((size_t)b * LOSS_GX + blockIdx.x) * SSI_PW;
    r[0] = a00; r[1] = a01; r[2] = a11; r[3] = b0; r[4] = b1;
  }
}
__global__ void ssi_solve_kernel(double* __restrict__ scratch, int B, int gx) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double* S = ssi_S(scratch, b);
  {
    double t[5] = {0, 0, 0, 0, 0};
    const double* r = ssi_P(scratch, B) + (size_t)b * LOSS_GX * SSI_PW;
    for (int x = 0; x < gx; ++x)
      for (int k = 0; k < 5; ++k) t[k] += r[x * SSI_PW + k];
    for (int k = 0; k < 5; ++k) S[k] = t[k];
  }
  // the reference evaluates this in fp32 (losses.py:18-23); keep its `det != 0` selection on the fp32 value
  const float a00 = (float)S[0], a01 = (float)S[1], a11 = (float)S[2], b0 = (float)S[3], b1 = (float)S[4];
  const float det = a00 * a11 - a01 * a01;
  if (det != 0.f) {
    const double dd = S[0] * S[2] - S[1] * S[1];
    S[5] = (S[2] * S[3] - S[1] * S[4]) / dd;
    S[6] = (-S[1] * S[3] + S[0] * S[4]) / dd;
    S[7] = 1.0;
  } else { S[5] = 0.0; S[6] = 0.0; S[7] = 0.0; }
  (void)b0; (void)b1;
}
__device__ __forceinline__ float ssi_diff(const float* p, const float* t, long long i, float s, float sh) {
  const float tv = t[i];
  return tv > 0.f ? s * p[i] + sh - tv : 0.f;
}
__global__ void __launch_bounds__(256) ssi_terms_kernel(const float* __restrict__ pred, const float* __restrict__ targ,
                                                        double* __restrict__ scratch, float* __restrict__ ssi_out,
                                                        int B, int H, int W, int scales) {
  __shared__ double red[8];
  const int b = blockIdx.y, HW = H * W;
  const float s = (float)ssi_S(scratch, b)[5], sh = (float)ssi_S(scratch, b)[6];
  const float* p = pred + (long long)b * HW;
  const float* t = targ + (long long)b * HW;
  double mse = 0, gn[SSI_MAX_SCALES], mk[SSI_MAX_SCALES];
#pragma unroll
  for (int k = 0; k < SSI_MAX_SCALES; ++k) { gn[k] = 0; mk[k] = 0; }
  for (int i = blockIdx.x * 256 + threadIdx.x; i < HW; i += gridDim.x * 256) {
    const int y = i / W, x = i % W;
    if (ssi_out) ssi_out[(long long)b * HW + i] = s * p[i] + sh;
    const bool m = t[i] > 0.f;
    const float d = ssi_diff(p, t, i, s, sh);
    mse += (double)(d * d);
#pragma unroll
    for (int k = 0; k < SSI_MAX_SCALES; ++k) {
      if (k >= scales) break;
      const int step = 1 << k;
      if ((y & (step - 1)) || (x & (step - 1))) break;
      if (m) mk[k] += 1.0;
      if (x + step < W && m && t[i + step] > 0.f) gn[k] += fabsf(ssi_diff(p, t, i + step, s, sh) - d);
      if (y + step < H && m && t[i + step * W] > 0.f) gn[k] += fabsf(ssi_diff(p, t, i + step * W, s, sh) - d);
    }
  }
  double* r = ssi_P(scratch, B) + ((size_t)b * LOSS_GX + blockIdx.x) * SSI_PW;
  mse = block_sum_d(mse, red);
  if (threadIdx.x == 0) r[0] = mse;
#pragma unroll
  for (int k = 0; k < SSI_MAX_SCALES; ++k) {
    if (k >= scales) break;
    const double a = block_sum_d(gn[k], red), c = block_sum_d(mk[k], red);
    if (threadIdx.x == 0) { r[1 + k] = a; r[1 + SSI_MAX_SCALES + k] = c; }
  }
}
// Combines the per-block rows into per-image totals (a thread owns whole images), then one thread per component adds the
// images in index order.  image_based != 0: reduction="image-based" (Depth_estimation/Metrics/losses.py:40-48, 82-85,
// 99-102): every image's term is divided by its own mask count and the mean over images is taken; otherwise
// reduction_batch_based (:28-37).  Also writes the per-image weights the backward pass multiplies with.
__global__ void __launch_bounds__(256) ssi_finalize_kernel(double* __restrict__ scratch, float* __restrict__ loss, int B, int gx,
                                                           float alpha, int scales, int image_based) {
  __shared__ double tot[256][SSI_PW];
  __shared__ double run[SSI_PW], img[SSI_PW];
  __shared__ double s_mt;
  double* G = ssi_G(scratch, B);
  if (threadIdx.x < SSI_PW) { run[threadIdx.x] = 0.0; img[threadIdx.x] = 0.0; }
  if (threadIdx.x == 0) s_mt = 0.0;
  __syncthreads();
  for (int b0 = 0; b0 < B; b0 += 256) {
    const int b = b0 + threadIdx.x;
    if (b < B) {
      const double* r = ssi_P(scratch, B) + (size_t)b * LOSS_GX * SSI_PW;
      for (int c = 0; c < SSI_PW; ++c) {
        const bool used = c == 0 || ((c - 1) % SSI_MAX_SCALES) < scales;
        double t = 0.0;
        if (used) for (int x = 0; x < gx; ++x) t += r[x * SSI_PW + c];
        tot[threadIdx.x][c] = t;
      }
    }
    __syncthreads();
    const int nb = min(256, B - b0);
    if (threadIdx.x < SSI_PW) {
      const int c = threadIdx.x;
      double a = run[c], ia = img[c];
      for (int i = 0; i < nb; ++i) {
        const double v = tot[i][c];
        a += v;
        if (c == 0) { const double m = ssi_S(scratch, b0 + i)[2]; ia += m > 0.0 ? v / (2.0 * m) : v; }
        else if (c <= SSI_MAX_SCALES) { const double m = tot[i][c + SSI_MAX_SCALES]; ia += m > 0.0 ? v / m : v; }
      }
      run[c] = a; img[c] = ia;
    } else if (threadIdx.x == 32) {
      double mt = s_mt;
      for (int i = 0; i < nb; ++i) mt += ssi_S(scratch, b0 + i)[2];
      s_mt = mt;
    }
    __syncthreads();
    if (b < B && image_based) {   // the weights of image-based reduction only need the image's own counts
      double* w = ssi_Wt(scratch, B, b);
      const double m = ssi_S(scratch, b)[2];
      w[0] = (m > 0.0 ? 1.0 / m : 1.0) / B;
      for (int k = 0; k < SSI_MAX_SCALES; ++k) {
        const double mk = tot[threadIdx.x][1 + SSI_MAX_SCALES + k];
        w[1 + k] = (k < scales && alpha > 0.f) ? (double)alpha * (mk > 0.0 ? 1.0 / mk : 1.0) / B : 0.0;
      }
    }
    __syncthreads();
  }
  const double mt = s_mt;
  if (threadIdx.x == 0) {
    G[0] = run[0]; G[1] = mt;
    for (int k = 0; k < SSI_MAX_SCALES; ++k) { G[2 + k] = run[1 + k]; G[10 + k] = run[1 + SSI_MAX_SCALES + k]; }
    double total;
    if (image_based) {
      total = img[0] / B;
      if (alpha > 0.f) for (int k = 0; k < scales; ++k) total += (double)alpha * img[1 + k] / B;
    } else {
      total = mt > 0.0 ? run[0] / (2.0 * mt) : 0.0;
      if (alpha > 0.f) for (int k = 0; k < scales; ++k) total += run[1 + SSI_MAX_SCALES + k] > 0.0 ? (double)alpha * run[1 + k] / run[1 + SSI_MAX_SCALES + k] : 0.0;
    }
    *loss = (float)total;
  }
  if (!image_based) {
    for (int b = threadIdx.x; b < B; b += 256) {
      double* w = ssi_Wt(scratch, B, b);
      w[0] = mt > 0.0 ? 1.0 / mt : 0.0;
      for (int k = 0; k < SSI_MAX_SCALES; ++k) {
        const double mk = run[1 + SSI_MAX_SCALES + k];
        w[1 + k] = (k < scales && alpha > 0.f && mk > 0.0) ? (double)alpha / mk : 0.0;
      }
    }
  }
}
// pass A of backward: q_i = d total / d ssi_i -> dpred (temporarily), and per-image G1 = sum q p, G0 = sum q
__global__ void __launch_bounds__(256) ssi_bwd_q_kernel(const float* __restrict__ pred, const float* __restrict__ targ,
                                                        double* __restrict__ scratch, float* __restrict__ dpred,
                                                        int B, int H, int W, float alpha, int scales) {
  __shared__ double red[8];
  const int b = blockIdx.y, HW = H * W;
  const float s = (float)ssi_S(scratch, b)[5], sh = (float)ssi_S(scratch, b)[6];
  const double* wt = ssi_Wt(scratch, B, b);
  const float inv_m = (float)wt[0];
  float inv_k[SSI_MAX_SCALES];
#pragma unroll
  for (int k = 0; k < SSI_MAX_SCALES; ++k) inv_k[k] = (float)wt[1 + k];
  (void)alpha;
  const float* p = pred + (long long)b * HW;
  const float* t = targ + (long long)b * HW;
  double g1 = 0, g0 = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < HW; i += gridDim.x * 256) {
    const int y = i / W, x = i % W;
    float q = 0.f;
    if (t[i] > 0.f) {
      const float d = ssi_diff(p, t, i, s, sh);
      q = d * inv_m;
#pragma unroll
      for (int k = 0; k < SSI_MAX_SCALES; ++k) {
        if (k >= scales) break;
        const int step = 1 << k;
        if ((y & (step - 1)) || (x & (step - 1))) break;
        float sg = 0.f;
        auto edge = [&](int j) {
          if (t[j] > 0.f) { const float e = d - ssi_diff(p, t, j, s, sh); sg += (e > 0.f) - (e < 0.f); }
        };
        if (x + step < W) edge(i + step);
        if (x - step >= 0) edge(i - step);
        if (y + step < H) edge(i + step * W);
        if (y - step >= 0) edge(i - step * W);
        q += sg * inv_k[k];
      }
    }
    dpred[(long long)b * HW + i] = q;
    g1 += (double)q * p[i]; g0 += (double)q;
  }
  g1 = block_sum_d(g1, red); g0 = block_sum_d(g0, red);
  if (threadIdx.x == 0) { double* r = ssi_P(scratch, B) + ((size_t)b * LOSS_GX + blockIdx.x) * SSI_PW; r[0] = g1; r[1] = g0; }
}
__global__ void ssi_bwd_sum_kernel(double* __restrict__ scratch, int B, int gx) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double* r = ssi_P(scratch, B) + (size_t)b * LOSS_GX * SSI_PW;
  double g1 = 0.0, g0 = 0.0;
  for (int x = 0; x < gx; ++x) { g1 += r[x * SSI_PW]; g0 += r[x * SSI_PW + 1]; }
  ssi_A(scratch, B, b)[0] = g1; ssi_A(scratch, B, b)[1] = g0;
}
// pass B: dpred_i = gout * (q_i * s + m_i (2 p_i A00 + A01 + t_i B0)) — the terms through d(scale,shift)/d(pred)
__global__ void ssi_bwd_final_kernel(const float* __restrict__ pred, const float* __restrict__ targ,
                                     double* __restrict__ scratch, const float* __restrict__ gout,
                                     float* __restrict__ dpred, int B, int HW) {
  const long long n = (long long)B * HW;
  const float g = gout ? *gout : 1.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(i / HW);
    const double* S = ssi_S(scratch, b);
    float out = 0.f;
    if (S[7] != 0.0) {
      const double a00 = S[0], a01 = S[1], a11 = S[2], b0 = S[3], b1 = S[4], s = S[5], sh = S[6];
      const double det = a00 * a11 - a01 * a01;
      const double G1 = ssi_A(scratch, B, b)[0], G0 = ssi_A(scratch, B, b)[1];
      const double A00 = (G1 * (-s * a11) + G0 * (b1 - sh * a11)) / det;
      const double A01 = (G1 * (-b1 + 2.0 * a01 * s) + G0 * (-b0 + 2.0 * a01 * sh)) / det;
      const double B0 = (G1 * a11 + G0 * (-a01)) / det;
      const float p = pred[i], t = targ[i];
      double v = (double)dpred[i] * s;
      if (t > 0.f) v += 2.0 * p * A00 + A01 + t * B0;
      out = (float)v;
    }
    dpred[i] = g * out;
  }
}
}  // namespace

size_t ssi_scratch_floats(int B) { return 2 * ((size_t)B * (11 + SSI_MAX_SCALES) + 24 + (size_t)B * LOSS_GX * SSI_PW); }  // doubles, expressed in floats
size_t dice_scratch_doubles(int B) { return (size_t)B * 3 * (1 + LOSS_GX); }
static inline int loss_gx(int HW) { return max(1, min(cdiv(HW, 2048), LOSS_GX)); }

int k_dice_fwd(const float* logits, const float* targets, float* sums, float* loss, int B, int HW, float smooth, cudaStream_t st) {
  double* s = reinterpret_cast<double*>(sums);
  double* part = s + (size_t)B * 3;
  const int gx = loss_gx(HW);
  dice_sums_kernel<<<dim3(gx, B), 256, 0, st>>>(logits, targets, part, HW);
  S4G_KERNEL_CHECK();
  dice_finalize_kernel<<<1, 256, 0, st>>>(s, part, loss, B, gx, smooth);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_dice_bwd(const float* logits, const float* targets, const float* sums, const float* gout, float* dlogits, int B, int HW, float smooth, cudaStream_t st) {
  dice_bwd_kernel<<<min(cdiv((long long)B * HW, 256), 148 * 16), 256, 0, st>>>(logits, targets, reinterpret_cast<const double*>(sums), gout, dlogits, B, HW, smooth);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_ssi_fwd(const float* pred, const float* targ, float* scratch, float* loss, float* ssi_out, int B, int H, int W, float alpha, int scales, int image_based, cudaStream_t st) {
  if (scales > SSI_MAX_SCALES || scales < 0) { s4g_set_error("ssi: scales must be in [0, 8]"); return S4G_ERR_SHAPE; }
  double* s = reinterpret_cast<double*>(scratch);
  const int HW = H * W;
  const int gx = loss_gx(HW);
  dim3 grid(gx, B);
  ssi_sums_kernel<<<grid, 256, 0, st>>>(pred, targ, s, B, HW);
  S4G_KERNEL_CHECK();
  ssi_solve_kernel<<<cdiv(B, 128), 128, 0, st>>>(s, B, gx);
  S4G_KERNEL_CHECK();
  ssi_terms_kernel<<<grid, 256, 0, st>>>(pred, targ, s, ssi_out, B, H, W, alpha > 0.f ? scales : 0);
  S4G_KERNEL_CHECK();
  ssi_finalize_kernel<<<1, 256, 0, st>>>(s, loss, B, gx, alpha, alpha > 0.f ? scales : 0, image_based);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_ssi_bwd(const float* pred, const float* targ, const float* scratch, const float* gout, float* dpred, int B, int H, int W, float alpha, int scales, cudaStream_t st) {
  if (scales > SSI_MAX_SCALES || scales < 0) { s4g_set_error("ssi: scales must be in [0, 8]"); return S4G_ERR_SHAPE; }
  double* s = reinterpret_cast<double*>(const_cast<float*>(scratch));
  const int HW = H * W;
  const int gx = loss_gx(HW);
  dim3 grid(gx, B);
  ssi_bwd_q_kernel<<<grid, 256, 0, st>>>(pred, targ, s, dpred, B, H, W, alpha, alpha > 0.f ? scales : 0);
  S4G_KERNEL_CHECK();
  ssi_bwd_sum_kernel<<<cdiv(B, 128), 128, 0, st>>>(s, B, gx);
  S4G_KERNEL_CHECK();
  ssi_bwd_final_kernel<<<min(cdiv((long long)B * HW, 256), 148 * 16), 256, 0, st>>>(pred, targ, s, gout, dpred, B, HW);
  S4G_KERNEL_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------ weighted cross-entropy
// nn.CrossEntropyLoss(weight) as used at Classification/train_classification.py:278:
//   loss = sum_i w[y_i] * (logsumexp(x_i) - x_i[y_i]) / sum_i w[y_i];  dL/dx_i = w[y_i] * (softmax(x_i) - onehot(y_i)) / sum w
// One block (B <= a few hundred rows, N ~ 23 classes): a warp per row, block reduction of the two sums, and in the same
// launch the gradient (scaled by grad_out = 1 when called from forward; backward only rescales).
namespace {
__global__ void __launch_bounds__(256) ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels,
                                                 const float* __restrict__ weight, float* __restrict__ loss,
                                                 float* __restrict__ dlogits, int B, int N) {
  __shared__ float s_num[8], s_den[8];
  __shared__ float s_inv;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float num = 0.f, den = 0.f;
  for (int b = warp; b < B; b += 8) {
    const float* x = logits + (long long)b * N;
    float mx = -INFINITY;
    for (int c = lane; c < N; c += 32) mx = fmaxf(mx, x[c]);
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < N; c += 32) se += __expf(x[c] - mx);
    se = warp_sum(se);
    const int y = (int)labels[b];
    const float w = weight ? weight[y] : 1.f;
    const float lse = mx + __logf(se);
    num += w * (lse - x[y]);
    den += w;
    for (int c = lane; c < N; c += 32) dlogits[(long long)b * N + c] = w * (__expf(x[c] - lse) - (c == y ? 1.f : 0.f));
  }
  if (lane == 0) { s_num[warp] = num; s_den[warp] = den; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, d = 0.f;
    for (int k = 0; k < 8; ++k) { a += s_num[k]; d += s_den[k]; }
    *loss = a / d;
    s_inv = 1.f / d;
  }
  __syncthreads();
  const float inv = s_inv;
  for (int i = threadIdx.x; i < B * N; i += 256) dlogits[i] *= inv;
}
__global__ void scale_kernel(const float* __restrict__ in, const float* __restrict__ g, float* __restrict__ out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i] * g[0];
}
}  // namespace
int k_ce_fwd(const float* logits, const long long* labels, const float* weight, float* loss, float* dlogits_unit, int B, int N,
             cudaStream_t st) {
  ce_kernel<<<1, 256, 0, st>>>(logits, labels, weight, loss, dlogits_unit, B, N);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_ce_bwd(const float* dlogits_unit, const float* gout, float* dlogits, int B, int N, cudaStream_t st) {
  scale_kernel<<<cdiv((long long)B * N, 256), 256, 0, st>>>(dlogits_unit, gout, dlogits, B * N);
  S4G_KERNEL_CHECK();
  return 0;
}
