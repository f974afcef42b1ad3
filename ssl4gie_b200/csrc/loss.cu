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
// sums (double) per image: [0] sum p*t, [1] sum p*p, [2] sum t*t
__global__ void __launch_bounds__(256) dice_sums_kernel(const float* __restrict__ logits, const float* __restrict__ targets,
                                                        double* __restrict__ sums, int HW) {
  __shared__ double red[8];
  const int b = blockIdx.y;
  double s0 = 0, s1 = 0, s2 = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < HW; i += gridDim.x * 256) {
    const float p = 1.f / (1.f + expf(-logits[(long long)b * HW + i]));
    const float t = targets[(long long)b * HW + i];
    s0 += p * t; s1 += p * p; s2 += t * t;
  }
  s0 = block_sum_d(s0, red); s1 = block_sum_d(s1, red); s2 = block_sum_d(s2, red);
  if (threadIdx.x == 0) { atomicAdd(&sums[b * 3], s0); atomicAdd(&sums[b * 3 + 1], s1); atomicAdd(&sums[b * 3 + 2], s2); }
}
__global__ void dice_finalize_kernel(const double* __restrict__ sums, float* __restrict__ loss, int B, float smooth) {
  if (threadIdx.x || blockIdx.x) return;
  double acc = 0.0;
  for (int b = 0; b < B; ++b) {
    const float inter = (float)sums[b * 3], pp = (float)sums[b * 3 + 1], tt = (float)sums[b * 3 + 2];
    acc += 2.0f * (inter + smooth) / (pp + tt + smooth);
  }
  *loss = 1.f - (float)(acc / B);
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
//   then global G[0]=mse_num, G[1]=Mtot, G[2+k]=grad_num_k, G[6+k]=M_k, then per image adjoint sums A[b*2+{0:G1,1:G0}]
__device__ __forceinline__ double* ssi_S(double* s, int b) { return s + b * 8; }
__device__ __forceinline__ double* ssi_G(double* s, int B) { return s + B * 8; }
__device__ __forceinline__ double* ssi_A(double* s, int B, int b) { return s + B * 8 + 16 + b * 2; }

__global__ void __launch_bounds__(256) ssi_sums_kernel(const float* __restrict__ pred, const float* __restrict__ targ,
                                                       double* __restrict__ scratch, int HW) {
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
    double* S = ssi_S(scratch, b);
    atomicAdd(&S[0], a00); atomicAdd(&S[1], a01); atomicAdd(&S[2], a11); atomicAdd(&S[3], b0); atomicAdd(&S[4], b1);
  }
}
__global__ void ssi_solve_kernel(double* __restrict__ scratch, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double* S = ssi_S(scratch, b);
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
  double mse = 0, gn[4] = {0, 0, 0, 0}, mk[4] = {0, 0, 0, 0};
  for (int i = blockIdx.x * 256 + threadIdx.x; i < HW; i += gridDim.x * 256) {
    const int y = i / W, x = i % W;
    if (ssi_out) ssi_out[(long long)b * HW + i] = s * p[i] + sh;
    const bool m = t[i] > 0.f;
    const float d = ssi_diff(p, t, i, s, sh);
    mse += (double)(d * d);
    for (int k = 0; k < scales; ++k) {
      const int step = 1 << k;
      if ((y & (step - 1)) || (x & (step - 1))) break;
      if (m) mk[k] += 1.0;
      if (x + step < W && m && t[i + step] > 0.f) gn[k] += fabsf(ssi_diff(p, t, i + step, s, sh) - d);
      if (y + step < H && m && t[i + step * W] > 0.f) gn[k] += fabsf(ssi_diff(p, t, i + step * W, s, sh) - d);
    }
  }
  double* G = ssi_G(scratch, B);
  mse = block_sum_d(mse, red);
  if (threadIdx.x == 0) atomicAdd(&G[0], mse);
  for (int k = 0; k < scales; ++k) {
    const double a = block_sum_d(gn[k], red), c = block_sum_d(mk[k], red);
    if (threadIdx.x == 0) { atomicAdd(&G[2 + k], a); atomicAdd(&G[6 + k], c); }
  }
}
__global__ void ssi_finalize_kernel(double* __restrict__ scratch, float* __restrict__ loss, int B, float alpha, int scales) {
  if (threadIdx.x || blockIdx.x) return;
  double* G = ssi_G(scratch, B);
  double mt = 0.0;
  for (int b = 0; b < B; ++b) mt += ssi_S(scratch, b)[2];
  G[1] = mt;
  double total = mt > 0.0 ? G[0] / (2.0 * mt) : 0.0;
  if (alpha > 0.f)
    for (int k = 0; k < scales; ++k) total += G[6 + k] > 0.0 ? alpha * G[2 + k] / G[6 + k] : 0.0;
  *loss = (float)total;
}
// pass A of backward: q_i = d total / d ssi_i -> dpred (temporarily), and per-image G1 = sum q p, G0 = sum q
__global__ void __launch_bounds__(256) ssi_bwd_q_kernel(const float* __restrict__ pred, const float* __restrict__ targ,
                                                        double* __restrict__ scratch, float* __restrict__ dpred,
                                                        int B, int H, int W, float alpha, int scales) {
  __shared__ double red[8];
  const int b = blockIdx.y, HW = H * W;
  const double* G = ssi_G(scratch, B);
  const float s = (float)ssi_S(scratch, b)[5], sh = (float)ssi_S(scratch, b)[6];
  const float inv_m = G[1] > 0.0 ? (float)(1.0 / G[1]) : 0.f;
  float inv_k[4];
  for (int k = 0; k < 4; ++k) inv_k[k] = (k < scales && alpha > 0.f && G[6 + k] > 0.0) ? (float)(alpha / G[6 + k]) : 0.f;
  const float* p = pred + (long long)b * HW;
  const float* t = targ + (long long)b * HW;
  double g1 = 0, g0 = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < HW; i += gridDim.x * 256) {
    const int y = i / W, x = i % W;
    float q = 0.f;
    if (t[i] > 0.f) {
      const float d = ssi_diff(p, t, i, s, sh);
      q = d * inv_m;
      for (int k = 0; k < scales; ++k) {
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
  if (threadIdx.x == 0) { atomicAdd(&ssi_A(scratch, B, b)[0], g1); atomicAdd(&ssi_A(scratch, B, b)[1], g0); }
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

size_t ssi_scratch_floats(int B) { return 2 * ((size_t)B * 10 + 16); }  // doubles, expressed in floats

int k_dice_fwd(const float* logits, const float* targets, float* sums, float* loss, int B, int HW, float smooth, cudaStream_t st) {
  double* s = reinterpret_cast<double*>(sums);
  S4G_CUDA(cudaMemsetAsync(s, 0, sizeof(double) * 3 * B, st));
  dice_sums_kernel<<<dim3(max(1, min(cdiv(HW, 2048), 64)), B), 256, 0, st>>>(logits, targets, s, HW);
  S4G_KERNEL_CHECK();
  dice_finalize_kernel<<<1, 32, 0, st>>>(s, loss, B, smooth);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_dice_bwd(const float* logits, const float* targets, const float* sums, const float* gout, float* dlogits, int B, int HW, float smooth, cudaStream_t st) {
  dice_bwd_kernel<<<min(cdiv((long long)B * HW, 256), 148 * 16), 256, 0, st>>>(logits, targets, reinterpret_cast<const double*>(sums), gout, dlogits, B, HW, smooth);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_ssi_fwd(const float* pred, const float* targ, float* scratch, float* loss, float* ssi_out, int B, int H, int W, float alpha, int scales, cudaStream_t st) {
  if (scales > 4) { s4g_set_error("ssi: scales must be <= 4"); return S4G_ERR_SHAPE; }
  double* s = reinterpret_cast<double*>(scratch);
  S4G_CUDA(cudaMemsetAsync(s, 0, sizeof(float) * ssi_scratch_floats(B), st));
  const int HW = H * W;
  dim3 grid(max(1, min(cdiv(HW, 2048), 64)), B);
  ssi_sums_kernel<<<grid, 256, 0, st>>>(pred, targ, s, HW);
  S4G_KERNEL_CHECK();
  ssi_solve_kernel<<<cdiv(B, 128), 128, 0, st>>>(s, B);
  S4G_KERNEL_CHECK();
  ssi_terms_kernel<<<grid, 256, 0, st>>>(pred, targ, s, ssi_out, B, H, W, alpha > 0.f ? scales : 0);
  S4G_KERNEL_CHECK();
  ssi_finalize_kernel<<<1, 32, 0, st>>>(s, loss, B, alpha, scales);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_ssi_bwd(const float* pred, const float* targ, const float* scratch, const float* gout, float* dpred, int B, int H, int W, float alpha, int scales, cudaStream_t st) {
  double* s = reinterpret_cast<double*>(const_cast<float*>(scratch));
  const int HW = H * W;
  S4G_CUDA(cudaMemsetAsync(s + (size_t)B * 8 + 16, 0, sizeof(double) * 2 * B, st));
  dim3 grid(max(1, min(cdiv(HW, 2048), 64)), B);
  ssi_bwd_q_kernel<<<grid, 256, 0, st>>>(pred, targ, s, dpred, B, H, W, alpha, alpha > 0.f ? scales : 0);
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
