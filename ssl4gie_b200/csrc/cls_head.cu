// Classification branch tail (reference Models/models.py:338,348-354): after the final LayerNorm, pool the tokens
// (x[:, 0] for out_token == "cls", x[:, 1:].mean(1) for "spatial") and apply lin_head (nn.Linear(768, num_classes),
// Models/models.py:294-295).  Sizes are tiny (B x 768 -> B x ~23): plain fp32 kernels on the fp32 master weights, fp32
// accumulation; the cost of the branch is the encoder.
#include "kernels.h"

namespace {

constexpr int NTOK = 197;

// pooled[b, c] (fp32) from hn [B*197, C]
template <typename T>
__global__ void pool_tokens_kernel(const T* __restrict__ hn, float* __restrict__ pooled, int B, int C, int spatial) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * C) return;
  const int b = idx / C, c = idx % C;
  const T* p = hn + (long long)b * NTOK * C + c;
  if (!spatial) { pooled[idx] = to_f(p[0]); return; }
  float s = 0.f;
  for (int t = 1; t < NTOK; ++t) s += to_f(p[(long long)t * C]);
  pooled[idx] = s * (1.0f / (NTOK - 1));
}

// out[b, n] = pooled[b, :] . W[n, :] + bias[n]; one warp per output
__global__ void head_fwd_kernel(const float* __restrict__ pooled, const float* __restrict__ W, const float* __restrict__ bias,
                                float* __restrict__ out, int B, int N, int C) {
  const int o = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (o >= B * N) return;
  const int b = o / N, n = o % N;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += pooled[(long long)b * C + c] * W[(long long)n * C + c];
  s = warp_sum(s);
  if (lane == 0) out[o] = s + (bias ? bias[n] : 0.f);
}

// dW[n, c] = sum_b dout[b, n] * pooled[b, c]; db[n] = sum_b dout[b, n]
__global__ void head_bwd_w_kernel(const float* __restrict__ dout, const float* __restrict__ pooled, float* __restrict__ dW,
                                  float* __restrict__ db, int B, int N, int C) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= N * C) return;
  const int n = idx / C, c = idx % C;
  float s = 0.f, sb = 0.f;
  for (int b = 0; b < B; ++b) {
    const float d = dout[b * N + n];
    s += d * pooled[(long long)b * C + c];
    sb += d;
  }
  dW[idx] = s;
  if (c == 0 && db) db[n] = sb;
}

// dpooled[b, c] = sum_n dout[b, n] * W[n, c]  (no head: dpooled = dout), scattered to the token rows of dhn:
// "cls": row 0 gets dpooled, rows 1.. get 0; "spatial": row 0 gets 0, rows 1.. get dpooled / 196
template <typename T>
__global__ void head_bwd_x_kernel(const float* __restrict__ dout, const float* __restrict__ W, T* __restrict__ dhn, int B,
                                  int N, int C, int spatial) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * C) return;
  const int b = idx / C, c = idx % C;
  float s = 0.f;
  if (W) {
    for (int n = 0; n < N; ++n) s += dout[b * N + n] * W[(long long)n * C + c];
  } else {
    s = dout[idx];
  }
  T* p = dhn + (long long)b * NTOK * C + c;
  const T first = from_f<T>(spatial ? 0.f : s);
  const T rest = from_f<T>(spatial ? s * (1.0f / (NTOK - 1)) : 0.f);
  p[0] = first;
  for (int t = 1; t < NTOK; ++t) p[(long long)t * C] = rest;
}

}  // namespace

template <typename T>
int k_cls_head_fwd(const T* hn, const float* W, const float* bias, float* pooled, float* out, int B, int N, int C,
                   int spatial, cudaStream_t st) {
  pool_tokens_kernel<T><<<cdiv((long long)B * C, 256), 256, 0, st>>>(hn, W ? pooled : out, B, C, spatial);
  S4G_KERNEL_CHECK();
  if (W) {
    head_fwd_kernel<<<cdiv((long long)B * N, 8), 256, 0, st>>>(pooled, W, bias, out, B, N, C);
    S4G_KERNEL_CHECK();
  }
  return 0;
}
template <typename T>
int k_cls_head_bwd(const float* dout, const float* pooled, const float* W, float* dW, float* db, T* dhn, int B, int N,
                   int C, int spatial, cudaStream_t st) {
  if (W && dW) {
    head_bwd_w_kernel<<<cdiv((long long)N * C, 256), 256, 0, st>>>(dout, pooled, dW, db, B, N, C);
    S4G_KERNEL_CHECK();
  }
  if (dhn) {
    head_bwd_x_kernel<T><<<cdiv((long long)B * C, 256), 256, 0, st>>>(dout, W, dhn, B, N, C, spatial);
    S4G_KERNEL_CHECK();
  }
  return 0;
}

template int k_cls_head_fwd<float>(const float*, const float*, const float*, float*, float*, int, int, int, int, cudaStream_t);
template int k_cls_head_fwd<bf16>(const bf16*, const float*, const float*, float*, float*, int, int, int, int, cudaStream_t);
template int k_cls_head_bwd<float>(const float*, const float*, const float*, float*, float*, float*, int, int, int, int, cudaStream_t);
template int k_cls_head_bwd<bf16>(const float*, const float*, const float*, float*, float*, bf16*, int, int, int, int, cudaStream_t);
