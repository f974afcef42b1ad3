// Fused AdamW over one flat fp32 parameter arena (reference: torch.optim.AdamW(model.parameters(), lr) at
// Binary_segmentation/train_segmentation.py:204; defaults betas (0.9, 0.999), eps 1e-8, weight_decay 0.01).
// HBM-bound: 28 B/param (read p,g,m,v; write p,m,v).
#include "kernels.h"

__global__ void __launch_bounds__(256) adamw_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m,
                                                    float4* __restrict__ v, long long n4, float lr, float b1, float b2, float eps,
                                                    float wd, float bc1, float bc2_sqrt, float gs,
                                                    const float* __restrict__ wd_slot) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pv = p[i], gv = g[i], mv = m[i], vv = v[i];
    if (wd_slot) wd = wd_slot[i >> 4];   // one weight decay per 64-float arena slot (parameter groups)
    float* pp = &pv.x; float* gp = &gv.x; float* mp = &mv.x; float* vp = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = gp[k] * gs;
      pp[k] *= 1.f - lr * wd;
      mp[k] = b1 * mp[k] + (1.f - b1) * gr;
      vp[k] = b2 * vp[k] + (1.f - b2) * gr * gr;
      const float denom = sqrtf(vp[k]) / bc2_sqrt + eps;
      pp[k] -= (lr / bc1) * (mp[k] / denom);
    }
    p[i] = pv; m[i] = mv; v[i] = vv;
  }
}
int k_adamw(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2, float eps, float wd,
            int step, float grad_scale, cudaStream_t st, const float* wd_slot) {
  if (n % 4) { s4g_set_error("adamw: arena length must be a multiple of 4"); return S4G_ERR_SHAPE; }
  const float bc1 = 1.f - powf(b1, (float)step), bc2 = 1.f - powf(b2, (float)step);
  adamw_kernel<<<148 * 8, 256, 0, st>>>(reinterpret_cast<float4*>(p), reinterpret_cast<const float4*>(g),
                                        reinterpret_cast<float4*>(m), reinterpret_cast<float4*>(v), n / 4, lr, b1, b2, eps, wd,
                                        bc1, sqrtf(bc2), grad_scale, wd_slot);
  S4G_KERNEL_CHECK();
  return 0;
}

// ---- device-resident optimizer state (CUDA-graph replays: by-value lr / step / seed arguments would be baked into the graph) ----
// hyper[8] = {lr, beta1, beta2, eps, weight_decay, grad_scale, -, -}; step counts COMPLETED steps (the kernel applies step + 1)
__global__ void __launch_bounds__(256) adamw_dev_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m,
                                                        float4* __restrict__ v, long long n4, const float* __restrict__ hyper,
                                                        const int* __restrict__ step_dev, const float* __restrict__ wd_slot) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], gs = hyper[5];
  float wd = hyper[4];
  const float step = (float)(*step_dev + 1);
  const float bc1 = 1.f - powf(b1, step), bc2_sqrt = sqrtf(1.f - powf(b2, step));
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pv = p[i], gv = g[i], mv = m[i], vv = v[i];
    if (wd_slot) wd = wd_slot[i >> 4];
    float* pp = &pv.x; float* gp = &gv.x; float* mp = &mv.x; float* vp = &vv.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gr = gp[k] * gs;
      pp[k] *= 1.f - lr * wd;
      mp[k] = b1 * mp[k] + (1.f - b1) * gr;
      vp[k] = b2 * vp[k] + (1.f - b2) * gr * gr;
      const float denom = sqrtf(vp[k]) / bc2_sqrt + eps;
      pp[k] -= (lr / bc1) * (mp[k] / denom);
    }
    p[i] = pv; m[i] = mv; v[i] = vv;
  }
}
__global__ void device_hyper_set_kernel(float* hyper, float lr, float b1, float b2, float eps, float wd, float gs) {
  hyper[0] = lr; hyper[1] = b1; hyper[2] = b2; hyper[3] = eps; hyper[4] = wd; hyper[5] = gs;
}
__global__ void device_counter_advance_kernel(int* step_dev, unsigned long long* seed_dev) {
  if (step_dev) *step_dev += 1;
  if (seed_dev) *seed_dev += 1;
}
int k_adamw_dev(float* p, const float* g, float* m, float* v, long long n, const float* hyper, const int* step_dev, const float* wd_slot,
                cudaStream_t st) {
  if (n % 4 || (wd_slot && n % 64)) { s4g_set_error("adamw: arena length must be a multiple of 4 (64 with weight-decay slots)"); return S4G_ERR_SHAPE; }
  adamw_dev_kernel<<<148 * 8, 256, 0, st>>>(reinterpret_cast<float4*>(p), reinterpret_cast<const float4*>(g), reinterpret_cast<float4*>(m),
                                            reinterpret_cast<float4*>(v), n / 4, hyper, step_dev, wd_slot);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_device_hyper_set(float* hyper, float lr, float b1, float b2, float eps, float wd, float gs, cudaStream_t st) {
  device_hyper_set_kernel<<<1, 1, 0, st>>>(hyper, lr, b1, b2, eps, wd, gs);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_device_counter_advance(int* step_dev, unsigned long long* seed_dev, cudaStream_t st) {
  device_counter_advance_kernel<<<1, 1, 0, st>>>(step_dev, seed_dev);
  S4G_KERNEL_CHECK();
  return 0;
}
