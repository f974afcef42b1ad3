// GPU-side training input pipeline: what the reference's 8 DataLoader workers do per sample with Pillow and
// torchvision (Binary_segmentation/Data/dataset.py:34-64, Data/dataloaders.py:62-76), on a batch of decoded 8-bit images:
//   Image.resize((224, 224))            Pillow's 8-bit two-pass bicubic resampler (fixed-point coefficients, 22 bits)
//   ColorJitter(0.4, 0.5, 0.25, 0.01)   ImageEnhance blends + Pillow's RGB <-> HSV, re-quantised to 8 bits after every step
//   GaussianBlur((25, 25), sigma)       float32 convolution on reflect-padded input, round half even
//   ToTensor + Normalize / Grayscale    planar float32
//   hflip / vflip / affine(nearest)     one gather with the inverse matrix torchvision builds, fill -1 (image) / 0 (mask)
// Integer and byte work, HBM-bound: nothing here is shaped for tensor cores.  The uint8 stages reproduce Pillow bit for bit
// (double-precision coefficient arithmetic uses explicit non-fused intrinsics: the host libraries are built without FMA).
#include <algorithm>
#include <vector>

#include "kernels.h"

namespace {

constexpr int OUT = 224;
constexpr int PRECISION_BITS = 32 - 8 - 2;   // Pillow Resample.c
constexpr int KMAX = 96;                     // coefficients per output index: sources up to ~5000 px per side

// ------------------------------------------------------------------------------------------------ resize coefficients
__device__ __forceinline__ double bicubic_filter(double x) {
  // ((a + 2) x - (a + 3)) x x + 1 for |x| < 1; (((x - 5) x + 8) x - 4) a for |x| < 2, a = -0.5  -- evaluated unfused
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return __dadd_rn(__dmul_rn(__dmul_rn(__dadd_rn(__dmul_rn(a + 2.0, x), -(a + 3.0)), x), x), 1.0);
  if (x < 2.0) return __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(x, -5.0), x), 8.0), x), -4.0), a);
  return 0.0;
}

// table[(img * 2 + axis) * OUT + xx] = {xmin, count, k[KMAX]} : precompute_coeffs + normalize_coeffs_8bpc
struct CoefRow { int xmin, count; int k[KMAX]; };

__global__ void resize_coeffs_kernel(const long long* __restrict__ desc, CoefRow* __restrict__ table, int B) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * 2 * OUT) return;
  const int xx = idx % OUT, axis = (idx / OUT) % 2, img = idx / (2 * OUT);
  const int in_size = (int)desc[img * 5 + (axis == 0 ? 2 : 1)];   // axis 0: width, axis 1: height
  const double scale = (double)in_size / (double)OUT;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = __dmul_rn(2.0, filterscale);
  const double ss = 1.0 / filterscale;
  const double center = __dmul_rn((double)xx + 0.5, scale);
  int xmin = (int)__dadd_rn(__dadd_rn(center, -support), 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)__dadd_rn(__dadd_rn(center, support), 0.5);
  if (xmax > in_size) xmax = in_size;
  xmax -= xmin;
  CoefRow& row = table[idx];
  row.xmin = xmin;
  if (xmax > KMAX) xmax = KMAX;   // rejected on the host (source too large); never taken for accepted inputs
  row.count = xmax;
  double w[KMAX];
  double ww = 0.0;
  for (int x = 0; x < xmax; ++x) {
    w[x] = bicubic_filter(__dmul_rn(__dadd_rn(__dadd_rn((double)(x + xmin), -center), 0.5), ss));
    ww = __dadd_rn(ww, w[x]);
  }
  for (int x = 0; x < xmax; ++x) {
    double v = w[x];
    if (ww != 0.0) v = v / ww;
    row.k[x] = v < 0 ? (int)__dadd_rn(-0.5, __dmul_rn(v, (double)(1 << PRECISION_BITS)))
                     : (int)__dadd_rn(0.5, __dmul_rn(v, (double)(1 << PRECISION_BITS)));
  }
}

__device__ __forceinline__ uint8_t clip8(long long acc) {
  const long long v = acc >> PRECISION_BITS;
  return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// horizontal pass: tmp[img][y][xx][c] from src[img][y][x][c]; one thread per (y, xx), all channels
template <int C>
__global__ void resize_h_kernel(const uint8_t* __restrict__ src, const long long* __restrict__ desc,
                                const CoefRow* __restrict__ table, uint8_t* __restrict__ tmp) {
  const int img = blockIdx.y;
  const long long off = desc[img * 5];
  const int h = (int)desc[img * 5 + 1], w = (int)desc[img * 5 + 2];
  const CoefRow* rows = table + (size_t)(img * 2) * OUT;
  uint8_t* dst = tmp + desc[img * 5 + 4];
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < (long long)h * OUT; t += (long long)gridDim.x * blockDim.x) {
    const int xx = (int)(t % OUT), y = (int)(t / OUT);
    const CoefRow& r = rows[xx];
    const uint8_t* p = src + off + ((long long)y * w + r.xmin) * C;
    long long acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = 1 << (PRECISION_BITS - 1);
    for (int x = 0; x < r.count; ++x) {
      const int k = r.k[x];
#pragma unroll
      for (int c = 0; c < C; ++c) acc[c] += (long long)p[x * C + c] * k;
    }
#pragma unroll
    for (int c = 0; c < C; ++c) dst[t * C + c] = clip8(acc[c]);
  }
}
// vertical pass: out[img][yy][x][c] from tmp[img][y][x][c]
template <int C>
__global__ void resize_v_kernel(const uint8_t* __restrict__ tmp, const long long* __restrict__ desc,
                                const CoefRow* __restrict__ table, uint8_t* __restrict__ out) {
  const int img = blockIdx.y;
  const CoefRow* rows = table + (size_t)(img * 2 + 1) * OUT;
  const uint8_t* s = tmp + desc[img * 5 + 4];
  uint8_t* dst = out + (size_t)img * OUT * OUT * C;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < OUT * OUT * C; t += gridDim.x * blockDim.x) {
    const int xc = t % (OUT * C), yy = t / (OUT * C);
    const CoefRow& r = rows[yy];
    long long acc = 1 << (PRECISION_BITS - 1);
    for (int y = 0; y < r.count; ++y) acc += (long long)s[(size_t)(r.xmin + y) * OUT * C + xc] * r.k[y];
    dst[t] = clip8(acc);
  }
}

// ------------------------------------------------------------------------------------------------ ColorJitter (8-bit)
__device__ __forceinline__ int luma(int r, int g, int b) { return (r * 19595 + g * 38470 + b * 7471 + 0x8000) >> 16; }   // convert("L")

// Image.blend(degenerate, image, f) on one channel: float32, unfused; truncation inside [0, 1], clip + truncation outside
__device__ __forceinline__ int blend8(int deg, int v, float f, bool interp) {
  const float t = __fadd_rn((float)deg, __fmul_rn(f, (float)(v - deg)));
  if (interp) return (int)(uint8_t)t;
  return t <= 0.f ? 0 : (t >= 255.f ? 255 : (int)t);
}

// Pillow Convert.c rgb2hsv / hsv2rgb (float variables, double literals: promotions as in oracle/aug_port.py)
__device__ __forceinline__ void rgb2hsv8(int r, int g, int b, int& uh, int& us, int& uv) {
  const int maxc = max(r, max(g, b)), minc = min(r, min(g, b));
  uv = maxc;
  if (minc == maxc) { uh = 0; us = 0; return; }
  const float cr = (float)(maxc - minc);
  const float s = __fdiv_rn(cr, (float)maxc);
  const float rc = __fdiv_rn((float)(maxc - r), cr), gc = __fdiv_rn((float)(maxc - g), cr), bc = __fdiv_rn((float)(maxc - b), cr);
  float h;
  if (r == maxc) h = __fadd_rn(bc, -gc);
  else if (g == maxc) h = (float)__dadd_rn(__dadd_rn(2.0, (double)rc), -(double)bc);
  else h = (float)__dadd_rn(__dadd_rn(4.0, (double)gc), -(double)rc);
  h = (float)fmod(__dadd_rn(__ddiv_rn((double)h, 6.0), 1.0), 1.0);
  const int ih = (int)__dmul_rn((double)h, 255.0), is = (int)__dmul_rn((double)s, 255.0);
  uh = min(max(ih, 0), 255);
  us = min(max(is, 0), 255);
}
__device__ __forceinline__ int round_pos(double x) { const int v = (int)floor(__dadd_rn(x, 0.5)); return min(max(v, 0), 255); }
__device__ __forceinline__ void hsv2rgb8(int h, int s, int v, int& r, int& g, int& b) {
  if (s == 0) { r = g = b = v; return; }
  const double hd = __ddiv_rn(__dmul_rn((double)h, 6.0), 255.0);
  const int i = (int)floor(hd);
  const float f = (float)__dadd_rn(hd, -(double)i);
  const float fs = (float)__ddiv_rn((double)s, 255.0);
  const double vf = (double)v;
  const int p = round_pos(__dmul_rn(vf, __dadd_rn(1.0, -(double)fs)));
  const int q = round_pos(__dmul_rn(vf, __dadd_rn(1.0, -(double)__fmul_rn(fs, f))));
  const int t = round_pos(__dmul_rn(vf, __dadd_rn(1.0, -__dmul_rn((double)fs, __dadd_rn(1.0, -(double)f)))));
  switch (i % 6) {
    case 0: r = v; g = t; b = p; break;
    case 1: r = q; g = v; b = p; break;
    case 2: r = p; g = v; b = t; break;
    case 3: r = p; g = q; b = v; break;
    case 4: r = t; g = p; b = v; break;
    default: r = v; g = p; b = q; break;
  }
}

struct JitterParams { int order[4]; int hue_shift; float factor[3]; };   // factor: brightness, contrast, saturation

// one CTA per image: the 147 KB image lives in shared memory across the four adjustments (contrast needs the mean luma of
// the image as left by the preceding steps: an exact integer block reduction)
__global__ void __launch_bounds__(1024) color_jitter_kernel(uint8_t* __restrict__ img, const JitterParams* __restrict__ params) {
  extern __shared__ __align__(16) uint8_t px[];
  __shared__ unsigned long long wsum[32];
  __shared__ int mean_s;
  constexpr int NPIX = OUT * OUT;
  uint8_t* g = img + (size_t)blockIdx.x * NPIX * 3;
  const JitterParams prm = params[blockIdx.x];
  for (int i = threadIdx.x; i < NPIX * 3 / 16; i += blockDim.x) reinterpret_cast<uint4*>(px)[i] = reinterpret_cast<const uint4*>(g)[i];
  __syncthreads();
  for (int step = 0; step < 4; ++step) {
    const int op = prm.order[step];
    if (op == 1) {   // contrast: blend with the rounded mean luma
      unsigned long long s = 0;
      for (int i = threadIdx.x; i < NPIX; i += blockDim.x) s += (unsigned)luma(px[3 * i], px[3 * i + 1], px[3 * i + 2]);
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = s;
      __syncthreads();
      if (threadIdx.x == 0) {
        unsigned long long tot = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += wsum[w];
        mean_s = (int)__dadd_rn(__ddiv_rn((double)tot, (double)NPIX), 0.5);
      }
      __syncthreads();
    }
    const float f = op < 3 ? prm.factor[op] : 0.f;
    const bool interp = f >= 0.f && f <= 1.f;
    const int mean = mean_s;
    for (int i = threadIdx.x; i < NPIX; i += blockDim.x) {
      int r = px[3 * i], gg = px[3 * i + 1], b = px[3 * i + 2];
      if (op == 0) { r = blend8(0, r, f, interp); gg = blend8(0, gg, f, interp); b = blend8(0, b, f, interp); }
      else if (op == 1) { r = blend8(mean, r, f, interp); gg = blend8(mean, gg, f, interp); b = blend8(mean, b, f, interp); }
      else if (op == 2) { const int l = luma(r, gg, b); r = blend8(l, r, f, interp); gg = blend8(l, gg, f, interp); b = blend8(l, b, f, interp); }
      else { int h, s, v; rgb2hsv8(r, gg, b, h, s, v); h = (h + prm.hue_shift) & 255; hsv2rgb8(h, s, v, r, gg, b); }
      px[3 * i] = (uint8_t)r; px[3 * i + 1] = (uint8_t)gg; px[3 * i + 2] = (uint8_t)b;
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < NPIX * 3 / 16; i += blockDim.x) reinterpret_cast<uint4*>(g)[i] = reinterpret_cast<const uint4*>(px)[i];
}

// ------------------------------------------------------------------------------------------------ GaussianBlur + ToTensor + Normalize
constexpr int BLUR_K = 25, BLUR_R = 12, BLUR_ROWS = 32, ROWB = OUT * 3;
constexpr int BLUR_SMEM = (BLUR_ROWS + 2 * BLUR_R) * ROWB + BLUR_ROWS * ROWB * (int)sizeof(float);
__device__ __forceinline__ int reflect(int i) { return i < 0 ? -i : (i >= OUT ? 2 * (OUT - 1) - i : i); }

// grid (7 row tiles, B).  sigma[b] <= 0: no blur (evaluation transform).  out [B, 3, 224, 224] float32:
// (round(blur(u8)) / 255 - mean[c]) / std[c]
__global__ void __launch_bounds__(256) blur_normalize_kernel(const uint8_t* __restrict__ img, const float* __restrict__ sigma,
                                                             float* __restrict__ out, float m0, float m1, float m2, float s0,
                                                             float s1, float s2) {
  extern __shared__ __align__(16) uint8_t sm[];
  uint8_t* px = sm;                                                       // [32 + 24][672]
  float* vb = reinterpret_cast<float*>(sm + (BLUR_ROWS + 2 * BLUR_R) * ROWB);   // [32][672]
  __shared__ float k1[BLUR_K];
  __shared__ float ksum;
  const int b = blockIdx.y, y0 = blockIdx.x * BLUR_ROWS;
  const float sg = sigma ? sigma[b] : 0.f;
  const uint8_t* g = img + (size_t)b * OUT * ROWB;
  if (threadIdx.x < BLUR_K) {   // torchvision _get_gaussian_kernel1d, float32
    const float x = __fdiv_rn((float)((int)threadIdx.x - BLUR_R), sg > 0.f ? sg : 1.f);
    k1[threadIdx.x] = expf(__fmul_rn(-0.5f, __fmul_rn(x, x)));
  }
  for (int i = threadIdx.x; i < (BLUR_ROWS + 2 * BLUR_R) * (ROWB / 16); i += 256) {
    const int r = i / (ROWB / 16), v = i % (ROWB / 16);
    reinterpret_cast<uint4*>(px + r * ROWB)[v] = reinterpret_cast<const uint4*>(g + (size_t)reflect(y0 - BLUR_R + r) * ROWB)[v];
  }
  __syncthreads();
  if (threadIdx.x == 0) { float t = 0.f; for (int i = 0; i < BLUR_K; ++i) t = __fadd_rn(t, k1[i]); ksum = t; }
  __syncthreads();
  if (threadIdx.x < BLUR_K) k1[threadIdx.x] = __fdiv_rn(k1[threadIdx.x], ksum);
  __syncthreads();
  if (sg > 0.f) {
    for (int i = threadIdx.x; i < BLUR_ROWS * ROWB; i += 256) {
      const int r = i / ROWB, xc = i % ROWB;
      float a = 0.f;
#pragma unroll
      for (int t = 0; t < BLUR_K; ++t) a = fmaf(k1[t], (float)px[(r + t) * ROWB + xc], a);
      vb[i] = a;
    }
    __syncthreads();
  }
  const float mean[3] = {m0, m1, m2}, stdv[3] = {s0, s1, s2};
  for (int i = threadIdx.x; i < 3 * BLUR_ROWS * OUT; i += 256) {
    const int x = i % OUT, r = (i / OUT) % BLUR_ROWS, c = i / (OUT * BLUR_ROWS);
    float v;
    if (sg > 0.f) {
      float a = 0.f;
#pragma unroll
      for (int t = 0; t < BLUR_K; ++t) a = fmaf(k1[t], vb[r * ROWB + reflect(x + t - BLUR_R) * 3 + c], a);
      v = fminf(fmaxf(rintf(a), 0.f), 255.f);
    } else {
      v = (float)px[(r + BLUR_R) * ROWB + x * 3 + c];
    }
    out[(((size_t)b * 3 + c) * OUT + y0 + r) * OUT + x] = __fdiv_rn(__fadd_rn(__fdiv_rn(v, 255.f), -mean[c]), stdv[c]);
  }
}

// ToTensor + Grayscale of the 8-bit mask [B, 224, 224, C] -> [B, 1, 224, 224] (torchvision rgb_to_grayscale weights)
template <int C>
__global__ void mask_to_tensor_kernel(const uint8_t* __restrict__ img, float* __restrict__ out, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    if (C == 1) { out[i] = __fdiv_rn((float)img[i], 255.f); continue; }
    const float r = __fdiv_rn((float)img[i * 3], 255.f), g = __fdiv_rn((float)img[i * 3 + 1], 255.f), b = __fdiv_rn((float)img[i * 3 + 2], 255.f);
    out[i] = __fadd_rn(__fadd_rn(__fmul_rn(0.2989f, r), __fmul_rn(0.587f, g)), __fmul_rn(0.114f, b));
  }
}

// ------------------------------------------------------------------------------------------------ flips + affine (nearest, fill)
struct GeoParams { int hflip, vflip, affine; float m[6]; };   // m: torchvision _get_inverse_affine_matrix (center 0, 0)

// x [B,3,224,224], y [B,1,224,224] -> xo, yo: TF.hflip, TF.vflip, TF.affine(NEAREST, fill) (dataset.py:46-63); the sampling grid
// is _gen_affine_grid + grid_sample(nearest, zeros, align_corners=False) in unfused float32
__global__ void geometric_kernel(const float* __restrict__ x, const float* __restrict__ y, float* __restrict__ xo, float* __restrict__ yo,
                                 const GeoParams* __restrict__ params, float fill_x, float fill_y, int B) {
  const long long n = (long long)B * OUT * OUT;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(t % OUT), i = (int)((t / OUT) % OUT), b = (int)(t / (OUT * OUT));
    const GeoParams& p = params[b];
    int si = i, sj = j;
    bool inside = true;
    if (p.affine) {
      const float half = 0.5f * OUT;
      const float xs = (float)j - (half - 0.5f), ys = (float)i - (half - 0.5f);
      const float gx = __fadd_rn(__fadd_rn(__fmul_rn(xs, __fdiv_rn(p.m[0], half)), __fmul_rn(ys, __fdiv_rn(p.m[1], half))), __fdiv_rn(p.m[2], half));
      const float gy = __fadd_rn(__fadd_rn(__fmul_rn(xs, __fdiv_rn(p.m[3], half)), __fmul_rn(ys, __fdiv_rn(p.m[4], half))), __fdiv_rn(p.m[5], half));
      const float fx = __fdiv_rn(__fadd_rn(__fmul_rn(__fadd_rn(gx, 1.f), (float)OUT), -1.f), 2.f);
      const float fy = __fdiv_rn(__fadd_rn(__fmul_rn(__fadd_rn(gy, 1.f), (float)OUT), -1.f), 2.f);
      const float rx = rintf(fx), ry = rintf(fy);
      inside = rx >= 0.f && rx < (float)OUT && ry >= 0.f && ry < (float)OUT;
      sj = (int)rx; si = (int)ry;
    }
    if (inside) {
      if (p.vflip) si = OUT - 1 - si;
      if (p.hflip) sj = OUT - 1 - sj;
    }
    const size_t src = (size_t)si * OUT + sj, dst = (size_t)i * OUT + j;
#pragma unroll
    for (int c = 0; c < 3; ++c) xo[((size_t)b * 3 + c) * OUT * OUT + dst] = inside ? x[((size_t)b * 3 + c) * OUT * OUT + src] : fill_x;
    if (y) yo[(size_t)b * OUT * OUT + dst] = inside ? y[(size_t)b * OUT * OUT + src] : fill_y;
  }
}

// ------------------------------------------------------------------------------------------------ depth target (I;16 -> float)
// Depth_estimation/Data/dataset.py:47-53 on the (already square-padded) 16-bit depth image: nearest resize as Pillow 9.1.0 does
// for "I;16" (source index floor((i + 0.5) * in / 224)), then / 65535 in double, stored as float: out [B, 1, 224, 224]
__global__ void depth_target_kernel(const uint16_t* __restrict__ src, const long long* __restrict__ desc, float* __restrict__ out, int B) {
  const long long n = (long long)B * OUT * OUT;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(t % OUT), y = (int)((t / OUT) % OUT), b = (int)(t / (OUT * OUT));
    const long long off = desc[b * 5];
    const int h = (int)desc[b * 5 + 1], w = (int)desc[b * 5 + 2];
    const int sy = min((int)floor(__ddiv_rn(__dmul_rn((double)y + 0.5, (double)h), (double)OUT)), h - 1);
    const int sx = min((int)floor(__ddiv_rn(__dmul_rn((double)x + 0.5, (double)w), (double)OUT)), w - 1);
    const uint16_t v = *reinterpret_cast<const uint16_t*>(reinterpret_cast<const uint8_t*>(src) + off + ((size_t)sy * w + sx) * 2);
    out[t] = (float)__ddiv_rn((double)v, 65535.0);
  }
}

}  // namespace

// ================================================================================================ host entry points
// desc [B][5] = {byte offset into src, height, width, channels, byte offset of the [h][224][C] intermediate inside tmp}
static size_t aug_table_bytes(int B) { return (sizeof(CoefRow) * (size_t)B * 2 * OUT + 255) & ~(size_t)255; }
size_t aug_resize_scratch_bytes(const long long* desc_host, int B, int C) {
  size_t tmp = 0;
  for (int b = 0; b < B; ++b) tmp = std::max(tmp, (size_t)desc_host[b * 5 + 4] + (size_t)desc_host[b * 5 + 1] * OUT * C);
  return aug_table_bytes(B) + tmp + 16;
}
int k_aug_resize(const uint8_t* src, const long long* desc_dev, const long long* desc_host, int B, int C, void* scratch,
                 size_t scratch_bytes, uint8_t* out, cudaStream_t st) {
  if (C != 1 && C != 3) { s4g_set_error("aug_resize: 1 or 3 channels"); return S4G_ERR_SHAPE; }
  long long max_h = 1;
  for (int b = 0; b < B; ++b) {
    const long long h = desc_host[b * 5 + 1], w = desc_host[b * 5 + 2];
    if (desc_host[b * 5 + 3] != C || h < 1 || w < 1 || desc_host[b * 5 + 4] < 0) { s4g_set_error("aug_resize: bad image descriptor"); return S4G_ERR_SHAPE; }
    // coefficients per output index: ceil(2 * scale) * 2 + 1 <= KMAX
    if (std::max(h, w) > (long long)OUT * (KMAX - 3) / 4) { s4g_set_error("aug_resize: source side too long for the coefficient table"); return S4G_ERR_SHAPE; }
    max_h = std::max(max_h, h);
  }
  if (scratch_bytes < aug_resize_scratch_bytes(desc_host, B, C)) { s4g_set_error("aug_resize: scratch too small"); return S4G_ERR_WORKSPACE; }
  CoefRow* table = reinterpret_cast<CoefRow*>(scratch);
  uint8_t* tmp = reinterpret_cast<uint8_t*>(scratch) + aug_table_bytes(B);
  resize_coeffs_kernel<<<cdiv((long long)B * 2 * OUT, 128), 128, 0, st>>>(desc_dev, table, B);
  S4G_KERNEL_CHECK();
  const dim3 gh((unsigned)std::min<long long>(cdiv(max_h * OUT, 256), 148 * 4), B), gv(cdiv((long long)OUT * OUT * C, 256), B);
  if (C == 3) {
    resize_h_kernel<3><<<gh, 256, 0, st>>>(src, desc_dev, table, tmp);
    resize_v_kernel<3><<<gv, 256, 0, st>>>(tmp, desc_dev, table, out);
  } else {
    resize_h_kernel<1><<<gh, 256, 0, st>>>(src, desc_dev, table, tmp);
    resize_v_kernel<1><<<gv, 256, 0, st>>>(tmp, desc_dev, table, out);
  }
  S4G_KERNEL_CHECK();
  return 0;
}
int k_aug_color_jitter(uint8_t* img, const void* params, int B, cudaStream_t st) {
  static bool attr = false;
  const int smem = OUT * OUT * 3;
  if (!attr) { S4G_CUDA(cudaFuncSetAttribute(color_jitter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); attr = true; }
  color_jitter_kernel<<<B, 1024, smem, st>>>(img, reinterpret_cast<const JitterParams*>(params));
  S4G_KERNEL_CHECK();
  return 0;
}
int k_aug_blur_normalize(const uint8_t* img, const float* sigma, float* out, const float* mean3, const float* std3, int B, cudaStream_t st) {
  static bool attr = false;
  if (!attr) { S4G_CUDA(cudaFuncSetAttribute(blur_normalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BLUR_SMEM)); attr = true; }
  blur_normalize_kernel<<<dim3(OUT / BLUR_ROWS, B), 256, BLUR_SMEM, st>>>(img, sigma, out, mean3[0], mean3[1], mean3[2], std3[0], std3[1], std3[2]);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_aug_mask_to_tensor(const uint8_t* img, float* out, int B, int C, cudaStream_t st) {
  const long long n = (long long)B * OUT * OUT;
  if (C == 3) mask_to_tensor_kernel<3><<<(unsigned)std::min<long long>(cdiv(n, 256), 148 * 8), 256, 0, st>>>(img, out, n);
  else if (C == 1) mask_to_tensor_kernel<1><<<(unsigned)std::min<long long>(cdiv(n, 256), 148 * 8), 256, 0, st>>>(img, out, n);
  else { s4g_set_error("aug_mask_to_tensor: 1 or 3 channels"); return S4G_ERR_SHAPE; }
  S4G_KERNEL_CHECK();
  return 0;
}
int k_aug_geometric(const float* x, const float* y, float* xo, float* yo, const void* params, float fill_x, float fill_y, int B,
                    cudaStream_t st) {
  const long long n = (long long)B * OUT * OUT;
  geometric_kernel<<<(unsigned)std::min<long long>(cdiv(n, 256), 148 * 8), 256, 0, st>>>(x, y, xo, yo, reinterpret_cast<const GeoParams*>(params),
                                                                                          fill_x, fill_y, B);
  S4G_KERNEL_CHECK();
  return 0;
}
int k_aug_depth_target(const void* src, const long long* desc_dev, float* out, int B, cudaStream_t st) {
  const long long n = (long long)B * OUT * OUT;
  depth_target_kernel<<<(unsigned)std::min<long long>(cdiv(n, 256), 148 * 8), 256, 0, st>>>(reinterpret_cast<const uint16_t*>(src), desc_dev, out, B);
  S4G_KERNEL_CHECK();
  return 0;
}
