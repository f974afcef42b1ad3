// Fused attention for the 197-token ViT-B/16 sequence on bf16 tensor cores (mma.sync m16n8k16, fp32 accumulate).
// One CTA per (image, head); Q, K, V (and dO in backward) live in XOR-swizzled shared memory for the whole kernel,
// the 197x197 score matrix never leaves registers, softmax is single-pass (the whole row fits), only LSE is saved.
//   forward : S = Q K^T -> softmax -> O = P V
//   backward: phase A per 16-query tile : S, dP = dO V^T, dS = P (dP - delta) s ; dQ = dS K
//             phase B per 16-key tile   : S^T = K Q^T, dP^T = V dO^T, dS^T ; dV = P^T dO, dK = dS^T Q
//             (scores are recomputed in both phases: no atomics, no cross-warp reduction, deterministic)
// Reference semantics: timm 0.6.12 Attention.forward (called from Models/models.py:333-334), scale = 64^-0.5.
#include "kernels.h"

namespace {

constexpr int D = 64;        // head dim
constexpr int NP = 208;      // 197 padded to 13 x 16
constexpr int NT = NP / 8;   // 26 n-tiles of 8
constexpr int WARPS = 7;     // 13 row tiles over 7 warps (2 rounds)
constexpr int THREADS = WARPS * 32;
constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ uint32_t sw_off(int row, int chunk) { return (uint32_t)(row * 128 + ((chunk ^ (row & 7)) << 4)); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// global [tokens, ld] slice (64 contiguous bf16 per row at `base + row*ld`) -> swizzled smem [NP][64]; rows >= n zeroed
__device__ __forceinline__ void load_tile(uint8_t* dst, const bf16* base, long long ld, int n) {
  for (int i = threadIdx.x; i < NP * 8; i += THREADS) {
    const int row = i >> 3, chunk = i & 7;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (row < n) v = *reinterpret_cast<const uint4*>(base + (long long)row * ld + chunk * 8);
    *reinterpret_cast<uint4*>(dst + sw_off(row, chunk)) = v;
  }
}

// A fragments (4 k-steps) of the 16-row tile starting at row0 of a swizzled [NP][64] matrix
__device__ __forceinline__ void load_a_frags(uint32_t smem, int row0, uint32_t (&a)[4][4]) {
  const int lane = threadIdx.x & 31;
  const int row = row0 + (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) ldsm_x4(smem + sw_off(row, ks * 2 + (lane >> 4)), a[ks][0], a[ks][1], a[ks][2], a[ks][3]);
}

// acc[j] (16 x 8 tile j) = A(16x64) . B[n0 + 8j .. +8][0..64]^T  for j in [0, NTILES)   (B row-major, contraction over d)
template <int NTILES>
__device__ __forceinline__ void gemm_abt(float (&acc)[NTILES][4], const uint32_t (&a)[4][4], uint32_t smemB, int n0) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < NTILES; ++j) {
    acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    const int row = n0 + j * 8 + (lane & 7);
    uint32_t b[8];
    ldsm_x4(smemB + sw_off(row, lane >> 3), b[0], b[1], b[2], b[3]);        // d 0..31
    ldsm_x4(smemB + sw_off(row, 4 + (lane >> 3)), b[4], b[5], b[6], b[7]);  // d 32..63
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) mma_bf16(acc[j], a[ks], b[2 * ks], b[2 * ks + 1]);
  }
}

// out[jn] (16 x 8 over d) += P(16 x 16*KSTEPS, as A fragments) . B[k0 .. k0+16*KSTEPS][0..64]   (B row-major, contraction over rows)
template <int KSTEPS>
__device__ __forceinline__ void gemm_pb(float (&out)[8][4], const uint32_t (&p)[KSTEPS][4], uint32_t smemB, int k0) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int ks = 0; ks < KSTEPS; ++ks) {
    const int row = k0 + ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
    for (int jp = 0; jp < 4; ++jp) {  // pairs of d n-tiles
      uint32_t b0, b1, b2, b3;
      ldsm_x4_t(smemB + sw_off(row, jp * 2 + (lane >> 4)), b0, b1, b2, b3);
      mma_bf16(out[2 * jp], p[ks], b0, b1);
      mma_bf16(out[2 * jp + 1], p[ks], b2, b3);
    }
  }
}

// ------------------------------------------------------------------------------------------ forward
__global__ void __launch_bounds__(THREADS, 1) attn_fwd_mma_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ o,
                                                                  float* __restrict__ lse, int N, int H) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint8_t* sQ = sm;
  uint8_t* sK = sm + NP * 128;
  uint8_t* sV = sm + 2 * NP * 128;
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int C = H * D, C3 = 3 * C;
  const bf16* base = qkv + (long long)b * N * C3 + h * D;
  load_tile(sQ, base, C3, N);
  load_tile(sK, base + C, C3, N);
  load_tile(sV, base + 2 * C, C3, N);
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t uQ = (uint32_t)__cvta_generic_to_shared(sQ), uK = (uint32_t)__cvta_generic_to_shared(sK),
                 uV = (uint32_t)__cvta_generic_to_shared(sV);
  const float sc = 0.125f * LOG2E;
  for (int tile = warp; tile < NP / 16; tile += WARPS) {
    uint32_t a[4][4];
    load_a_frags(uQ, tile * 16, a);
    float s[NT][4];
    gemm_abt<NT>(s, a, uK, 0);
    float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int col = j * 8 + 2 * c;
      if (col >= N) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
      if (col + 1 >= N) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
      m0 = fmaxf(m0, fmaxf(s[j][0], s[j][1]));
      m1 = fmaxf(m1, fmaxf(s[j][2], s[j][3]));
    }
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
    float l0 = 0.f, l1 = 0.f;
    uint32_t p[NT / 2][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const float e0 = exp2f((s[j][0] - m0) * sc), e1 = exp2f((s[j][1] - m0) * sc);
      const float e2 = exp2f((s[j][2] - m1) * sc), e3 = exp2f((s[j][3] - m1) * sc);
      l0 += e0 + e1; l1 += e2 + e3;
      p[j >> 1][(j & 1) * 2] = pack_bf16(e0, e1);
      p[j >> 1][(j & 1) * 2 + 1] = pack_bf16(e2, e3);
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    float acc[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    gemm_pb<NT / 2>(acc, p, uV, 0);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    const int r0 = tile * 16 + g, r1 = r0 + 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int d = j * 8 + 2 * c;
      if (r0 < N) *reinterpret_cast<uint32_t*>(o + ((long long)b * N + r0) * C + h * D + d) = pack_bf16(acc[j][0] * i0, acc[j][1] * i0);
      if (r1 < N) *reinterpret_cast<uint32_t*>(o + ((long long)b * N + r1) * C + h * D + d) = pack_bf16(acc[j][2] * i1, acc[j][3] * i1);
    }
    if (c == 0) {
      if (r0 < N) lse[((long long)b * H + h) * N + r0] = m0 * 0.125f + logf(l0);
      if (r1 < N) lse[((long long)b * H + h) * N + r1] = m1 * 0.125f + logf(l1);
    }
  }
}

// ------------------------------------------------------------------------------------------ backward
// one half (NH n-tiles starting at column n0) of: P = exp(S s - lse_row), dS = P (dP - delta_row) s  -> bf16 A fragments
//   row-stat variant (phase A): statistics indexed by the accumulator ROW (queries are rows)
template <int NH>
__device__ __forceinline__ void ds_rows(const float (&s)[NH][4], const float (&dp)[NH][4], int n0, int N, float lse0, float lse1,
                                        float dl0, float dl1, uint32_t (&ds)[NH / 2][4]) {
  const int c = threadIdx.x & 3;
  const float sc = 0.125f * LOG2E;
#pragma unroll
  for (int j = 0; j < NH; ++j) {
    const int col = n0 + j * 8 + 2 * c;
    float p0 = exp2f(s[j][0] * sc - lse0), p1 = exp2f(s[j][1] * sc - lse0);
    float p2 = exp2f(s[j][2] * sc - lse1), p3 = exp2f(s[j][3] * sc - lse1);
    if (col >= N) { p0 = 0.f; p2 = 0.f; }
    if (col + 1 >= N) { p1 = 0.f; p3 = 0.f; }
    ds[j >> 1][(j & 1) * 2] = pack_bf16(p0 * (dp[j][0] - dl0) * 0.125f, p1 * (dp[j][1] - dl0) * 0.125f);
    ds[j >> 1][(j & 1) * 2 + 1] = pack_bf16(p2 * (dp[j][2] - dl1) * 0.125f, p3 * (dp[j][3] - dl1) * 0.125f);
  }
}
//   column-stat variant (phase B): queries are COLUMNS; statistics come from shared memory (padded queries: lse = +big)
template <int NH>
__device__ __forceinline__ void ds_cols(const float (&s)[NH][4], const float (&dp)[NH][4], int n0, const float* slse,
                                        const float* sdelta, uint32_t (&pt)[NH / 2][4], uint32_t (&ds)[NH / 2][4]) {
  const int c = threadIdx.x & 3;
  const float sc = 0.125f * LOG2E;
#pragma unroll
  for (int j = 0; j < NH; ++j) {
    const int col = n0 + j * 8 + 2 * c;
    const float la = slse[col], lb = slse[col + 1], da = sdelta[col], db = sdelta[col + 1];
    const float p0 = exp2f(s[j][0] * sc - la), p1 = exp2f(s[j][1] * sc - lb);
    const float p2 = exp2f(s[j][2] * sc - la), p3 = exp2f(s[j][3] * sc - lb);
    pt[j >> 1][(j & 1) * 2] = pack_bf16(p0, p1);
    pt[j >> 1][(j & 1) * 2 + 1] = pack_bf16(p2, p3);
    ds[j >> 1][(j & 1) * 2] = pack_bf16(p0 * (dp[j][0] - da) * 0.125f, p1 * (dp[j][1] - db) * 0.125f);
    ds[j >> 1][(j & 1) * 2 + 1] = pack_bf16(p2 * (dp[j][2] - da) * 0.125f, p3 * (dp[j][3] - db) * 0.125f);
  }
}

__global__ void __launch_bounds__(THREADS, 1) attn_bwd_mma_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ o,
                                                                  const bf16* __restrict__ dout, const float* __restrict__ lse,
                                                                  bf16* __restrict__ dqkv, int N, int H) {
  extern __shared__ __align__(128) uint8_t sm[];
  uint8_t* sQ = sm;
  uint8_t* sK = sm + NP * 128;
  uint8_t* sV = sm + 2 * NP * 128;
  uint8_t* sdO = sm + 3 * NP * 128;
  float* slse = reinterpret_cast<float*>(sm + 4 * NP * 128);  // lse * log2(e); padded queries: +big
  float* sdelta = slse + NP;
  const int b = blockIdx.x / H, h = blockIdx.x % H;
  const int C = H * D, C3 = 3 * C;
  const bf16* base = qkv + (long long)b * N * C3 + h * D;
  const bf16* dob = dout + (long long)b * N * C + h * D;
  const bf16* ob = o + (long long)b * N * C + h * D;
  bf16* dbase = dqkv + (long long)b * N * C3 + h * D;
  load_tile(sQ, base, C3, N);
  load_tile(sK, base + C, C3, N);
  load_tile(sV, base + 2 * C, C3, N);
  load_tile(sdO, dob, C, N);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // delta_i = sum_d dO[i,d] * O[i,d]; 4 lanes per row
  for (int i = threadIdx.x; i < NP * 4; i += THREADS) {
    const int row = i >> 2, part = i & 3;
    float acc = 0.f;
    if (row < N) {
      Vec8<bf16> x, y;
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        x.load(dob + (long long)row * C + part * 16 + k * 8);
        y.load(ob + (long long)row * C + part * 16 + k * 8);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc += x.v[e] * y.v[e];
      }
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (part == 0) {
      sdelta[row] = acc;
      slse[row] = row < N ? lse[((long long)b * H + h) * N + row] * LOG2E : 1e30f;
    }
  }
  __syncthreads();
  const int g = lane >> 2, c = lane & 3;
  const uint32_t uQ = (uint32_t)__cvta_generic_to_shared(sQ), uK = (uint32_t)__cvta_generic_to_shared(sK),
                 uV = (uint32_t)__cvta_generic_to_shared(sV), udO = (uint32_t)__cvta_generic_to_shared(sdO);
  constexpr int NH0 = 14, NH1 = 12;  // column halves: 112 + 96 = 208

  // ---------------- phase A: dQ for each 16-query tile
  for (int tile = warp; tile < NP / 16; tile += WARPS) {
    uint32_t aq[4][4], ado[4][4];
    load_a_frags(uQ, tile * 16, aq);
    load_a_frags(udO, tile * 16, ado);
    const int r0 = tile * 16 + g, r1 = r0 + 8;
    const float lse0 = slse[r0], lse1 = slse[r1], dl0 = sdelta[r0], dl1 = sdelta[r1];
    float acc[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    {
      float s[NH0][4], dp[NH0][4];
      uint32_t ds[NH0 / 2][4];
      gemm_abt<NH0>(s, aq, uK, 0);
      gemm_abt<NH0>(dp, ado, uV, 0);
      ds_rows<NH0>(s, dp, 0, N, lse0, lse1, dl0, dl1, ds);
      gemm_pb<NH0 / 2>(acc, ds, uK, 0);
    }
    {
      float s[NH1][4], dp[NH1][4];
      uint32_t ds[NH1 / 2][4];
      gemm_abt<NH1>(s, aq, uK, NH0 * 8);
      gemm_abt<NH1>(dp, ado, uV, NH0 * 8);
      ds_rows<NH1>(s, dp, NH0 * 8, N, lse0, lse1, dl0, dl1, ds);
      gemm_pb<NH1 / 2>(acc, ds, uK, NH0 * 8);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int d = j * 8 + 2 * c;
      if (r0 < N) *reinterpret_cast<uint32_t*>(dbase + (long long)r0 * C3 + d) = pack_bf16(acc[j][0], acc[j][1]);
      if (r1 < N) *reinterpret_cast<uint32_t*>(dbase + (long long)r1 * C3 + d) = pack_bf16(acc[j][2], acc[j][3]);
    }
  }
  // ---------------- phase B: dK, dV for each 16-key tile (rows = keys, columns = queries)
  for (int tile = warp; tile < NP / 16; tile += WARPS) {
    uint32_t ak[4][4], av[4][4];
    load_a_frags(uK, tile * 16, ak);
    load_a_frags(uV, tile * 16, av);
    float dk[8][4], dv[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dk[j][0] = dk[j][1] = dk[j][2] = dk[j][3] = 0.f;
      dv[j][0] = dv[j][1] = dv[j][2] = dv[j][3] = 0.f;
    }
    {
      float s[NH0][4], dp[NH0][4];
      uint32_t pt[NH0 / 2][4], ds[NH0 / 2][4];
      gemm_abt<NH0>(s, ak, uQ, 0);
      gemm_abt<NH0>(dp, av, udO, 0);
      ds_cols<NH0>(s, dp, 0, slse, sdelta, pt, ds);
      gemm_pb<NH0 / 2>(dv, pt, udO, 0);
      gemm_pb<NH0 / 2>(dk, ds, uQ, 0);
    }
    {
      float s[NH1][4], dp[NH1][4];
      uint32_t pt[NH1 / 2][4], ds[NH1 / 2][4];
      gemm_abt<NH1>(s, ak, uQ, NH0 * 8);
      gemm_abt<NH1>(dp, av, udO, NH0 * 8);
      ds_cols<NH1>(s, dp, NH0 * 8, slse, sdelta, pt, ds);
      gemm_pb<NH1 / 2>(dv, pt, udO, NH0 * 8);
      gemm_pb<NH1 / 2>(dk, ds, uQ, NH0 * 8);
    }
    const int r0 = tile * 16 + g, r1 = r0 + 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int d = j * 8 + 2 * c;
      if (r0 < N) {
        *reinterpret_cast<uint32_t*>(dbase + (long long)r0 * C3 + C + d) = pack_bf16(dk[j][0], dk[j][1]);
        *reinterpret_cast<uint32_t*>(dbase + (long long)r0 * C3 + 2 * C + d) = pack_bf16(dv[j][0], dv[j][1]);
      }
      if (r1 < N) {
        *reinterpret_cast<uint32_t*>(dbase + (long long)r1 * C3 + C + d) = pack_bf16(dk[j][2], dk[j][3]);
        *reinterpret_cast<uint32_t*>(dbase + (long long)r1 * C3 + 2 * C + d) = pack_bf16(dv[j][2], dv[j][3]);
      }
    }
  }
}

}  // namespace

int attn_fwd_mma(const bf16* qkv, bf16* o, float* lse, int B, int N, int H, cudaStream_t st) {
  if (N > NP || N < 1) { s4g_set_error("attn_fwd_mma: specialised for sequences of at most 208 tokens"); return S4G_ERR_SHAPE; }
  const size_t smem = 3 * NP * 128;
  S4G_CUDA(cudaFuncSetAttribute(attn_fwd_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  attn_fwd_mma_kernel<<<B * H, THREADS, smem, st>>>(qkv, o, lse, N, H);
  S4G_KERNEL_CHECK();
  return 0;
}
int attn_bwd_mma(const bf16* qkv, const bf16* o, const bf16* dout, const float* lse, bf16* dqkv, int B, int N, int H,
                 cudaStream_t st) {
  if (N > NP || N < 1) { s4g_set_error("attn_bwd_mma: specialised for sequences of at most 208 tokens"); return S4G_ERR_SHAPE; }
  const size_t smem = 4 * NP * 128 + 2 * NP * sizeof(float);
  S4G_CUDA(cudaFuncSetAttribute(attn_bwd_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  attn_bwd_mma_kernel<<<B * H, THREADS, smem, st>>>(qkv, o, dout, lse, dqkv, N, H);
  S4G_KERNEL_CHECK();
  return 0;
}
