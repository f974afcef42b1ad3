// Fused attention for the 197-token ViT-B/16 sequence on bf16 tensor cores (mma.sync m16n8k16, fp32 accumulate).
// Persistent CTAs (one per SM, 13 warps = one 16-row tile each) loop over (image, head) pairs; Q, K, V (and dO in
// backward) of the NEXT pair stream into the other half of a double-buffered, XOR-swizzled shared-memory stage with
// cp.async while the current pair is computed.  The 197x197 score matrix never leaves registers; only LSE is saved.
//   forward : S = Q K^T -> softmax -> O = P V
//   backward: phase A per 16-query tile : S, dP = dO V^T, dS = P (dP - delta) s ; dQ = dS K
//             phase B per 16-key tile   : S^T = K Q^T, dP^T = V dO^T, dS^T ; dV = P^T dO, dK = dS^T Q
//             (scores are recomputed in both phases: no atomics, no cross-warp reduction, deterministic)
// Reference semantics: timm 0.6.12 Attention.forward (called from Models/models.py:333-334), scale = 64^-0.5.
#include <cstdlib>
#include "kernels.h"

namespace {

constexpr int D = 64;        // head dim
constexpr int NP = 208;      // 197 padded to 13 x 16
constexpr int NT = NP / 8;   // 26 n-tiles of 8
constexpr int WARPS = 13;    // one 16-row tile per warp
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

constexpr int TILE_BYTES = NP * 128;

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// global [tokens, ld] slice (64 contiguous bf16 per row at `base + row*ld`) -> swizzled smem [NP][64], rows < n only
// (the padding rows n..NP-1 are zeroed once per stage by zero_pad_rows)
__device__ __forceinline__ void load_tile_async(uint32_t dst, const bf16* base, long long ld, int n) {
  for (int i = threadIdx.x; i < n * 8; i += THREADS) {
    const int row = i >> 3, chunk = i & 7;
    cp_async16(dst + sw_off(row, chunk), base + (long long)row * ld + chunk * 8);
  }
}
__device__ __forceinline__ void zero_pad_rows(uint8_t* tile, int n) {
  for (int i = n * 8 + threadIdx.x; i < NP * 8; i += THREADS)
    *reinterpret_cast<uint4*>(tile + sw_off(i >> 3, i & 7)) = make_uint4(0, 0, 0, 0);
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
  static_assert(NTILES % 2 == 0, "n-tiles are processed in pairs");
#pragma unroll
  for (int j = 0; j < NTILES; j += 2) {
    acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    acc[j + 1][0] = acc[j + 1][1] = acc[j + 1][2] = acc[j + 1][3] = 0.f;
    const int row = n0 + j * 8 + (lane & 7);
    uint32_t b[8], e[8];
    ldsm_x4(smemB + sw_off(row, lane >> 3), b[0], b[1], b[2], b[3]);            // d 0..31
    ldsm_x4(smemB + sw_off(row, 4 + (lane >> 3)), b[4], b[5], b[6], b[7]);      // d 32..63
    ldsm_x4(smemB + sw_off(row + 8, lane >> 3), e[0], e[1], e[2], e[3]);
    ldsm_x4(smemB + sw_off(row + 8, 4 + (lane >> 3)), e[4], e[5], e[6], e[7]);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      mma_bf16(acc[j], a[ks], b[2 * ks], b[2 * ks + 1]);
      mma_bf16(acc[j + 1], a[ks], e[2 * ks], e[2 * ks + 1]);
    }
  }
}

// one k-step (16 contraction rows starting at row0 of the row-major [NP][64] matrix B): out[8 d-tiles] += P(16x16) . B
__device__ __forceinline__ void pb_step(float (&out)[8][4], const uint32_t (&p)[4], uint32_t smemB, int row0) {
  const int lane = threadIdx.x & 31;
  const int row = row0 + (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
  for (int jp = 0; jp < 4; ++jp) {
    uint32_t b0, b1, b2, b3;
    ldsm_x4_t(smemB + sw_off(row, jp * 2 + (lane >> 4)), b0, b1, b2, b3);
    mma_bf16(out[2 * jp], p, b0, b1);
    mma_bf16(out[2 * jp + 1], p, b2, b3);
  }
}

// ------------------------------------------------------------------------------------------ forward
// online softmax over one block of NH key n-tiles starting at key n0
template <int NH>
__device__ __forceinline__ void fwd_block(const uint32_t (&a)[4][4], uint32_t uK, uint32_t uV, int n0, int N, float (&acc)[8][4],
                                          float& m0, float& m1, float& l0, float& l1) {
  const int c = threadIdx.x & 3;
  const float sc = 0.125f * LOG2E;
  float s[NH][4];
  gemm_abt<NH>(s, a, uK, n0);
  float x0 = -INFINITY, x1 = -INFINITY;
#pragma unroll
  for (int j = 0; j < NH; ++j) {
    const int col = n0 + j * 8 + 2 * c;
    if (col >= N) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
    if (col + 1 >= N) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
    x0 = fmaxf(x0, fmaxf(s[j][0], s[j][1]));
    x1 = fmaxf(x1, fmaxf(s[j][2], s[j][3]));
  }
  x0 = fmaxf(x0, __shfl_xor_sync(0xffffffffu, x0, 1)); x0 = fmaxf(x0, __shfl_xor_sync(0xffffffffu, x0, 2));
  x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, 1)); x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, 2));
  const float n0m = fmaxf(m0, x0), n1m = fmaxf(m1, x1);   // finite: key 0 is never masked
  const float c0 = exp2f((m0 - n0m) * sc), c1 = exp2f((m1 - n1m) * sc);
  m0 = n0m; m1 = n1m;
  l0 *= c0; l1 *= c1;
#pragma unroll
  for (int j = 0; j < 8; ++j) { acc[j][0] *= c0; acc[j][1] *= c0; acc[j][2] *= c1; acc[j][3] *= c1; }
  const float b0 = m0 * sc, b1 = m1 * sc;
#pragma unroll
  for (int ks = 0; ks < NH / 2; ++ks) {
    uint32_t p[4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int j = 2 * ks + h;
      const float e0 = exp2f(fmaf(s[j][0], sc, -b0)), e1 = exp2f(fmaf(s[j][1], sc, -b0));
      const float e2 = exp2f(fmaf(s[j][2], sc, -b1)), e3 = exp2f(fmaf(s[j][3], sc, -b1));
      l0 += e0 + e1; l1 += e2 + e3;
      p[h * 2] = pack_bf16(e0, e1);
      p[h * 2 + 1] = pack_bf16(e2, e3);
    }
    pb_step(acc, p, uV, n0 + ks * 16);
  }
}

__global__ void __launch_bounds__(THREADS, 1) attn_fwd_mma_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ o,
                                                                  float* __restrict__ lse, int N, int H, int total) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  extern __shared__ __align__(128) uint8_t sm[];
  const int C = H * D, C3 = 3 * C;
  const uint32_t sm_u = (uint32_t)__cvta_generic_to_shared(sm);
  for (int t = 0; t < 6; ++t) zero_pad_rows(sm + t * TILE_BYTES, N);
  auto issue = [&](int work, int stage) {
    const bf16* base = qkv + (long long)(work / H) * N * C3 + (work % H) * D;
    for (int t = 0; t < 3; ++t) load_tile_async(sm_u + (stage * 3 + t) * TILE_BYTES, base + t * C, C3, N);
    cp_async_commit();
  };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  int stage = 0;
  if ((int)blockIdx.x < total) issue(blockIdx.x, 0);
  for (int work = blockIdx.x; work < total; work += gridDim.x, stage ^= 1) {
    const int next = work + gridDim.x;
    if (next < total) { issue(next, stage ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const int b = work / H, h = work % H;
    const uint32_t uQ = sm_u + stage * 3 * TILE_BYTES, uK = uQ + TILE_BYTES, uV = uK + TILE_BYTES;
    uint32_t a[4][4];
    load_a_frags(uQ, warp * 16, a);
    float acc[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    fwd_block<14>(a, uK, uV, 0, N, acc, m0, m1, l0, l1);
    fwd_block<12>(a, uK, uV, 112, N, acc, m0, m1, l0, l1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    const int r0 = warp * 16 + g, r1 = r0 + 8;
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
    __syncthreads();  // every warp is done with this stage before it is refilled
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

// phase A chunk: NH key n-tiles starting at key n0 for this warp's 16 queries
template <int NH>
__device__ __forceinline__ void bwd_a_chunk(const uint32_t (&aq)[4][4], const uint32_t (&ado)[4][4], uint32_t uK, uint32_t uV, int n0,
                                            int N, float lse0, float lse1, float dl0, float dl1, float (&acc)[8][4]) {
  float s[NH][4], dp[NH][4];
  uint32_t ds[NH / 2][4];
  gemm_abt<NH>(s, aq, uK, n0);
  gemm_abt<NH>(dp, ado, uV, n0);
  ds_rows<NH>(s, dp, n0, N, lse0, lse1, dl0, dl1, ds);
#pragma unroll
  for (int ks = 0; ks < NH / 2; ++ks) pb_step(acc, ds[ks], uK, n0 + ks * 16);
}
// phase B chunk: NH query n-tiles starting at query n0 for this warp's 16 keys
template <int NH>
__device__ __forceinline__ void bwd_b_chunk(const uint32_t (&ak)[4][4], const uint32_t (&av)[4][4], uint32_t uQ, uint32_t udO, int n0,
                                            const float* slse, const float* sdelta, float (&dk)[8][4], float (&dv)[8][4]) {
  float s[NH][4], dp[NH][4];
  uint32_t pt[NH / 2][4], ds[NH / 2][4];
  gemm_abt<NH>(s, ak, uQ, n0);
  gemm_abt<NH>(dp, av, udO, n0);
  ds_cols<NH>(s, dp, n0, slse, sdelta, pt, ds);
#pragma unroll
  for (int ks = 0; ks < NH / 2; ++ks) {
    pb_step(dv, pt[ks], udO, n0 + ks * 16);
    pb_step(dk, ds[ks], uQ, n0 + ks * 16);
  }
}

constexpr int BWD_STAGE_BYTES = 4 * TILE_BYTES + 2 * NP * (int)sizeof(float);

__global__ void __launch_bounds__(THREADS, 1) attn_bwd_mma_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ o,
                                                                  const bf16* __restrict__ dout, const float* __restrict__ lse,
                                                                  bf16* __restrict__ dqkv, int N, int H, int total) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  extern __shared__ __align__(128) uint8_t sm[];
  const int C = H * D, C3 = 3 * C;
  const uint32_t sm_u = (uint32_t)__cvta_generic_to_shared(sm);
  for (int st = 0; st < 2; ++st)
    for (int t = 0; t < 4; ++t) zero_pad_rows(sm + st * BWD_STAGE_BYTES + t * TILE_BYTES, N);
  auto issue = [&](int work, int stage) {
    const int b = work / H, h = work % H;
    const bf16* base = qkv + (long long)b * N * C3 + h * D;
    const uint32_t dst = sm_u + stage * BWD_STAGE_BYTES;
    for (int t = 0; t < 3; ++t) load_tile_async(dst + t * TILE_BYTES, base + t * C, C3, N);
    load_tile_async(dst + 3 * TILE_BYTES, dout + (long long)b * N * C + h * D, C, N);
    cp_async_commit();
  };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, c = lane & 3;
  int stage = 0;
  if ((int)blockIdx.x < total) issue(blockIdx.x, 0);
  for (int work = blockIdx.x; work < total; work += gridDim.x, stage ^= 1) {
    const int next = work + gridDim.x;
    if (next < total) issue(next, stage ^ 1);
    const int b = work / H, h = work % H;
    uint8_t* stg = sm + stage * BWD_STAGE_BYTES;
    float* slse = reinterpret_cast<float*>(stg + 4 * TILE_BYTES);  // lse * log2(e); padded queries: +big
    float* sdelta = slse + NP;
    {
      // delta_i = sum_d dO[i,d] * O[i,d] straight from global memory (overlaps the cp.async traffic); 4 lanes per row
      const bf16* dob = dout + (long long)b * N * C + h * D;
      const bf16* ob = o + (long long)b * N * C + h * D;
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
    }
    if (next < total) { cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const uint32_t uQ = sm_u + stage * BWD_STAGE_BYTES, uK = uQ + TILE_BYTES, uV = uK + TILE_BYTES, udO = uV + TILE_BYTES;
    bf16* dbase = dqkv + (long long)b * N * C3 + h * D;
    const int r0 = warp * 16 + g, r1 = r0 + 8;
    {
      // ---------------- phase A: dQ of this warp's 16 queries
      uint32_t aq[4][4], ado[4][4];
      load_a_frags(uQ, warp * 16, aq);
      load_a_frags(udO, warp * 16, ado);
      const float lse0 = slse[r0], lse1 = slse[r1], dl0 = sdelta[r0], dl1 = sdelta[r1];
      float acc[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
#pragma unroll 1
      for (int n0 = 0; n0 < 192; n0 += 32) bwd_a_chunk<4>(aq, ado, uK, uV, n0, N, lse0, lse1, dl0, dl1, acc);
      bwd_a_chunk<2>(aq, ado, uK, uV, 192, N, lse0, lse1, dl0, dl1, acc);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int d = j * 8 + 2 * c;
        if (r0 < N) *reinterpret_cast<uint32_t*>(dbase + (long long)r0 * C3 + d) = pack_bf16(acc[j][0], acc[j][1]);
        if (r1 < N) *reinterpret_cast<uint32_t*>(dbase + (long long)r1 * C3 + d) = pack_bf16(acc[j][2], acc[j][3]);
      }
    }
    {
      // ---------------- phase B: dK, dV of this warp's 16 keys (rows = keys, columns = queries)
      uint32_t ak[4][4], av[4][4];
      load_a_frags(uK, warp * 16, ak);
      load_a_frags(uV, warp * 16, av);
      float dk[8][4], dv[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dk[j][0] = dk[j][1] = dk[j][2] = dk[j][3] = 0.f;
        dv[j][0] = dv[j][1] = dv[j][2] = dv[j][3] = 0.f;
      }
#pragma unroll 1
      for (int n0 = 0; n0 < 192; n0 += 32) bwd_b_chunk<4>(ak, av, uQ, udO, n0, slse, sdelta, dk, dv);
      bwd_b_chunk<2>(ak, av, uQ, udO, 192, slse, sdelta, dk, dv);
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
    __syncthreads();  // every warp is done with this stage before it is refilled
  }
}

}  // namespace

static int g_attn_sms = 0;
static int attn_grid(int total) {
  if (!g_attn_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_attn_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_attn_sms <= 0) g_attn_sms = 148;
    if (const char* lim = getenv("S4G_SM_LIMIT")) { const int v = atoi(lim); if (v >= 2 && v < g_attn_sms) g_attn_sms = v; }
  }
  return total < g_attn_sms ? total : g_attn_sms;
}

// S4G_ATTN=mma keeps this mma.sync kernel (A/B measurements); the default tensor-core attention is attention_umma.cu
static bool use_umma_attention() {
  static const bool v = [] { const char* e = getenv("S4G_ATTN"); return !(e && std::string(e) == "mma"); }();
  return v;
}

int attn_fwd_mma(const bf16* qkv, bf16* o, float* lse, int B, int N, int H, cudaStream_t st) {
  if (use_umma_attention()) return attn_fwd_umma(qkv, o, lse, B, N, H, st);
  if (N > NP || N < 1) { s4g_set_error("attn_fwd_mma: specialised for sequences of at most 208 tokens"); return S4G_ERR_SHAPE; }
  const size_t smem = 2 * 3 * TILE_BYTES;
  static bool attr = false;
  if (!attr) { S4G_CUDA(cudaFuncSetAttribute(attn_fwd_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
  S4G_CUDA(launch_dep(attn_fwd_mma_kernel, dim3(attn_grid(B * H)), dim3(THREADS), smem, st, qkv, o, lse, N, H, B * H));
  S4G_KERNEL_CHECK();
  return 0;
}
bool attn_uses_umma() { return use_umma_attention(); }

int attn_bwd_mma(const bf16* qkv, const bf16* o, const bf16* dout, const float* lse, bf16* dqkv, int B, int N, int H,
                 cudaStream_t st, const float* delta) {
  if (use_umma_attention()) return attn_bwd_umma(qkv, o, dout, lse, dqkv, B, N, H, st, delta);
  if (N > NP || N < 1) { s4g_set_error("attn_bwd_mma: specialised for sequences of at most 208 tokens"); return S4G_ERR_SHAPE; }
  const size_t smem = 2 * BWD_STAGE_BYTES;
  static bool attr = false;
  if (!attr) { S4G_CUDA(cudaFuncSetAttribute(attn_bwd_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
  S4G_CUDA(launch_dep(attn_bwd_mma_kernel, dim3(attn_grid(B * H)), dim3(THREADS), smem, st, qkv, o, dout, lse, dqkv, N, H, B * H));
  S4G_KERNEL_CHECK();
  return 0;
}
