// Fused attention for the 197-token ViT-B/16 sequence on the 5th-generation tensor cores (tcgen05.mma, accumulators and
// the probability matrix in tensor memory, operands by TMA).  Reference semantics: timm 0.6.12 Attention.forward as called
// from Models/models.py:333-336 (12 heads, d = 64, scale = 64^-0.5, softmax over all 197 keys, no dropout).
//
// The whole key range of a head fits one MMA (N = 208 = 197 padded to 13 x 16), so the softmax is single pass: no running
// maximum, no rescaling of the output accumulator.
//
// forward, one persistent CTA per SM, 10 warps:
//   warp 0      TMA producer: Q, K, V of the next (image, head) into a 2-stage ring (3-D tensor map over qkv, the rows
//               197..207 of every box are out of bounds and arrive as zeros)
//   warp 1      MMA issuer: S_t = Q_t K^T (M = 128 query rows, N = 208, K = 64) into TMEM slot t, then O_t = P_t V with
//               P_t read straight from tensor memory (A operand in TMEM) and V as an MN-major B operand
//   warps 2-5   softmax + epilogue of query tile 0 (rows 0..127), warps 6-9 of query tile 1 (rows 128..196): a thread owns
//               one query row (= one TMEM lane): row max, exp2, row sum, bf16 P written over the first half of S,
//               later O / sum -> global.  The two tiles run half a phase apart, so the MMAs of one overlap the exponentials
//               of the other.
// backward, one persistent CTA per SM, 10 warps, keys on the TMEM lanes (S^T = K Q^T, dP^T = V dO^T), so that P^T and
// dS^T are A operands in tensor memory for dV = P^T dO and dK = dS^T Q; dS^T additionally goes through a 32 KB shared
// memory tile as the MN-major A operand of dQ = dS K.  Four steps per head (2 key tiles x 2 query halves), accumulators
// of dV, dK (per key tile) and dQ (both query tiles) stay in tensor memory: 512 columns exactly.  No atomics anywhere:
// the result is bit-reproducible.
#include <cuda.h>

#include <cstdlib>

#include "kernels.h"
#include "umma_ptx.cuh"

namespace {

using namespace s4g_ptx;

constexpr int D = 64;                    // head dim
constexpr int NP = 208;                  // tokens padded to 13 x 16
constexpr int TILE_BYTES = NP * 128;     // one of Q / K / V / dO in shared memory: rows of 128 B, SWIZZLE_128B
constexpr float LOG2E = 1.4426950408889634f;
constexpr float SCALE = 0.125f;          // 64^-0.5
constexpr int THREADS = 320;

__device__ __forceinline__ uint8_t* align_1024(uint8_t* p) {
  return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~(uintptr_t)1023);
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// =========================================================================================== forward
constexpr int FWD_HALF_ROWS = 104;                       // TMA box rows: two boxes per tensor
constexpr int FWD_STAGE_BYTES = 3 * TILE_BYTES;
constexpr int FWD_STAGES = 2;
constexpr int SLOT_COLS = 256;                           // TMEM columns per query tile: S [0,208), P [0,104), O [128,192)
constexpr int O_COL = 128;

struct FwdCtl {
  uint64_t full[FWD_STAGES], empty[FWD_STAGES], s_full[2], p_full[2], o_full[2], slot_free[2];
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(THREADS, 1) attn_fwd_umma_kernel(const __grid_constant__ CUtensorMap tmQKV, bf16* __restrict__ o,
                                                                   float* __restrict__ lse, int N, int H, int total) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  FwdCtl* ctl = reinterpret_cast<FwdCtl*>(smem + FWD_STAGES * FWD_STAGE_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = H * D;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    for (int i = 0; i < FWD_STAGES; ++i) { mbar_init(&ctl->full[i], 1); mbar_init(&ctl->empty[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&ctl->s_full[i], 1); mbar_init(&ctl->p_full[i], 4); mbar_init(&ctl->o_full[i], 1); mbar_init(&ctl->slot_free[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&ctl->tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();     // everything above overlapped the previous kernel's tail; global memory is touched only below
  pdl_trigger();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int i = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x, ++i) {
        const int stage = i & 1;
        mbar_wait(&ctl->empty[stage], ((i >> 1) & 1) ^ 1);
        mbar_expect_tx(&ctl->full[stage], FWD_STAGE_BYTES);
        const int b = w / H, h = w % H;
        uint8_t* s = smem + stage * FWD_STAGE_BYTES;
        for (int t = 0; t < 3; ++t)
          for (int half = 0; half < 2; ++half)
            tma_load_3d(&tmQKV, &ctl->full[stage], s + t * TILE_BYTES + half * FWD_HALF_ROWS * 128, t * C + h * D, half * FWD_HALF_ROWS, b);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t idesc_s = make_idesc(128, NP, 0, 0);     // S = Q K^T: both operands K-major
    const uint32_t idesc_o = make_idesc(128, D, 0, 1);      // O = P V: A in TMEM, V MN-major (rows = keys)
    // O_t of head ip: needs P_t complete; the last reader of the head's shared memory stage
    auto issue_pv = [&](int ip, int t) {
      mbar_wait(&ctl->p_full[t], ip & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sv = smem_u32(smem + (ip & 1) * FWD_STAGE_BYTES + 2 * TILE_BYTES);
        const uint32_t slot = tmem_base + t * SLOT_COLS;
#pragma unroll
        for (int kk = 0; kk < NP / 16; ++kk)
          umma_bf16_ts(slot + O_COL, slot + kk * 8, make_desc(sv + kk * 2048, TILE_BYTES, 1024), idesc_o, kk != 0);
        umma_commit(&ctl->o_full[t]);
        if (t == 1) umma_commit(&ctl->empty[ip & 1]);
      }
      __syncwarp();
    };
    int i = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++i) {
      const int stage = i & 1;
      mbar_wait(&ctl->full[stage], (i >> 1) & 1);
      tc_fence_after();
      for (int t = 0; t < 2; ++t) {
        mbar_wait(&ctl->slot_free[t], (i & 1) ^ 1);   // the epilogue of the previous head has read O_t out of this slot
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sq = smem_u32(smem + stage * FWD_STAGE_BYTES) + t * 128 * 128;
          const uint32_t sk = smem_u32(smem + stage * FWD_STAGE_BYTES + TILE_BYTES);
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk)
            umma_bf16(tmem_base + t * SLOT_COLS, make_desc(sq + kk * 32, 16, 1024), make_desc(sk + kk * 32, 16, 1024), idesc_s, kk != 0);
          umma_commit(&ctl->s_full[t]);
        }
        __syncwarp();
        // the P V product of the job issued before this one (tile 1 of the previous head, or tile 0 of this head)
        if (t == 0) { if (i > 0) issue_pv(i - 1, 1); } else issue_pv(i, 0);
      }
    }
    if (i > 0) issue_pv(i - 1, 1);
  } else {
    // ------------------------------------------------------------------ softmax + epilogue: warps 2-5 tile 0, 6-9 tile 1
    const int t = (warp - 2) >> 2;
    const int q = warp & 3;                    // TMEM lane quarter this warp may touch
    const int row = t * 128 + q * 32 + lane;   // query row of this thread
    const bool warp_active = t * 128 + q * 32 < N;
    const uint32_t taddr = tmem_base + t * SLOT_COLS + ((uint32_t)(q * 32) << 16);
    const float sl2 = SCALE * LOG2E;
    int i = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++i) {
      const int b = w / H, h = w % H;
      mbar_wait(&ctl->s_full[t], i & 1);
      tc_fence_after();
      float mx = -INFINITY, sum = 0.f;
      if (warp_active) {
        // pass 1: row maximum over the valid keys
#pragma unroll 1
        for (int c0 = 0; c0 < NP; c0 += 32) {
          if (c0 + 32 <= NP) {
            uint32_t v[32];
            tmem_ld32(taddr + c0, v);
            tmem_ld_wait();
            if (c0 + 32 <= N) {
#pragma unroll
              for (int k = 0; k < 32; ++k) mx = fmaxf(mx, __uint_as_float(v[k]));
            } else {
#pragma unroll
              for (int k = 0; k < 32; ++k) if (c0 + k < N) mx = fmaxf(mx, __uint_as_float(v[k]));
            }
          } else {
            uint32_t v[16];
            tmem_ld16(taddr + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 16; ++k) if (c0 + k < N) mx = fmaxf(mx, __uint_as_float(v[k]));
          }
        }
        const float nb = -mx * sl2;
        // pass 2: p = exp2((s - max) * scale * log2 e) -> bf16 pairs written over the first half of S (a thread only ever
        // touches its own lane, and P chunk c lands on columns this thread has already read)
#pragma unroll 1
        for (int c0 = 0; c0 < NP; c0 += 32) {
          if (c0 + 32 <= NP) {
            uint32_t v[32], pk[16];
            tmem_ld32(taddr + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 32; k += 2) {
              float p0 = ex2_approx(fmaf(__uint_as_float(v[k]), sl2, nb)), p1 = ex2_approx(fmaf(__uint_as_float(v[k + 1]), sl2, nb));
              if (c0 + k >= N) p0 = 0.f;
              if (c0 + k + 1 >= N) p1 = 0.f;
              sum += p0 + p1;
              pk[k >> 1] = pack_bf16(p0, p1);
            }
            tmem_st16(taddr + (c0 >> 1), pk);
          } else {
            uint32_t v[16], pk[8];
            tmem_ld16(taddr + c0, v);
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 16; k += 2) {
              float p0 = ex2_approx(fmaf(__uint_as_float(v[k]), sl2, nb)), p1 = ex2_approx(fmaf(__uint_as_float(v[k + 1]), sl2, nb));
              if (c0 + k >= N) p0 = 0.f;
              if (c0 + k + 1 >= N) p1 = 0.f;
              sum += p0 + p1;
              pk[k >> 1] = pack_bf16(p0, p1);
            }
            tmem_st8(taddr + (c0 >> 1), pk);
          }
        }
        tmem_st_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl->p_full[t]);
      // epilogue: O / sum -> bf16 -> global
      mbar_wait(&ctl->o_full[t], i & 1);
      tc_fence_after();
      uint32_t acc[2][32];
      if (warp_active) {
        tmem_ld32(taddr + O_COL, acc[0]);
        tmem_ld32(taddr + O_COL + 32, acc[1]);
        tmem_ld_wait();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ctl->slot_free[t]);   // the slot may take the next head's S while the stores drain
      if (warp_active && row < N) {
        const float inv = 1.f / sum;
        bf16* dst = o + ((long long)b * N + row) * C + h * D;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const uint32_t* a = acc[g >> 2] + (g & 3) * 8;
          uint4 pk;
          pk.x = pack_bf16(__uint_as_float(a[0]) * inv, __uint_as_float(a[1]) * inv);
          pk.y = pack_bf16(__uint_as_float(a[2]) * inv, __uint_as_float(a[3]) * inv);
          pk.z = pack_bf16(__uint_as_float(a[4]) * inv, __uint_as_float(a[5]) * inv);
          pk.w = pack_bf16(__uint_as_float(a[6]) * inv, __uint_as_float(a[7]) * inv);
          *reinterpret_cast<uint4*>(dst + g * 8) = pk;
        }
        lse[((long long)b * H + h) * N + row] = mx * SCALE + logf(sum);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// =========================================================================================== backward
// shared memory: Q | K | V | dO (each [lo 128 rows | hi 80 rows] x 128 B), dS^T staging (2 query groups x 128 keys x 128 B),
// lse * log2(e) and delta of the 208 (padded) queries, barriers
constexpr int LO_ROWS = 128, HI_ROWS = NP - LO_ROWS;    // 128 + 80
constexpr int STG_BYTES = 2 * 128 * 128;                 // 32 KB
constexpr int BWD_SMEM_TENSORS = 4 * TILE_BYTES;
// TMEM columns
constexpr int COL_S = 0, COL_DP = 128, COL_DV = 256, COL_DK = 320, COL_DQ = 384;

struct BwdCtl {
  uint64_t full[3], empty[3];       // input groups: 0 = {K_lo, V_lo}, 1 = {Q_lo, dO_lo}, 2 = {K_hi, V_hi, Q_hi, dO_hi}
  uint64_t sdp_full, p_ready, stg_free, dvdk_full, dvdk_free, dq_full, dq_free;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(THREADS, 1) attn_bwd_umma_kernel(const __grid_constant__ CUtensorMap tmQKVlo,
                                                                   const __grid_constant__ CUtensorMap tmQKVhi,
                                                                   const __grid_constant__ CUtensorMap tmDOlo,
                                                                   const __grid_constant__ CUtensorMap tmDOhi,
                                                                   const bf16* __restrict__ o, const bf16* __restrict__ dout,
                                                                   const float* __restrict__ lse, bf16* __restrict__ dqkv,
                                                                   int N, int H, int total) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  uint8_t* sQ = smem;
  uint8_t* sK = smem + TILE_BYTES;
  uint8_t* sV = smem + 2 * TILE_BYTES;
  uint8_t* sDO = smem + 3 * TILE_BYTES;
  uint8_t* stg = smem + BWD_SMEM_TENSORS;
  float* s_lse = reinterpret_cast<float*>(stg + STG_BYTES);   // lse * log2(e); padded queries: +big  (=> P = 0)
  float* s_delta = s_lse + NP;
  BwdCtl* ctl = reinterpret_cast<BwdCtl*>(s_delta + NP);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = H * D, C3 = 3 * C;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKVlo); tma_prefetch_desc(&tmQKVhi); tma_prefetch_desc(&tmDOlo); tma_prefetch_desc(&tmDOhi);
    for (int i = 0; i < 3; ++i) { mbar_init(&ctl->full[i], 1); mbar_init(&ctl->empty[i], 1); }
    mbar_init(&ctl->sdp_full, 1); mbar_init(&ctl->p_ready, 8); mbar_init(&ctl->stg_free, 1);
    mbar_init(&ctl->dvdk_full, 1); mbar_init(&ctl->dvdk_free, 8); mbar_init(&ctl->dq_full, 1); mbar_init(&ctl->dq_free, 8);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&ctl->tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();     // everything above overlapped the previous kernel's tail; global memory is touched only below
  pdl_trigger();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (single-buffered, three groups)
    if (elect_one()) {
      int i = 0;
      for (int w = blockIdx.x; w < total; w += gridDim.x, ++i) {
        const int b = w / H, h = w % H;
        const uint32_t ph = (i & 1) ^ 1;
        mbar_wait(&ctl->empty[0], ph);
        mbar_expect_tx(&ctl->full[0], 2 * LO_ROWS * 128);
        tma_load_3d(&tmQKVlo, &ctl->full[0], sK, C + h * D, 0, b);
        tma_load_3d(&tmQKVlo, &ctl->full[0], sV, 2 * C + h * D, 0, b);
        mbar_wait(&ctl->empty[1], ph);
        mbar_expect_tx(&ctl->full[1], 2 * LO_ROWS * 128);
        tma_load_3d(&tmQKVlo, &ctl->full[1], sQ, h * D, 0, b);
        tma_load_3d(&tmDOlo, &ctl->full[1], sDO, h * D, 0, b);
        mbar_wait(&ctl->empty[2], ph);
        mbar_expect_tx(&ctl->full[2], 4 * HI_ROWS * 128);
        tma_load_3d(&tmQKVhi, &ctl->full[2], sK + LO_ROWS * 128, C + h * D, LO_ROWS, b);
        tma_load_3d(&tmQKVhi, &ctl->full[2], sV + LO_ROWS * 128, 2 * C + h * D, LO_ROWS, b);
        tma_load_3d(&tmQKVhi, &ctl->full[2], sQ + LO_ROWS * 128, h * D, LO_ROWS, b);
        tma_load_3d(&tmDOhi, &ctl->full[2], sDO + LO_ROWS * 128, h * D, LO_ROWS, b);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t uQ = smem_u32(sQ), uK = smem_u32(sK), uV = smem_u32(sV), uDO = smem_u32(sDO), uStg = smem_u32(stg);
    int i = 0, step = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++i) {
      const uint32_t ph = i & 1;
      for (int t = 0; t < 2; ++t) {
        for (int qh = 0; qh < 2; ++qh, ++step) {
          const int nq = qh ? HI_ROWS : LO_ROWS;       // query columns of this step
          const int nk = t ? HI_ROWS : LO_ROWS;        // keys that take part in the dQ contraction
          if (t == 0 && qh == 0) { mbar_wait(&ctl->full[0], ph); mbar_wait(&ctl->full[1], ph); }
          if (t == 0 && qh == 1) mbar_wait(&ctl->full[2], ph);
          tc_fence_after();
          if (elect_one()) {
            // S^T = K_t Q_qh^T and dP^T = V_t dO_qh^T (M = 128 keys, N = nq queries, K = 64)
            const uint32_t idesc = make_idesc(128, nq, 0, 0);
            const uint32_t ak = uK + t * 128 * 128, av = uV + t * 128 * 128, bq = uQ + qh * 128 * 128, bdo = uDO + qh * 128 * 128;
#pragma unroll
            for (int kk = 0; kk < D / 16; ++kk)
              umma_bf16(tmem_base + COL_S, make_desc(ak + kk * 32, 16, 1024), make_desc(bq + kk * 32, 16, 1024), idesc, kk != 0);
#pragma unroll
            for (int kk = 0; kk < D / 16; ++kk)
              umma_bf16(tmem_base + COL_DP, make_desc(av + kk * 32, 16, 1024), make_desc(bdo + kk * 32, 16, 1024), idesc, kk != 0);
            umma_commit(&ctl->sdp_full);
          }
          __syncwarp();
          // P^T, dS^T (tensor memory) and dS^T (staging) of this step are complete
          mbar_wait(&ctl->p_ready, step & 1);
          if (qh == 0 && (t == 1 || i > 0)) mbar_wait(&ctl->dvdk_free, (2 * i + t - 1) & 1);   // previous key tile's dV / dK read out
          if (t == 0 && qh == 0 && i > 0) mbar_wait(&ctl->dq_free, (i - 1) & 1);               // previous head's dQ read out
          tc_fence_after();
          if (elect_one()) {
            const uint32_t idesc_kv = make_idesc(128, D, 0, 1);    // A in TMEM (K-major), B = dO / Q rows, MN-major
            const uint32_t idesc_q = make_idesc(128, D, 1, 1);     // A = dS^T staging (MN-major), B = K rows (MN-major)
            const int n0 = nq == LO_ROWS ? 64 : 48;                // columns owned by the first compute warp group
            for (int kk = 0; kk < nq / 16; ++kk) {
              const int qc = kk * 16;                               // query offset inside this half
              const uint32_t acol = qc < n0 ? (qc >> 1) : n0 + ((qc - n0) >> 1);   // where that group put its packed columns
              umma_bf16_ts(tmem_base + COL_DV, tmem_base + COL_S + acol, make_desc(uDO + (qh * 128 + qc) * 128, TILE_BYTES, 1024), idesc_kv,
                           (qh | kk) != 0);
            }
            for (int kk = 0; kk < nq / 16; ++kk) {
              const int qc = kk * 16;
              const uint32_t acol = qc < n0 ? (qc >> 1) : n0 + ((qc - n0) >> 1);
              umma_bf16_ts(tmem_base + COL_DK, tmem_base + COL_DP + acol, make_desc(uQ + (qh * 128 + qc) * 128, TILE_BYTES, 1024), idesc_kv,
                           (qh | kk) != 0);
            }
            for (int kk = 0; kk < nk / 16; ++kk)
              umma_bf16(tmem_base + COL_DQ + qh * 64, make_desc(uStg + kk * 2048, 128 * 128, 1024),
                        make_desc(uK + (t * 128 + kk * 16) * 128, TILE_BYTES, 1024), idesc_q, (t | kk) != 0);
            umma_commit(&ctl->stg_free);
            if (qh == 1) umma_commit(&ctl->dvdk_full);
            if (t == 1 && qh == 1) umma_commit(&ctl->dq_full);
            // input groups whose last reader has been issued
            if (t == 0 && qh == 1) umma_commit(&ctl->empty[0]);
            if (t == 1 && qh == 0) umma_commit(&ctl->empty[1]);
            if (t == 1 && qh == 1) umma_commit(&ctl->empty[2]);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ compute warps: thread = key row (TMEM lane)
    const int wg = (warp - 2) >> 2;
    const int q = warp & 3;
    const int ctid = threadIdx.x - 64;     // 0..255
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const float sl2 = SCALE * LOG2E;
    int i = 0, step = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++i) {
      const int b = w / H, h = w % H;
      // per-query statistics of this head: delta_i = sum_d dO[i,d] O[i,d] (straight from global memory), lse_i * log2(e)
      asm volatile("bar.sync 1, 256;" ::: "memory");   // every compute warp is done with the previous head's statistics
      {
        const bf16* dob = dout + (long long)b * N * C + h * D;
        const bf16* ob = o + (long long)b * N * C + h * D;
        for (int r = ctid; r < NP; r += 256) {
          float acc = 0.f;
          if (r < N) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              Vec8<bf16> x, y;
              x.load(dob + (long long)r * C + k * 8);
              y.load(ob + (long long)r * C + k * 8);
#pragma unroll
              for (int e = 0; e < 8; ++e) acc += x.v[e] * y.v[e];
            }
          }
          s_delta[r] = acc;
          s_lse[r] = r < N ? lse[((long long)b * H + h) * N + r] * LOG2E : 1e30f;
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      for (int t = 0; t < 2; ++t) {
        const bool warp_active = t * 128 + q * 32 < NP;    // key rows of this warp exist in shared memory (incl. zero padding)
        const int krow = q * 32 + lane;                     // key row inside the tile
        for (int qh = 0; qh < 2; ++qh, ++step) {
          const int nq = qh ? HI_ROWS : LO_ROWS;
          const int n0 = nq == LO_ROWS ? 64 : 48;
          const int c_begin = wg ? n0 : 0, c_end = wg ? nq : n0;
          mbar_wait(&ctl->sdp_full, step & 1);
          if (step > 0) mbar_wait(&ctl->stg_free, (step - 1) & 1);   // the previous step's dQ MMAs have read the staging tile
          tc_fence_after();
          if (warp_active) {
#pragma unroll 1
            for (int c0 = c_begin; c0 < c_end; c0 += 16) {
              uint32_t sv[16], dv[16], pp[8], ds[8];
              tmem_ld16(tmem_base + lane_off + COL_S + c0, sv);
              tmem_ld16(tmem_base + lane_off + COL_DP + c0, dv);
              tmem_ld_wait();
              const int qi = qh * 128 + c0;   // global query index of column c0
#pragma unroll
              for (int k = 0; k < 16; k += 4) {
                const float4 l4 = *reinterpret_cast<const float4*>(s_lse + qi + k);
                const float4 d4 = *reinterpret_cast<const float4*>(s_delta + qi + k);
                const float p0 = ex2_approx(fmaf(__uint_as_float(sv[k]), sl2, -l4.x)), p1 = ex2_approx(fmaf(__uint_as_float(sv[k + 1]), sl2, -l4.y));
                const float p2 = ex2_approx(fmaf(__uint_as_float(sv[k + 2]), sl2, -l4.z)), p3 = ex2_approx(fmaf(__uint_as_float(sv[k + 3]), sl2, -l4.w));
                pp[k >> 1] = pack_bf16(p0, p1);
                pp[(k >> 1) + 1] = pack_bf16(p2, p3);
                ds[k >> 1] = pack_bf16(p0 * (__uint_as_float(dv[k]) - d4.x) * SCALE, p1 * (__uint_as_float(dv[k + 1]) - d4.y) * SCALE);
                ds[(k >> 1) + 1] = pack_bf16(p2 * (__uint_as_float(dv[k + 2]) - d4.z) * SCALE, p3 * (__uint_as_float(dv[k + 3]) - d4.w) * SCALE);
              }
              // packed columns go to this warp group's own part of the S / dP regions (already read by this thread)
              const uint32_t pc = c_begin + ((c0 - c_begin) >> 1);
              tmem_st8(tmem_base + lane_off + COL_S + pc, pp);
              tmem_st8(tmem_base + lane_off + COL_DP + pc, ds);
              // dS^T staging: row = key, 64-query groups 16 KB apart, 16 B chunks XOR-swizzled by the row (SWIZZLE_128B)
              uint8_t* srow = stg + (c0 >> 6) * (128 * 128) + krow * 128;
              const int ch = (c0 & 63) >> 3;
              *reinterpret_cast<uint4*>(srow + (((ch) ^ (krow & 7)) << 4)) = make_uint4(ds[0], ds[1], ds[2], ds[3]);
              *reinterpret_cast<uint4*>(srow + (((ch + 1) ^ (krow & 7)) << 4)) = make_uint4(ds[4], ds[5], ds[6], ds[7]);
            }
            tmem_st_wait();
          }
          fence_proxy_async();   // the staging stores must be visible to the tensor core's (async proxy) reads
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&ctl->p_ready);
          if (qh == 1) {
            // dV (warp group 0) / dK (warp group 1) of key tile t
            mbar_wait(&ctl->dvdk_full, (2 * i + t) & 1);
            tc_fence_after();
            uint32_t acc[2][32];
            if (warp_active) {
              tmem_ld32(tmem_base + lane_off + (wg ? COL_DK : COL_DV), acc[0]);
              tmem_ld32(tmem_base + lane_off + (wg ? COL_DK : COL_DV) + 32, acc[1]);
              tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->dvdk_free);
            const int key = t * 128 + krow;
            if (warp_active && key < N) {
              bf16* dst = dqkv + ((long long)b * N + key) * C3 + (wg ? C : 2 * C) + h * D;
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                const uint32_t* a = acc[g >> 2] + (g & 3) * 8;
                *reinterpret_cast<uint4*>(dst + g * 8) =
                    make_uint4(pack_bf16(__uint_as_float(a[0]), __uint_as_float(a[1])), pack_bf16(__uint_as_float(a[2]), __uint_as_float(a[3])),
                               pack_bf16(__uint_as_float(a[4]), __uint_as_float(a[5])), pack_bf16(__uint_as_float(a[6]), __uint_as_float(a[7])));
              }
            }
          }
          if (t == 1 && qh == 1) {
            // dQ: warp group g reads query tile g
            mbar_wait(&ctl->dq_full, i & 1);
            tc_fence_after();
            const bool rows_here = wg * 128 + q * 32 < N;
            uint32_t acc[2][32];
            if (rows_here) {
              tmem_ld32(tmem_base + lane_off + COL_DQ + wg * 64, acc[0]);
              tmem_ld32(tmem_base + lane_off + COL_DQ + wg * 64 + 32, acc[1]);
              tmem_ld_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->dq_free);
            const int qrow = wg * 128 + q * 32 + lane;
            if (rows_here && qrow < N) {
              bf16* dst = dqkv + ((long long)b * N + qrow) * C3 + h * D;
#pragma unroll
              for (int g = 0; g < 8; ++g) {
                const uint32_t* a = acc[g >> 2] + (g & 3) * 8;
                *reinterpret_cast<uint4*>(dst + g * 8) =
                    make_uint4(pack_bf16(__uint_as_float(a[0]), __uint_as_float(a[1])), pack_bf16(__uint_as_float(a[2]), __uint_as_float(a[3])),
                               pack_bf16(__uint_as_float(a[4]), __uint_as_float(a[5])), pack_bf16(__uint_as_float(a[6]), __uint_as_float(a[7])));
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

int make_rows_map(CUtensorMap* map, const bf16* base, int cols, int N, int B, int box_rows) {
  // [B, N, cols] row-major as a 3-D tensor {cols, N, B}: rows past N of a box are out of bounds and are zero-filled
  uint64_t dims[3] = {(uint64_t)cols, (uint64_t)N, (uint64_t)B};
  uint64_t strides[2] = {(uint64_t)cols * 2, (uint64_t)N * cols * 2};
  uint32_t box[3] = {64, (uint32_t)box_rows, 1};
  return umma_make_map_bf16(map, base, 3, dims, strides, box);
}

}  // namespace

static int attn_umma_grid(int total) {
  const int sms = umma_num_sms();
  return total < sms ? total : sms;
}

int attn_fwd_umma(const bf16* qkv, bf16* o, float* lse, int B, int N, int H, cudaStream_t st) {
  if (N > NP || N < 1) { s4g_set_error("attn_fwd_umma: specialised for sequences of at most 208 tokens"); return S4G_ERR_SHAPE; }
  CUtensorMap tm;
  S4G_TRY(make_rows_map(&tm, qkv, 3 * H * D, N, B, FWD_HALF_ROWS));
  const size_t smem = 1024 + FWD_STAGES * FWD_STAGE_BYTES + sizeof(FwdCtl);
  static bool attr = false;
  if (!attr) { S4G_CUDA(cudaFuncSetAttribute(attn_fwd_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
  S4G_CUDA(launch_dep(attn_fwd_umma_kernel, dim3(attn_umma_grid(B * H)), dim3(THREADS), smem, st, tm, o, lse, N, H, B * H));
  S4G_KERNEL_CHECK();
  return 0;
}

int attn_bwd_umma(const bf16* qkv, const bf16* o, const bf16* dout, const float* lse, bf16* dqkv, int B, int N, int H,
                  cudaStream_t st) {
  if (N > NP || N < 1) { s4g_set_error("attn_bwd_umma: specialised for sequences of at most 208 tokens"); return S4G_ERR_SHAPE; }
  CUtensorMap qlo, qhi, dlo, dhi;
  S4G_TRY(make_rows_map(&qlo, qkv, 3 * H * D, N, B, LO_ROWS));
  S4G_TRY(make_rows_map(&qhi, qkv, 3 * H * D, N, B, HI_ROWS));
  S4G_TRY(make_rows_map(&dlo, dout, H * D, N, B, LO_ROWS));
  S4G_TRY(make_rows_map(&dhi, dout, H * D, N, B, HI_ROWS));
  const size_t smem = 1024 + BWD_SMEM_TENSORS + STG_BYTES + 2 * NP * sizeof(float) + sizeof(BwdCtl);
  static bool attr = false;
  if (!attr) { S4G_CUDA(cudaFuncSetAttribute(attn_bwd_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); attr = true; }
  S4G_CUDA(launch_dep(attn_bwd_umma_kernel, dim3(attn_umma_grid(B * H)), dim3(THREADS), smem, st, qlo, qhi, dlo, dhi, o, dout, lse, dqkv, N, H,
                      B * H));
  S4G_KERNEL_CHECK();
  return 0;
}
