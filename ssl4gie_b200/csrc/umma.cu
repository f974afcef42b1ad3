// tcgen05 / TMEM / TMA contraction engine for sm_100a (bf16 operands, fp32 accumulation in tensor memory).
//
//   umma_conv_fprop : D[M, N] = sum_{tap, c} A_tap[M, c] * W[N, tap*Cin + c]   (+ fused epilogue)
//        A_tap is the NHWC activation shifted by the tap offset, fetched by 4-D TMA boxes whose
//        out-of-bounds elements the hardware zero-fills (= the conv's zero padding, no im2col buffer).
//        Linear layers / 1x1 convs are the 1-tap case.  Both operands K-major, SWIZZLE_128B.
//   umma_conv_wgrad : dW[Cout, tap*Cin + c] = sum_pixels dY[pix, Cout] * X_tap[pix, c]
//        contraction over pixels: both operands are MN-major (channel-contiguous) in shared memory,
//        the same TMA boxes as fprop, split over pixel ranges across CTAs, fp32 red.add epilogue.
//
// Kernel anatomy (one CTA per SM, persistent over tiles, 192 threads):
//   warp 0      TMA producer (one elected lane): global -> smem ring, mbarrier complete_tx
//   warp 1      tcgen05.mma issuer (one elected lane) + TMEM allocator; tcgen05.commit frees smem slots
//   warps 2..5  epilogue: tcgen05.ld 32x32b from TMEM (two accumulator stages ping-pong with the main loop),
//               bias / GELU / GELU' / ReLU / residual, vectorised global stores
#include <cuda.h>

#include <cstdlib>
#include "kernels.h"
#include "umma_ptx.cuh"

namespace {

constexpr int BLOCK_K = 64;           // bf16 elements = 128 B = one swizzle row
constexpr int MAX_STAGES = 8;          // smem ring depth: as many stages as fit, at most this
constexpr int A_STAGE_BYTES = 128 * 128;  // 128 rows x 128 B
constexpr int NUM_THREADS = 192;          // wgrad kernel: producer, MMA issuer, 4 epilogue warps
constexpr int EPI_WARPS = 8;              // fprop kernel: two epilogue warps per TMEM lane quarter (column slabs alternate)
constexpr int FPROP_THREADS = 64 + 32 * EPI_WARPS;

using namespace s4g_ptx;

// ------------------------------------------------------------------------------------------ shared layout
constexpr int TQ = 4;   // tile queue depth (dynamic scheduling)
struct SmemCtl {
  uint64_t full[MAX_STAGES], empty[MAX_STAGES], tfull[2], tempty[2];
  uint64_t qfull[TQ], qempty[TQ];   // dynamic tile scheduler: queue of tile indices, filled by the leader CTA's producer thread
  int qtile[TQ];
  uint32_t tmem_base;
};

// Dynamic tile scheduling.  With a static partition (tile = group + k * groups) a CTA that cannot be placed -- another
// kernel (an NCCL collective on the communication stream) holds its SM -- keeps its whole share of tiles waiting while the
// resident CTAs go idle.  Here the leader CTA's producer thread draws tile indices from a global counter and publishes each
// one to the MMA / epilogue warps of the pair and to the peer CTA's producer through a small shared-memory queue
// (st.async + mbarrier complete_tx, so the value is visible with the barrier phase); late CTAs simply find no work left.
// tile -> output mapping is unchanged, so results do not depend on who computed what.  counter[0] = next tile,
// counter[1] = groups that have finished; the last group to finish resets both for the next launch.
struct TileSched {
  SmemCtl* ctl;
  int* counter;      // null: static partition
  int group, num_groups, num_tiles, rank, cg, consumers;
  __device__ __forceinline__ bool dyn() const { return counter != nullptr; }
  __device__ __forceinline__ int fetch() const { return atomicAdd(counter, 1); }
  // leader producer thread: hand tile `t` (or -1 = no more work) of step k to every consumer of the pair
  __device__ __forceinline__ void publish(int k, int t) const {
    const int slot = k & (TQ - 1);
    mbar_wait(&ctl->qempty[slot], ((k / TQ) & 1) ^ 1);
    if (cg == 1) {   // no cluster: an ordinary store, published by the barrier arrival (release) / wait (acquire)
      *reinterpret_cast<volatile int*>(&ctl->qtile[slot]) = t;
      mbar_arrive(&ctl->qfull[slot]);
      return;
    }
    for (int c = 0; c < cg; ++c) st_async_u32(map_to_cta(&ctl->qtile[slot], c), (uint32_t)t, map_to_cta(&ctl->qfull[slot], c));
    mbar_expect_tx(&ctl->qfull[slot], 4);   // this CTA's arrival for the phase (the leader's producer does not read the queue)
  }
  // peer producer thread: arm this CTA's barrier, then read like any consumer
  __device__ __forceinline__ int take_armed(int k) const {
    mbar_expect_tx(&ctl->qfull[k & (TQ - 1)], 4);
    const int t = take(k);
    release(k);
    return t;
  }
  // read step k (any number of threads); ONE thread per consumer then calls release(k) once all of them have read
  __device__ __forceinline__ int take(int k) const {
    if (!dyn()) { const int t = group + k * num_groups; return t < num_tiles ? t : -1; }
    const int slot = k & (TQ - 1);
    mbar_wait(&ctl->qfull[slot], (k / TQ) & 1);
    return *reinterpret_cast<volatile int*>(&ctl->qtile[slot]);
  }
  __device__ __forceinline__ void release(int k) const {
    if (!dyn()) return;
    if (cg == 1) mbar_arrive(&ctl->qempty[k & (TQ - 1)]);
    else mbar_arrive_leader(&ctl->qempty[k & (TQ - 1)]);
  }
  __device__ __forceinline__ void init_barriers() const {
    for (int i = 0; i < TQ; ++i) { mbar_init(&ctl->qfull[i], 1); mbar_init(&ctl->qempty[i], consumers); }
  }
  // after the kernel's final barrier, one thread of the leader CTA
  __device__ __forceinline__ void finish() const {
    if (!dyn()) return;
    if (atomicAdd(counter + 1, 1) == num_groups - 1) { counter[0] = 0; counter[1] = 0; __threadfence(); }
  }
};

struct TileGeom {  // how a row tile maps onto the NHWC tensor: rows = (dn, dy, dx) boxes
  int bw, bh, bn;  // box extents (pixels)
  int tw, th, tn;  // tiles along w, h, n
  int H, W, Nimg;
};
__device__ __forceinline__ void tile_origin(const TileGeom& t, int tile, int& n0, int& h0, int& w0) {
  const int iw = tile % t.tw, ih = (tile / t.tw) % t.th, in = tile / (t.tw * t.th);
  w0 = iw * t.bw; h0 = ih * t.bh; n0 = in * t.bn;
}

struct FpropParams {
  TileGeom tg;
  int taps, S, pad, chunks;  // chunks = ceil(Cin / 64)
  int tiles_m, tiles_n, block_n, N;   // tiles_m: row tiles per CTA group step (pairs of row tiles when CG == 2)
  int stages;                          // smem ring depth
  int tma_epi;       // 1: bf16 outputs staged in swizzled smem and written with TMA stores
  int f32_epi;       // 1: fp32 outputs (+ fp32 residual) of a plain matrix, transposed through the staging slab (coalesced)
  long long M;       // rows of the output matrix (for the aux read guard)
  int* tile_counter; // dynamic tile scheduler (null: static partition)
  EpiParams e;
};

constexpr int EPI_WARP_BYTES = 32 * 128;  // per-warp staging: 32 rows x 64 bf16 columns, SWIZZLE_128B

// 32 consecutive columns of one output row
__device__ __forceinline__ void epi_row32(const EpiParams& e, long long row, int col0, int ncols, const uint32_t (&acc)[32]) {
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(acc[i]) * e.alpha;
  if (e.bias) {
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      if (i < ncols) {
        const float4 b = *reinterpret_cast<const float4*>(e.bias + col0 + i);
        v[i] += b.x; v[i + 1] += b.y; v[i + 2] += b.z; v[i + 3] += b.w;
      }
    }
  }
  const long long o = row * e.ldo + col0;
  if (e.mode == EPI_GELU) {
    if (e.out2) {
      bf16* p2 = reinterpret_cast<bf16*>(e.out2) + o;
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        if (i < ncols) {
          Vec8<bf16> t;
#pragma unroll
          for (int k = 0; k < 8; ++k) t.v[k] = v[i + k];
          t.store(p2 + i);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = gelu_f(v[i]);
  } else if (e.mode == EPI_DGELU) {
    const bf16* pa = reinterpret_cast<const bf16*>(e.aux) + o;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      if (i < ncols) {
        Vec8<bf16> t;
        t.load(pa + i);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[i + k] *= dgelu_f(t.v[k]);
      }
    }
  } else if (e.mode == EPI_RELU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
  } else if (e.mode == EPI_MUL) {
    const bf16* pa = reinterpret_cast<const bf16*>(e.aux) + o;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      if (i < ncols) {
        Vec8<bf16> t;
        t.load(pa + i);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[i + k] *= t.v[k];
      }
    }
  }
  if (e.res) {
    const long long r = row * e.ldr + col0;
    if (e.res_f32) {
      const float* pr = reinterpret_cast<const float*>(e.res) + r;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        if (i < ncols) {
          const float4 t = *reinterpret_cast<const float4*>(pr + i);
          v[i] += t.x; v[i + 1] += t.y; v[i + 2] += t.z; v[i + 3] += t.w;
        }
      }
    } else {
      const bf16* pr = reinterpret_cast<const bf16*>(e.res) + r;
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        if (i < ncols) {
          Vec8<bf16> t;
          t.load(pr + i);
#pragma unroll
          for (int k = 0; k < 8; ++k) v[i + k] += t.v[k];
        }
      }
    }
  }
  if (e.out_f32) {
    float* po = reinterpret_cast<float*>(e.out) + o;
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      if (i < ncols) {
        float4 t = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        if (e.accumulate) {
          const float4 q = *reinterpret_cast<const float4*>(po + i);
          t.x += q.x; t.y += q.y; t.z += q.z; t.w += q.w;
        }
        *reinterpret_cast<float4*>(po + i) = t;
      }
    }
  } else {
    bf16* po = reinterpret_cast<bf16*>(e.out) + o;
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
      if (i < ncols) {
        Vec8<bf16> t;
#pragma unroll
        for (int k = 0; k < 8; ++k) t.v[k] = v[i + k];
        t.store(po + i);
      }
    }
  }
}

__device__ __forceinline__ uint8_t* align_1024(uint8_t* p) {
  return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~(uintptr_t)1023);
}

// =========================================================================================== fprop
// CG = 1: one CTA per 128 x N tile.  CG = 2: a CTA pair (cluster of 2, one TPC) computes a 256 x N tile with
// tcgen05.mma.cta_group::2 -- each CTA stages its own 128 rows of A and HALF of the B tile, so the L2 -> smem operand
// traffic per FLOP drops by a third and the smaller stages give a deeper ring for the same shared memory.
template <int CG, int STATS = 0>
__global__ void __launch_bounds__(FPROP_THREADS, 1) umma_fprop_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                    const __grid_constant__ CUtensorMap tmB,
                                                                    const __grid_constant__ CUtensorMap tmO,
                                                                    const __grid_constant__ CUtensorMap tmO2, const FpropParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  const int bn_cta = p.block_n / CG;                // B rows staged by this CTA
  const int b_stage_bytes = bn_cta * 128;
  const int stage_bytes = A_STAGE_BYTES + b_stage_bytes;
  const int STG = p.stages;
  uint8_t* epi_smem = smem + STG * stage_bytes;  // EPI_WARPS x EPI_WARP_BYTES (1024 B aligned: stage sizes are multiples of 1 KB)
  SmemCtl* ctl = reinterpret_cast<SmemCtl*>(epi_smem + EPI_WARPS * EPI_WARP_BYTES);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int k_iters = p.taps * p.chunks;
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;
  const int group = blockIdx.x / CG, num_groups = gridDim.x / CG;
  // queue consumers that release a slot: the epilogue warps of every CTA of the group, the MMA warp, the peer's producer
  const TileSched sched{ctl, p.tile_counter, group, num_groups, num_tiles, rank, CG, CG * EPI_WARPS + 1 + (CG - 1)};

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.tma_epi) { tma_prefetch_desc(&tmO); if (p.e.out2) tma_prefetch_desc(&tmO2); }
    for (int i = 0; i < STG; ++i) { mbar_init(&ctl->full[i], 1); mbar_init(&ctl->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&ctl->tfull[i], 1); mbar_init(&ctl->tempty[i], EPI_WARPS * CG); }
    sched.init_barriers();
    fence_barrier_init();
  }
  if (warp == 1) { if (CG == 2) tmem_alloc_pair(&ctl->tmem_base, 512); else tmem_alloc(&ctl->tmem_base, 512); }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  pdl_wait();     // everything above overlapped the previous kernel's tail; global memory is touched only below
  pdl_trigger();
  const uint32_t tmem_base = ctl->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (+ tile scheduler in the leader CTA)
    if (elect_one()) {
      int stage = 0; uint32_t phase = 0;
      int next = (sched.dyn() && rank == 0) ? sched.fetch() : 0;
      for (int k = 0;; ++k) {
        int tile;
        if (sched.dyn() && rank == 0) {
          tile = next < num_tiles ? next : -1;
          sched.publish(k, tile);
          if (tile >= 0) next = sched.fetch();   // in flight while this tile's loads are issued
        } else if (sched.dyn()) {
          tile = sched.take_armed(k);
        } else {
          tile = sched.take(k);
        }
        if (tile < 0) break;
        const int tm = (tile / p.tiles_n) * CG + rank, tn = tile % p.tiles_n;
        int n0, h0, w0;
        tile_origin(p.tg, tm, n0, h0, w0);
        for (int tap = 0; tap < p.taps; ++tap) {
          const int dy = tap / p.S - p.pad, dx = tap % p.S - p.pad;
          for (int ch = 0; ch < p.chunks; ++ch) {
            mbar_wait(&ctl->empty[stage], phase ^ 1);
            uint8_t* sa = smem + stage * stage_bytes;
            if (CG == 2) {
              // the leader's barrier collects the bytes of both CTAs' loads
              if (rank == 0) mbar_expect_tx(&ctl->full[stage], 2 * stage_bytes);
              tma_load_4d_pair(&tmA, &ctl->full[stage], sa, ch * BLOCK_K, w0 + dx, h0 + dy, n0);
              tma_load_3d_pair(&tmB, &ctl->full[stage], sa + A_STAGE_BYTES, ch * BLOCK_K, tap, tn * p.block_n + rank * bn_cta);
            } else {
              mbar_expect_tx(&ctl->full[stage], stage_bytes);
              tma_load_4d(&tmA, &ctl->full[stage], sa, ch * BLOCK_K, w0 + dx, h0 + dy, n0);
              tma_load_3d(&tmB, &ctl->full[stage], sa + A_STAGE_BYTES, ch * BLOCK_K, tap, tn * p.block_n);
            }
            if (++stage == STG) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (CG == 2: the leader CTA only)
    const uint32_t idesc = make_idesc(128 * CG, p.block_n, 0, 0);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    for (int k = 0; rank == 0; ++k) {
      const int tile = sched.take(k);
      __syncwarp();
      if (lane == 0) sched.release(k);
      if (tile < 0) break;
      mbar_wait(&ctl->tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * 256;
      for (int ki = 0; ki < k_iters; ++ki) {
        mbar_wait(&ctl->full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * stage_bytes);
          const uint32_t sb = sa + A_STAGE_BYTES;
#pragma unroll
          for (int kk = 0; kk < BLOCK_K / 16; ++kk) {
            const uint64_t da = make_desc(sa + kk * 32, 16, 1024);
            const uint64_t db = make_desc(sb + kk * 32, 16, 1024);
            if (CG == 2) umma_bf16_pair(tmem_d, da, db, idesc, (ki | kk) != 0);
            else umma_bf16(tmem_d, da, db, idesc, (ki | kk) != 0);
          }
          if (CG == 2) {
            umma_commit_pair(&ctl->empty[stage]);
            if (ki == k_iters - 1) umma_commit_pair(&ctl->tfull[acc]);
          } else {
            umma_commit(&ctl->empty[stage]);
            if (ki == k_iters - 1) umma_commit(&ctl->tfull[acc]);
          }
        }
        __syncwarp();
        if (++stage == STG) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..5)
    const int q = warp & 3;  // TMEM lane quarter this warp may touch
    const int r = q * 32 + lane;
    int acc = 0; uint32_t acc_phase = 0;
    const TileGeom& t = p.tg;
    float st_acc[STATS ? 8 : 1] = {};   // [slab of this warp (2)][sum col 2l, sum col 2l+1, sumsq col 2l, sumsq col 2l+1]
    for (int k = 0;; ++k) {
      const int tile = sched.take(k);
      __syncwarp();
      if (lane == 0) sched.release(k);
      if (tile < 0) break;
      const int tm = (tile / p.tiles_n) * CG + rank, tn = tile % p.tiles_n;
      int n0, h0, w0;
      tile_origin(t, tm, n0, h0, w0);
      const int dx = r % t.bw, dy = (r / t.bw) % t.bh, dn = r / (t.bw * t.bh);
      const int n = n0 + dn, h = h0 + dy, w = w0 + dx;
      const bool valid = dn < t.bn && n < t.Nimg && h < t.H && w < t.W;
      const long long row = ((long long)n * t.H + h) * t.W + w;
      mbar_wait(&ctl->tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * 256 + ((uint32_t)(q * 32) << 16);
      const int half = (warp - 2) >> 2;  // which alternate column slabs this warp owns
      bool released = false;
      if (p.tma_epi) {
        // bf16 outputs: each warp stages 32 rows x 64 columns in its private 128B-swizzled smem slab and writes it with
        // its own TMA store (which clips the M / N tails) -- no cross-warp synchronisation in the epilogue
        const int npass = p.e.out2 ? 2 : 1;
        const int nslabs = (p.block_n + 63) >> 6;
        uint8_t* buf = epi_smem + (warp - 2) * EPI_WARP_BYTES;
        for (int sl = half; sl < nslabs; sl += 2) {
          float rowdot = 0.f;
          for (int pass = 0; pass < npass; ++pass) {
            if (lane == 0) tma_store_wait_read<0>();  // the previous store is done reading the staging slab
            __syncwarp();
            const bool pre = p.e.out2 && pass == 0;  // pre-activation copy
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
              uint32_t v[32];
              tmem_ld32(taddr + sl * 64 + hf * 32, v);
              tmem_ld_wait();
              if (sl + 2 >= nslabs && pass == npass - 1 && hf == 1) {
                // accumulator fully read by this warp: hand the TMEM stage back to the MMA warp before the stores
                tc_fence_before();
                __syncwarp();
                if (lane == 0) { if (CG == 2) mbar_arrive_leader(&ctl->tempty[acc]); else mbar_arrive(&ctl->tempty[acc]); }
                released = true;
              }
              const int col0 = tn * p.block_n + sl * 64 + hf * 32;
              const int ncols = p.N - col0;
              float f[32];
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]) * p.e.alpha;
              if (p.e.bias) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                  if (i < ncols) {
                    const float4 bb = *reinterpret_cast<const float4*>(p.e.bias + col0 + i);
                    f[i] += bb.x; f[i + 1] += bb.y; f[i + 2] += bb.z; f[i + 3] += bb.w;
                  }
                }
              }
              if (!pre) {
                if (p.e.mode == EPI_GELU) {
#pragma unroll
                  for (int i = 0; i < 32; ++i) f[i] = gelu_fast(f[i]);
                } else if (p.e.mode == EPI_RELU) {
#pragma unroll
                  for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.f);
                } else if (p.e.mode == EPI_DGELU) {
                  if (valid) {
                    const bf16* pa = reinterpret_cast<const bf16*>(p.e.aux) + row * p.e.ldo + col0;
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
                      if (i < ncols) {
                        Vec8<bf16> t8;
                        t8.load(pa + i);
#pragma unroll
                        for (int k = 0; k < 8; ++k) f[i + k] *= dgelu_fast(t8.v[k]);
                      }
                    }
                  }
                } else if (p.e.mode == EPI_MUL) {
                  if (valid) {
                    const bf16* pa = reinterpret_cast<const bf16*>(p.e.aux) + row * p.e.ldo + col0;
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
                      if (i < ncols) {
                        Vec8<bf16> t8;
                        t8.load(pa + i);
#pragma unroll
                        for (int k = 0; k < 8; ++k) f[i + k] *= t8.v[k];
                      }
                    }
                  }
                }
              }
              if (p.e.dot_aux && valid) {
                const bf16* pa = reinterpret_cast<const bf16*>(p.e.dot_aux) + row * p.e.ldo + col0;
#pragma unroll
                for (int i = 0; i < 32; i += 8) {
                  if (i < ncols) {
                    Vec8<bf16> t8;
                    t8.load(pa + i);
#pragma unroll
                    for (int k = 0; k < 8; ++k) rowdot = fmaf(f[i + k], t8.v[k], rowdot);
                  }
                }
              }
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint4 pk;
                __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                for (int k = 0; k < 4; ++k) h2[k] = __floats2bfloat162_rn(f[j * 8 + 2 * k], f[j * 8 + 2 * k + 1]);
                const int chunk = hf * 4 + j;
                *reinterpret_cast<uint4*>(buf + lane * 128 + ((chunk ^ (lane & 7)) << 4)) = pk;
              }
            }
            fence_proxy_async();
            __syncwarp();
            if (p.e.dot_aux && valid) {
              const int g = (tn * p.block_n + sl * 64) >> 6;   // 64-column group = attention head
              if (g < p.e.dot_groups)
                p.e.dot_out[(row / p.e.dot_rows) * (long long)p.e.dot_groups * p.e.dot_rows + (long long)g * p.e.dot_rows + row % p.e.dot_rows] = rowdot;
            }
            if (lane == 0) {
              const int cc = tn * p.block_n + sl * 64;
              if (cc < p.N) {
                const int r0 = q * 32;  // origin of this warp's 32-row sub-box inside the tile
                tma_store_4d(pre ? &tmO2 : &tmO, buf, cc, w0 + r0 % t.bw, h0 + (r0 / t.bw) % t.bh, n0 + r0 / (t.bw * t.bh));
                tma_store_commit();
              }
            }
            if (STATS) {
              // column sums of the staged bf16 slab (32 rows x 64 columns): lane l owns columns 2l, 2l + 1 -- conflict-free 4-byte
              // reads of the swizzled rows; rows outside the tensor are skipped
              const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
              float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
              for (int rr = 0; rr < 32; ++rr) {
                if ((vmask >> rr) & 1) {
                  const uint32_t w = *reinterpret_cast<const uint32_t*>(buf + rr * 128 + (((lane >> 2) ^ (rr & 7)) << 4) + (lane & 3) * 4);
                  const float a = __uint_as_float(w << 16), b = __uint_as_float(w & 0xffff0000u);
                  s0 += a; s1 += b;
                  q0 = fmaf(a, a, q0); q1 = fmaf(b, b, q1);
                }
              }
              if (((sl - half) >> 1) == 0) { st_acc[0] += s0; st_acc[1] += s1; st_acc[2] += q0; st_acc[3] += q1; }
              else { st_acc[4] += s0; st_acc[5] += s1; st_acc[6] += q0; st_acc[7] += q1; }
            }
          }
        }
      } else if (p.f32_epi) {
        // fp32 output (+ fp32 residual, the token residual stream): a thread owns a ROW of the accumulator, but rows are
        // 3 KB apart in memory, so the 32 x 32 chunk goes through the warp's swizzled staging slab and is added to the
        // residual / written with each warp instruction covering 4 rows x 128 contiguous bytes.  The residual loads
        // are issued before the accumulator is read, so their latency hides behind tcgen05.ld.
        const EpiParams& e = p.e;
        float* outp = reinterpret_cast<float*>(e.out);
        const float* resp = reinterpret_cast<const float*>(e.res);
        uint8_t* buf = epi_smem + (warp - 2) * EPI_WARP_BYTES;
        const long long row_base = (long long)tm * 128 + q * 32;
        const int sub = lane >> 3, c4 = lane & 7;
        for (int c0 = half * 32; c0 < p.block_n; c0 += 64) {
          const int col0 = tn * p.block_n + c0;
          const int col = col0 + c4 * 4;
          float4 rv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const long long grow = row_base + i * 4 + sub;
            rv[i] = (resp && grow < p.M && col < p.N) ? *reinterpret_cast<const float4*>(resp + grow * e.ldr + col)
                                                      : make_float4(0.f, 0.f, 0.f, 0.f);
          }
          uint32_t v[32];
          tmem_ld32(taddr + c0, v);
          tmem_ld_wait();
          if (c0 + 64 >= p.block_n) {   // last chunk of this warp: hand the TMEM stage back before the stores
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG == 2) mbar_arrive_leader(&ctl->tempty[acc]); else mbar_arrive(&ctl->tempty[acc]); }
            released = true;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 f = make_float4(__uint_as_float(v[4 * j]) * e.alpha, __uint_as_float(v[4 * j + 1]) * e.alpha,
                                   __uint_as_float(v[4 * j + 2]) * e.alpha, __uint_as_float(v[4 * j + 3]) * e.alpha);
            if (e.bias && col0 + 4 * j < p.N) {
              const float4 bb = *reinterpret_cast<const float4*>(e.bias + col0 + 4 * j);
              f.x += bb.x; f.y += bb.y; f.z += bb.z; f.w += bb.w;
            }
            if (e.mode == EPI_RELU) { f.x = fmaxf(f.x, 0.f); f.y = fmaxf(f.y, 0.f); f.z = fmaxf(f.z, 0.f); f.w = fmaxf(f.w, 0.f); }
            *reinterpret_cast<float4*>(buf + lane * 128 + ((j ^ (lane & 7)) << 4)) = f;
          }
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = i * 4 + sub;
            const long long grow = row_base + rr;
            float4 a = *reinterpret_cast<const float4*>(buf + rr * 128 + ((c4 ^ (rr & 7)) << 4));
            a.x += rv[i].x; a.y += rv[i].y; a.z += rv[i].z; a.w += rv[i].w;
            if (grow < p.M && col < p.N) *reinterpret_cast<float4*>(outp + grow * e.ldo + col) = a;
          }
          __syncwarp();
        }
      } else if (half == 0) {
        for (int c0 = 0; c0 < p.block_n; c0 += 32) {
          uint32_t v[32];
          tmem_ld32(taddr + c0, v);
          tmem_ld_wait();
          const int col0 = tn * p.block_n + c0;
          const int ncols = min(32, p.N - col0);
          if (valid && ncols > 0) epi_row32(p.e, row, col0, ncols, v);
        }
      }
      if (!released) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { if (CG == 2) mbar_arrive_leader(&ctl->tempty[acc]); else mbar_arrive(&ctl->tempty[acc]); }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (p.tma_epi && lane == 0) tma_store_wait_all();
    if (STATS) {
      // park this warp's sums in its (now idle) staging slab: [8 values][32 lanes]; combined across the warps below
      __syncwarp();
      float* mine = reinterpret_cast<float*>(epi_smem + (warp - 2) * EPI_WARP_BYTES);
#pragma unroll
      for (int k = 0; k < 8; ++k) mine[k * 32 + lane] = st_acc[k];
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();   // the peer may still signal this CTA's barriers until here
  if (warp == 1) { if (CG == 2) tmem_dealloc_pair(tmem_base, 512); else tmem_dealloc(tmem_base, 512); }
  if (STATS) {
    // one row [2 N] per CTA: column c lives in slab c / 64, owned by the four warps (row quarters 0..3) of half (c / 64) & 1;
    // fixed summation order => deterministic
    float* row = p.e.stat_part + (size_t)blockIdx.x * 2 * p.N;
    for (int i = threadIdx.x; i < 2 * p.N; i += blockDim.x) {
      const int sq = i >= p.N, c = sq ? i - p.N : i;
      const int sl = c >> 6, hf = sl & 1, j = sl >> 1, ln = (c & 63) >> 1, k = j * 4 + sq * 2 + (c & 1);
      float tsum = 0.f;
#pragma unroll
      for (int qq = 0; qq < 4; ++qq) tsum += reinterpret_cast<const float*>(epi_smem + (hf * 4 + qq) * EPI_WARP_BYTES)[k * 32 + ln];
      row[i] = tsum;
    }
  }
  if (threadIdx.x == 0 && rank == 0) sched.finish();
}

// =========================================================================================== wgrad
struct WgradParams {
  TileGeom tg;      // 64-pixel chunks
  int taps, S, pad;
  int Cout, Cin, K;  // K = taps * Cin (row length of dW)
  int tiles_co, tiles_ci, block_n, splits;   // tiles_co: 128*CG-row output tiles
  int stages;        // smem ring depth
  int chunks_total;  // pixel chunks
  int chunks_per_split;
  float* part;       // split-K partials: [split][tap][tci][tco][block_n columns (ci)][128 rows (co)] fp32
  float* dW;         // direct mode (splits == 1, linear layer): the epilogue writes dW [Cout][Cin] itself
  int direct;
  int* tile_counter; // dynamic work scheduler (null: static partition)
};

// CG = 2: a CTA pair owns a 256 (co) x block_n (ci) output tile (cta_group::2): each CTA stages its own 128 dY channels
// and HALF of the X channels of every pixel chunk -- a third less L2 -> smem traffic per FLOP and a deeper ring.
template <int CG>
__global__ void __launch_bounds__(NUM_THREADS, 1) umma_wgrad_kernel(const __grid_constant__ CUtensorMap tmY,
                                                                    const __grid_constant__ CUtensorMap tmX, const WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_1024(smem_raw);
  const int a_bytes = 2 * BLOCK_K * 128;                   // two 64-channel groups of dY, 64 pixels each
  const int ngroups = p.block_n / 64 / CG;                 // 64-channel groups of X staged by this CTA
  const int stage_bytes = a_bytes + ngroups * BLOCK_K * 128;
  const int STG = p.stages;
  SmemCtl* ctl = reinterpret_cast<SmemCtl*>(smem + STG * stage_bytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int out_tiles = p.tiles_co * p.tiles_ci * p.taps;
  const int num_work = out_tiles * p.splits;
  const int chunks_per_split = p.chunks_per_split;
  const int rank = CG == 2 ? (int)cluster_ctarank() : 0;
  const int group = blockIdx.x / CG, num_groups = gridDim.x / CG;
  const TileSched sched{ctl, p.tile_counter, group, num_groups, num_work, rank, CG, CG * 4 + 1 + (CG - 1)};

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmY);
    tma_prefetch_desc(&tmX);
    for (int i = 0; i < STG; ++i) { mbar_init(&ctl->full[i], 1); mbar_init(&ctl->empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&ctl->tfull[i], 1); mbar_init(&ctl->tempty[i], 4 * CG); }
    sched.init_barriers();
    fence_barrier_init();
  }
  if (warp == 1) { if (CG == 2) tmem_alloc_pair(&ctl->tmem_base, 512); else tmem_alloc(&ctl->tmem_base, 512); }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  pdl_wait();     // everything above overlapped the previous kernel's tail; global memory is touched only below
  pdl_trigger();
  const uint32_t tmem_base = ctl->tmem_base;

  // work item -> (split, tap, ci tile, co tile); consecutive CTAs share the pixel range (L2 reuse of dY / X)
  auto decode = [&](int wi, int& split, int& tap, int& tci, int& tco) {
    tco = wi % p.tiles_co; wi /= p.tiles_co;
    tci = wi % p.tiles_ci; wi /= p.tiles_ci;
    tap = wi % p.taps; split = wi / p.taps;
  };

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0; uint32_t phase = 0;
      int next = (sched.dyn() && rank == 0) ? sched.fetch() : 0;
      for (int k = 0;; ++k) {
        int wi;
        if (sched.dyn() && rank == 0) {
          wi = next < num_work ? next : -1;
          sched.publish(k, wi);
          if (wi >= 0) next = sched.fetch();
        } else if (sched.dyn()) {
          wi = sched.take_armed(k);
        } else {
          wi = sched.take(k);
        }
        if (wi < 0) break;
        int split, tap, tci, tco;
        decode(wi, split, tap, tci, tco);
        const int co0 = (tco * CG + rank) * 128;
        const int ci0 = tci * p.block_n + rank * (p.block_n / CG);
        const int dy = tap / p.S - p.pad, dx = tap % p.S - p.pad;
        const int c_begin = split * chunks_per_split, c_end = min(p.chunks_total, c_begin + chunks_per_split);
        for (int c = c_begin; c < c_end; ++c) {
          int n0, h0, w0;
          tile_origin(p.tg, c, n0, h0, w0);
          mbar_wait(&ctl->empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * stage_bytes;
          if (CG == 2) {
            if (rank == 0) mbar_expect_tx(&ctl->full[stage], 2 * stage_bytes);   // the leader's barrier collects both CTAs' bytes
            for (int g = 0; g < 2; ++g) tma_load_4d_pair(&tmY, &ctl->full[stage], sa + g * BLOCK_K * 128, co0 + g * 64, w0, h0, n0);
            for (int g = 0; g < ngroups; ++g)
              tma_load_4d_pair(&tmX, &ctl->full[stage], sa + a_bytes + g * BLOCK_K * 128, ci0 + g * 64, w0 + dx, h0 + dy, n0);
          } else {
            mbar_expect_tx(&ctl->full[stage], stage_bytes);
            for (int g = 0; g < 2; ++g) tma_load_4d(&tmY, &ctl->full[stage], sa + g * BLOCK_K * 128, co0 + g * 64, w0, h0, n0);
            for (int g = 0; g < ngroups; ++g)
              tma_load_4d(&tmX, &ctl->full[stage], sa + a_bytes + g * BLOCK_K * 128, ci0 + g * 64, w0 + dx, h0 + dy, n0);
          }
          if (++stage == STG) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = make_idesc(128 * CG, p.block_n, 1, 1);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    for (int k = 0; rank == 0; ++k) {
      const int wi = sched.take(k);
      __syncwarp();
      if (lane == 0) sched.release(k);
      if (wi < 0) break;
      int split, tap, tci, tco;
      decode(wi, split, tap, tci, tco);
      const int c_begin = split * chunks_per_split, c_end = min(p.chunks_total, c_begin + chunks_per_split);
      if (c_begin >= c_end) continue;
      mbar_wait(&ctl->tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * 256;
      for (int c = c_begin; c < c_end; ++c) {
        mbar_wait(&ctl->full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * stage_bytes);
          const uint32_t sb = sa + a_bytes;
#pragma unroll
          for (int kk = 0; kk < BLOCK_K / 16; ++kk) {
            // MN-major: 16 pixels = two 8-row atoms (SBO = 1024 B); 64-channel groups are LBO = 64 px * 128 B apart
            const uint64_t da = make_desc(sa + kk * 2048, BLOCK_K * 128, 1024);
            const uint64_t db = make_desc(sb + kk * 2048, BLOCK_K * 128, 1024);
            if (CG == 2) umma_bf16_pair(tmem_d, da, db, idesc, (c != c_begin) || kk != 0);
            else umma_bf16(tmem_d, da, db, idesc, (c != c_begin) || kk != 0);
          }
          if (CG == 2) {
            umma_commit_pair(&ctl->empty[stage]);
            if (c == c_end - 1) umma_commit_pair(&ctl->tfull[acc]);
          } else {
            umma_commit(&ctl->empty[stage]);
            if (c == c_end - 1) umma_commit(&ctl->tfull[acc]);
          }
        }
        __syncwarp();
        if (++stage == STG) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    const int q = warp & 3;
    int acc = 0; uint32_t acc_phase = 0;
    for (int k = 0;; ++k) {
      const int wi = sched.take(k);
      __syncwarp();
      if (lane == 0) sched.release(k);
      if (wi < 0) break;
      int split, tap, tci, tco;
      decode(wi, split, tap, tci, tco);
      const int c_begin = split * chunks_per_split, c_end = min(p.chunks_total, c_begin + chunks_per_split);
      if (c_begin >= c_end) continue;
      // partial tile of this CTA's 128 rows, column-major (co contiguous): each store instruction of a warp writes 128
      // contiguous bytes.  Scratch layout [split][tap][tci][128-row tile][column][128]: the reduce kernels see 128-row tiles.
      const size_t ptile = (((size_t)split * p.taps + tap) * p.tiles_ci + tci) * (p.tiles_co * CG) + tco * CG + rank;
      float* part = p.part + ptile * 128 * p.block_n + q * 32 + lane;
      mbar_wait(&ctl->tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * 256 + ((uint32_t)(q * 32) << 16);
      const int co = (tco * CG + rank) * 128 + q * 32 + lane;
      float* drow = p.dW + (size_t)co * p.Cin + tci * p.block_n;   // direct mode: this thread's row of dW
      for (int c0 = 0; c0 < p.block_n; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(taddr + c0, v);
        tmem_ld_wait();
        if (p.direct) {
          // 128 contiguous bytes per thread: whole lines, no partials, no reduce launch
          if (co < p.Cout) {
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              if (tci * p.block_n + c0 + i < p.Cin)
                *reinterpret_cast<float4*>(drow + c0 + i) = make_float4(__uint_as_float(v[i]), __uint_as_float(v[i + 1]),
                                                                         __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
          }
        } else if (tci * p.block_n + c0 < p.Cin) {
#pragma unroll
          for (int i = 0; i < 32; ++i) part[(size_t)(c0 + i) * 128] = __uint_as_float(v[i]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (CG == 2) mbar_arrive_leader(&ctl->tempty[acc]); else mbar_arrive(&ctl->tempty[acc]); }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) { if (CG == 2) tmem_dealloc_pair(tmem_base, 512); else tmem_dealloc(tmem_base, 512); }
  if (threadIdx.x == 0 && rank == 0) sched.finish();
}

// Deterministic split-K reduction + layout change: dW[co*so + ci*si + tap*st] = sum_split part[split][tap][tci][tco][ci][co]
struct WgradReduceParams {
  const float* part;
  float* dW;
  int Cout, Cin, taps, tiles_co, tiles_ci, block_n, splits;
  long long so, si, st;
  int groups;     // linear form: the split range is cut into `groups` parts (gridDim.y), combined in a second stage (see the kernel)
  int* tickets;   // one arrival counter per blockIdx.x (zero between launches)
};
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const WgradReduceParams p) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  __shared__ float tile[32][33];
  const int tx = threadIdx.x, ty = threadIdx.y;   // 32 x 8
  const int co0 = blockIdx.x * 32, ci0 = blockIdx.y * 32, tap = blockIdx.z;
  const int tco = co0 >> 7, tci = ci0 / p.block_n;
  const size_t tile_elems = (size_t)128 * p.block_n;
  const size_t split_stride = (size_t)p.taps * p.tiles_ci * p.tiles_co * tile_elems;
  const float* base = p.part + (((size_t)tap * p.tiles_ci + tci) * p.tiles_co + tco) * tile_elems + (co0 & 127) + tx;
  // the four columns of a thread advance together through the splits: 4 (x 2 splits) independent loads per trip
  float acc4[4] = {0.f, 0.f, 0.f, 0.f};
  const float* src4[4];
  bool ok[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int cil = ty + j * 8;                       // ci within this 32-block
    const int col = ci0 - tci * p.block_n + cil;      // column inside the tile
    ok[j] = ci0 + cil < p.Cin;
    src4[j] = base + (size_t)(ok[j] ? col : 0) * 128;
  }
  int s = 0;
  for (; s + 2 <= p.splits; s += 2) {
    float v0[4], v1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { v0[j] = src4[j][(size_t)s * split_stride]; v1[j] = src4[j][(size_t)(s + 1) * split_stride]; }
#pragma unroll
    for (int j = 0; j < 4; ++j) { acc4[j] += v0[j]; acc4[j] += v1[j]; }
  }
  if (s < p.splits) {
#pragma unroll
    for (int j = 0; j < 4; ++j) acc4[j] += src4[j][(size_t)s * split_stride];
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) tile[ty + j * 8][tx] = ok[j] ? acc4[j] : 0.f;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int col = ty + j * 8;                       // co within the 32-block
    const int co = co0 + col, ci = ci0 + tx;
    if (co < p.Cout && ci < p.Cin) p.dW[co * p.so + ci * p.si + tap * p.st] = tile[tx][col];
  }
}

// Linear-layer form of the reduction (taps == 1, dW row-major [Cout][Cin]): one block per 128 (co) x 32 (ci) panel of an
// output tile, 16 independent 128 B-coalesced loads per thread and split, transposed through a conflict-free smem
// panel so that every warp store is a full 128 B row segment of dW.
__global__ void __launch_bounds__(256) wgrad_reduce_linear_kernel(const WgradReduceParams p) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  __shared__ float tile[32][129];
  __shared__ int is_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int subs = p.block_n >> 5;
  const int t = blockIdx.x / subs, sub = blockIdx.x % subs;
  const int tco = t % p.tiles_co, tci = t / p.tiles_co;
  const size_t tile_elems = (size_t)128 * p.block_n;
  const size_t split_stride = (size_t)p.tiles_ci * p.tiles_co * tile_elems;
  float* base = const_cast<float*>(p.part) + ((size_t)tci * p.tiles_co + tco) * tile_elems + (size_t)sub * 32 * 128;
  // A layer with few output tiles is split many ways (the 56 x 56 out_conv: 74 splits of ONE 256 x 256 tile), and 16 blocks
  // walking 74 splits each are bound by memory latency (104 us for 19 MB).  The split range is therefore cut into gridDim.y
  // groups: every block sums its group (in split order) and leaves the sum in the slot of the group's first split; the block
  // that arrives last (ticket counter, integer) adds the group sums in group order.  Fixed partition and order: deterministic.
  const int G = p.groups, per = (p.splits + G - 1) / G;
  const int s_begin = blockIdx.y * per, s_end = min(p.splits, s_begin + per);
  float acc[16];
#pragma unroll
  for (int it = 0; it < 16; ++it) acc[it] = 0.f;
  int s = s_begin;
  for (; s + 2 <= s_end; s += 2) {   // two splits per trip: 32 independent loads in flight per thread
    float v[2][16];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const float* src = base + (size_t)(s + u) * split_stride;
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int item = it * 8 + warp;            // 128 items = 32 columns x 4 groups of 32 rows
        v[u][it] = src[(item >> 2) * 128 + (item & 3) * 32 + lane];
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int it = 0; it < 16; ++it) acc[it] += v[u][it];
  }
  if (s < s_end) {
    const float* src = base + (size_t)s * split_stride;
#pragma unroll
    for (int it = 0; it < 16; ++it) {
      const int item = it * 8 + warp;
      acc[it] += src[(item >> 2) * 128 + (item & 3) * 32 + lane];
    }
  }
  if (G > 1) {
    if (s_begin < p.splits) {
      float* dst = base + (size_t)s_begin * split_stride;   // only this block reads or writes the slots of its group
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int item = it * 8 + warp;
        dst[(item >> 2) * 128 + (item & 3) * 32 + lane] = acc[it];
      }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const int ticket = atomicAdd(p.tickets + blockIdx.x, 1);
      is_last = ticket == G - 1;
      if (is_last) p.tickets[blockIdx.x] = 0;   // ready for the next launch
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
#pragma unroll
    for (int it = 0; it < 16; ++it) acc[it] = 0.f;
    for (int g = 0; g * per < p.splits; ++g) {
      const float* src = base + (size_t)(g * per) * split_stride;
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int item = it * 8 + warp;
        acc[it] += __ldcg(src + (item >> 2) * 128 + (item & 3) * 32 + lane);
      }
    }
  }
#pragma unroll
  for (int it = 0; it < 16; ++it) {
    const int item = it * 8 + warp;
    tile[item >> 2][(item & 3) * 32 + lane] = acc[it];
  }
  __syncthreads();
  const int ci = tci * p.block_n + sub * 32 + lane;
  if (ci >= p.Cin) return;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int col = warp * 16 + k;
    const int co = tco * 128 + col;
    if (co < p.Cout) p.dW[(long long)co * p.so + ci] = tile[lane][col];
  }
}

// ------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
int g_num_sms = 148;
// dynamic tile scheduler: a ring of {next tile, finished groups} counter pairs, one per launch; each kernel leaves its pair
// zeroed (TileSched::finish), so the ring never needs clearing again.  OFF by default (S4G_DYN_TILES=1 turns it on): measured
// on B200 it costs 0.3 ms per step alone (one atomic round trip before the first load of each of ~230 launches) and, against
// expectation, 2 ms per step next to NCCL at N = 2 -- kept as an experiment knob with its parity tests.
constexpr int TILE_COUNTER_RING = 256;
constexpr int MAX_DEV = 64;
int* g_tile_counters[MAX_DEV] = {};
constexpr int REDUCE_TICKETS = 1024;   // arrival counters of the two-level split-K reduce (wgrad_reduce_linear_kernel), zero at rest
int* g_reduce_tickets[MAX_DEV] = {};
unsigned g_tile_counter_next = 0;
int g_dyn_tiles = 0;
int* next_tile_counter() {
  if (!g_dyn_tiles) return nullptr;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV || !g_tile_counters[dev]) return nullptr;
  return g_tile_counters[dev] + 2 * (g_tile_counter_next++ % TILE_COUNTER_RING);
}
int g_pdl = 1;  // programmatic dependent launch of the tcgen05 kernels (S4G_PDL=0 turns it off)

int make_map(CUtensorMap* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
             const uint32_t* box) {
  cuuint64_t gd[4]; cuuint64_t gs[3]; cuuint32_t bx[4]; cuuint32_t es[4] = {1, 1, 1, 1};
  for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) gs[i] = strides_bytes[i];
  if (reinterpret_cast<uintptr_t>(base) & 15) { s4g_set_error("tcgen05 path: operand base must be 16 B aligned"); return S4G_ERR_ALIGN; }
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    char buf[256];
    snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed (%d): rank %d dims %llu %llu %llu %llu box %u %u %u %u", (int)r, rank,
             (unsigned long long)dims[0], (unsigned long long)dims[1], rank > 2 ? (unsigned long long)dims[2] : 0ull,
             rank > 3 ? (unsigned long long)dims[3] : 0ull, box[0], box[1], rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    s4g_set_error(buf);
    return S4G_ERR_CUDA;
  }
  return 0;
}

int pow2_divisor(int v, int cap) {
  int p = 1;
  while (p * 2 <= cap && v % (p * 2) == 0) p *= 2;
  return p;
}
// rows-per-tile decomposition: bw | W and bh | H powers of two, bn fills the rest (bw*bh*bn == rows exactly: the
// TMA transaction byte count of a stage depends on it)
TileGeom make_tiles(int Nimg, int H, int W, int rows) {
  TileGeom t;
  t.H = H; t.W = W; t.Nimg = Nimg;
  t.bw = pow2_divisor(W, rows);
  t.bh = pow2_divisor(H, rows / t.bw);
  t.bn = rows / (t.bw * t.bh);
  t.tw = (W + t.bw - 1) / t.bw; t.th = (H + t.bh - 1) / t.bh; t.tn = (Nimg + t.bn - 1) / t.bn;
  return t;
}
// plain [M, C] matrix: rows run along "w"
TileGeom matrix_tiles(long long M, int rows) {
  TileGeom t;
  t.H = 1; t.W = (int)M; t.Nimg = 1;
  t.bw = rows; t.bh = 1; t.bn = 1;
  t.tw = (int)((M + rows - 1) / rows); t.th = 1; t.tn = 1;
  return t;
}
// activation tensor map: NHWC [Nimg,H,W,C] (row stride ld) -> dims {C, W, H, N}; plain matrix [M, C] -> {C, M, 1, 1}
// rows_div = 4: the box covers a quarter (32 consecutive rows) of the tile -- the per-warp epilogue store box
int make_act_map(CUtensorMap* map, const bf16* base, const ConvGeom& g, int C, int ld, const TileGeom& t, bool matrix, int H, int W,
                 int rows_div = 1) {
  uint64_t dims[4], strides[3]; uint32_t box[4];
  int bw = t.bw, bh = t.bh, bn = t.bn;
  if (rows_div > 1) {
    const int rows = bw * bh * bn / rows_div;
    bw = std::min(bw, rows);
    bh = std::min(bh, rows / bw);
    bn = rows / (bw * bh);
  }
  if (matrix) {
    dims[0] = C; dims[1] = (uint64_t)g.Nimg * H * W; dims[2] = 1; dims[3] = 1;
    strides[0] = (uint64_t)ld * 2; strides[1] = strides[0] * dims[1]; strides[2] = strides[1];
    box[0] = 64; box[1] = bw; box[2] = 1; box[3] = 1;
  } else {
    dims[0] = C; dims[1] = W; dims[2] = H; dims[3] = g.Nimg;
    strides[0] = (uint64_t)ld * 2; strides[1] = strides[0] * W; strides[2] = strides[1] * H;
    box[0] = 64; box[1] = bw; box[2] = bh; box[3] = bn;
  }
  return make_map(map, base, 4, dims, strides, box);
}
// Output-tile width.  Measured (profiles/r01_d): the token GEMMs are bound by the L2 -> smem operand fill (about 12 TB/s
// chip-wide), not by tcgen05 issue, so the cost of a persistent launch is waves x bytes staged per K step (16 KB of A
// rows plus this CTA's share of the B tile).  With CTA pairs N = 768 runs fastest as 4 x 192 columns (2.7 waves of
// 28 KB steps instead of 2.03 -> 3 waves of 32 KB), N = 2304 / 3072 as 256-column tiles.  S4G_FPROP_BN overrides.
int pick_block_n(int N, int row_tiles, int cg) {
  if (N <= 256) return (N + 15) / 16 * 16;
  static const int forced = getenv("S4G_FPROP_BN") ? atoi(getenv("S4G_FPROP_BN")) : 0;
  if (forced) return forced;
  const int groups = g_num_sms / cg;
  int best = 256;
  double best_cost = 1e30;
  for (int bn = 256; bn >= 128; bn -= 64) {
    const long long work = (long long)cdiv(row_tiles, cg) * cdiv(N, bn);
    const double cost = (double)((work + groups - 1) / groups) * (A_STAGE_BYTES + bn * 128.0 / cg);
    if (cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

}  // namespace

// bf16 tiled tensor map, SWIZZLE_128B (for the other tcgen05 kernels of the library)
int umma_make_map_bf16(void* map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box) {
  S4G_TRY(umma_init());
  return make_map(reinterpret_cast<CUtensorMap*>(map), base, rank, dims, strides_bytes, box);
}
int umma_num_sms() { return umma_init() == 0 ? g_num_sms : 148; }

int umma_init() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  S4G_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess) { s4g_set_error("cuTensorMapEncodeTiled is not available in this driver"); return S4G_ERR_CUDA; }
  g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  int dev = 0;
  S4G_CUDA(cudaGetDevice(&dev));
  S4G_CUDA(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  if (const char* e = getenv("S4G_PDL")) g_pdl = atoi(e) != 0;
  if (const char* e = getenv("S4G_DYN_TILES")) g_dyn_tiles = atoi(e) != 0;
  if (dev >= 0 && dev < MAX_DEV && !g_tile_counters[dev]) {
    S4G_CUDA(cudaMalloc(&g_tile_counters[dev], sizeof(int) * 2 * TILE_COUNTER_RING));
    S4G_CUDA(cudaMemset(g_tile_counters[dev], 0, sizeof(int) * 2 * TILE_COUNTER_RING));
  }
  if (dev >= 0 && dev < MAX_DEV && !g_reduce_tickets[dev]) {
    S4G_CUDA(cudaMalloc(&g_reduce_tickets[dev], sizeof(int) * REDUCE_TICKETS));
    S4G_CUDA(cudaMemset(g_reduce_tickets[dev], 0, sizeof(int) * REDUCE_TICKETS));
  }
  // S4G_SM_LIMIT: persistent grids leave SMs free for a concurrently running collective (experiment knob)
  if (const char* lim = getenv("S4G_SM_LIMIT")) { const int v = atoi(lim); if (v >= 2 && v < g_num_sms) g_num_sms = v & ~1; }
  S4G_CUDA(cudaFuncSetAttribute(umma_fprop_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  S4G_CUDA(cudaFuncSetAttribute(umma_fprop_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  S4G_CUDA(cudaFuncSetAttribute(umma_fprop_kernel<1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  S4G_CUDA(cudaFuncSetAttribute(umma_fprop_kernel<2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  S4G_CUDA(cudaFuncSetAttribute(umma_wgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  S4G_CUDA(cudaFuncSetAttribute(umma_wgrad_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  return 0;
}

// number of partial rows umma_conv_fprop writes into EpiParams::stat_part for this geometry (= its grid size)
int umma_fprop_stat_rows(const ConvGeom& g) {
  if (umma_init() != 0) return 0;
  const bool matrix = g.Hi == 1 && g.Wi == 1;
  const TileGeom tg = matrix ? matrix_tiles(g.Nimg, 128) : make_tiles(g.Nimg, g.Hi, g.Wi, 128);
  const int row_tiles = tg.tw * tg.th * tg.tn;
  static const int forced_cg = getenv("S4G_FPROP_CG") ? atoi(getenv("S4G_FPROP_CG")) : 0;
  int cg = forced_cg ? forced_cg : (row_tiles >= 2 ? 2 : 1);
  const int block_n = pick_block_n(g.Cout, row_tiles, cg);
  if (block_n % 32) cg = 1;
  const int tiles = cdiv(row_tiles, cg) * cdiv(g.Cout, block_n);
  return cg == 2 ? 2 * std::min(tiles, g_num_sms / 2) : std::min(tiles, g_num_sms);
}

int umma_conv_fprop(const ConvGeom& g, const bf16* X, const bf16* W, const EpiParams& e, cudaStream_t st) {
  S4G_TRY(umma_init());
  if (g.stride != 1 || g.Ho != g.Hi || g.Wo != g.Wi) { s4g_set_error("umma_conv_fprop: stride-1 'same' convolutions only"); return S4G_ERR_SHAPE; }
  if (g.Cin % 8 || g.ldx % 8 || g.Cout % 8) { s4g_set_error("umma_conv_fprop: channel counts must be multiples of 8"); return S4G_ERR_SHAPE; }
  const bool matrix = g.Hi == 1 && g.Wi == 1;
  FpropParams p{};
  p.tg = matrix ? matrix_tiles(g.Nimg, 128) : make_tiles(g.Nimg, g.Hi, g.Wi, 128);
  p.taps = g.R * g.S; p.S = g.S; p.pad = g.pad;
  p.chunks = cdiv(g.Cin, BLOCK_K);
  const int row_tiles = p.tg.tw * p.tg.th * p.tg.tn;
  // CTA pairs (cta_group::2) whenever there are at least two row tiles; S4G_FPROP_CG=1 forces the single-CTA kernel
  static const int forced_cg = getenv("S4G_FPROP_CG") ? atoi(getenv("S4G_FPROP_CG")) : 0;
  int cg = forced_cg ? forced_cg : (row_tiles >= 2 ? 2 : 1);
  p.block_n = pick_block_n(g.Cout, row_tiles, cg);
  if (p.block_n % 32) cg = 1;
  p.N = g.Cout;
  p.tiles_n = cdiv(g.Cout, p.block_n);
  p.e = e;
  p.tiles_m = cdiv(row_tiles, cg);
  const int stage_bytes = A_STAGE_BYTES + (p.block_n / cg) * 128;
  const size_t fixed = 1024 + EPI_WARPS * EPI_WARP_BYTES + sizeof(SmemCtl);
  p.stages = (int)std::min<size_t>(MAX_STAGES, (227 * 1024 - fixed) / stage_bytes);
  CUtensorMap tmA, tmB;
  S4G_TRY(make_act_map(&tmA, X, g, g.Cin, g.ldx, p.tg, matrix, g.Hi, g.Wi));
  {
    // weights [Cout][taps][Cin] as a 3-D tensor {Cin, taps, Cout}: the channel tail of a tap is zero-filled by
    // TMA instead of running into the next tap
    uint64_t dims[3] = {(uint64_t)g.Cin, (uint64_t)p.taps, (uint64_t)g.Cout};
    uint64_t strides[2] = {(uint64_t)g.Cin * 2, (uint64_t)g.Cin * p.taps * 2};
    uint32_t box[3] = {64, 1, (uint32_t)(p.block_n / cg)};
    S4G_TRY(make_map(&tmB, W, 3, dims, strides, box));
  }
  p.M = (long long)g.Nimg * g.Hi * g.Wi;
  p.tile_counter = next_tile_counter();
  p.tma_epi = !(e.out_f32 || e.res || e.accumulate) && e.ldo % 8 == 0 && (reinterpret_cast<uintptr_t>(e.out) & 15) == 0 &&
              (!e.out2 || (reinterpret_cast<uintptr_t>(e.out2) & 15) == 0);
  p.f32_epi = matrix && e.out_f32 && !e.accumulate && (!e.res || e.res_f32) && !e.out2 &&
              (e.mode == EPI_NONE || e.mode == EPI_RELU) && e.ldo % 4 == 0 && (!e.res || e.ldr % 4 == 0) && g.Cout % 4 == 0 &&
              (reinterpret_cast<uintptr_t>(e.out) & 15) == 0 && (reinterpret_cast<uintptr_t>(e.res) & 15) == 0 &&
              (!e.bias || (reinterpret_cast<uintptr_t>(e.bias) & 15) == 0);
  if (e.stat_part && (!p.tma_epi || p.tiles_n != 1 || g.Cout > 256 || g.Cout % 64 || e.out2 || e.mode != EPI_NONE)) {
    s4g_set_error("umma_conv_fprop: fused column statistics need the bf16 TMA-store epilogue, one column tile and Cout a multiple of 64 (<= 256)");
    return S4G_ERR_SHAPE;
  }
  if (e.dot_aux && (!p.tma_epi || p.block_n % 64 || !matrix || e.out2)) {
    s4g_set_error("umma_conv_fprop: the fused row-dot needs the bf16 TMA-store epilogue of a plain matrix with 64-column slabs");
    return S4G_ERR_SHAPE;
  }
  CUtensorMap tmO = tmA, tmO2 = tmA;
  if (p.tma_epi) {
    S4G_TRY(make_act_map(&tmO, reinterpret_cast<const bf16*>(e.out), g, g.Cout, e.ldo, p.tg, matrix, g.Hi, g.Wi, 4));
    if (e.out2) S4G_TRY(make_act_map(&tmO2, reinterpret_cast<const bf16*>(e.out2), g, g.Cout, e.ldo, p.tg, matrix, g.Hi, g.Wi, 4));
  }
  const size_t smem = fixed + (size_t)p.stages * stage_bytes;
  if (cg == 2) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * std::min(p.tiles_m * p.tiles_n, g_num_sms / 2));
    cfg.blockDim = dim3(FPROP_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = g_pdl ? 2 : 1;
    if (e.stat_part) S4G_CUDA(cudaLaunchKernelEx(&cfg, umma_fprop_kernel<2, 1>, tmA, tmB, tmO, tmO2, p));
    else S4G_CUDA(cudaLaunchKernelEx(&cfg, umma_fprop_kernel<2>, tmA, tmB, tmO, tmO2, p));
  } else {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(std::min(p.tiles_m * p.tiles_n, g_num_sms));
    cfg.blockDim = dim3(FPROP_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = g_pdl ? 1 : 0;
    if (e.stat_part) S4G_CUDA(cudaLaunchKernelEx(&cfg, umma_fprop_kernel<1, 1>, tmA, tmB, tmO, tmO2, p));
    else S4G_CUDA(cudaLaunchKernelEx(&cfg, umma_fprop_kernel<1>, tmA, tmB, tmO, tmO2, p));
  }
  S4G_KERNEL_CHECK();
  return 0;
}

namespace {
struct WgradPlan {
  TileGeom tg;
  int block_n, tiles_co, tiles_ci, out_tiles, chunks_total, splits, cps;   // tiles_co: 128*cg-row tiles
  int cg, stages;
  bool direct;
  size_t scratch_bytes, smem_bytes;
};
// One candidate: tile width block_n, best pixel split for it.  direct = the whole pixel range in one work item of a
// linear layer: the epilogue writes dW itself, no partials and no reduce launch.
WgradPlan plan_wgrad_bn(const ConvGeom& g, int block_n, double& cost) {
  WgradPlan w;
  const bool matrix = g.Hi == 1 && g.Wi == 1;
  w.tg = matrix ? matrix_tiles(g.Nimg, BLOCK_K) : make_tiles(g.Nimg, g.Hi, g.Wi, BLOCK_K);
  const int taps = g.R * g.S;
  w.block_n = block_n;
  // CTA pairs when a pair tile (256 co) is not mostly padding and each CTA's X share is whole 64-channel groups
  static const int forced_cg = getenv("S4G_WGRAD_CG") ? atoi(getenv("S4G_WGRAD_CG")) : 0;
  w.cg = forced_cg ? forced_cg : ((g.Cout > 128 && w.block_n % 128 == 0) ? 2 : 1);
  if (w.block_n % 128) w.cg = 1;
  const int stage_bytes = 2 * BLOCK_K * 128 + (w.block_n / 64 / w.cg) * BLOCK_K * 128;
  w.stages = (int)std::min<size_t>(MAX_STAGES, (227 * 1024 - 1024 - sizeof(SmemCtl)) / stage_bytes);
  w.smem_bytes = 1024 + (size_t)w.stages * stage_bytes + sizeof(SmemCtl);
  const int groups = g_num_sms / w.cg;
  w.tiles_co = cdiv(g.Cout, 128 * w.cg);
  w.tiles_ci = cdiv(g.Cin, w.block_n);
  w.out_tiles = w.tiles_co * w.tiles_ci * taps;
  w.chunks_total = w.tg.tw * w.tg.th * w.tg.tn;
  const double t_chunk = 0.0014 * w.block_n;                       // us per 64-pixel chunk (4 MMAs of 128 x block_n x 16)
  const double tile_bytes = 128.0 * w.block_n * 4;
  double best = 1e30;
  w.splits = 1; w.cps = w.chunks_total;
  for (int s = 1; s <= std::min(w.chunks_total, 256); ++s) {
    const int cps = cdiv(w.chunks_total, s);
    const int se = cdiv(w.chunks_total, cps);
    if (se != s) continue;
    const long long items = (long long)w.out_tiles * se;
    const long long rounds = (items + groups - 1) / groups;
    const bool direct = se == 1 && taps == 1;
    // partials are written once and read once by the reduce kernel, whose launch costs ~8 us on top
    const double t = rounds * (cps * t_chunk + 2.0) + (direct ? 0.0 : items * w.cg * tile_bytes * 2 / 3.0e6 + 8.0);
    if (t < best) { best = t; w.splits = se; w.cps = cps; }
  }
  w.direct = w.splits == 1 && taps == 1;
  w.scratch_bytes = w.direct ? 0 : (size_t)w.out_tiles * w.cg * w.splits * 128 * w.block_n * sizeof(float);
  cost = best;
  return w;
}
// Split-K plan: enough (output tile x pixel range) work items to fill the SMs in whole waves, few enough that the
// fp32 partials stay cheap.  Tiles are as wide as Cin allows (<= 256 columns): measured, 128-column tiles run the
// channel-major (MN-major) operands at less than half the rate per FLOP (fc1 weight gradient as 72 unsplit 256 x 128
// pair tiles: 83 us, against 48 + 9 us for 2-way split 256 x 256 tiles plus the reduce).  S4G_WGRAD_BN overrides.
WgradPlan plan_wgrad(const ConvGeom& g) {
  const int bn0 = std::min(256, (g.Cin + 63) / 64 * 64);
  static const int forced_bn = getenv("S4G_WGRAD_BN") ? atoi(getenv("S4G_WGRAD_BN")) : 0;
  double cost;
  return plan_wgrad_bn(g, forced_bn && forced_bn <= bn0 ? forced_bn : bn0, cost);
}
}  // namespace

size_t umma_wgrad_scratch_bytes(const ConvGeom& g) {
  if (umma_init() != 0) return 0;
  return plan_wgrad(g).scratch_bytes;
}

int umma_conv_wgrad(const ConvGeom& g, const bf16* dY, const bf16* X, float* dW, float* scratch, size_t scratch_bytes,
                    cudaStream_t st, int torch_layout) {
  S4G_TRY(umma_init());
  if (g.stride != 1 || g.Ho != g.Hi || g.Wo != g.Wi) { s4g_set_error("umma_conv_wgrad: stride-1 'same' convolutions only"); return S4G_ERR_SHAPE; }
  if (g.Cin % 8 || g.ldx % 8 || g.Cout % 8 || g.ldy % 8) { s4g_set_error("umma_conv_wgrad: channel counts must be multiples of 8"); return S4G_ERR_SHAPE; }
  const bool matrix = g.Hi == 1 && g.Wi == 1;
  const WgradPlan w = plan_wgrad(g);
  if (w.scratch_bytes && (!scratch || scratch_bytes < w.scratch_bytes)) { s4g_set_error("umma_conv_wgrad: split-K scratch too small (see s4g_wgrad_scratch_bytes)"); return S4G_ERR_WORKSPACE; }
  WgradParams p{};
  p.tg = w.tg;
  p.taps = g.R * g.S; p.S = g.S; p.pad = g.pad;
  p.Cout = g.Cout; p.Cin = g.Cin; p.K = p.taps * g.Cin;
  p.block_n = w.block_n;
  p.tiles_co = w.tiles_co;
  p.tiles_ci = w.tiles_ci;
  p.chunks_total = w.chunks_total;
  p.splits = w.splits;
  p.tile_counter = next_tile_counter();
  p.stages = w.stages;
  p.chunks_per_split = w.cps;
  p.part = scratch;
  p.dW = dW;
  p.direct = w.direct;
  if (w.direct && (reinterpret_cast<uintptr_t>(dW) & 15)) { s4g_set_error("umma_conv_wgrad: dW must be 16 B aligned"); return S4G_ERR_ALIGN; }
  CUtensorMap tmY, tmX;
  S4G_TRY(make_act_map(&tmY, dY, g, g.Cout, g.ldy, p.tg, matrix, g.Hi, g.Wi));
  S4G_TRY(make_act_map(&tmX, X, g, g.Cin, g.ldx, p.tg, matrix, g.Hi, g.Wi));
  if (w.cg == 2) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * std::min(w.out_tiles * p.splits, g_num_sms / 2));
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = w.smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = g_pdl ? 2 : 1;
    S4G_CUDA(cudaLaunchKernelEx(&cfg, umma_wgrad_kernel<2>, tmY, tmX, p));
  } else {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(std::min(w.out_tiles * p.splits, g_num_sms));
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = w.smem_bytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = g_pdl ? 1 : 0;
    S4G_CUDA(cudaLaunchKernelEx(&cfg, umma_wgrad_kernel<1>, tmY, tmX, p));
  }
  S4G_KERNEL_CHECK();
  if (w.direct) return 0;
  WgradReduceParams r{};
  r.part = scratch; r.dW = dW;
  r.Cout = g.Cout; r.Cin = g.Cin; r.taps = p.taps; r.tiles_co = p.tiles_co * w.cg; r.tiles_ci = p.tiles_ci; r.block_n = p.block_n;
  r.splits = p.splits;
  if (torch_layout) { r.so = (long long)g.Cin * p.taps; r.si = p.taps; r.st = 1; }   // [Cout][Cin][R][S]
  else { r.so = p.K; r.si = 1; r.st = g.Cin; }                                        // packed [Cout][tap][Cin]
  if (p.taps == 1 && r.si == 1) {
    const int bx = r.tiles_co * r.tiles_ci * (r.block_n / 32);
    // enough blocks to cover the memory latency: ~4 splits per block once a layer is split more than 8 ways
    int dev = 0;
    S4G_CUDA(cudaGetDevice(&dev));
    r.tickets = (dev >= 0 && dev < MAX_DEV) ? g_reduce_tickets[dev] : nullptr;
    r.groups = (r.splits > 8 && bx <= REDUCE_TICKETS && r.tickets) ? std::min(16, cdiv(r.splits, 4)) : 1;
    S4G_CUDA(launch_dep(wgrad_reduce_linear_kernel, dim3(bx, r.groups), dim3(256), 0, st, r));
  }
  else
    S4G_CUDA(launch_dep(wgrad_reduce_kernel, dim3(dim3(cdiv(g.Cout, 32), cdiv(g.Cin, 32), p.taps)), dim3(dim3(32, 8)), 0, st, r));
  S4G_KERNEL_CHECK();
  return 0;
}
