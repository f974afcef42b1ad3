// Deterministic cross-block reductions.
//
// Every kernel of the path that reduces over rows / pixels (LayerNorm and BatchNorm backward, bias-gradient column sums,
// BatchNorm statistics, the head and loss sums) writes ONE partial row per block -- combined inside the block in a fixed
// order -- and the rows are then summed by finalize_kernel in a fixed partition and order.  No floating-point atomics
// anywhere: two runs of the same step on the same device give bit-identical results, as torch's LayerNorm / BatchNorm
// backward and the reference's losses do (Binary_segmentation/train_segmentation.py:41-52 under a fixed seed).
//
// The combine step of several producers can be queued (Reducer::defer) and run as ONE grouped launch: inside the model
// backward the parameter gradients it produces are only needed at the next gradient-bucket boundary.
#include <mutex>

#include "kernels.h"

namespace {

__global__ void __launch_bounds__(256) finalize_kernel(const __grid_constant__ FinalizeJobs jobs) {
  pdl_wait();  // launched programmatically dependent: nothing in global memory is touched before this
  pdl_trigger();
  __shared__ double red[8][33];
  int lo = 0, hi = jobs.n - 1;
  while (lo < hi) {  // last job whose first block is <= blockIdx.x
    const int mid = (lo + hi + 1) >> 1;
    if (jobs.job[mid].block0 <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const FinalizeJob& j = jobs.job[lo];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = ((int)blockIdx.x - j.block0) * 32 + lane;
  // warp w owns rows w, w + 8, ...; four running sums per thread keep loads in flight.  Partition and order are fixed.
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  if (col < j.ncols) {
    const size_t ld = (size_t)j.ld;
    int r = warp;
    if (j.in_f64) {
      const double* p = reinterpret_cast<const double*>(j.part) + col;
      for (; r + 24 < j.nrows; r += 32) { a0 += p[r * ld]; a1 += p[(r + 8) * ld]; a2 += p[(r + 16) * ld]; a3 += p[(r + 24) * ld]; }
      for (; r < j.nrows; r += 8) a0 += p[r * ld];
    } else {
      const float* p = reinterpret_cast<const float*>(j.part) + col;
      for (; r + 24 < j.nrows; r += 32) { a0 += (double)p[r * ld]; a1 += (double)p[(r + 8) * ld]; a2 += (double)p[(r + 16) * ld]; a3 += (double)p[(r + 24) * ld]; }
      for (; r < j.nrows; r += 8) a0 += (double)p[r * ld];
    }
  }
  red[warp][lane] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  if (warp == 0 && col < j.ncols) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[k][lane];
    if (j.out_f64) reinterpret_cast<double*>(j.out)[col] = s;
    else reinterpret_cast<float*>(j.out)[col] = (float)s;
  }
}

constexpr size_t LIB_SCRATCH_BYTES = 32u << 20;
constexpr int MAX_DEVICES = 64;
void* g_scratch[MAX_DEVICES] = {};
std::mutex g_scratch_mutex;

}  // namespace

int reduce_init() {
  int dev = 0;
  S4G_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) { s4g_set_error("reduce_init: device index out of range"); return S4G_ERR_STATE; }
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch[dev]) S4G_CUDA(cudaMalloc(&g_scratch[dev], LIB_SCRATCH_BYTES));
  return 0;
}

void* reduce_scratch(const Reducer* r, size_t bytes) {
  if (r && r->scratch) {
    if (bytes > r->scratch_bytes) { s4g_set_error("internal: reduction scratch too small"); return nullptr; }
    return r->scratch;
  }
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES || !g_scratch[dev]) {
    s4g_set_error("reduction scratch missing: call s4g_init(device) first");
    return nullptr;
  }
  if (bytes > LIB_SCRATCH_BYTES) { s4g_set_error("reduction too large for the library scratch"); return nullptr; }
  return g_scratch[dev];
}

int k_finalize_flush(FinalizeJobs* jobs, cudaStream_t st) {
  if (!jobs || jobs->n == 0) return 0;
  S4G_CUDA(launch_dep(finalize_kernel, dim3(jobs->total_blocks), dim3(256), 0, st, *jobs));
  S4G_KERNEL_CHECK();
  jobs->n = 0; jobs->total_blocks = 0;
  jobs->launches++;
  return 0;
}

int k_finalize(const Reducer* r, const void* part, void* out, int nrows, int ncols, int ld, int in_f64, int out_f64, cudaStream_t st) {
  if (!out || ncols <= 0) return 0;
  // queued only when the partials live in caller-owned scratch (the library scratch is reused by the next producer)
  FinalizeJobs* q = (r && r->scratch && r->defer) ? r->defer : nullptr;
  FinalizeJobs local;
  local.n = 0; local.total_blocks = 0; local.launches = 0;
  FinalizeJobs* jobs = q ? q : &local;
  if (jobs->n == MAX_FIN_JOBS) S4G_TRY(k_finalize_flush(jobs, st));
  FinalizeJob& j = jobs->job[jobs->n++];
  j.part = part; j.out = out; j.nrows = nrows; j.ncols = ncols; j.ld = ld; j.in_f64 = in_f64; j.out_f64 = out_f64;
  j.block0 = jobs->total_blocks;
  jobs->total_blocks += cdiv(ncols, 32);
  if (!q) return k_finalize_flush(jobs, st);
  return 0;
}
