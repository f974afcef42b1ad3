// SyncBatchNorm statistic exchange over NVLink peer memory (reference: torch.nn.SyncBatchNorm under DDP,
// Binary_segmentation/train_segmentation.py:199-203 -- an all_gather of per-rank mean/invstd/count in forward and an
// all_reduce of sum_dy / sum_dy_xmu in backward, one NCCL collective each, 29 per training step).
//
// Here the exchange is ONE single-CTA kernel per statistic vector (<= 512 doubles): every rank stores its vector into
// its slot of every peer's mailbox with plain NVLink P2P stores, publishes a sequence flag with a system-scope release,
// spins on its own mailbox until all ranks' flags for this exchange have arrived, and sums the slots in rank order --
// so every rank computes bit-identical totals, with no host callback and no NCCL launch on the critical path.
//
// Mailbox layout (one per rank, cudaMalloc'ed, exported with cudaIpc): double slots[2][MAXW][512]; u64 flags[2][MAXW].
// Slots are double-buffered by the parity of the exchange counter: a peer can only write parity p again (exchange
// s + 2) after it has finished exchange s + 1, which needed this rank's flag for s + 1, which this rank sends only
// after the kernel of exchange s (the reader of parity p) has completed on its stream.
#include "kernels.h"

namespace {

constexpr int MAXW = S4G_PEER_MAX_WORLD;
constexpr int SLOT = 512;

struct PeerBoxes { void* p[MAXW]; };

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(SLOT) peer_sum_f64_kernel(double* __restrict__ data, int count, PeerBoxes boxes, int rank,
                                                            int world, unsigned long long seq) {
  const int tid = threadIdx.x;
  const int par = (int)(seq & 1ull);
  if (tid < count) {
    const double v = data[tid];
    for (int r = 0; r < world; ++r) {
      double* slots = reinterpret_cast<double*>(boxes.p[r]);
      slots[((size_t)par * MAXW + rank) * SLOT + tid] = v;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (tid < world) {
    unsigned long long* flags = reinterpret_cast<unsigned long long*>(reinterpret_cast<double*>(boxes.p[tid]) + 2 * MAXW * SLOT);
    st_release_sys(flags + par * MAXW + rank, seq + 1);
    const unsigned long long* mine =
        reinterpret_cast<const unsigned long long*>(reinterpret_cast<const double*>(boxes.p[rank]) + 2 * MAXW * SLOT) + par * MAXW + tid;
    const unsigned long long t0 = globaltimer_ns();
    while (ld_acquire_sys(mine) < seq + 1) {
      if (globaltimer_ns() - t0 > 20000000000ull) {  // 20 s: a peer died or the ranks' schedules diverged
        printf("ssl4gie_b200 peer exchange %llu: rank %d timed out waiting for rank %d\n", seq, rank, tid);
        __trap();
      }
    }
  }
  __syncthreads();
  if (tid < count) {
    const volatile double* slots = reinterpret_cast<const volatile double*>(boxes.p[rank]) + (size_t)par * MAXW * SLOT;
    double s = 0.0;
    for (int r = 0; r < world; ++r) s += slots[(size_t)r * SLOT + tid];
    data[tid] = s;
  }
}

}  // namespace

size_t k_peer_mailbox_bytes() { return sizeof(double) * 2 * MAXW * SLOT + sizeof(unsigned long long) * 2 * MAXW; }

int k_peer_sum_f64(double* data, int count, void* const* boxes, int rank, int world, unsigned long long seq, cudaStream_t st) {
  if (count > SLOT || world > MAXW || world < 1) { s4g_set_error("peer_sum_f64: count > 512 or world > 8"); return S4G_ERR_SHAPE; }
  PeerBoxes b{};
  for (int r = 0; r < world; ++r) b.p[r] = boxes[r];
  peer_sum_f64_kernel<<<1, SLOT, 0, st>>>(data, count, b, rank, world, seq);
  S4G_KERNEL_CHECK();
  return 0;
}
