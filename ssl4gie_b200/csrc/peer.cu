// SyncBatchNorm statistic exchange over NVLink peer memory (reference: torch.nn.SyncBatchNorm under DDP,
// Binary_segmentation/train_segmentation.py:199-203 -- an all_gather of per-rank mean/invstd/count in forward and an
// all_reduce of sum_dy / sum_dy_xmu in backward, one NCCL collective each, 29 per training step).
//
// Here the exchange is ONE single-CTA kernel per statistic vector (<= 512 doubles): every rank stores its vector into
// its slot of every peer's mailbox with plain NVLink P2P stores, publishes a sequence flag with a system-scope release,
// spins on its own mailbox until all ranks' flags for this exchange have arrived, and sums the slots in rank order --
// so every rank computes bit-identical totals, with no host callback and no NCCL launch on the critical path.
//
// Mailbox layout (one per rank, cudaMalloc'ed, exported with cudaIpc): double slots[2][MAXW][512]; u64 flags[2][MAXW];
// u64 seq; u64 error.  Slots are double-buffered by the parity of the exchange counter: a peer can only write parity p
// again (exchange s + 2) after it has finished exchange s + 1, which needed this rank's flag for s + 1, which this rank
// sends only after the kernel of exchange s (the reader of parity p) has completed on its stream.
// The exchange counter `seq` lives in the rank's OWN mailbox (read at kernel start, incremented at kernel end; kernels
// of one rank are stream-ordered): every model handle of a process shares one mailbox set and one counter, nothing
// about the exchange is a kernel argument that changes per call (CUDA-graph replayable), and a handle created on one
// rank only (rank-0 evaluation) cannot desynchronise it.
// A peer that does not answer within S4G_PEER_TIMEOUT_S seconds (default 600) sets `error` in the own mailbox and the
// kernel returns without trapping the context; s4g_peer_check() reports it on the host.
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

__device__ __forceinline__ unsigned long long* box_flags(void* box) {
  return reinterpret_cast<unsigned long long*>(reinterpret_cast<double*>(box) + 2 * MAXW * SLOT);
}

__global__ void __launch_bounds__(SLOT) peer_sum_f64_kernel(double* __restrict__ data, int count, PeerBoxes boxes, int rank,
                                                            int world, unsigned long long timeout_ns) {
  const int tid = threadIdx.x;
  unsigned long long* own_ctl = box_flags(boxes.p[rank]) + 2 * MAXW;   // [0] = seq, [1] = error
  const unsigned long long seq = *reinterpret_cast<volatile unsigned long long*>(own_ctl);
  const int par = (int)(seq & 1ull);
  __syncthreads();   // every thread has read seq before thread 0 advances it below
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
    st_release_sys(box_flags(boxes.p[tid]) + par * MAXW + rank, seq + 1);
    const unsigned long long* mine = box_flags(boxes.p[rank]) + par * MAXW + tid;
    const unsigned long long t0 = globaltimer_ns();
    while (ld_acquire_sys(mine) < seq + 1) {
      if (globaltimer_ns() - t0 > timeout_ns) {  // a peer died or the ranks' schedules diverged: flag it, do not trap
        own_ctl[1] = ((seq + 1) << 8) | (unsigned long long)(tid + 1);
        break;
      }
    }
  }
  __syncthreads();
  if (tid == 0) *reinterpret_cast<volatile unsigned long long*>(own_ctl) = seq + 1;
  if (tid < count) {
    const volatile double* slots = reinterpret_cast<const volatile double*>(boxes.p[rank]) + (size_t)par * MAXW * SLOT;
    double s = 0.0;
    for (int r = 0; r < world; ++r) s += slots[(size_t)r * SLOT + tid];
    data[tid] = s;
  }
}

}  // namespace

size_t k_peer_mailbox_bytes() { return sizeof(double) * 2 * MAXW * SLOT + sizeof(unsigned long long) * (2 * MAXW + 2); }
// byte offset of the {seq, error} pair inside a mailbox
size_t k_peer_ctl_offset() { return sizeof(double) * 2 * MAXW * SLOT + sizeof(unsigned long long) * 2 * MAXW; }

int k_peer_sum_f64(double* data, int count, void* const* boxes, int rank, int world, cudaStream_t st) {
  static const unsigned long long timeout_ns = [] {
    const char* e = getenv("S4G_PEER_TIMEOUT_S");
    const double s = e ? atof(e) : 600.0;
    return (unsigned long long)((s > 0 ? s : 600.0) * 1e9);
  }();
  if (count > SLOT || world > MAXW || world < 1) { s4g_set_error("peer_sum_f64: count > 512 or world > 8"); return S4G_ERR_SHAPE; }
  PeerBoxes b{};
  for (int r = 0; r < world; ++r) b.p[r] = boxes[r];
  peer_sum_f64_kernel<<<1, SLOT, 0, st>>>(data, count, b, rank, world, timeout_ns);
  S4G_KERNEL_CHECK();
  return 0;
}
