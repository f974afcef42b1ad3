// Tensor-core attention entry points.  Until the mma kernels land these forward to the FFMA kernels so the
// tensor backend is complete end to end.
#include "kernels.h"
int attn_fwd_mma(const bf16* qkv, bf16* o, float* lse, int B, int N, int H, cudaStream_t st) { return attn_fwd_simt<bf16>(qkv, o, lse, B, N, H, st); }
int attn_bwd_mma(const bf16* qkv, const bf16* o, const bf16* dout, const float* lse, bf16* dqkv, int B, int N, int H, cudaStream_t st) {
  return attn_bwd_simt<bf16>(qkv, o, dout, lse, dqkv, B, N, H, st);
}
