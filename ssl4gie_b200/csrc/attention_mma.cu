#include "kernels.h"
int attn_fwd_mma(const bf16*, bf16*, float*, int, int, int, cudaStream_t) { s4g_set_error("mma attention not built yet"); return S4G_ERR_STATE; }
int attn_bwd_mma(const bf16*, const bf16*, const bf16*, const float*, bf16*, int, int, int, cudaStream_t) { s4g_set_error("mma attention not built yet"); return S4G_ERR_STATE; }
