#include "kernels.h"
int umma_init() { return 0; }
int umma_conv_fprop(const ConvGeom&, const bf16*, const bf16*, const EpiParams&, cudaStream_t) { s4g_set_error("tcgen05 path not built yet"); return S4G_ERR_STATE; }
int umma_conv_wgrad(const ConvGeom&, const bf16*, const bf16*, float*, float*, size_t, cudaStream_t) { s4g_set_error("tcgen05 path not built yet"); return S4G_ERR_STATE; }
size_t umma_wgrad_scratch_bytes(const ConvGeom&) { return 0; }
