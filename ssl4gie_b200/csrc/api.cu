// C-ABI plumbing: error state, device check, operator-level entry points declared in include/ssl4gie_b200.h.
#include <cstring>
#include "kernels.h"

static thread_local std::string g_err;
void s4g_set_error(const std::string& msg) { g_err = msg; }
int s4g_cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  g_err = buf;
  return S4G_ERR_CUDA;
}

#define ST(s) reinterpret_cast<cudaStream_t>(s)
#define DISPATCH(dtype, expr_f, expr_h)                      \
  do {                                                       \
    if ((dtype) == S4G_DTYPE_F32) { typedef float T; return (expr_f); } \
    if ((dtype) == S4G_DTYPE_BF16) { typedef bf16 T; return (expr_h); } \
    s4g_set_error("bad dtype");                              \
    return S4G_ERR_SHAPE;                                    \
  } while (0)

extern "C" {

int s4g_version(void) { return 100; }
const char* s4g_last_error(void) { return g_err.c_str(); }

int s4g_init(int device) {
  cudaDeviceProp prop;
  S4G_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    char buf[128];
    snprintf(buf, sizeof(buf), "libssl4gie_b200 is built for sm_100a only; device %d is sm_%d%d", device, prop.major, prop.minor);
    g_err = buf;
    return S4G_ERR_ARCH;
  }
  S4G_CUDA(cudaSetDevice(device));
  S4G_TRY(reduce_init());   // the only device allocation the library ever makes: 32 MB of reduction scratch per device
  return umma_init();
}

static int check_backend(int dtype, int backend) {
  if (backend == S4G_BACKEND_TENSOR && dtype != S4G_DTYPE_BF16) {
    s4g_set_error("tensor-core backend needs bf16");
    return S4G_ERR_STATE;
  }
  return 0;
}

int s4g_linear_fwd(const void* A, const void* W, const float* bias, const float* residual, void* out, long long M, int N,
                   int K, int act, int out_f32, int dtype, int backend, void* stream) {
  S4G_TRY(check_backend(dtype, backend));
  EpiParams e{};
  e.bias = bias; e.out = out; e.ldo = N; e.mode = act; e.out_f32 = out_f32; e.alpha = 1.f;
  e.res = residual; e.res_f32 = 1; e.ldr = N;
  const ConvGeom g = linear_geom(M, N, K);
  if (backend == S4G_BACKEND_TENSOR) return umma_conv_fprop(g, (const bf16*)A, (const bf16*)W, e, ST(stream));
  DISPATCH(dtype, simt_conv_fprop<T>(g, (const T*)A, (const T*)W, e, ST(stream)), simt_conv_fprop<T>(g, (const T*)A, (const T*)W, e, ST(stream)));
}
int s4g_linear_wgrad(const void* dY, const void* X, float* dW, long long M, int N, int K, int dtype, int backend,
                     void* scratch, size_t scratch_bytes, void* stream) {
  S4G_TRY(check_backend(dtype, backend));
  const ConvGeom g = linear_geom(M, N, K);
  if (backend == S4G_BACKEND_TENSOR) return umma_conv_wgrad(g, (const bf16*)dY, (const bf16*)X, dW, (float*)scratch, scratch_bytes, ST(stream));
  DISPATCH(dtype, simt_conv_wgrad<T>(g, (const T*)dY, (const T*)X, dW, ST(stream)), simt_conv_wgrad<T>(g, (const T*)dY, (const T*)X, dW, ST(stream)));
}
int s4g_conv3x3_fwd(const void* X, const void* Wp, const float* bias, void* out, int B, int H, int W, int Cin, int Cout,
                    int dtype, int backend, void* stream) {
  S4G_TRY(check_backend(dtype, backend));
  EpiParams e{};
  e.bias = bias; e.out = out; e.ldo = Cout; e.alpha = 1.f;
  const ConvGeom g = conv3_geom(B, H, W, Cin, Cout);
  if (backend == S4G_BACKEND_TENSOR) return umma_conv_fprop(g, (const bf16*)X, (const bf16*)Wp, e, ST(stream));
  DISPATCH(dtype, simt_conv_fprop<T>(g, (const T*)X, (const T*)Wp, e, ST(stream)), simt_conv_fprop<T>(g, (const T*)X, (const T*)Wp, e, ST(stream)));
}
int s4g_conv3x3_wgrad(const void* dY, const void* X, float* dWp, int B, int H, int W, int Cin, int Cout, int dtype,
                      int backend, void* scratch, size_t scratch_bytes, void* stream) {
  S4G_TRY(check_backend(dtype, backend));
  const ConvGeom g = conv3_geom(B, H, W, Cin, Cout);
  if (backend == S4G_BACKEND_TENSOR) return umma_conv_wgrad(g, (const bf16*)dY, (const bf16*)X, dWp, (float*)scratch, scratch_bytes, ST(stream));
  DISPATCH(dtype, simt_conv_wgrad<T>(g, (const T*)dY, (const T*)X, dWp, ST(stream)), simt_conv_wgrad<T>(g, (const T*)dY, (const T*)X, dWp, ST(stream)));
}
size_t s4g_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout, int taps) {
  ConvGeom g = taps == 9 ? conv3_geom(B, H, W, Cin, Cout) : linear_geom((long long)B * H * W, Cout, Cin);
  return umma_wgrad_scratch_bytes(g);
}
int s4g_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd, int rows,
                      int C, float eps, int dtype, void* stream) {
  DISPATCH(dtype, k_layernorm_fwd<T>(x, gamma, beta, (T*)y, mean, rstd, rows, C, eps, ST(stream)),
           k_layernorm_fwd<T>(x, gamma, beta, (T*)y, mean, rstd, rows, C, eps, ST(stream)));
}
int s4g_layernorm_bwd(const void* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                      const float* dres, float* dx, float* dgamma, float* dbeta, int rows, int C, int dtype, void* stream) {
  DISPATCH(dtype, k_layernorm_bwd<T>((const T*)dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, rows, C, ST(stream)),
           k_layernorm_bwd<T>((const T*)dy, x, gamma, mean, rstd, dres, dx, dgamma, dbeta, rows, C, ST(stream)));
}
int s4g_attention_fwd(const void* qkv, void* o, float* lse, int B, int N, int H, int dtype, int backend, void* stream) {
  S4G_TRY(check_backend(dtype, backend));
  if (backend == S4G_BACKEND_TENSOR) return attn_fwd_mma((const bf16*)qkv, (bf16*)o, lse, B, N, H, ST(stream));
  DISPATCH(dtype, attn_fwd_simt<T>((const T*)qkv, (T*)o, lse, B, N, H, ST(stream)), attn_fwd_simt<T>((const T*)qkv, (T*)o, lse, B, N, H, ST(stream)));
}
int s4g_attention_bwd(const void* qkv, const void* o, const void* dout, const float* lse, void* dqkv, int B, int N, int H,
                      int dtype, int backend, void* stream) {
  S4G_TRY(check_backend(dtype, backend));
  if (backend == S4G_BACKEND_TENSOR) return attn_bwd_mma((const bf16*)qkv, (const bf16*)o, (const bf16*)dout, lse, (bf16*)dqkv, B, N, H, ST(stream));
  DISPATCH(dtype, attn_bwd_simt<T>((const T*)qkv, (const T*)o, (const T*)dout, lse, (T*)dqkv, B, N, H, ST(stream)),
           attn_bwd_simt<T>((const T*)qkv, (const T*)o, (const T*)dout, lse, (T*)dqkv, B, N, H, ST(stream)));
}
int s4g_bilinear2x_fwd(const void* x, void* y, int B, int H, int W, int C, int dtype, void* stream) {
  DISPATCH(dtype, k_bilinear2x_fwd<T>((const T*)x, (T*)y, B, H, W, C, ST(stream)), k_bilinear2x_fwd<T>((const T*)x, (T*)y, B, H, W, C, ST(stream)));
}
int s4g_bilinear2x_bwd(const void* dy, void* dx, int B, int H, int W, int C, int dtype, void* stream) {
  DISPATCH(dtype, k_bilinear2x_bwd<T>((const T*)dy, (T*)dx, B, H, W, C, ST(stream)), k_bilinear2x_bwd<T>((const T*)dy, (T*)dx, B, H, W, C, ST(stream)));
}
size_t s4g_dice_scratch_bytes(int B) { return sizeof(double) * dice_scratch_doubles(B); }
int s4g_dice_fwd(const float* logits, const float* targets, void* scratch, float* loss, int B, int HW, float smooth, void* stream) {
  return k_dice_fwd(logits, targets, (float*)scratch, loss, B, HW, smooth, ST(stream));
}
int s4g_dice_bwd(const float* logits, const float* targets, const void* scratch, const float* grad_out, float* dlogits, int B,
                 int HW, float smooth, void* stream) {
  return k_dice_bwd(logits, targets, (const float*)scratch, grad_out, dlogits, B, HW, smooth, ST(stream));
}
size_t s4g_ssi_scratch_bytes(int B) { return sizeof(float) * ssi_scratch_floats(B); }
int s4g_ssi_fwd(const float* pred, const float* target, void* scratch, float* loss, float* ssi_out, int B, int H, int W,
                float alpha, int scales, int image_based, void* stream) {
  return k_ssi_fwd(pred, target, (float*)scratch, loss, ssi_out, B, H, W, alpha, scales, image_based, ST(stream));
}
int s4g_ssi_bwd(const float* pred, const float* target, void* scratch, const float* grad_out, float* dpred, int B, int H,
                int W, float alpha, int scales, void* stream) {
  return k_ssi_bwd(pred, target, (const float*)scratch, grad_out, dpred, B, H, W, alpha, scales, ST(stream));
}
int s4g_ce_fwd(const float* logits, const long long* labels, const float* class_weight, float* loss, float* dlogits_unit, int B,
               int num_classes, void* stream) {
  if (!logits || !labels || !loss || !dlogits_unit || B <= 0 || num_classes <= 0) { s4g_set_error("ce_fwd: bad argument"); return S4G_ERR_SHAPE; }
  return k_ce_fwd(logits, labels, class_weight, loss, dlogits_unit, B, num_classes, ST(stream));
}
int s4g_ce_bwd(const float* dlogits_unit, const float* grad_out, float* dlogits, int B, int num_classes, void* stream) {
  return k_ce_bwd(dlogits_unit, grad_out, dlogits, B, num_classes, ST(stream));
}
int s4g_adamw_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr, float beta1,
                   float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream) {
  return k_adamw(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step, grad_scale, ST(stream));
}
int s4g_adamw_step_groups(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                          float beta1, float beta2, float eps, const float* weight_decay_per_slot, int step, float grad_scale,
                          void* stream) {
  if (!weight_decay_per_slot || n % 64) { s4g_set_error("adamw_step_groups: need one weight decay per 64-float slot"); return S4G_ERR_SHAPE; }
  return k_adamw(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, 0.f, step, grad_scale, ST(stream), weight_decay_per_slot);
}

size_t s4g_peer_mailbox_bytes(void) { return k_peer_mailbox_bytes(); }
int s4g_peer_alloc(size_t bytes, void** dptr, unsigned char handle[S4G_PEER_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == S4G_PEER_HANDLE_BYTES, "cudaIpcMemHandle_t size");
  if (!dptr || !handle) { s4g_set_error("peer_alloc: null argument"); return S4G_ERR_STATE; }
  S4G_CUDA(cudaMalloc(dptr, bytes));
  S4G_CUDA(cudaMemset(*dptr, 0, bytes));
  S4G_CUDA(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  S4G_CUDA(cudaIpcGetMemHandle(&h, *dptr));
  memcpy(handle, &h, sizeof(h));
  return 0;
}
int s4g_peer_open(const unsigned char handle[S4G_PEER_HANDLE_BYTES], void** dptr) {
  if (!dptr || !handle) { s4g_set_error("peer_open: null argument"); return S4G_ERR_STATE; }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  S4G_CUDA(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int s4g_peer_close(void* dptr) { S4G_CUDA(cudaIpcCloseMemHandle(dptr)); return 0; }
int s4g_peer_free(void* dptr) { S4G_CUDA(cudaFree(dptr)); return 0; }
int s4g_peer_sum_f64(double* data, int count, void* const* mailboxes, int rank, int world, void* stream) {
  return k_peer_sum_f64(data, count, mailboxes, rank, world, ST(stream));
}
int s4g_peer_check(void* own_mailbox) {
  if (!own_mailbox) { s4g_set_error("peer_check: null mailbox"); return S4G_ERR_STATE; }
  unsigned long long ctl[2] = {0, 0};
  S4G_CUDA(cudaDeviceSynchronize());
  S4G_CUDA(cudaMemcpy(ctl, reinterpret_cast<char*>(own_mailbox) + k_peer_ctl_offset(), sizeof(ctl), cudaMemcpyDeviceToHost));
  if (ctl[1]) {
    char buf[160];
    snprintf(buf, sizeof(buf), "peer exchange %llu timed out waiting for rank %d (S4G_PEER_TIMEOUT_S)", ctl[1] >> 8, (int)(ctl[1] & 255) - 1);
    s4g_set_error(buf);
    return S4G_ERR_STATE;
  }
  return 0;
}

size_t s4g_mae_loss_scratch_bytes(int B) { return mae_loss_scratch_bytes(B); }
int s4g_mae_loss_fwd(const float* pred, const float* images, const float* mask, void* scratch, float* loss, int B,
                     int norm_pix, void* stream) {
  if (!pred || !images || !mask || !scratch || !loss || B < 1) { s4g_set_error("mae_loss_fwd: bad argument"); return S4G_ERR_STATE; }
  return k_mae_loss_fwd(pred, images, mask, scratch, loss, B, norm_pix, ST(stream));
}
int s4g_mae_loss_bwd(const float* pred, const float* images, const float* mask, const void* scratch,
                     const float* grad_out, float* dpred, int B, int norm_pix, void* stream) {
  if (!pred || !images || !mask || !scratch || !grad_out || !dpred || B < 1) { s4g_set_error("mae_loss_bwd: bad argument"); return S4G_ERR_STATE; }
  return k_mae_loss_bwd(pred, images, mask, scratch, grad_out, dpred, B, norm_pix, ST(stream));
}

size_t s4g_aug_resize_scratch_bytes(const long long* desc_host, int B, int C) {
  if (!desc_host || B < 1) return 0;
  return aug_resize_scratch_bytes(desc_host, B, C);
}
int s4g_aug_resize(const unsigned char* src, const long long* desc_dev, const long long* desc_host, int B, int C, void* scratch,
                   size_t scratch_bytes, unsigned char* out, void* stream) {
  if (!src || !desc_dev || !desc_host || !scratch || !out || B < 1) { s4g_set_error("aug_resize: bad argument"); return S4G_ERR_STATE; }
  return k_aug_resize(src, desc_dev, desc_host, B, C, scratch, scratch_bytes, out, ST(stream));
}
int s4g_aug_color_jitter(unsigned char* img, const void* params, int B, void* stream) {
  if (!img || !params || B < 1) { s4g_set_error("aug_color_jitter: bad argument"); return S4G_ERR_STATE; }
  return k_aug_color_jitter(img, params, B, ST(stream));
}
int s4g_aug_blur_normalize(const unsigned char* img, const float* sigma, float* out, const float* mean3_host,
                           const float* std3_host, int B, void* stream) {
  if (!img || !out || !mean3_host || !std3_host || B < 1) { s4g_set_error("aug_blur_normalize: bad argument"); return S4G_ERR_STATE; }
  return k_aug_blur_normalize(img, sigma, out, mean3_host, std3_host, B, ST(stream));
}
int s4g_aug_mask_to_tensor(const unsigned char* img, float* out, int B, int C, void* stream) {
  if (!img || !out || B < 1) { s4g_set_error("aug_mask_to_tensor: bad argument"); return S4G_ERR_STATE; }
  return k_aug_mask_to_tensor(img, out, B, C, ST(stream));
}
int s4g_aug_depth_target(const void* src, const long long* desc_dev, float* out, int B, void* stream) {
  if (!src || !desc_dev || !out || B < 1) { s4g_set_error("aug_depth_target: bad argument"); return S4G_ERR_STATE; }
  return k_aug_depth_target(src, desc_dev, out, B, ST(stream));
}
int s4g_aug_geometric(const float* x, const float* y, float* xo, float* yo, const void* params, float fill_x, float fill_y,
                      int B, void* stream) {
  if (!x || !xo || !params || (y && !yo) || B < 1) { s4g_set_error("aug_geometric: bad argument"); return S4G_ERR_STATE; }
  return k_aug_geometric(x, y, xo, yo, params, fill_x, fill_y, B, ST(stream));
}

int s4g_adamw_step_device(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, const float* hyper_dev,
                          const int* step_dev, const float* weight_decay_per_slot, void* stream) {
  if (!params || !grads || !exp_avg || !exp_avg_sq || !hyper_dev || !step_dev) { s4g_set_error("adamw_step_device: null argument"); return S4G_ERR_STATE; }
  return k_adamw_dev(params, grads, exp_avg, exp_avg_sq, n, hyper_dev, step_dev, weight_decay_per_slot, ST(stream));
}
int s4g_device_hyper_set(float* hyper_dev, float lr, float beta1, float beta2, float eps, float weight_decay, float grad_scale, void* stream) {
  if (!hyper_dev) { s4g_set_error("device_hyper_set: null argument"); return S4G_ERR_STATE; }
  return k_device_hyper_set(hyper_dev, lr, beta1, beta2, eps, weight_decay, grad_scale, ST(stream));
}
int s4g_device_counter_advance(int* step_dev, unsigned long long* seed_dev, void* stream) {
  return k_device_counter_advance(step_dev, seed_dev, ST(stream));
}

}  // extern "C"
