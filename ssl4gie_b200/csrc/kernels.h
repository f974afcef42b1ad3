// Internal launcher declarations (host side). Every launcher enqueues on `st`, never synchronises,
// never allocates, and returns 0 or a negative S4G_ERR_* code.
#pragma once
#include "../../include/ssl4gie_b200.h"
#include "common.cuh"

struct ConvGeom {
  int Nimg, Hi, Wi, Cin, Ho, Wo, Cout, R, S, stride, pad;
  int ldx;  // channel stride of X rows (elements), normally Cin
  int ldy;  // channel stride of dY rows (wgrad), normally Cout
};
static inline ConvGeom linear_geom(long long M, int N, int K) {
  ConvGeom g{(int)M, 1, 1, K, 1, 1, N, 1, 1, 1, 0, K, N};
  return g;
}
static inline ConvGeom conv3_geom(int B, int H, int W, int Cin, int Cout) {
  ConvGeom g{B, H, W, Cin, H, W, Cout, 3, 3, 1, 1, Cin, Cout};
  return g;
}

// ---- deterministic cross-block reductions: reduce.cu
// A producer kernel writes one partial row per block into scratch; finalize sums the rows in a fixed order.
struct FinalizeJob { const void* part; void* out; int nrows, ncols, ld, in_f64, out_f64, block0; };
constexpr int MAX_FIN_JOBS = 64;
struct FinalizeJobs { FinalizeJob job[MAX_FIN_JOBS]; int n, total_blocks, launches; };
struct Reducer {
  void* scratch = nullptr;        // where the partial rows go; null: the library's per-device scratch (s4g_init)
  size_t scratch_bytes = 0;
  FinalizeJobs* defer = nullptr;  // non-null (with own scratch): queue the combine step instead of launching it
};
int reduce_init();
void* reduce_scratch(const Reducer* r, size_t bytes);  // null + error set when too small / not initialised
// out[c] = sum_{row < nrows} part[row * ld + c], c < ncols (part: float or double rows; out: float or double)
int k_finalize(const Reducer* r, const void* part, void* out, int nrows, int ncols, int ld, int in_f64, int out_f64, cudaStream_t st);
int k_finalize_flush(FinalizeJobs* jobs, cudaStream_t st);
// scratch a producer needs (upper bounds, bytes)
static inline size_t ln_bwd_scratch_bytes(int C) { return (size_t)148 * 3 * C * sizeof(float); }
static inline size_t colsum_scratch_bytes(int N) { return (size_t)((148 * 4) / ((N + 255) / 256) + 2) * N * sizeof(float); }   // k_colsum / k_dgelu_mul row blocks
static inline size_t bn_scratch_bytes(int C) { return (size_t)(148 * 4 + 8) * 2 * C * sizeof(double); }
static inline size_t seghead_scratch_bytes() { return (size_t)(148 * 4) * 776 * sizeof(double); }
static inline size_t depthhead_scratch_bytes() { return (size_t)(148 * 4) * 33 * sizeof(float); }

// ---- validation (SIMT) contractions: simt.cu
template <typename T> int simt_conv_fprop(const ConvGeom& g, const T* X, const T* W, const EpiParams& e, cudaStream_t st);
template <typename T> int simt_conv_wgrad(const ConvGeom& g, const T* dY, const T* X, float* dW, cudaStream_t st);

// ---- tcgen05 contractions (bf16 in, fp32 accumulate in TMEM): umma.cu
int umma_conv_fprop(const ConvGeom& g, const bf16* X, const bf16* W, const EpiParams& e, cudaStream_t st);
int umma_fprop_stat_rows(const ConvGeom& g);   // rows of EpiParams::stat_part that call writes (its grid size)
// dW layout: packed [Cout][tap][Cin] (torch_layout = 0) or torch's [Cout][Cin][R][S] (torch_layout = 1)
int umma_conv_wgrad(const ConvGeom& g, const bf16* dY, const bf16* X, float* dW, float* scratch, size_t scratch_bytes,
                    cudaStream_t st, int torch_layout = 0);
size_t umma_wgrad_scratch_bytes(const ConvGeom& g);
int umma_init();
int umma_make_map_bf16(void* cu_tensor_map, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box);
int umma_num_sms();

// ---- elementwise / normalisation / resampling: elementwise.cu
template <typename T> int k_patchify(const float* img, T* out, int B, cudaStream_t st);
int k_assemble_tokens(const float* pe, const float* cls, const float* pos, float* x, int B, cudaStream_t st);
int k_assemble_tokens_bwd(const float* dx, float* dpe, float* dcls, float* dpos, int B, cudaStream_t st);
template <typename T>
int k_layernorm_fwd(const float* x, const float* gamma, const float* beta, T* y, float* mean, float* rstd, int rows,
                    int C, float eps, cudaStream_t st);
// dx = (dres ? dres : 0) + LN-backward(dy); dgamma / dbeta are overwritten (deterministic two-stage reduction)
template <typename T>
int k_layernorm_bwd(const T* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                    const float* dres, float* dx, float* dgamma, float* dbeta, int rows, int C, cudaStream_t st,
                    T* dx_cast = nullptr, float* dxsum = nullptr, const Reducer* red = nullptr);
// dx_cast: optional T copy of dx (the next GEMM's operand); dxsum: optional [C] = column sums of dx
template <typename T> int k_dgelu_mul(const T* dy, const T* u, T* out, long long n, cudaStream_t st, float* colsum = nullptr, int N = 0, const Reducer* red = nullptr);
template <typename T> int k_gelu(const T* u, T* out, long long n, cudaStream_t st, T* gprime = nullptr);   // gprime (may alias u) = gelu'(u)
template <typename T> int k_colsum(const T* dy, float* out, long long M, int N, int ld, cudaStream_t st, const Reducer* red = nullptr);  // out overwritten
template <typename T> int k_tokens_to_nhwc(const float* x, T* out, int B, cudaStream_t st);                // drop cls, cast
template <typename T> int k_tap_grad(const T* dtap, float* dx, int B, int accumulate, cudaStream_t st, T* cast_out = nullptr);  // inverse of above (+ optional T copy of the result)
// GPU input pipeline (augment.cu; Binary_segmentation/Data/dataset.py:34-64, dataloaders.py:62-76)
size_t aug_resize_scratch_bytes(const long long* desc_host, int B, int C);
int k_aug_resize(const uint8_t* src, const long long* desc_dev, const long long* desc_host, int B, int C, void* scratch, size_t scratch_bytes, uint8_t* out, cudaStream_t st);
int k_aug_color_jitter(uint8_t* img, const void* params, int B, cudaStream_t st);
int k_aug_blur_normalize(const uint8_t* img, const float* sigma, float* out, const float* mean3, const float* std3, int B, cudaStream_t st);
int k_aug_mask_to_tensor(const uint8_t* img, float* out, int B, int C, cudaStream_t st);
int k_aug_depth_target(const void* src, const long long* desc_dev, float* out, int B, cudaStream_t st);
int k_aug_geometric(const float* x, const float* y, float* xo, float* yo, const void* params, float fill_x, float fill_y, int B, cudaStream_t st);
// MAE pretraining step (mae.cu; Models/mae/models_mae.py:123-214)
int k_mae_masking(const float* noise, int* ids_restore, int* ids_keep, float* mask, int B, int L, int keep, cudaStream_t st);
template <typename T> int k_mae_patchify_kept(const float* img, const int* ids_keep, T* out, int B, int keep, cudaStream_t st);
int k_mae_assemble(const float* pe, const float* cls, const float* pos, const int* ids_keep, float* x, int B, int keep, int C, cudaStream_t st);
template <typename T> int k_mae_assemble_bwd(const float* dx, T* dpe, float* dcls, int B, int keep, int C, cudaStream_t st);
int k_mae_unshuffle(const float* de, const float* mask_token, const float* dpos, const int* ids_restore, float* y, int B, int L, int keep, int C, cudaStream_t st);
size_t mae_unshuffle_scratch_bytes(int B, int C);
template <typename T>
int k_mae_unshuffle_bwd(const float* dy, const int* ids_restore, T* dde, float* dmask_token, int B, int L, int keep, int C, cudaStream_t st, const Reducer* r);
// [rows, heads*32] <-> [rows, heads*64] (zero upper half); the first `scaled` heads are multiplied by `scale`
template <typename T> int k_head_pad(const T* in, T* out, long long rows, int heads, int scaled, float scale, cudaStream_t st);
template <typename T> int k_head_unpad(const T* in, T* out, long long rows, int heads, int scaled, float scale, cudaStream_t st);
int k_drop_cls(const float* full, float* out, int B, int L, int C, cudaStream_t st);
template <typename T> int k_add_cls_row(const float* dout, T* dfull, int B, int L, int C, cudaStream_t st);
size_t mae_loss_scratch_bytes(int B);
int k_mae_loss_fwd(const float* pred, const float* img, const float* mask, void* scratch, float* loss, int B, int norm_pix, cudaStream_t st);
int k_mae_loss_bwd(const float* pred, const float* img, const float* mask, const void* scratch, const float* grad_out, float* dpred, int B, int norm_pix, cudaStream_t st);
// classification tail (cls_head.cu): token pooling + lin_head on fp32 master weights; W == nullptr: no head (out = pooled)
template <typename T>
int k_cls_head_fwd(const T* hn, const float* W, const float* bias, float* pooled, float* out, int B, int N, int C,
                   int spatial, cudaStream_t st);
template <typename T>
int k_cls_head_bwd(const float* dout, const float* pooled, const float* W, float* dW, float* db, T* dhn, int B, int N,
                   int C, int spatial, cudaStream_t st);
// grouped weight packing: every job is one [N, K] fp32 weight -> bf16/fp32 copy wp [N, K] and (optional) transpose wt [K, N]
constexpr int PACK_TILE = 64;   // k_pack_linear_grouped works in PACK_TILE x PACK_TILE tiles: tile0 counts those
struct PackLinearJob { const float* src; void* wp; void* wt; int N, K, tile0, pad; };
struct PackLinearJobs { PackLinearJob job[80]; int n; };
template <typename T> int k_pack_linear_grouped(const PackLinearJobs& jobs, int total_tiles, cudaStream_t st);
// 3x3 conv weights [Co][Ci][3][3] -> wf [Co][tap][Ci] and wd [Ci][8 - tap][Co]; job: N = Co, K = Ci (<= 768)
template <typename T> int k_pack_conv3_grouped(const PackLinearJobs& jobs, int total_blocks, cudaStream_t st);
template <typename T> int k_cast_f32(const float* in, T* out, long long n, cudaStream_t st);
template <typename T> int k_relu(const T* x, T* y, long long n, cudaStream_t st);
// dx = (dskip ? dskip : 0) + (x > 0 ? dr : 0)
template <typename T> int k_relu_bwd_add(const T* dr, const T* x, const T* dskip, T* dx, long long n, cudaStream_t st);
template <typename T> int k_add(const T* a, const T* b, T* y, long long n, cudaStream_t st);
template <typename T>
int k_pixel_shuffle(const T* ycol, const float* bias, T* out, int B, int H, int W, int C, int k, cudaStream_t st);
template <typename T>
int k_pixel_unshuffle(const T* dout, T* dycol, int B, int H, int W, int C, int k, cudaStream_t st);
template <typename T> int k_im2col_s2(const T* x, T* col, int B, int H, int W, int C, cudaStream_t st);    // 3x3 s2 p1
template <typename T> int k_col2im_s2(const T* dcol, T* dx, int B, int H, int W, int C, cudaStream_t st);
template <typename T> int k_bilinear2x_fwd(const T* x, T* y, int B, int H, int W, int C, cudaStream_t st);
template <typename T> int k_bilinear2x_bwd(const T* dy, T* dx, int B, int H, int W, int C, cudaStream_t st);

// BatchNorm2d(train): stats[0..C) = sum, [C..2C) = sumsq as double (overwritten)
template <typename T> int k_bn_stats(const T* x, double* stats, long long M, int C, cudaStream_t st, const Reducer* red = nullptr);
// mean/rstd from (possibly all-reduced) sums over `count` rows; running stats update like torch
int k_bn_finalize(const double* stats, double count, float* mean, float* rstd, float* running_mean, float* running_var,
                  long long* num_batches_tracked, float momentum, float eps, int C, cudaStream_t st);
int k_bn_eval_prepare(const float* running_mean, const float* running_var, float* mean, float* rstd, float eps, int C,
                      cudaStream_t st);
// y = bn(x) [relu] [+ skip]
template <typename T>
int k_bn_apply(const T* x, const float* mean, const float* rstd, const float* gamma, const float* beta, const T* skip,
               T* y, int relu, long long M, int C, cudaStream_t st);
// sums[0..C) = sum(dy'), [C..2C) = sum(dy' * xhat) where dy' = relu ? dy*(bn(x)>0) : dy   (double, overwritten)
template <typename T>
int k_bn_bwd_reduce(const T* dy, const T* x, const float* mean, const float* rstd, const float* gamma,
                    const float* beta, double* sums, int relu, long long M, int C, cudaStream_t st, const Reducer* red = nullptr);
template <typename T>
int k_bn_bwd_apply(const T* dy, const T* x, const float* mean, const float* rstd, const float* gamma, const float* beta,
                   const double* sums, double count, T* dx, float* dgamma, float* dbeta, int relu, long long M, int C,
                   cudaStream_t st);

// seg head tail: BN -> ReLU -> Dropout(p) -> 1x1 (C->1) + b  => small [M] fp32
template <typename T>
int k_seghead_fwd(const T* x, const float* mean, const float* rstd, const float* gamma, const float* beta,
                  const float* w, const float* b, const uint8_t* mask, float p, uint64_t seed, float* small,
                  long long M, int C, cudaStream_t st, const unsigned long long* seed_dev = nullptr);   // seed_dev: device-resident seed, overrides `seed`
// pass 1: reductions (BN sums + dw[C] + db) ; pass 2: dx
template <typename T>
int k_seghead_bwd_reduce(const float* dsmall, const T* x, const float* mean, const float* rstd, const float* gamma,
                         const float* beta, const float* w, const uint8_t* mask, float p, uint64_t seed, double* sums,
                         float* dw, float* db, long long M, int C, cudaStream_t st, const Reducer* red = nullptr,
                         const unsigned long long* seed_dev = nullptr);
template <typename T>
int k_seghead_bwd_apply(const float* dsmall, const T* x, const float* mean, const float* rstd, const float* gamma,
                        const float* beta, const float* w, const uint8_t* mask, float p, uint64_t seed,
                        const double* sums, double count, T* dx, float* dgamma, float* dbeta, long long M, int C,
                        cudaStream_t st, const unsigned long long* seed_dev = nullptr);
// depth head tail: (already ReLU'd) h[M,C] . w + b -> sigmoid => out [M] fp32 ; bwd gives dh and dw/db
template <typename T>
int k_depthhead_fwd(const T* h, const float* w, const float* b, float* out, long long M, int C, cudaStream_t st);
template <typename T>
int k_depthhead_bwd(const float* dout, const float* out, const T* h, const float* w, T* dh, float* dw, float* db,
                    long long M, int C, cudaStream_t st, const Reducer* red = nullptr);
// single-channel fp32 bilinear x2 (seg logits)
int k_bilinear2x_f32_fwd(const float* x, float* y, int B, int H, int W, cudaStream_t st);
int k_bilinear2x_f32_bwd(const float* dy, float* dx, int B, int H, int W, cudaStream_t st);

// ---- weight packing (fp32 torch layout -> T kernel layout): elementwise.cu
template <typename T> int k_pack_linear(const float* w, T* wp, T* wt, int N, int K, cudaStream_t st);  // wp[N,K], wt[K,N]
template <typename T> int k_pack_conv(const float* w, T* wf, T* wd, int Co, int Ci, int R, cudaStream_t st);
template <typename T> int k_pack_convT(const float* w, T* wf, T* wd, int Ci, int Co, int k, cudaStream_t st);
int k_unpack_conv_grad(const float* dwp, float* dw, int Co, int Ci, int R, cudaStream_t st);
int k_unpack_convT_grad(const float* dwp, float* dw, int Ci, int Co, int k, cudaStream_t st);

// ---- attention: attention.cu   qkv [B*197, 2304] (q|k|v, head-major inside), o [B*197, 768]
template <typename T> int attn_fwd_simt(const T* qkv, T* o, float* lse, int B, int Ntok, int H, cudaStream_t st);
template <typename T>
int attn_bwd_simt(const T* qkv, const T* o, const T* dout, const float* lse, T* dqkv, int B, int Ntok, int H,
                  cudaStream_t st);
// tcgen05 attention (attention_umma.cu): S and P in tensor memory, Q/K/V by TMA; head dim 64, Ntok <= 208
int attn_fwd_umma(const bf16* qkv, bf16* o, float* lse, int B, int Ntok, int H, cudaStream_t st);
// delta (optional): precomputed rowsum(dO o O) [B, H, Ntok]; null = computed by a pass over o and dout
int attn_bwd_umma(const bf16* qkv, const bf16* o, const bf16* dout, const float* lse, bf16* dqkv, int B, int Ntok, int H,
                  cudaStream_t st, const float* delta = nullptr);
bool attn_uses_umma();   // false under S4G_ATTN=mma
int attn_fwd_mma(const bf16* qkv, bf16* o, float* lse, int B, int Ntok, int H, cudaStream_t st);
int attn_bwd_mma(const bf16* qkv, const bf16* o, const bf16* dout, const float* lse, bf16* dqkv, int B, int Ntok, int H,
                 cudaStream_t st, const float* delta = nullptr);

// ---- losses: loss.cu   (scratch holds the per-block partial rows and the combined sums; nothing to zero)
size_t dice_scratch_doubles(int B);
int k_dice_fwd(const float* logits, const float* targets, float* sums, float* loss, int B, int HW, float smooth,
               cudaStream_t st);
int k_dice_bwd(const float* logits, const float* targets, const float* sums, const float* gout, float* dlogits, int B,
               int HW, float smooth, cudaStream_t st);
int k_ssi_fwd(const float* pred, const float* targ, float* scratch, float* loss, float* ssi_out, int B, int H, int W,
              float alpha, int scales, int image_based, cudaStream_t st);
int k_ce_fwd(const float* logits, const long long* labels, const float* weight, float* loss, float* dlogits_unit, int B, int N,
             cudaStream_t st);
int k_ce_bwd(const float* dlogits_unit, const float* gout, float* dlogits, int B, int N, cudaStream_t st);
int k_ssi_bwd(const float* pred, const float* targ, const float* scratch, const float* gout, float* dpred, int B, int H,
              int W, float alpha, int scales, cudaStream_t st);
size_t ssi_scratch_floats(int B);

// ---- optimizer: optim.cu
int k_adamw(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2, float eps, float wd,
            int step, float grad_scale, cudaStream_t st, const float* wd_slot = nullptr);
int k_adamw_dev(float* p, const float* g, float* m, float* v, long long n, const float* hyper, const int* step_dev, const float* wd_slot, cudaStream_t st);
int k_device_hyper_set(float* hyper, float lr, float b1, float b2, float eps, float wd, float gs, cudaStream_t st);
int k_device_counter_advance(int* step_dev, unsigned long long* seed_dev, cudaStream_t st);

// peer.cu: SyncBN statistic exchange over NVLink peer memory
size_t k_peer_mailbox_bytes();
size_t k_peer_ctl_offset();
int k_peer_sum_f64(double* data, int count, void* const* boxes, int rank, int world, cudaStream_t st);
