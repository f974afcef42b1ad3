/* libssl4gie_b200 — C ABI of the B200-native (sm_100a) ViT-B/16 + DPT training hot path.
 *
 * The reference (ESandML/SSL4GIE) has no FFI/plugin interface: the path sits behind ordinary PyTorch
 * nn.Modules.  This header is therefore the boundary we define UNDER those modules; each entry
 * point names the reference code it replaces (paths relative to the reference repository root).
 * The Python side (ssl4gie_b200/_C.py, ctypes) passes tensor.data_ptr() and
 * torch.cuda.current_stream().cuda_stream; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless stated otherwise
 *   - every call enqueues on `stream` (a cudaStream_t passed as void*), never synchronises, never allocates
 *     device memory, never frees caller memory; the caller owns every buffer
 *   - return 0 on success, negative S4G_ERR_* otherwise; message via s4g_last_error() (thread-local)
 *   - token matrices are row-major [rows, C]; feature maps are NHWC; dtype 0 = fp32 validation mode,
 *     dtype 1 = bf16 activations / fp32 accumulate (parameters, gradients and the residual stream are fp32 in both)
 */
#ifndef SSL4GIE_B200_H
#define SSL4GIE_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define S4G_OK 0
#define S4G_ERR_SHAPE (-1)
#define S4G_ERR_ALIGN (-2)
#define S4G_ERR_WORKSPACE (-3)
#define S4G_ERR_ARCH (-4)
#define S4G_ERR_CUDA (-5)
#define S4G_ERR_STATE (-6)
#define S4G_ERR_NAME (-7)

#define S4G_DTYPE_F32 0
#define S4G_DTYPE_BF16 1

#define S4G_DENSE_NONE 0  /* classification branch, Models/models.py:348-354 */
#define S4G_DENSE_SEG 1   /* dense="seg"   */
#define S4G_DENSE_DEPTH 2 /* dense="depth" */
#define S4G_DENSE_MAE 3   /* MaskedAutoencoderViT (mae_vit_base_patch16): out = pred [B,196,768], Models/mae/models_mae.py:150-196 */

#define S4G_BACKEND_SIMT 0    /* FFMA kernels: fp32 validation mode / cross-check */
#define S4G_BACKEND_TENSOR 1  /* tcgen05 + TMEM + TMA contractions, tensor-core attention (bf16 only) */

int s4g_version(void);
const char* s4g_last_error(void);
/* Checks the device is sm_100 and prepares per-device state. */
int s4g_init(int device);

/* ------------------------------------------------------------------------------------------------------------
 * Whole-model schedule: replaces VisionTransformer_from_Any.forward / ViT_from_MAE.forward /
 * ViT_from_MoCoV3.forward (Models/models.py:340-357, :458-475, :560-578), DPT_decoder.forward
 * (Models/DPT_decoder.py:501-539) and their autograd backward (train_segmentation.py:45-47).
 * ---------------------------------------------------------------------------------------------------------- */
typedef struct S4GModel S4GModel;

typedef struct {
  int batch;          /* images per call on this GPU */
  int dense;          /* S4G_DENSE_* */
  int dtype;          /* S4G_DTYPE_* */
  int backend;        /* S4G_BACKEND_* */
  int frozen;         /* encoder under no_grad (Models/models.py:341-343) */
  int world_size;     /* ranks sharing BatchNorm statistics (SyncBatchNorm, train_segmentation.py:199); 1 = local BN */
  int num_classes;    /* lin_head width for S4G_DENSE_NONE (0 = no head) */
  int out_token;      /* 0 = "cls", 1 = "spatial" (Models/models.py:349-352) */
  float dropout_p;    /* seg head nn.Dropout(0.1) (Models/DPT_decoder.py:494) */
} S4GModelDesc;

int s4g_model_create(const S4GModelDesc* desc, S4GModel** out);
void s4g_model_destroy(S4GModel* m);
/* Bind one state_dict entry by its reference name (e.g. "blocks.0.attn.qkv.weight",
 * "decoder.refinenet1.resConfUnit2.bn1.running_mean").  `grad` may be NULL (buffers, frozen parameters).
 * All float entries are fp32; num_batches_tracked is int64.  May be called again to re-point storage. */
int s4g_model_bind(S4GModel* m, const char* name, void* data, void* grad, long long numel);
/* Optional: tells the library that all bound gradients live in [grads, grads + bytes) so backward can clear them
 * with one memset. */
int s4g_model_set_grad_arena(S4GModel* m, void* grads, size_t bytes);
size_t s4g_model_workspace_bytes(S4GModel* m);
int s4g_model_set_workspace(S4GModel* m, void* workspace, size_t bytes);
/* images [B,3,224,224] fp32 NCHW -> out: seg logits / depth sigmoid [B,1,224,224] fp32, or [B,num_classes].
 * training != 0: batch statistics + running-stat update + dropout(seed) and activations kept for backward.
 * dropout_mask: optional [B*112*112*256] bytes (NHWC) overriding the RNG (parity tests); NULL = RNG. */
int s4g_model_forward(S4GModel* m, const float* images, float* out, int training, uint64_t seed,
                      const uint8_t* dropout_mask, void* stream);
/* S4G_DENSE_MAE only, before s4g_model_forward: noise [B,196] = the uniform noise random_masking draws
 * (Models/mae/models_mae.py:132; smallest 49 per image are kept); mask_out [B,196] float (0 keep, 1 remove, :141-145) and
 * ids_restore_out [B,196] int32 (:136) are optional.  The pointers are read / written by every later forward. */
int s4g_model_set_mae_io(S4GModel* m, const float* noise, float* mask_out, int* ids_restore_out);
/* CUDA-graph support: read the seg head's dropout seed (Models/DPT_decoder.py:494) from device memory instead of the by-value
 * `seed` of s4g_model_forward, which a captured graph would replay unchanged.  NULL restores the by-value seed. */
int s4g_model_set_seed_device(S4GModel* m, const unsigned long long* seed_dev);
/* dout: gradient w.r.t. `out`.  Writes (overwrites) the gradient of every bound parameter that has one. */
int s4g_model_backward(S4GModel* m, const float* dout, void* stream);
/* Cross-rank sum used by SyncBatchNorm statistics (forward: sum/sumsq, backward: sum dy / sum dy*xhat).  Called
 * on the host while enqueueing; must enqueue an in-place sum-all-reduce of `count` doubles on `stream`. */
typedef int (*s4g_allreduce_fn)(void* user, void* device_doubles, int count, void* stream);
int s4g_model_set_allreduce(S4GModel* m, s4g_allreduce_fn fn, void* user);
/* The same cross-rank sum without the host callback: a one-CTA kernel that exchanges the statistic vector through
 * NVLink peer memory (P2P stores into every rank's mailbox + system-scope flags; csrc/peer.cu).  Replaces the NCCL
 * all_gather / all_reduce pair inside torch.nn.SyncBatchNorm (train_segmentation.py:199).  Every rank allocates ONE
 * mailbox per process group, shared by all of its model handles (s4g_peer_alloc: zero-filled device memory + a 64-byte cudaIpc handle to send to the peers
 * by any host channel), opens the peers' (s4g_peer_open), and passes the `world` device pointers in rank order
 * (mailboxes[rank] = its own).  All ranks must have set the pointers (host barrier) before the first forward. */
#define S4G_PEER_MAX_WORLD 8
#define S4G_PEER_HANDLE_BYTES 64
size_t s4g_peer_mailbox_bytes(void);
int s4g_peer_alloc(size_t bytes, void** dptr, unsigned char handle[S4G_PEER_HANDLE_BYTES]);
int s4g_peer_open(const unsigned char handle[S4G_PEER_HANDLE_BYTES], void** dptr);
int s4g_peer_close(void* dptr);
int s4g_peer_free(void* dptr);
int s4g_model_set_peer_exchange(S4GModel* m, void* const* mailboxes, int rank, int world);
/* Stand-alone exchange (tests): in-place sum over ranks of `count` <= 512 doubles.  The exchange counter lives in the
 * rank's own mailbox (device memory), so every handle sharing the mailboxes stays in step and the launch is
 * CUDA-graph replayable. */
int s4g_peer_sum_f64(double* data, int count, void* const* mailboxes, int rank, int world, void* stream);
/* Synchronises the device and reports a peer that did not answer within S4G_PEER_TIMEOUT_S (default 600 s):
 * 0 = healthy, S4G_ERR_STATE + s4g_last_error() otherwise (the exchange kernel flags the time-out, it never traps). */
int s4g_peer_check(void* own_mailbox);
/* Gradient-bucket boundaries for overlapping the data-parallel all-reduce with backward (DDP reducer,
 * train_segmentation.py:203).  The library records events[i] (cudaEvent_t) on `stream` during backward once every
 * parameter gradient of bucket i is final: bucket 0 = DPT decoder, buckets 1..n-1 = encoder blocks from the last
 * to the first, split evenly, the final bucket also covering patch_embed / cls_token / pos_embed. */
int s4g_model_set_bucket_events(S4GModel* m, void** events, int n);
/* Name list of the parameters whose gradients belong to bucket i, as a '\n'-separated string (host memory owned by
 * the library, valid until the next call). */
const char* s4g_model_bucket_params(S4GModel* m, int bucket, int n_buckets);
/* number of kernels the last forward / backward call enqueued */
int s4g_model_last_launches(S4GModel* m);

/* ------------------------------------------------------------------------------------------------------------
 * Operator-level entry points (used by the parity tests and by the loss / optimizer modules)
 * ---------------------------------------------------------------------------------------------------------- */
/* nn.Linear / 1x1 Conv2d: out[M,N] = epilogue(A[M,K] . W[N,K]^T + bias).  act: 0 none, 1 GELU(erf) (timm Mlp),
 * 3 ReLU.  residual: optional fp32 [M,N] added after the activation (timm Block residual).  A, W are `dtype`;
 * out is fp32 if out_f32 else `dtype`. */
int s4g_linear_fwd(const void* A, const void* W, const float* bias, const float* residual, void* out, long long M,
                   int N, int K, int act, int out_f32, int dtype, int backend, void* stream);
/* dW[N,K] (fp32) = dY[M,N]^T . X[M,K] */
int s4g_linear_wgrad(const void* dY, const void* X, float* dW, long long M, int N, int K, int dtype, int backend,
                     void* scratch, size_t scratch_bytes, void* stream);
/* nn.Conv2d 3x3 stride 1 pad 1 on NHWC: X [B,H,W,Cin], Wp [Cout][3*3][Cin] (packed), out [B,H,W,Cout] */
int s4g_conv3x3_fwd(const void* X, const void* Wp, const float* bias, void* out, int B, int H, int W, int Cin,
                    int Cout, int dtype, int backend, void* stream);
/* dWp [Cout][9][Cin] fp32 */
int s4g_conv3x3_wgrad(const void* dY, const void* X, float* dWp, int B, int H, int W, int Cin, int Cout, int dtype,
                      int backend, void* scratch, size_t scratch_bytes, void* stream);
size_t s4g_wgrad_scratch_bytes(int B, int H, int W, int Cin, int Cout, int taps);
/* nn.LayerNorm(C, eps) on fp32 rows -> `dtype` output; saves mean/rstd */
int s4g_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                      int rows, int C, float eps, int dtype, void* stream);
int s4g_layernorm_bwd(const void* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                      const float* dres, float* dx, float* dgamma, float* dbeta, int rows, int C, int dtype,
                      void* stream);
/* Allocation-free variant of s4g_layernorm_bwd (same arguments, same results).  It has no counterpart in the reference: it
 * was added for callers that cannot call cudaMalloc (CUDA graph capture, serving loops,
 * harnesses that forbid allocation).  The partial rows of the dgamma / dbeta reduction go into a file-scope __device__ array of
 * 1 363 968 bytes (148 blocks x 3 x 768 floats) that every GPU loading the library holds until exit, so neither the call nor
 * any set-up it needs (s4g_init is not required) allocates.  Limit: C <= 768 (kernels exist for C = 512, 768), any row count;
 * larger shapes return S4G_ERR_WORKSPACE.  All calls on a device share the array (it keeps the last call's rows; nothing has to
 * start at zero): calls that overlap on the GPU -- different streams without an ordering between them -- race. */
int s4g_layernorm_bwd_static(const void* dy, const float* x, const float* gamma, const float* mean, const float* rstd,
                             const float* dres, float* dx, float* dgamma, float* dbeta, int rows, int C, int dtype,
                             void* stream);
/* timm Attention core: qkv [B*N, 3*H*64] -> o [B*N, H*64], lse [B,H,N] */
int s4g_attention_fwd(const void* qkv, void* o, float* lse, int B, int N, int H, int dtype, int backend, void* stream);
int s4g_attention_bwd(const void* qkv, const void* o, const void* dout, const float* lse, void* dqkv, int B, int N,
                      int H, int dtype, int backend, void* stream);
/* F.interpolate(scale_factor=2, mode="bilinear", align_corners=True) on NHWC (Models/DPT_decoder.py:295-297) */
int s4g_bilinear2x_fwd(const void* x, void* y, int B, int H, int W, int C, int dtype, void* stream);
int s4g_bilinear2x_bwd(const void* dy, void* dx, int B, int H, int W, int C, int dtype, void* stream);

/* SoftDiceLoss.forward (Binary_segmentation/Metrics/losses.py:10-24) and its gradient.
 * scratch: >= s4g_dice_scratch_bytes(B) bytes, kept between fwd and bwd. */
size_t s4g_dice_scratch_bytes(int B);
int s4g_dice_fwd(const float* logits, const float* targets, void* scratch, float* loss, int B, int HW, float smooth,
                 void* stream);
int s4g_dice_bwd(const float* logits, const float* targets, const void* scratch, const float* grad_out,
                 float* dlogits, int B, int HW, float smooth, void* stream);
/* ScaleAndShiftInvariantLoss.forward (Depth_estimation/Metrics/losses.py:130-146); image_based = 0:
 * reduction_batch_based (:28-37), else reduction_image_based (:40-48); scales <= 8.
 * ssi_out: optional [B,H,W] prediction_ssi (losses.py:138,148-151).  All sums are two-stage, fixed-order (no atomics). */
size_t s4g_ssi_scratch_bytes(int B);
int s4g_ssi_fwd(const float* pred, const float* target, void* scratch, float* loss, float* ssi_out, int B, int H,
                int W, float alpha, int scales, int image_based, void* stream);
int s4g_ssi_bwd(const float* pred, const float* target, void* scratch, const float* grad_out, float* dpred, int B,
                int H, int W, float alpha, int scales, void* stream);
/* MaskedAutoencoderViT.forward_loss (Models/mae/models_mae.py:198-214): target = patchify(images) (:95-107), per-patch
 * (t - mean) / sqrt(var + 1e-6) with the unbiased variance when norm_pix != 0, loss = sum(mean_e (pred - t)^2 * mask) /
 * sum(mask).  pred / dpred [B,196,768], images [B,3,224,224], mask [B,196]; scratch kept between fwd and bwd. */
size_t s4g_mae_loss_scratch_bytes(int B);
int s4g_mae_loss_fwd(const float* pred, const float* images, const float* mask, void* scratch, float* loss, int B,
                     int norm_pix, void* stream);
int s4g_mae_loss_bwd(const float* pred, const float* images, const float* mask, const void* scratch,
                     const float* grad_out, float* dpred, int B, int norm_pix, void* stream);
/* nn.CrossEntropyLoss(weight) of the classification fine-tune (Classification/train_classification.py:278):
 * loss = sum_i w[y_i] * -log softmax(x_i)[y_i] / sum_i w[y_i].  labels int64 [B]; class_weight [num_classes] or NULL.
 * dlogits_unit [B,num_classes] receives dloss/dlogits for grad_out = 1 (kept between fwd and bwd). */
int s4g_ce_fwd(const float* logits, const long long* labels, const float* class_weight, float* loss, float* dlogits_unit,
               int B, int num_classes, void* stream);
int s4g_ce_bwd(const float* dlogits_unit, const float* grad_out, float* dlogits, int B, int num_classes, void* stream);

/* The same step with its scalars in device memory, for capture in a CUDA graph (train_segmentation.py:204 + the per-iteration LR
 * of Models/mae/engine_pretrain.py:42-43): hyper_dev[8] = {lr, beta1, beta2, eps, weight_decay, grad_scale, -, -} is written by
 * s4g_device_hyper_set (a 1-thread kernel with by-value arguments: stream-ordered, no host buffer to race with); step_dev counts
 * completed steps (the kernel applies step_dev + 1) and is advanced, together with an optional dropout seed, by
 * s4g_device_counter_advance.  weight_decay_per_slot as in s4g_adamw_step_groups, or NULL. */
int s4g_adamw_step_device(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n,
                          const float* hyper_dev, const int* step_dev, const float* weight_decay_per_slot, void* stream);
int s4g_device_hyper_set(float* hyper_dev, float lr, float beta1, float beta2, float eps, float weight_decay,
                         float grad_scale, void* stream);
int s4g_device_counter_advance(int* step_dev, unsigned long long* seed_dev, void* stream);

/* ---- GPU-side input pipeline: the per-sample Pillow / torchvision work of the reference's DataLoader workers
 * (Binary_segmentation/Data/dataset.py:34-64, Data/dataloaders.py:62-76) on a batch of decoded 8-bit images.  Every entry
 * point works on B images of 224 x 224 except s4g_aug_resize, which produces them.
 * s4g_aug_resize: Image.resize((224, 224)) (Pillow 8-bit bicubic, bit-exact).  src = packed images [h][w][C]; desc [B][5] =
 *   {byte offset into src, height, width, channels (== C, 1 or 3), byte offset of this image's [h][224][C] intermediate inside
 *   the scratch (caller-chosen, non-overlapping)} given both as a device and a host array; out [B,224,224,C].  No host
 *   synchronisation. */
size_t s4g_aug_resize_scratch_bytes(const long long* desc_host, int B, int C);
int s4g_aug_resize(const unsigned char* src, const long long* desc_dev, const long long* desc_host, int B, int C,
                   void* scratch, size_t scratch_bytes, unsigned char* out, void* stream);
/* transforms.ColorJitter on 8-bit RGB [B,224,224,3] in place (ImageEnhance blends, Pillow RGB<->HSV; bit-exact).
 * params (device) [B] of {int order[4] (permutation of 0 brightness, 1 contrast, 2 saturation, 3 hue); int hue_shift
 * (= uint8(hue_factor * 255)); float factor[3] (brightness, contrast, saturation)}. */
int s4g_aug_color_jitter(unsigned char* img, const void* params, int B, void* stream);
/* transforms.GaussianBlur((25, 25), sigma[b]) + ToTensor + Normalize(mean, std): [B,224,224,3] u8 -> [B,3,224,224] f32.
 * sigma (device, [B]) may be NULL or <= 0 per image: no blur (the evaluation transform, dataloaders.py:104-111). */
int s4g_aug_blur_normalize(const unsigned char* img, const float* sigma, float* out, const float* mean3_host,
                           const float* std3_host, int B, void* stream);
/* ToTensor + Grayscale of the masks (dataloaders.py:72-74): [B,224,224,C] u8 -> [B,1,224,224] f32. */
int s4g_aug_mask_to_tensor(const unsigned char* img, float* out, int B, int C, void* stream);
/* Depth target (Depth_estimation/Data/dataset.py:47-53): packed 16-bit depth images (already padded to squares, make_square :9-16;
 * desc as for s4g_aug_resize, byte offsets even) -> nearest resize to 224 x 224 as the pinned Pillow 9.1.0 does for "I;16"
 * (source index floor((i + 0.5) * in / 224)) -> / 65535: out [B,1,224,224] f32. */
int s4g_aug_depth_target(const void* src, const long long* desc_dev, float* out, int B, void* stream);
/* TF.hflip, TF.vflip, TF.affine(interpolation=NEAREST, fill) applied jointly to x [B,3,224,224] and y [B,1,224,224] (y, yo
 * may be NULL) (dataset.py:46-63).  params (device) [B] of {int hflip, vflip, affine; float m[6]} with m = torchvision's
 * _get_inverse_affine_matrix((0, 0), angle, translate, scale, (shear, 0)). */
int s4g_aug_geometric(const float* x, const float* y, float* xo, float* yo, const void* params, float fill_x, float fill_y,
                      int B, void* stream);

/* torch.optim.AdamW step over a flat fp32 arena (train_segmentation.py:204). grad_scale multiplies g first
 * (1/world_size after a sum-all-reduce). step counts from 1. */
int s4g_adamw_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                   float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale, void* stream);
/* The same with parameter groups (timm add_weight_decay + AdamW(betas=(0.9, 0.95)), Models/mae/main_pretrain.py:179-180):
 * weight_decay_per_slot[n / 64] holds the weight decay of each 64-float arena slot (parameters start on slot boundaries). */
int s4g_adamw_step_groups(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                          float beta1, float beta2, float eps, const float* weight_decay_per_slot, int step,
                          float grad_scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif
