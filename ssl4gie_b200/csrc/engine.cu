// Static forward/backward schedule of ViT-B/16 (+ DPT decoder | classification head).
//
// The schedule is built once per (batch, dense, dtype, backend): every activation gets a fixed offset in one
// caller-owned workspace, every op becomes a closure that enqueues kernels on the caller's stream, and
// backward is the reverse list of the closures pushed alongside.  No allocation, no host sync, no autograd
// graph walk at run time => the whole step is CUDA-graph capturable.
//
// Reference semantics followed (paths relative to the reference repo):
//   encoder   Models/models.py:325-338 (== :427-456, :545-558) + timm 0.6.12 Block/Attention/Mlp
//   decoder   Models/DPT_decoder.py:501-539, RCU :212-233, fusion :281-301, heads :469-497
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "kernels.h"

namespace {

struct Param {
  float* p = nullptr;
  float* g = nullptr;
  long long n = 0;
};

struct Act {  // NHWC activation (or [rows, C] matrix with H = W = 1)
  size_t v = 0, g = 0;
  int B = 0, H = 0, W = 0, C = 0;
  long long rows() const { return (long long)B * H * W; }
  long long numel() const { return rows() * C; }
};

typedef std::function<int(cudaStream_t)> Op;

}  // namespace

struct S4GModel {
  S4GModelDesc d;
  std::map<std::string, Param> params;
  char* ws = nullptr;
  size_t ws_bytes = 0, ws_need = 0;
  bool built = false;
  std::vector<Op> pack, fwd, bwd_dec, bwd_enc;  // bwd lists are stored in forward order and run reversed
  std::vector<Act> acts;
  std::vector<std::pair<int, std::string>> bucket_of;  // (bwd position class, param name)
  void* grad_arena = nullptr;
  size_t grad_arena_bytes = 0;
  s4g_allreduce_fn allreduce = nullptr;
  void* allreduce_user = nullptr;
  void* peer_boxes[S4G_PEER_MAX_WORLD] = {};  // SyncBN exchange through NVLink peer memory (peer.cu) when peer_world > 1
  int peer_rank = 0, peer_world = 0;
  // BatchNorm accumulators of all layers live in two contiguous pools, zeroed by ONE memset at the top of forward
  // (sum / sumsq) and ONE at the top of backward (sum dy / sum dy*xhat) instead of one memset per layer
  size_t bn_stats_pool = 0, bn_sums_pool = 0;
  int bn_count = 0;
  // in-place sum over ranks of `count` doubles: the peer-memory kernel when mailboxes are set, else the host callback
  int exchange(double* v, int count, cudaStream_t st) {
    if (peer_world > 1) { launches++; return k_peer_sum_f64(v, count, peer_boxes, peer_rank, peer_world, st); }
    if (!allreduce) { s4g_set_error("SyncBatchNorm needs s4g_model_set_peer_exchange or s4g_model_set_allreduce"); return S4G_ERR_STATE; }
    return allreduce(allreduce_user, v, count, st);
  }
  std::vector<cudaEvent_t> events;
  // combine steps of the deterministic two-stage reductions (reduce.cu) queued by backward closures; run as one grouped
  // launch at every gradient-bucket boundary and at the end of backward
  FinalizeJobs pending{};
  int flush_pending(cudaStream_t st) {
    if (pending.n == 0) return 0;
    launches++;
    return k_finalize_flush(&pending, st);
  }
  std::string bucket_str;
  int launches = 0;
  // per-call state read by closures
  const float* images = nullptr;
  float* out = nullptr;
  const float* dout = nullptr;
  int training = 1;
  uint64_t seed = 0;
  const unsigned long long* seed_dev = nullptr;   // device-resident dropout seed (s4g_model_set_seed_device): overrides `seed`
  const uint8_t* dropout_mask = nullptr;
  // MAE pretraining step (S4G_DENSE_MAE): masking noise in, mask / ids_restore out (s4g_model_set_mae_io)
  const float* mae_noise = nullptr;
  float* mae_mask_out = nullptr;
  int* mae_ids_restore_out = nullptr;
  size_t mae_mask_off = 0, mae_ids_restore_off = 0;
  std::vector<std::string> enc_param_order;  // names in backward completion order (for buckets)
  std::vector<std::string> dec_param_names;
  std::vector<size_t> enc_block_mark;         // bwd_enc index at which block i's grads are complete

  template <typename U> U* at(size_t off) const { return reinterpret_cast<U*>(ws + off); }
};

namespace {

template <typename T>
struct Builder {
  S4GModel* m;
  size_t top = 0;
  const int B;
  const bool tensor;  // tcgen05 / tensor-core kernels
  size_t dwp_scratch = 0, dwp_bytes = 0;      // packed weight-gradient staging
  size_t wg_scratch = 0, wg_bytes = 0;        // split-K partials for the tcgen05 wgrad
  std::vector<Op>* bwd;                        // current backward list

  explicit Builder(S4GModel* mm) : m(mm), B(mm->d.batch), tensor(mm->d.backend == S4G_BACKEND_TENSOR) {}

  size_t alloc(size_t bytes) {
    size_t o = (top + 255) & ~(size_t)255;
    top = o + bytes;
    return o;
  }
  // scratch of one two-stage reduction call site (partial rows, reduce.cu).  Deferred sites own their slot: the rows must
  // survive until the grouped combine at the next bucket boundary.  Immediate sites share `shared_red`.
  struct RedSlot { size_t off = 0, bytes = 0; };
  RedSlot red_alloc(size_t bytes) { RedSlot r; r.off = alloc(bytes); r.bytes = bytes; return r; }
  Reducer red(const RedSlot& sl, bool defer) const {
    Reducer r;
    r.scratch = m->ws + sl.off; r.scratch_bytes = sl.bytes; r.defer = defer ? &m->pending : nullptr;
    return r;
  }
  RedSlot shared_red;
  int act(int Bn, int H, int W, int C, bool grad = true) {
    Act a;
    a.B = Bn; a.H = H; a.W = W; a.C = C;
    a.v = alloc(sizeof(T) * a.numel());
    if (grad) a.g = alloc(sizeof(T) * a.numel());
    m->acts.push_back(a);
    return (int)m->acts.size() - 1;
  }
  Param* P(const std::string& name) {
    auto it = m->params.find(name);
    if (it == m->params.end()) { missing.push_back(name); static Param dummy; return &dummy; }
    return &it->second;
  }
  Param* Popt(const std::string& name) {
    auto it = m->params.find(name);
    return it == m->params.end() ? nullptr : &it->second;
  }
  std::vector<std::string> missing;
  T* V(int a) const { return m->template at<T>(m->acts[a].v); }
  T* G(int a) const { return m->template at<T>(m->acts[a].g); }

  // ---- contraction dispatch -----------------------------------------------------------------------------
  int fprop(const ConvGeom& g, const T* X, const T* W, const EpiParams& e, cudaStream_t st) const {
    m->launches++;
    if constexpr (std::is_same<T, bf16>::value) {
      if (tensor) return umma_conv_fprop(g, X, W, e, st);
    }
    return simt_conv_fprop<T>(g, X, W, e, st);
  }
  // torch_layout (tcgen05 path only): write torch's [Cout][Cin][R][S] instead of the packed [Cout][tap][Cin]
  int wgrad(const ConvGeom& g, const T* dY, const T* X, float* dW, cudaStream_t st, int torch_layout = 0) const {
    m->launches++;
    if constexpr (std::is_same<T, bf16>::value) {
      if (tensor) {
        m->launches++;  // split-K reduce
        return umma_conv_wgrad(g, dY, X, dW, m->template at<float>(wg_scratch), wg_bytes, st, torch_layout);
      }
    }
    return simt_conv_wgrad<T>(g, dY, X, dW, st);
  }
  bool umma_path() const { return std::is_same<T, bf16>::value && tensor; }
  static bool gelu_bwd_fused() {
    static const bool v = [] { const char* e = getenv("S4G_GELU_BWD"); return !(e && std::string(e) == "pass"); }();
    return v;
  }
  static EpiParams epi(void* out, int ldo, const float* bias = nullptr, int mode = EPI_NONE, int out_f32 = 0) {
    EpiParams e{};
    e.bias = bias; e.out = out; e.ldo = ldo; e.mode = mode; e.out_f32 = out_f32; e.alpha = 1.f;
    return e;
  }

  // ---- linear layers (nn.Linear and 1x1 Conv2d share the [N,K] weight layout) ------------------------------
  struct Lin { Param* w; Param* b; size_t wf, wd; int N, K; RedSlot bias_red; };
  Lin linear(const std::string& name, int N, int K, bool dgrad = true) {
    Lin l{P(name + ".weight"), Popt(name + ".bias"), 0, 0, N, K, RedSlot()};
    l.wf = alloc(sizeof(T) * (size_t)N * K);
    l.wd = dgrad ? alloc(sizeof(T) * (size_t)N * K) : 0;
    if (l.b) l.bias_red = red_alloc(colsum_scratch_bytes(N));
    lin_jobs.push_back(l);   // packed by ONE grouped launch (finish())
    return l;
  }
  std::vector<Lin> lin_jobs;
  // dW, db of a linear layer given dY [M,N], X [M,K]
  int linear_param_grads(const Lin& l, const T* dY, const T* X, long long M, cudaStream_t st, bool bias_done = false) const {
    if (l.w->g) S4G_TRY(wgrad(linear_geom(M, l.N, l.K), dY, X, l.w->g, st));
    if (l.b && l.b->g && !bias_done) {
      m->launches++;
      const Reducer r = red(l.bias_red, true);
      S4G_TRY(k_colsum<T>(dY, l.b->g, M, l.N, l.N, st, &r));
    }
    return 0;
  }

  // ---- 3x3 convolutions --------------------------------------------------------------------------------------
  struct Conv { Param* w; Param* b; size_t wf, wd; int Co, Ci; RedSlot bias_red; };
  Conv conv3(const std::string& name, int Co, int Ci) {
    Conv c{P(name + ".weight"), Popt(name + ".bias"), 0, 0, Co, Ci, RedSlot()};
    if (c.b) c.bias_red = red_alloc(colsum_scratch_bytes(Co));
    c.wf = alloc(sizeof(T) * (size_t)Co * Ci * 9);
    c.wd = alloc(sizeof(T) * (size_t)Co * Ci * 9);
    dwp_bytes = std::max(dwp_bytes, sizeof(float) * (size_t)Co * Ci * 9);
    conv_jobs.push_back(c);   // packed by ONE grouped launch (finish())
    return c;
  }
  std::vector<Conv> conv_jobs;
  // out = conv3x3(in) (+bias) (epilogue mode); pushes forward and backward closures
  // stats_into (tcgen05 path, training): offset of a BatchNorm's [2 Co] double accumulator; the conv epilogue then also produces
  // the per-channel sum / sum of squares of its (bf16-rounded) output, so the BatchNorm that follows needs no statistics pass
  static bool bn_stats_fused() {
    static const bool v = [] { const char* e = getenv("S4G_BN_STATS"); return !(e && std::string(e) == "pass"); }();
    return v;
  }
  void conv3_layer(const Conv& c, int in, int out, int mode = EPI_NONE, bool in_grad = true, size_t stats_into = 0) {
    Builder* self = this;
    S4GModel* mm = m;
    const Act ai = m->acts[in];
    note_wgrad(conv3_geom(ai.B, ai.H, ai.W, c.Ci, c.Co));
    m->fwd.push_back([=](cudaStream_t st) {
      const Act& a = mm->acts[in];
      const ConvGeom g = conv3_geom(a.B, a.H, a.W, c.Ci, c.Co);
      EpiParams e = epi(self->V(out), c.Co, c.b ? c.b->p : nullptr, mode);
      if (stats_into && mm->training && self->umma_path() && bn_stats_fused()) {
        const Reducer r = self->red(self->shared_red, false);
        const int rows = umma_fprop_stat_rows(g);
        e.stat_part = reinterpret_cast<float*>(reduce_scratch(&r, sizeof(float) * (size_t)rows * 2 * c.Co));
        if (!e.stat_part) return (int)S4G_ERR_WORKSPACE;
        S4G_TRY(self->fprop(g, self->V(in), mm->template at<T>(c.wf), e, st));
        mm->launches++;
        return k_finalize(&r, e.stat_part, mm->template at<double>(stats_into), rows, 2 * c.Co, 2 * c.Co, 0, 1, st);
      }
      return self->fprop(g, self->V(in), mm->template at<T>(c.wf), e, st);
    });
    bwd->push_back([=](cudaStream_t st) {
      const Act& a = mm->acts[in];
      const ConvGeom g = conv3_geom(a.B, a.H, a.W, c.Ci, c.Co);
      if (in_grad) {
        EpiParams e = epi(self->G(in), c.Ci);
        S4G_TRY(self->fprop(conv3_geom(a.B, a.H, a.W, c.Co, c.Ci), self->G(out), mm->template at<T>(c.wd), e, st));
      }
      if (c.w->g) {
        if (self->umma_path()) {
          S4G_TRY(self->wgrad(g, self->G(out), self->V(in), c.w->g, st, 1));
        } else {
          float* dwp = mm->template at<float>(self->dwp_scratch);
          S4G_TRY(self->wgrad(g, self->G(out), self->V(in), dwp, st));
          mm->launches++;
          S4G_TRY(k_unpack_conv_grad(dwp, c.w->g, c.Co, c.Ci, 3, st));
        }
      }
      if (c.b && c.b->g) {
        mm->launches++;
        const Reducer r = self->red(c.bias_red, true);
        S4G_TRY(k_colsum<T>(self->G(out), c.b->g, a.rows(), c.Co, c.Co, st, &r));
      }
      return 0;
    });
  }
  void note_wgrad(const ConvGeom& g) {
    if (std::is_same<T, bf16>::value && tensor) wg_bytes = std::max(wg_bytes, umma_wgrad_scratch_bytes(g));
  }
  // out = in . W^T + b  (1x1 conv / linear on an Act)
  void lin_layer(const Lin& l, int in, int out, bool in_grad = true) {
    Builder* self = this;
    S4GModel* mm = m;
    note_wgrad(linear_geom(m->acts[in].rows(), l.N, l.K));
    m->fwd.push_back([=](cudaStream_t st) {
      EpiParams e = epi(self->V(out), l.N, l.b ? l.b->p : nullptr);
      return self->fprop(linear_geom(mm->acts[in].rows(), l.N, l.K), self->V(in), mm->template at<T>(l.wf), e, st);
    });
    bwd->push_back([=](cudaStream_t st) {
      const long long M = mm->acts[in].rows();
      if (in_grad) {
        EpiParams e = epi(self->G(in), l.K);
        S4G_TRY(self->fprop(linear_geom(M, l.K, l.N), self->G(out), mm->template at<T>(l.wd), e, st));
      }
      return self->linear_param_grads(l, self->G(out), self->V(in), M, st);
    });
  }

  // ---- BatchNorm2d (train / eval) ----------------------------------------------------------------------------
  struct BN { Param *w, *b, *rm, *rv, *nbt; size_t stats, sums, mean, rstd; };
  BN bn(const std::string& name) {
    BN n{P(name + ".weight"), P(name + ".bias"), P(name + ".running_mean"), P(name + ".running_var"),
         P(name + ".num_batches_tracked"), 0, 0, 0, 0};
    constexpr int MAX_BN = 17;  // DPT_decoder.py:204-206,492
    if (m->bn_count == 0) {
      m->bn_stats_pool = alloc(sizeof(double) * 512 * MAX_BN);
      m->bn_sums_pool = alloc(sizeof(double) * 512 * MAX_BN);
    }
    need_shared_red(std::max(bn_scratch_bytes(256), seghead_scratch_bytes()));
    n.stats = m->bn_stats_pool + sizeof(double) * 512 * m->bn_count;
    n.sums = m->bn_sums_pool + sizeof(double) * 512 * m->bn_count;
    m->bn_count++;
    n.mean = alloc(sizeof(float) * 256);
    n.rstd = alloc(sizeof(float) * 256);
    return n;
  }
  void need_shared_red(size_t bytes) { if (shared_red.bytes < bytes) shared_red = red_alloc(bytes); }
  // statistics of x (forward): fills mean/rstd, updates running stats when training
  int bn_forward_stats(const BN& n, const T* x, long long M, cudaStream_t st, bool stats_done = false) const {
    float* mean = m->template at<float>(n.mean);
    float* rstd = m->template at<float>(n.rstd);
    if (!m->training) { m->launches++; return k_bn_eval_prepare(n.rm->p, n.rv->p, mean, rstd, 1e-5f, 256, st); }
    double* stats = m->template at<double>(n.stats);
    m->launches += stats_done ? 1 : 3;
    const Reducer r = red(shared_red, false);
    // stats_done: the producing conv's epilogue already left sum / sum of squares in `stats` (conv3_layer, stats_into)
    if (!stats_done) S4G_TRY(k_bn_stats<T>(x, stats, M, 256, st, &r));
    if (m->d.world_size > 1) S4G_TRY(m->exchange(stats, 512, st));
    return k_bn_finalize(stats, (double)M * m->d.world_size, mean, rstd, n.rm->p, n.rv->p,
                         reinterpret_cast<long long*>(n.nbt->p), 0.1f, 1e-5f, 256, st);
  }
  int bn_backward(const BN& n, const T* dy, const T* x, T* dx, int relu, long long M, cudaStream_t st) const {
    double* sums = m->template at<double>(n.sums);
    const float* mean = m->template at<float>(n.mean);
    const float* rstd = m->template at<float>(n.rstd);
    m->launches += 3;
    const Reducer r = red(shared_red, false);
    S4G_TRY(k_bn_bwd_reduce<T>(dy, x, mean, rstd, n.w->p, n.b->p, sums, relu, M, 256, st, &r));
    if (m->d.world_size > 1) S4G_TRY(m->exchange(sums, 512, st));
    return k_bn_bwd_apply<T>(dy, x, mean, rstd, n.w->p, n.b->p, sums, (double)M * m->d.world_size, dx, n.w->g, n.b->g,
                             relu, M, 256, st);
  }

  // ---- ResidualConvUnit_custom (Models/DPT_decoder.py:170-233) ------------------------------------------------
  // returns the output act; x.g receives skip + ReLU-path gradient
  int rcu(const std::string& p, int x, bool use_bn) {
    const Act ax = m->acts[x];
    Builder* self = this;
    S4GModel* mm = m;
    const long long M = ax.rows(), n = ax.numel();
    const int r = act(ax.B, ax.H, ax.W, 256), c1 = act(ax.B, ax.H, ax.W, 256), r2 = act(ax.B, ax.H, ax.W, 256),
              c2 = act(ax.B, ax.H, ax.W, 256), out = act(ax.B, ax.H, ax.W, 256);
    Conv k1 = conv3(p + "conv1", 256, 256), k2 = conv3(p + "conv2", 256, 256);
    BN b1{}, b2{};
    if (use_bn) { b1 = bn(p + "bn1"); b2 = bn(p + "bn2"); }
    // r = relu(x);  backward: x.g = out.g + (x > 0) * r.g
    m->fwd.push_back([=](cudaStream_t st) { mm->launches++; return k_relu<T>(self->V(x), self->V(r), n, st); });
    bwd->push_back([=](cudaStream_t st) {
      mm->launches++;
      return k_relu_bwd_add<T>(self->G(r), self->V(x), self->G(out), self->G(x), n, st);
    });
    const bool fuse_stats = use_bn && umma_path() && bn_stats_fused();
    conv3_layer(k1, r, c1, EPI_NONE, true, fuse_stats ? b1.stats : 0);
    if (use_bn) {
      m->fwd.push_back([=](cudaStream_t st) {
        S4G_TRY(self->bn_forward_stats(b1, self->V(c1), M, st, fuse_stats));
        mm->launches++;
        return k_bn_apply<T>(self->V(c1), mm->template at<float>(b1.mean), mm->template at<float>(b1.rstd), b1.w->p, b1.b->p,
                             nullptr, self->V(r2), 1, M, 256, st);
      });
      bwd->push_back([=](cudaStream_t st) { return self->bn_backward(b1, self->G(r2), self->V(c1), self->G(c1), 1, M, st); });
    } else {
      m->fwd.push_back([=](cudaStream_t st) { mm->launches++; return k_relu<T>(self->V(c1), self->V(r2), n, st); });
      bwd->push_back([=](cudaStream_t st) {
        mm->launches++;
        return k_relu_bwd_add<T>(self->G(r2), self->V(c1), nullptr, self->G(c1), n, st);
      });
    }
    conv3_layer(k2, r2, c2, EPI_NONE, true, fuse_stats ? b2.stats : 0);
    if (use_bn) {
      m->fwd.push_back([=](cudaStream_t st) {
        S4G_TRY(self->bn_forward_stats(b2, self->V(c2), M, st, fuse_stats));
        mm->launches++;
        return k_bn_apply<T>(self->V(c2), mm->template at<float>(b2.mean), mm->template at<float>(b2.rstd), b2.w->p, b2.b->p,
                             self->V(x), self->V(out), 0, M, 256, st);
      });
      bwd->push_back([=](cudaStream_t st) { return self->bn_backward(b2, self->G(out), self->V(c2), self->G(c2), 0, M, st); });
    } else {
      m->fwd.push_back([=](cudaStream_t st) { mm->launches++; return k_add<T>(self->V(c2), self->V(x), self->V(out), n, st); });
      // d(c2) == d(out); resolved at run time because fusion() may re-alias out.g after this unit is built
      bwd->push_back([=](cudaStream_t) { mm->acts[c2].g = mm->acts[out].g; return 0; });
    }
    return out;
  }

  // ---- FeatureFusionBlock_custom (Models/DPT_decoder.py:238-301) ----------------------------------------------
  int fusion(const std::string& p, int x0, int x1, bool use_bn) {
    Builder* self = this;
    S4GModel* mm = m;
    int s = x0;
    if (x1 >= 0) {
      const int res = rcu(p + "resConfUnit1.", x1, use_bn);
      const Act a = m->acts[x0];
      s = act(a.B, a.H, a.W, 256);
      const long long n = a.numel();
      m->fwd.push_back([=](cudaStream_t st) { mm->launches++; return k_add<T>(self->V(x0), self->V(res), self->V(s), n, st); });
      bwd->push_back([](cudaStream_t) { return 0; });
      m->acts[x0].g = m->acts[s].g;   // d(x0) == d(res) == d(s)
      m->acts[res].g = m->acts[s].g;
    }
    const int t = rcu(p + "resConfUnit2.", s, use_bn);
    // The reference upsamples and then applies the 1x1 out_conv (Models/DPT_decoder.py:295-299).  Both are linear and
    // the bilinear weights of every output pixel sum to 1, so conv1x1(up(x)) + b == up(conv1x1(x) + b): the 1x1 conv
    // runs on the 4x smaller map (fprop, dgrad and wgrad) and the full-resolution intermediate is never materialised.
    const Act at = m->acts[t];
    const int v = act(at.B, at.H, at.W, 256);
    lin_layer(linear(p + "out_conv", 256, 256), t, v);
    const int out = act(at.B, 2 * at.H, 2 * at.W, 256);
    m->fwd.push_back([=](cudaStream_t st) {
      const Act& a = mm->acts[v];
      mm->launches++;
      return k_bilinear2x_fwd<T>(self->V(v), self->V(out), a.B, a.H, a.W, 256, st);
    });
    bwd->push_back([=](cudaStream_t st) {
      const Act& a = mm->acts[v];
      mm->launches++;
      return k_bilinear2x_bwd<T>(self->G(out), self->G(v), a.B, a.H, a.W, 256, st);
    });
    return out;
  }

  // ---- encoder -------------------------------------------------------------------------------------------------
  struct Block {
    Param *n1w, *n1b, *n2w, *n2b;
    Lin qkv, proj, fc1, fc2;
    size_t h1, mean1, rstd1, qkvb, ao, lse, h2, mean2, rstd2, u, a;
    RedSlot ln1_red, ln2_red, gelu_red;
  };
  std::vector<size_t> xs;  // residual stream snapshots: xs[2i] = block i input, xs[2i+1] = after attention
  std::vector<Block> blocks;
  size_t patches = 0, pe = 0;
  Lin pe_lin{};
  size_t dxa = 0, dxb = 0, dxT = 0, dh = 0, du = 0, dao = 0, dqkv = 0, attn_delta = 0;  // backward temporaries
  int tap_act[4] = {-1, -1, -1, -1};

  int attn_fwd(const T* qkv, T* o, float* lse, cudaStream_t st) const {
    m->launches++;
    if constexpr (std::is_same<T, bf16>::value) {
      if (tensor) return attn_fwd_mma(qkv, o, lse, B, 197, 12, st);
    }
    return attn_fwd_simt<T>(qkv, o, lse, B, 197, 12, st);
  }
  int attn_bwd(const T* qkv, const T* o, const T* dout, const float* lse, T* dqkv_, cudaStream_t st, const float* delta = nullptr) const {
    m->launches++;
    if constexpr (std::is_same<T, bf16>::value) {
      if (tensor) return attn_bwd_mma(qkv, o, dout, lse, dqkv_, B, 197, 12, st, delta);
    }
    return attn_bwd_simt<T>(qkv, o, dout, lse, dqkv_, B, 197, 12, st);
  }

  void build_encoder(bool dense) {
    Builder* self = this;
    S4GModel* mm = m;
    const long long Tk = (long long)B * 197, Pn = (long long)B * 196;
    const bool train_enc = !m->d.frozen;
    patches = alloc(sizeof(T) * Pn * 768);
    pe = alloc(sizeof(float) * Pn * 768);
    pe_lin = linear("patch_embed.proj", 768, 768, false);
    note_wgrad(linear_geom(Pn, 768, 768));
    Param* cls = P("cls_token");
    Param* pos = P("pos_embed");
    xs.resize(25);
    for (auto& x : xs) x = alloc(sizeof(float) * Tk * 768);
    m->fwd.push_back([=](cudaStream_t st) {
      mm->launches += 3;
      S4G_TRY(k_patchify<T>(mm->images, mm->template at<T>(self->patches), self->B, st));
      EpiParams e = epi(mm->template at<float>(self->pe), 768, self->pe_lin.b->p, EPI_NONE, 1);
      S4G_TRY(self->fprop(linear_geom(Pn, 768, 768), mm->template at<T>(self->patches), mm->template at<T>(self->pe_lin.wf), e, st));
      return k_assemble_tokens(mm->template at<float>(self->pe), cls->p, pos->p, mm->template at<float>(self->xs[0]), self->B, st);
    });
    for (int i = 0; i < 12; ++i) {
      const std::string p = "blocks." + std::to_string(i) + ".";
      Block b;
      b.n1w = P(p + "norm1.weight"); b.n1b = P(p + "norm1.bias");
      b.n2w = P(p + "norm2.weight"); b.n2b = P(p + "norm2.bias");
      b.qkv = linear(p + "attn.qkv", 2304, 768, train_enc);
      b.proj = linear(p + "attn.proj", 768, 768, train_enc);
      b.fc1 = linear(p + "mlp.fc1", 3072, 768, train_enc);
      b.fc2 = linear(p + "mlp.fc2", 768, 3072, train_enc);
      b.h1 = alloc(sizeof(T) * Tk * 768); b.h2 = alloc(sizeof(T) * Tk * 768);
      b.mean1 = alloc(sizeof(float) * Tk); b.rstd1 = alloc(sizeof(float) * Tk);
      b.mean2 = alloc(sizeof(float) * Tk); b.rstd2 = alloc(sizeof(float) * Tk);
      b.qkvb = alloc(sizeof(T) * Tk * 2304);
      b.ao = alloc(sizeof(T) * Tk * 768);
      b.lse = alloc(sizeof(float) * (size_t)B * 12 * 197);
      b.u = alloc(sizeof(T) * Tk * 3072); b.a = alloc(sizeof(T) * Tk * 3072);
      if (train_enc) {
        b.ln1_red = red_alloc(ln_bwd_scratch_bytes(768)); b.ln2_red = red_alloc(ln_bwd_scratch_bytes(768));
        b.gelu_red = red_alloc(colsum_scratch_bytes(3072));
      }
      note_wgrad(linear_geom(Tk, 2304, 768)); note_wgrad(linear_geom(Tk, 768, 768));
      note_wgrad(linear_geom(Tk, 3072, 768)); note_wgrad(linear_geom(Tk, 768, 3072));
      blocks.push_back(b);
      const size_t xin = xs[2 * i], xmid = xs[2 * i + 1], xout = xs[2 * i + 2];
      m->fwd.push_back([=](cudaStream_t st) {
        float* x0 = mm->template at<float>(xin);
        float* x1 = mm->template at<float>(xmid);
        float* x2 = mm->template at<float>(xout);
        T* h1 = mm->template at<T>(b.h1);
        T* h2 = mm->template at<T>(b.h2);
        T* qkv = mm->template at<T>(b.qkvb);
        T* ao = mm->template at<T>(b.ao);
        mm->launches += 2;
        S4G_TRY(k_layernorm_fwd<T>(x0, b.n1w->p, b.n1b->p, h1, mm->template at<float>(b.mean1), mm->template at<float>(b.rstd1), (int)Tk, 768, 1e-6f, st));
        S4G_TRY(self->fprop(linear_geom(Tk, 2304, 768), h1, mm->template at<T>(b.qkv.wf), epi(qkv, 2304, b.qkv.b->p), st));
        S4G_TRY(self->attn_fwd(qkv, ao, mm->template at<float>(b.lse), st));
        EpiParams e1 = epi(x1, 768, b.proj.b->p, EPI_NONE, 1);
        e1.res = x0; e1.res_f32 = 1; e1.ldr = 768;
        S4G_TRY(self->fprop(linear_geom(Tk, 768, 768), ao, mm->template at<T>(b.proj.wf), e1, st));
        S4G_TRY(k_layernorm_fwd<T>(x1, b.n2w->p, b.n2b->p, h2, mm->template at<float>(b.mean2), mm->template at<float>(b.rstd2), (int)Tk, 768, 1e-6f, st));
        if (self->umma_path()) {
          // K = 768 is too short to hide the GELU arithmetic behind the MMAs of a tile: u from the GEMM, a = gelu(u)
          // as its own HBM-bound pass
          S4G_TRY(self->fprop(linear_geom(Tk, 3072, 768), h2, mm->template at<T>(b.fc1.wf), epi(mm->template at<T>(b.u), 3072, b.fc1.b->p), st));
          mm->launches++;
          // the same pass leaves gelu'(u) in place of u (u has no other reader): backward multiplies by it in the fc2 dgrad
          // epilogue (S4G_GELU_BWD=pass keeps the separate GELU' pass over du)
          S4G_TRY(k_gelu<T>(mm->template at<T>(b.u), mm->template at<T>(b.a), Tk * 3072, st, gelu_bwd_fused() ? mm->template at<T>(b.u) : nullptr));
        } else {
          EpiParams e2 = epi(mm->template at<T>(b.a), 3072, b.fc1.b->p, EPI_GELU);
          e2.out2 = mm->template at<T>(b.u);
          S4G_TRY(self->fprop(linear_geom(Tk, 3072, 768), h2, mm->template at<T>(b.fc1.wf), e2, st));
        }
        EpiParams e3 = epi(x2, 768, b.fc2.b->p, EPI_NONE, 1);
        e3.res = x1; e3.res_f32 = 1; e3.ldr = 768;
        return self->fprop(linear_geom(Tk, 768, 3072), mm->template at<T>(b.a), mm->template at<T>(b.fc2.wf), e3, st);
      });
      if (dense && (i == 2 || i == 5 || i == 8 || i == 11)) {
        const int k = i / 3;
        tap_act[k] = act(B, 14, 14, 768, train_enc);
        const int ta = tap_act[k];
        m->fwd.push_back([=](cudaStream_t st) {
          mm->launches++;
          return k_tokens_to_nhwc<T>(mm->template at<float>(xout), self->V(ta), self->B, st);
        });
      }
    }
    if (!train_enc) return;
    // ---------------- encoder backward (stored in forward order; run reversed) ----------------
    dxa = alloc(sizeof(float) * Tk * 768); dxb = alloc(sizeof(float) * Tk * 768);
    dxT = alloc(sizeof(T) * Tk * 768); dh = alloc(sizeof(T) * Tk * 768);
    du = alloc(sizeof(T) * Tk * 3072); dao = alloc(sizeof(T) * Tk * 768); dqkv = alloc(sizeof(T) * Tk * 2304);
    attn_delta = alloc(sizeof(float) * (size_t)B * 12 * 197);
    // patch-embed / token-assembly backward: gradient arrives in dxa (12 blocks => even number of ping-pongs)
    m->bwd_enc.push_back([=](cudaStream_t st) {
      float* dx = mm->template at<float>(self->dxa);
      mm->launches += 2;
      if (cls->g || pos->g) S4G_TRY(k_assemble_tokens_bwd(dx, nullptr, cls->g, pos->g, self->B, st));
      T* dpe = mm->template at<T>(self->dxT);
      S4G_TRY(k_tokens_to_nhwc<T>(dx, dpe, self->B, st));
      return self->linear_param_grads(self->pe_lin, dpe, mm->template at<T>(self->patches), Pn, st);
    });
    for (int i = 0; i < 12; ++i) {
      const Block b = blocks[i];
      const size_t xin = xs[2 * i], xmid = xs[2 * i + 1];
      const int tapk = (dense && (i == 2 || i == 5 || i == 8 || i == 11)) ? i / 3 : -1;
      const int ta = tapk >= 0 ? tap_act[tapk] : -1;
      // bias gradients of proj / fc2 are column sums of the fp32 residual gradients the LayerNorm backward kernels
      // produce: LN2-bwd of this block yields proj's, LN1-bwd of this block yields fc2's of block i-1 (unless a tap
      // gradient is added in between)
      const bool fc2_bias_fused = i < 11 && ta < 0;
      const bool prev_fc2_fused = i > 0 && !(dense && (i - 1 == 2 || i - 1 == 5 || i - 1 == 8));
      float* prev_fc2_bias_g = i > 0 ? blocks[i - 1].fc2.b->g : nullptr;
      m->bwd_enc.push_back([=](cudaStream_t st) {
        // incoming gradient of the block output: dxa; result (block input gradient) also ends in dxa
        float* dxo = mm->template at<float>(self->dxa);
        float* dxm = mm->template at<float>(self->dxb);
        T* dxt = mm->template at<T>(self->dxT);
        T* dhb = mm->template at<T>(self->dh);
        T* dub = mm->template at<T>(self->du);
        T* daob = mm->template at<T>(self->dao);
        T* dqkvb = mm->template at<T>(self->dqkv);
        // dxt = bf16 copy of the incoming gradient: written by the tap-gradient kernel, or by the previous block's
        // LayerNorm backward (dx_cast)
        if (ta >= 0) { mm->launches++; S4G_TRY(k_tap_grad<T>(self->G(ta), dxo, self->B, i == 11 ? 0 : 1, st, dxt)); }
        else if (i == 11) { mm->launches++; S4G_TRY(k_cast_f32<T>(dxo, dxt, Tk * 768, st)); }
        mm->launches += 2;
        // mlp: x2 = x1 + fc2(gelu(fc1(ln2(x1))))
        if (self->umma_path() && gelu_bwd_fused()) {
          // du = (dx2 . W2) * gelu'(u): the stored derivative is one multiply in the dgrad epilogue; fc1's bias gradient is
          // a column sum of du
          EpiParams e = epi(dub, 3072, nullptr, EPI_MUL);
          e.aux = mm->template at<T>(b.u);
          S4G_TRY(self->fprop(linear_geom(Tk, 3072, 768), dxt, mm->template at<T>(b.fc2.wd), e, st));
          mm->launches++;
          const Reducer rg = self->red(b.gelu_red, true);
          if (b.fc1.b->g) S4G_TRY(k_colsum<T>(dub, b.fc1.b->g, Tk, 3072, 3072, st, &rg));
        } else if (self->umma_path()) {
          S4G_TRY(self->fprop(linear_geom(Tk, 3072, 768), dxt, mm->template at<T>(b.fc2.wd), epi(dub, 3072), st));
          mm->launches++;
          // fc1's bias gradient (column sums of du) comes out of the same pass
          const Reducer rg = self->red(b.gelu_red, true);
          S4G_TRY(k_dgelu_mul<T>(dub, mm->template at<T>(b.u), dub, Tk * 3072, st, b.fc1.b->g, 3072, &rg));
        } else {
          EpiParams e = epi(dub, 3072, nullptr, EPI_DGELU);
          e.aux = mm->template at<T>(b.u);
          S4G_TRY(self->fprop(linear_geom(Tk, 3072, 768), dxt, mm->template at<T>(b.fc2.wd), e, st));
        }
        S4G_TRY(self->linear_param_grads(b.fc2, dxt, mm->template at<T>(b.a), Tk, st, fc2_bias_fused));
        S4G_TRY(self->fprop(linear_geom(Tk, 768, 3072), dub, mm->template at<T>(b.fc1.wd), epi(dhb, 768), st));
        S4G_TRY(self->linear_param_grads(b.fc1, dub, mm->template at<T>(b.h2), Tk, st, self->umma_path()));
        const Reducer r2 = self->red(b.ln2_red, true), r1 = self->red(b.ln1_red, true);
        S4G_TRY(k_layernorm_bwd<T>(dhb, mm->template at<float>(xmid), b.n2w->p, mm->template at<float>(b.mean2), mm->template at<float>(b.rstd2), dxo, dxm, b.n2w->g, b.n2b->g, (int)Tk, 768, st, dxt, b.proj.b->g, &r2));
        // attention: x1 = x0 + proj(attn(qkv(ln1(x0))))
        // d(attention output) = dx1 . Wproj; on the tcgen05 path its epilogue also yields delta = rowsum(dO o O) per head, which
        // the attention backward would otherwise compute with its own pass over O and dO
        EpiParams ep = epi(daob, 768);
        const float* delta = nullptr;
        if (self->umma_path() && attn_uses_umma()) {
          ep.dot_aux = mm->template at<T>(b.ao); ep.dot_out = mm->template at<float>(self->attn_delta); ep.dot_rows = 197; ep.dot_groups = 12;
          delta = ep.dot_out;
        }
        S4G_TRY(self->fprop(linear_geom(Tk, 768, 768), dxt, mm->template at<T>(b.proj.wd), ep, st));
        S4G_TRY(self->linear_param_grads(b.proj, dxt, mm->template at<T>(b.ao), Tk, st, true));
        S4G_TRY(self->attn_bwd(mm->template at<T>(b.qkvb), mm->template at<T>(b.ao), daob, mm->template at<float>(b.lse), dqkvb, st, delta));
        S4G_TRY(self->fprop(linear_geom(Tk, 768, 2304), dqkvb, mm->template at<T>(b.qkv.wd), epi(dhb, 768), st));
        S4G_TRY(self->linear_param_grads(b.qkv, dqkvb, mm->template at<T>(b.h1), Tk, st));
        return k_layernorm_bwd<T>(dhb, mm->template at<float>(xin), b.n1w->p, mm->template at<float>(b.mean1), mm->template at<float>(b.rstd1), dxm, dxo, b.n1w->g, b.n1b->g, (int)Tk, 768, st, i > 0 ? dxt : nullptr,
                                  prev_fc2_fused ? prev_fc2_bias_g : nullptr, &r1);
      });
    }
  }

  // ---- decoder ---------------------------------------------------------------------------------------------------
  size_t head_small = 0, head_dsmall = 0;
  void build_decoder() {
    Builder* self = this;
    S4GModel* mm = m;
    bwd = &m->bwd_dec;
    const bool seg = m->d.dense == S4G_DENSE_SEG;
    const bool tap_grad = !m->d.frozen;
    const std::string d = "decoder.";
    // reassemble (Models/DPT_decoder.py:517-520)
    int l[4];
    {  // act_postprocess12: 1x1 768->96, ConvT k4 s4
      const int a = act(B, 14, 14, 96);
      lin_layer(linear(d + "act_postprocess12.0", 96, 768), tap_act[0], a, tap_grad);
      l[0] = convT_layer(d + "act_postprocess12.1", a, 96, 4);
    }
    {
      const int a = act(B, 14, 14, 192);
      lin_layer(linear(d + "act_postprocess22.0", 192, 768), tap_act[1], a, tap_grad);
      l[1] = convT_layer(d + "act_postprocess22.1", a, 192, 2);
    }
    {
      l[2] = act(B, 14, 14, 384);
      lin_layer(linear(d + "act_postprocess32.0", 384, 768), tap_act[2], l[2], tap_grad);
    }
    {
      const int a = act(B, 14, 14, 768);
      lin_layer(linear(d + "act_postprocess42.0", 768, 768), tap_act[3], a, tap_grad);
      l[3] = conv_s2_layer(d + "act_postprocess42.1", a);
    }
    const int feat[4] = {96, 192, 384, 768};
    int r[4];
    for (int k = 0; k < 4; ++k) {
      const Act a = m->acts[l[k]];
      r[k] = act(a.B, a.H, a.W, 256);
      conv3_layer(conv3(d + "layer" + std::to_string(k + 1) + "_rn", 256, feat[k]), l[k], r[k]);
    }
    const int p4 = fusion(d + "refinenet4.", r[3], -1, seg);
    const int p3 = fusion(d + "refinenet3.", p4, r[2], seg);
    const int p2 = fusion(d + "refinenet2.", p3, r[1], seg);
    const int p1 = fusion(d + "refinenet1.", p2, r[0], seg);
    const long long M1 = (long long)B * 112 * 112;
    head_small = alloc(sizeof(float) * M1);
    head_dsmall = alloc(sizeof(float) * M1);
    if (seg) {
      const int c = act(B, 112, 112, 256);
      BN n = bn(d + "output_conv.1");
      const bool fuse_stats = umma_path() && bn_stats_fused();
      conv3_layer(conv3(d + "output_conv.0", 256, 256), p1, c, EPI_NONE, true, fuse_stats ? n.stats : 0);
      Param* w4 = P(d + "output_conv.4.weight");
      Param* b4 = P(d + "output_conv.4.bias");
      m->fwd.push_back([=](cudaStream_t st) {
        S4G_TRY(self->bn_forward_stats(n, self->V(c), M1, st, fuse_stats));
        mm->launches += 2;
        float* small = mm->template at<float>(self->head_small);
        S4G_TRY(k_seghead_fwd<T>(self->V(c), mm->template at<float>(n.mean), mm->template at<float>(n.rstd), n.w->p, n.b->p, w4->p, b4->p,
                                 mm->training ? mm->dropout_mask : nullptr, mm->training ? mm->d.dropout_p : 0.f, mm->seed, small, M1, 256, st, mm->seed_dev));
        return k_bilinear2x_f32_fwd(small, mm->out, self->B, 112, 112, st);
      });
      bwd->push_back([=](cudaStream_t st) {
        float* ds = mm->template at<float>(self->head_dsmall);
        double* sums = mm->template at<double>(n.sums);
        mm->launches += 4;
        S4G_TRY(k_bilinear2x_f32_bwd(mm->dout, ds, self->B, 112, 112, st));
        const Reducer rs = self->red(self->shared_red, false);
        S4G_TRY(k_seghead_bwd_reduce<T>(ds, self->V(c), mm->template at<float>(n.mean), mm->template at<float>(n.rstd), n.w->p, n.b->p, w4->p,
                                        mm->dropout_mask, mm->d.dropout_p, mm->seed, sums, w4->g, b4->g, M1, 256, st, &rs, mm->seed_dev));
        if (mm->d.world_size > 1) S4G_TRY(mm->exchange(sums, 512, st));
        return k_seghead_bwd_apply<T>(ds, self->V(c), mm->template at<float>(n.mean), mm->template at<float>(n.rstd), n.w->p, n.b->p, w4->p,
                                      mm->dropout_mask, mm->d.dropout_p, mm->seed, sums, (double)M1 * mm->d.world_size, self->G(c),
                                      n.w->g, n.b->g, M1, 256, st, mm->seed_dev);
      });
    } else {
      const int c = act(B, 112, 112, 128);
      conv3_layer(conv3(d + "output_conv.0", 128, 256), p1, c);
      const int u = act(B, 224, 224, 128);
      m->fwd.push_back([=](cudaStream_t st) { mm->launches++; return k_bilinear2x_fwd<T>(self->V(c), self->V(u), self->B, 112, 112, 128, st); });
      bwd->push_back([=](cudaStream_t st) { mm->launches++; return k_bilinear2x_bwd<T>(self->G(u), self->G(c), self->B, 112, 112, 128, st); });
      const int h = act(B, 224, 224, 32);
      conv3_layer(conv3(d + "output_conv.2", 32, 128), u, h, EPI_RELU);
      Param* w4 = P(d + "output_conv.4.weight");
      Param* b4 = P(d + "output_conv.4.bias");
      const long long M2 = (long long)B * 224 * 224;
      m->fwd.push_back([=](cudaStream_t st) { mm->launches++; return k_depthhead_fwd<T>(self->V(h), w4->p, b4->p, mm->out, M2, 32, st); });
      // inserted BEFORE conv(…2)'s backward in run order because bwd runs reversed
      need_shared_red(depthhead_scratch_bytes());
      bwd->push_back([=](cudaStream_t st) {
        mm->launches += 3;
        const Reducer rd = self->red(self->shared_red, false);
        S4G_TRY(k_depthhead_bwd<T>(mm->dout, mm->out, self->V(h), w4->p, self->G(h), w4->g, b4->g, M2, 32, st, &rd));
        return k_relu_bwd_add<T>(self->G(h), self->V(h), nullptr, self->G(h), M2 * 32, st);  // ReLU mask (h = relu(.))
      });
    }
  }

  // ConvTranspose2d(C, C, k, stride k) (Models/DPT_decoder.py:341-350, 365-374)
  int convT_layer(const std::string& name, int in, int C, int k) {
    Builder* self = this;
    S4GModel* mm = m;
    Param* w = P(name + ".weight");
    Param* b = P(name + ".bias");
    const int kk = k * k;
    const size_t wf = alloc(sizeof(T) * (size_t)kk * C * C), wd = alloc(sizeof(T) * (size_t)kk * C * C);
    dwp_bytes = std::max(dwp_bytes, sizeof(float) * (size_t)kk * C * C);
    m->pack.push_back([=](cudaStream_t st) { mm->launches++; return k_pack_convT<T>(w->p, mm->template at<T>(wf), mm->template at<T>(wd), C, C, k, st); });
    const Act a = m->acts[in];
    const long long M = a.rows();
    const size_t ycol = alloc(sizeof(T) * M * kk * C), dycol = alloc(sizeof(T) * M * kk * C);
    const int out = act(a.B, a.H * k, a.W * k, C);
    note_wgrad(linear_geom(M, kk * C, C));
    const RedSlot bias_red = red_alloc(colsum_scratch_bytes(C));
    m->fwd.push_back([=](cudaStream_t st) {
      S4G_TRY(self->fprop(linear_geom(M, kk * C, C), self->V(in), mm->template at<T>(wf), epi(mm->template at<T>(ycol), kk * C), st));
      mm->launches++;
      return k_pixel_shuffle<T>(mm->template at<T>(ycol), b->p, self->V(out), a.B, a.H, a.W, C, k, st);
    });
    bwd->push_back([=](cudaStream_t st) {
      T* dy = mm->template at<T>(dycol);
      mm->launches += 3;
      S4G_TRY(k_pixel_unshuffle<T>(self->G(out), dy, a.B, a.H, a.W, C, k, st));
      S4G_TRY(self->fprop(linear_geom(M, C, kk * C), dy, mm->template at<T>(wd), epi(self->G(in), C), st));
      float* dwp = mm->template at<float>(self->dwp_scratch);
      S4G_TRY(self->wgrad(linear_geom(M, kk * C, C), dy, self->V(in), dwp, st));
      S4G_TRY(k_unpack_convT_grad(dwp, w->g, C, C, k, st));
      const Reducer r = self->red(bias_red, true);
      return k_colsum<T>(self->G(out), b->g, M * kk, C, C, st, &r);
    });
    return out;
  }

  // Conv2d(768, 768, 3, stride 2, pad 1) (Models/DPT_decoder.py:403-409): explicit im2col + GEMM
  int conv_s2_layer(const std::string& name, int in) {
    Builder* self = this;
    S4GModel* mm = m;
    Param* w = P(name + ".weight");
    Param* b = P(name + ".bias");
    const Act a = m->acts[in];
    const int C = a.C, Ho = (a.H + 1) / 2, Wo = (a.W + 1) / 2;
    const long long M = (long long)a.B * Ho * Wo;
    const size_t wf = alloc(sizeof(T) * (size_t)C * C * 9), wt = alloc(sizeof(T) * (size_t)C * C * 9);
    dwp_bytes = std::max(dwp_bytes, sizeof(float) * (size_t)C * C * 9);
    const size_t col = alloc(sizeof(T) * M * 9 * C), dcol = alloc(sizeof(T) * M * 9 * C);
    const size_t wstage = alloc(sizeof(float) * (size_t)C * C * 9);
    m->pack.push_back([=](cudaStream_t st) {
      // wf[co][tap][ci] ; wt = wf^T ([tap*C+ci][co]) via the fp32 staging copy of the packed layout
      mm->launches += 3;
      S4G_TRY(k_pack_conv<float>(w->p, mm->template at<float>(wstage), nullptr, C, C, 3, st));
      return k_pack_linear<T>(mm->template at<float>(wstage), mm->template at<T>(wf), mm->template at<T>(wt), C, 9 * C, st);
    });
    const int out = act(a.B, Ho, Wo, C);
    note_wgrad(linear_geom(M, C, 9 * C));
    const RedSlot bias_red = red_alloc(colsum_scratch_bytes(C));
    m->fwd.push_back([=](cudaStream_t st) {
      mm->launches++;
      S4G_TRY(k_im2col_s2<T>(self->V(in), mm->template at<T>(col), a.B, a.H, a.W, C, st));
      return self->fprop(linear_geom(M, C, 9 * C), mm->template at<T>(col), mm->template at<T>(wf), epi(self->V(out), C, b->p), st);
    });
    bwd->push_back([=](cudaStream_t st) {
      T* dc = mm->template at<T>(dcol);
      mm->launches += 3;
      S4G_TRY(self->fprop(linear_geom(M, 9 * C, C), self->G(out), mm->template at<T>(wt), epi(dc, 9 * C), st));
      S4G_TRY(k_col2im_s2<T>(dc, self->G(in), a.B, a.H, a.W, C, st));
      float* dwp = mm->template at<float>(self->dwp_scratch);
      S4G_TRY(self->wgrad(linear_geom(M, C, 9 * C), self->G(out), mm->template at<T>(col), dwp, st));
      S4G_TRY(k_unpack_conv_grad(dwp, w->g, C, C, 3, st));
      const Reducer r = self->red(bias_red, true);
      return k_colsum<T>(self->G(out), b->g, M, C, C, st, &r);
    });
    return out;
  }

  // ---- classification branch (Models/models.py:338,348-354) -------------------------------------------------------
  void build_cls_head() {
    Builder* self = this;
    S4GModel* mm = m;
    bwd = &m->bwd_dec;
    const long long Tk = (long long)B * 197;
    const bool train_enc = !m->d.frozen;
    const int nc = m->d.num_classes;
    const int spatial = m->d.out_token;
    Param* nw = P("norm.weight");
    Param* nb = P("norm.bias");
    Param* hw = nc > 0 ? P("lin_head.weight") : nullptr;
    Param* hb = nc > 0 ? P("lin_head.bias") : nullptr;
    const size_t hn = alloc(sizeof(T) * Tk * 768), mean = alloc(sizeof(float) * Tk), rstd = alloc(sizeof(float) * Tk);
    const size_t pooled = alloc(sizeof(float) * (size_t)B * 768);
    const size_t dhn = train_enc ? alloc(sizeof(T) * Tk * 768) : 0;
    const size_t xlast = xs[24];
    const RedSlot ln_red = red_alloc(ln_bwd_scratch_bytes(768));
    m->fwd.push_back([=](cudaStream_t st) {
      mm->launches += 3;
      S4G_TRY(k_layernorm_fwd<T>(mm->template at<float>(xlast), nw->p, nb->p, mm->template at<T>(hn), mm->template at<float>(mean),
                                 mm->template at<float>(rstd), (int)Tk, 768, 1e-6f, st));
      return k_cls_head_fwd<T>(mm->template at<T>(hn), hw ? hw->p : nullptr, hb ? hb->p : nullptr, mm->template at<float>(pooled),
                               mm->out, self->B, nc, 768, spatial, st);
    });
    bwd->push_back([=](cudaStream_t st) {
      mm->launches += 3;
      S4G_TRY(k_cls_head_bwd<T>(mm->dout, mm->template at<float>(pooled), hw ? hw->p : nullptr, hw ? hw->g : nullptr,
                                hb ? hb->g : nullptr, train_enc ? mm->template at<T>(dhn) : nullptr, self->B, nc, 768, spatial, st));
      if (!train_enc) return 0;
      // gradient of the last block's output lands in dxa, where the encoder backward expects it
      const Reducer r = self->red(ln_red, true);
      return k_layernorm_bwd<T>(mm->template at<T>(dhn), mm->template at<float>(xlast), nw->p, mm->template at<float>(mean),
                                mm->template at<float>(rstd), nullptr, mm->template at<float>(self->dxa), nw->g, nb->g, (int)Tk, 768, st,
                                nullptr, nullptr, &r);
    });
  }

  // ---- generic pre-LN transformer stack (timm 0.6.12 Block) for the MAE pretraining step ---------------------------------
  // The fine-tuning encoder above is specialised (taps, 197 tokens, fused delta); this one takes (tokens, C, heads) and is
  // used for the 50-token encoder and the 512-d / 16-head decoder of Models/mae/models_mae.py:41-59.  Head dims of 32 run on
  // the 64-dim attention kernels through a zero-padded copy of q | k | v: q is scaled by sqrt(2) in the pad pass, so the
  // kernels' 64^-1/2 becomes 32^-1/2, and dq by sqrt(2) in the un-pad pass.
  struct TBlock {
    Param *n1w, *n1b, *n2w, *n2b;
    Lin qkv, proj, fc1, fc2;
    size_t h1, mean1, rstd1, qkvb, qkvp, ao, aop, lse, h2, mean2, rstd2, u, a;
    RedSlot ln1_red, ln2_red, gelu_red;
  };
  struct TStack {
    int C = 0, H = 0, D = 0, N = 0, depth = 0;
    long long Tk = 0;
    std::vector<TBlock> blocks;
    std::vector<size_t> xs;  // fp32 residual stream: xs[2i] block i input, xs[2i+1] after attention, xs[2*depth] output
    size_t dxa = 0, dxb = 0, dxT = 0, dh = 0, du = 0, dao = 0, daop = 0, dqkv = 0, dqkvp = 0;
  };
  int attn_fwd_g(const T* qkv, T* o, float* lse, int N, int H, cudaStream_t st) const {
    m->launches++;
    if constexpr (std::is_same<T, bf16>::value) {
      if (tensor) return attn_fwd_mma(qkv, o, lse, B, N, H, st);
    }
    return attn_fwd_simt<T>(qkv, o, lse, B, N, H, st);
  }
  int attn_bwd_g(const T* qkv, const T* o, const T* dout, const float* lse, T* dqkv_, int N, int H, cudaStream_t st) const {
    m->launches++;
    if constexpr (std::is_same<T, bf16>::value) {
      if (tensor) return attn_bwd_mma(qkv, o, dout, lse, dqkv_, B, N, H, st, nullptr);
    }
    return attn_bwd_simt<T>(qkv, o, dout, lse, dqkv_, B, N, H, st);
  }
  // forward closures go to m->fwd, backward closures (one per block, forward order) to *bl
  TStack* tstack(const std::string& prefix, int depth, int C, int H, int N, std::vector<Op>* bl) {
    stacks.emplace_back(new TStack());
    TStack* s = stacks.back().get();
    s->C = C; s->H = H; s->D = C / H; s->N = N; s->depth = depth; s->Tk = (long long)B * N;
    Builder* self = this;
    S4GModel* mm = m;
    const long long Tk = s->Tk;
    const int C3 = 3 * C, C4 = 4 * C, HP = H * 64;
    const bool pad = s->D == 32;
    s->xs.resize(2 * depth + 1);
    for (auto& x : s->xs) x = alloc(sizeof(float) * Tk * C);
    for (int i = 0; i < depth; ++i) {
      const std::string p = prefix + std::to_string(i) + ".";
      TBlock b{};
      b.n1w = P(p + "norm1.weight"); b.n1b = P(p + "norm1.bias");
      b.n2w = P(p + "norm2.weight"); b.n2b = P(p + "norm2.bias");
      b.qkv = linear(p + "attn.qkv", C3, C); b.proj = linear(p + "attn.proj", C, C);
      b.fc1 = linear(p + "mlp.fc1", C4, C); b.fc2 = linear(p + "mlp.fc2", C, C4);
      b.h1 = alloc(sizeof(T) * Tk * C); b.h2 = alloc(sizeof(T) * Tk * C);
      b.mean1 = alloc(sizeof(float) * Tk); b.rstd1 = alloc(sizeof(float) * Tk);
      b.mean2 = alloc(sizeof(float) * Tk); b.rstd2 = alloc(sizeof(float) * Tk);
      b.qkvb = alloc(sizeof(T) * Tk * C3); b.ao = alloc(sizeof(T) * Tk * C);
      if (pad) { b.qkvp = alloc(sizeof(T) * Tk * 3 * HP); b.aop = alloc(sizeof(T) * Tk * HP); }
      b.lse = alloc(sizeof(float) * (size_t)B * H * N);
      b.u = alloc(sizeof(T) * Tk * C4); b.a = alloc(sizeof(T) * Tk * C4);
      b.ln1_red = red_alloc(ln_bwd_scratch_bytes(C)); b.ln2_red = red_alloc(ln_bwd_scratch_bytes(C));
      b.gelu_red = red_alloc(colsum_scratch_bytes(C4));
      note_wgrad(linear_geom(Tk, C3, C)); note_wgrad(linear_geom(Tk, C, C));
      note_wgrad(linear_geom(Tk, C4, C)); note_wgrad(linear_geom(Tk, C, C4));
      s->blocks.push_back(b);
      const size_t xin = s->xs[2 * i], xmid = s->xs[2 * i + 1], xout = s->xs[2 * i + 2];
      m->fwd.push_back([=](cudaStream_t st) {
        float* x0 = mm->template at<float>(xin);
        float* x1 = mm->template at<float>(xmid);
        float* x2 = mm->template at<float>(xout);
        T* h1 = mm->template at<T>(b.h1);
        T* h2 = mm->template at<T>(b.h2);
        T* qkv = mm->template at<T>(b.qkvb);
        T* ao = mm->template at<T>(b.ao);
        mm->launches += 2;
        S4G_TRY(k_layernorm_fwd<T>(x0, b.n1w->p, b.n1b->p, h1, mm->template at<float>(b.mean1), mm->template at<float>(b.rstd1), (int)Tk, C, 1e-6f, st));
        S4G_TRY(self->fprop(linear_geom(Tk, C3, C), h1, mm->template at<T>(b.qkv.wf), epi(qkv, C3, b.qkv.b->p), st));
        if (pad) {
          T* qp = mm->template at<T>(b.qkvp);
          T* op = mm->template at<T>(b.aop);
          mm->launches += 2;
          S4G_TRY(k_head_pad<T>(qkv, qp, Tk, 3 * H, H, 1.41421356237309505f, st));
          S4G_TRY(self->attn_fwd_g(qp, op, mm->template at<float>(b.lse), N, H, st));
          S4G_TRY(k_head_unpad<T>(op, ao, Tk, H, 0, 1.f, st));
        } else {
          S4G_TRY(self->attn_fwd_g(qkv, ao, mm->template at<float>(b.lse), N, H, st));
        }
        EpiParams e1 = epi(x1, C, b.proj.b->p, EPI_NONE, 1);
        e1.res = x0; e1.res_f32 = 1; e1.ldr = C;
        S4G_TRY(self->fprop(linear_geom(Tk, C, C), ao, mm->template at<T>(b.proj.wf), e1, st));
        S4G_TRY(k_layernorm_fwd<T>(x1, b.n2w->p, b.n2b->p, h2, mm->template at<float>(b.mean2), mm->template at<float>(b.rstd2), (int)Tk, C, 1e-6f, st));
        if (self->umma_path()) {
          S4G_TRY(self->fprop(linear_geom(Tk, C4, C), h2, mm->template at<T>(b.fc1.wf), epi(mm->template at<T>(b.u), C4, b.fc1.b->p), st));
          mm->launches++;
          S4G_TRY(k_gelu<T>(mm->template at<T>(b.u), mm->template at<T>(b.a), Tk * C4, st, mm->template at<T>(b.u)));
        } else {
          EpiParams e2 = epi(mm->template at<T>(b.a), C4, b.fc1.b->p, EPI_GELU);
          e2.out2 = mm->template at<T>(b.u);
          S4G_TRY(self->fprop(linear_geom(Tk, C4, C), h2, mm->template at<T>(b.fc1.wf), e2, st));
        }
        EpiParams e3 = epi(x2, C, b.fc2.b->p, EPI_NONE, 1);
        e3.res = x1; e3.res_f32 = 1; e3.ldr = C;
        return self->fprop(linear_geom(Tk, C, C4), mm->template at<T>(b.a), mm->template at<T>(b.fc2.wf), e3, st);
      });
    }
    s->dxa = alloc(sizeof(float) * Tk * C); s->dxb = alloc(sizeof(float) * Tk * C);
    s->dxT = alloc(sizeof(T) * Tk * C); s->dh = alloc(sizeof(T) * Tk * C);
    s->du = alloc(sizeof(T) * Tk * C4); s->dao = alloc(sizeof(T) * Tk * C); s->dqkv = alloc(sizeof(T) * Tk * C3);
    if (pad) { s->daop = alloc(sizeof(T) * Tk * HP); s->dqkvp = alloc(sizeof(T) * Tk * 3 * HP); }
    for (int i = 0; i < depth; ++i) {
      const TBlock b = s->blocks[i];
      const size_t xin = s->xs[2 * i], xmid = s->xs[2 * i + 1];
      float* prev_fc2_bias_g = i > 0 ? s->blocks[i - 1].fc2.b->g : nullptr;
      const bool last = i == depth - 1;
      bl->push_back([=](cudaStream_t st) {
        // incoming: fp32 gradient of the block output in dxa and its T copy in dxT (written by whoever produced dxa); the
        // gradient of the block input is left the same way
        float* dxo = mm->template at<float>(s->dxa);
        float* dxm = mm->template at<float>(s->dxb);
        T* dxt = mm->template at<T>(s->dxT);
        T* dhb = mm->template at<T>(s->dh);
        T* dub = mm->template at<T>(s->du);
        T* daob = mm->template at<T>(s->dao);
        T* dqkvb = mm->template at<T>(s->dqkv);
        mm->launches += 2;
        const Reducer rg = self->red(b.gelu_red, true);
        if (self->umma_path()) {
          EpiParams e = epi(dub, C4, nullptr, EPI_MUL);
          e.aux = mm->template at<T>(b.u);
          S4G_TRY(self->fprop(linear_geom(Tk, C4, C), dxt, mm->template at<T>(b.fc2.wd), e, st));
          mm->launches++;
          if (b.fc1.b->g) S4G_TRY(k_colsum<T>(dub, b.fc1.b->g, Tk, C4, C4, st, &rg));
        } else {
          EpiParams e = epi(dub, C4, nullptr, EPI_DGELU);
          e.aux = mm->template at<T>(b.u);
          S4G_TRY(self->fprop(linear_geom(Tk, C4, C), dxt, mm->template at<T>(b.fc2.wd), e, st));
        }
        S4G_TRY(self->linear_param_grads(b.fc2, dxt, mm->template at<T>(b.a), Tk, st, !last));
        S4G_TRY(self->fprop(linear_geom(Tk, C, C4), dub, mm->template at<T>(b.fc1.wd), epi(dhb, C), st));
        S4G_TRY(self->linear_param_grads(b.fc1, dub, mm->template at<T>(b.h2), Tk, st, self->umma_path()));
        const Reducer r2 = self->red(b.ln2_red, true), r1 = self->red(b.ln1_red, true);
        S4G_TRY(k_layernorm_bwd<T>(dhb, mm->template at<float>(xmid), b.n2w->p, mm->template at<float>(b.mean2), mm->template at<float>(b.rstd2), dxo, dxm, b.n2w->g, b.n2b->g, (int)Tk, C, st, dxt, b.proj.b->g, &r2));
        S4G_TRY(self->fprop(linear_geom(Tk, C, C), dxt, mm->template at<T>(b.proj.wd), epi(daob, C), st));
        S4G_TRY(self->linear_param_grads(b.proj, dxt, mm->template at<T>(b.ao), Tk, st, true));
        if (pad) {
          T* dop = mm->template at<T>(s->daop);
          T* dqp = mm->template at<T>(s->dqkvp);
          mm->launches += 2;
          S4G_TRY(k_head_pad<T>(daob, dop, Tk, H, 0, 1.f, st));
          S4G_TRY(self->attn_bwd_g(mm->template at<T>(b.qkvp), mm->template at<T>(b.aop), dop, mm->template at<float>(b.lse), dqp, N, H, st));
          S4G_TRY(k_head_unpad<T>(dqp, dqkvb, Tk, 3 * H, H, 1.41421356237309505f, st));
        } else {
          S4G_TRY(self->attn_bwd_g(mm->template at<T>(b.qkvb), mm->template at<T>(b.ao), daob, mm->template at<float>(b.lse), dqkvb, N, H, st));
        }
        S4G_TRY(self->fprop(linear_geom(Tk, C, C3), dqkvb, mm->template at<T>(b.qkv.wd), epi(dhb, C), st));
        S4G_TRY(self->linear_param_grads(b.qkv, dqkvb, mm->template at<T>(b.h1), Tk, st));
        return k_layernorm_bwd<T>(dhb, mm->template at<float>(xin), b.n1w->p, mm->template at<float>(b.mean1), mm->template at<float>(b.rstd1), dxm, dxo, b.n1w->g, b.n1b->g, (int)Tk, C, st, dxt, prev_fc2_bias_g, &r1);
      });
    }
    return s;
  }
  std::vector<std::unique_ptr<TStack>> stacks;

  // ---- MAE pretraining step: MaskedAutoencoderViT.forward up to `pred` (Models/mae/models_mae.py:150-196) ----------------
  // out = pred [B, 196, 768] (fp32); the masked norm-pix loss (:198-214) is its own entry point (k_mae_loss_*).
  void build_mae() {
    Builder* self = this;
    S4GModel* mm = m;
    constexpr int L = 196, KEEP = 49, CE = 768, CD = 512;   // int(L * (1 - 0.75)) = 49 (:130)
    const long long Pk = (long long)B * KEEP, Te = (long long)B * (KEEP + 1), Td = (long long)B * (L + 1);
    const size_t ids_restore = alloc(sizeof(int) * (size_t)B * L), ids_keep = alloc(sizeof(int) * (size_t)Pk);
    const size_t mask = alloc(sizeof(float) * (size_t)B * L);
    m->mae_mask_off = mask; m->mae_ids_restore_off = ids_restore;
    patches = alloc(sizeof(T) * Pk * CE);
    pe = alloc(sizeof(float) * Pk * CE);
    pe_lin = linear("patch_embed.proj", CE, CE, false);
    note_wgrad(linear_geom(Pk, CE, CE));
    Param* cls = P("cls_token");
    Param* pos = P("pos_embed");
    Param* dpos = P("decoder_pos_embed");
    Param* mtok = P("mask_token");
    // ---------------- forward ----------------
    const size_t fwd_head = m->fwd.size();
    m->fwd.push_back(nullptr);  // filled below once the encoder stack (its xs[0]) exists
    bwd = &m->bwd_enc;
    m->bwd_enc.push_back(nullptr);  // [0] = patch embed / token assembly backward, filled below
    TStack* enc = tstack("blocks.", 12, CE, 12, KEEP + 1, &m->bwd_enc);
    m->fwd[fwd_head] = [=](cudaStream_t st) {
      if (!mm->mae_noise) { s4g_set_error("MAE forward: s4g_model_set_mae_io(noise, ...) has not been called"); return (int)S4G_ERR_STATE; }
      int* keep = mm->template at<int>(ids_keep);
      mm->launches += 4;
      S4G_TRY(k_mae_masking(mm->mae_noise, mm->template at<int>(ids_restore), keep, mm->template at<float>(mask), self->B, L, KEEP, st));
      S4G_TRY(k_mae_patchify_kept<T>(mm->images, keep, mm->template at<T>(self->patches), self->B, KEEP, st));
      EpiParams e = epi(mm->template at<float>(self->pe), CE, self->pe_lin.b->p, EPI_NONE, 1);
      S4G_TRY(self->fprop(linear_geom(Pk, CE, CE), mm->template at<T>(self->patches), mm->template at<T>(self->pe_lin.wf), e, st));
      return k_mae_assemble(mm->template at<float>(self->pe), cls->p, pos->p, keep, mm->template at<float>(enc->xs[0]), self->B, KEEP, CE, st);
    };
    m->bwd_enc[0] = [=](cudaStream_t st) {
      T* dpe = mm->template at<T>(enc->dxT);   // [B*49, 768] fits in the [B*50, 768] buffer
      mm->launches++;
      S4G_TRY(k_mae_assemble_bwd<T>(mm->template at<float>(enc->dxa), dpe, cls->g, self->B, KEEP, CE, st));
      return self->linear_param_grads(self->pe_lin, dpe, mm->template at<T>(self->patches), Pk, st);
    };
    // encoder norm -> decoder_embed -> mask tokens + un-shuffle + decoder pos (:173-185)
    bwd = &m->bwd_dec;
    Param* nw = P("norm.weight");
    Param* nb = P("norm.bias");
    const Lin dembed = linear("decoder_embed", CD, CE);
    note_wgrad(linear_geom(Te, CD, CE));
    const size_t hn = alloc(sizeof(T) * Te * CE), nmean = alloc(sizeof(float) * Te), nrstd = alloc(sizeof(float) * Te);
    const size_t de = alloc(sizeof(float) * Te * CD), dde = alloc(sizeof(T) * Te * CD), dhn = alloc(sizeof(T) * Te * CE);
    const RedSlot n_red = red_alloc(ln_bwd_scratch_bytes(CE)), mt_red = red_alloc(mae_unshuffle_scratch_bytes(B, CD));
    const size_t fwd_mid = m->fwd.size();
    m->fwd.push_back(nullptr);
    const size_t bwd_mid = m->bwd_dec.size();
    m->bwd_dec.push_back(nullptr);
    TStack* dec = tstack("decoder_blocks.", 8, CD, 16, L + 1, &m->bwd_dec);
    const size_t x_enc_out = enc->xs[2 * enc->depth];
    m->fwd[fwd_mid] = [=](cudaStream_t st) {
      mm->launches += 2;
      S4G_TRY(k_layernorm_fwd<T>(mm->template at<float>(x_enc_out), nw->p, nb->p, mm->template at<T>(hn), mm->template at<float>(nmean),
                                 mm->template at<float>(nrstd), (int)Te, CE, 1e-6f, st));
      EpiParams e = epi(mm->template at<float>(de), CD, dembed.b->p, EPI_NONE, 1);
      S4G_TRY(self->fprop(linear_geom(Te, CD, CE), mm->template at<T>(hn), mm->template at<T>(dembed.wf), e, st));
      return k_mae_unshuffle(mm->template at<float>(de), mtok->p, dpos->p, mm->template at<int>(ids_restore),
                             mm->template at<float>(dec->xs[0]), self->B, L, KEEP, CD, st);
    };
    m->bwd_dec[bwd_mid] = [=](cudaStream_t st) {
      T* ddeb = mm->template at<T>(dde);
      T* dhnb = mm->template at<T>(dhn);
      mm->launches += 2;
      const Reducer rm = self->red(mt_red, true), rn = self->red(n_red, true);
      S4G_TRY(k_mae_unshuffle_bwd<T>(mm->template at<float>(dec->dxa), mm->template at<int>(ids_restore), ddeb, mtok->g, self->B, L, KEEP, CD, st, &rm));
      S4G_TRY(self->fprop(linear_geom(Te, CE, CD), ddeb, mm->template at<T>(dembed.wd), epi(dhnb, CE), st));
      S4G_TRY(self->linear_param_grads(dembed, ddeb, mm->template at<T>(hn), Te, st));
      // gradient of the encoder output: fp32 in enc->dxa with its T copy in enc->dxT, as the block backward expects
      return k_layernorm_bwd<T>(dhnb, mm->template at<float>(x_enc_out), nw->p, mm->template at<float>(nmean), mm->template at<float>(nrstd),
                                nullptr, mm->template at<float>(enc->dxa), nw->g, nb->g, (int)Te, CE, st, mm->template at<T>(enc->dxT), nullptr, &rn);
    };
    // decoder_norm -> decoder_pred -> drop cls (:190-194)
    Param* dnw = P("decoder_norm.weight");
    Param* dnb = P("decoder_norm.bias");
    const Lin pred = linear("decoder_pred", CE, CD);
    note_wgrad(linear_geom(Td, CE, CD));
    const size_t hd = alloc(sizeof(T) * Td * CD), dmean = alloc(sizeof(float) * Td), drstd = alloc(sizeof(float) * Td);
    const size_t pfull = alloc(sizeof(float) * Td * CE), dpf = alloc(sizeof(T) * Td * CE), dhd = alloc(sizeof(T) * Td * CD);
    const RedSlot dn_red = red_alloc(ln_bwd_scratch_bytes(CD));
    const size_t x_dec_out = dec->xs[2 * dec->depth];
    m->fwd.push_back([=](cudaStream_t st) {
      mm->launches += 2;
      S4G_TRY(k_layernorm_fwd<T>(mm->template at<float>(x_dec_out), dnw->p, dnb->p, mm->template at<T>(hd), mm->template at<float>(dmean),
                                 mm->template at<float>(drstd), (int)Td, CD, 1e-6f, st));
      EpiParams e = epi(mm->template at<float>(pfull), CE, pred.b->p, EPI_NONE, 1);
      S4G_TRY(self->fprop(linear_geom(Td, CE, CD), mm->template at<T>(hd), mm->template at<T>(pred.wf), e, st));
      return k_drop_cls(mm->template at<float>(pfull), mm->out, self->B, L, CE, st);
    });
    m->bwd_dec.push_back([=](cudaStream_t st) {
      T* dpfb = mm->template at<T>(dpf);
      T* dhdb = mm->template at<T>(dhd);
      mm->launches += 2;
      S4G_TRY(k_add_cls_row<T>(mm->dout, dpfb, self->B, L, CE, st));
      S4G_TRY(self->fprop(linear_geom(Td, CD, CE), dpfb, mm->template at<T>(pred.wd), epi(dhdb, CD), st));
      S4G_TRY(self->linear_param_grads(pred, dpfb, mm->template at<T>(hd), Td, st));
      const Reducer rn = self->red(dn_red, true);
      return k_layernorm_bwd<T>(dhdb, mm->template at<float>(x_dec_out), dnw->p, mm->template at<float>(dmean), mm->template at<float>(drstd),
                                nullptr, mm->template at<float>(dec->dxa), dnw->g, dnb->g, (int)Td, CD, st, mm->template at<T>(dec->dxT), nullptr, &rn);
    });
  }

  int finish() {
    if (!lin_jobs.empty()) {
      Builder* self = this;
      S4GModel* mm = m;
      // first in the pack list: conv_s2 re-packs through the same kernel family afterwards.  80 layers per grouped launch.
      m->pack.insert(m->pack.begin(), [self, mm](cudaStream_t st) {
        for (size_t first = 0; first < self->lin_jobs.size(); first += 80) {
          PackLinearJobs jobs;
          int tiles = 0, n = 0;
          for (size_t k = first; k < self->lin_jobs.size() && k < first + 80; ++k) {
            const Lin& l = self->lin_jobs[k];
            PackLinearJob& j = jobs.job[n++];
            j.src = l.w->p; j.wp = mm->template at<T>(l.wf); j.wt = l.wd ? (void*)mm->template at<T>(l.wd) : nullptr;
            j.N = l.N; j.K = l.K; j.tile0 = tiles; j.pad = 0;
            tiles += cdiv(l.N, PACK_TILE) * cdiv(l.K, PACK_TILE);
          }
          jobs.n = n;
          mm->launches++;
          S4G_TRY(k_pack_linear_grouped<T>(jobs, tiles, st));
        }
        return 0;
      });
    }
    if (!conv_jobs.empty()) {
      if (conv_jobs.size() > 80) { s4g_set_error("internal: more than 80 3x3 convolutions"); return S4G_ERR_STATE; }
      Builder* self = this;
      S4GModel* mm = m;
      m->pack.push_back([self, mm](cudaStream_t st) {
        PackLinearJobs jobs;
        int blocks = 0, n = 0;
        for (const Conv& c : self->conv_jobs) {
          PackLinearJob& j = jobs.job[n++];
          j.src = c.w->p; j.wp = mm->template at<T>(c.wf); j.wt = mm->template at<T>(c.wd);
          j.N = c.Co; j.K = c.Ci; j.tile0 = blocks; j.pad = 0;
          blocks += c.Co + c.Ci;
        }
        jobs.n = n;
        mm->launches++;
        return k_pack_conv3_grouped<T>(jobs, blocks, st);
      });
    }
    dwp_scratch = alloc(dwp_bytes);
    wg_scratch = alloc(wg_bytes);
    m->ws_need = top + 256;
    if (!missing.empty()) {
      std::string s = "unbound state_dict entries:";
      for (size_t i = 0; i < missing.size() && i < 8; ++i) s += " " + missing[i];
      s4g_set_error(s);
      return S4G_ERR_NAME;
    }
    return 0;
  }
};

// Builders are kept alive with the model (closures hold `self`).
struct BuilderHolder {
  std::unique_ptr<Builder<float>> f;
  std::unique_ptr<Builder<bf16>> h;
};
std::map<S4GModel*, BuilderHolder> g_builders;

template <typename T>
int build(S4GModel* m, std::unique_ptr<Builder<T>>& slot) {
  slot.reset(new Builder<T>(m));
  Builder<T>& b = *slot;
  b.bwd = &m->bwd_dec;
  if (m->d.dense == S4G_DENSE_MAE) {
    b.build_mae();
    return b.finish();
  }
  b.build_encoder(m->d.dense != S4G_DENSE_NONE);
  if (m->d.dense != S4G_DENSE_NONE) b.build_decoder();
  else b.build_cls_head();
  return b.finish();
}

int ensure_built(S4GModel* m) {
  if (m->built) return 0;
  m->pack.clear(); m->fwd.clear(); m->bwd_dec.clear(); m->bwd_enc.clear(); m->acts.clear();
  m->bn_count = 0;
  int r;
  if (m->d.dtype == S4G_DTYPE_F32) r = build<float>(m, g_builders[m].f);
  else r = build<bf16>(m, g_builders[m].h);
  if (r == 0) m->built = true;
  return r;
}

}  // namespace

extern "C" {

int s4g_model_create(const S4GModelDesc* desc, S4GModel** out) {
  if (!desc || !out || desc->batch <= 0) { s4g_set_error("model_create: bad descriptor"); return S4G_ERR_SHAPE; }
  if (desc->dtype == S4G_DTYPE_F32 && desc->backend == S4G_BACKEND_TENSOR) {
    s4g_set_error("model_create: the tensor-core backend is bf16 only (fp32 is the FFMA validation mode)");
    return S4G_ERR_STATE;
  }
  S4GModel* m = new S4GModel();
  m->d = *desc;
  if (m->d.world_size < 1) m->d.world_size = 1;
  *out = m;
  return 0;
}
void s4g_model_destroy(S4GModel* m) {
  if (!m) return;
  g_builders.erase(m);
  delete m;
}
int s4g_model_bind(S4GModel* m, const char* name, void* data, void* grad, long long numel) {
  if (!m || !name || !data) { s4g_set_error("model_bind: null argument"); return S4G_ERR_STATE; }
  Param& p = m->params[name];  // node address is stable; closures hold Param*
  p.p = reinterpret_cast<float*>(data);
  p.g = reinterpret_cast<float*>(grad);
  p.n = numel;
  return 0;
}
int s4g_model_set_grad_arena(S4GModel* m, void* grads, size_t bytes) {
  m->grad_arena = grads; m->grad_arena_bytes = bytes;
  return 0;
}
size_t s4g_model_workspace_bytes(S4GModel* m) {
  if (ensure_built(m) != 0) return 0;
  return m->ws_need;
}
int s4g_model_set_workspace(S4GModel* m, void* workspace, size_t bytes) {
  S4G_TRY(ensure_built(m));
  if (bytes < m->ws_need) { s4g_set_error("model_set_workspace: workspace too small"); return S4G_ERR_WORKSPACE; }
  if (reinterpret_cast<uintptr_t>(workspace) & 255) { s4g_set_error("model_set_workspace: need 256 B alignment"); return S4G_ERR_ALIGN; }
  m->ws = reinterpret_cast<char*>(workspace);
  m->ws_bytes = bytes;
  return 0;
}
int s4g_model_set_peer_exchange(S4GModel* m, void* const* mailboxes, int rank, int world) {
  if (!m || !mailboxes || world < 1 || world > S4G_PEER_MAX_WORLD || rank < 0 || rank >= world) {
    s4g_set_error("model_set_peer_exchange: bad argument (world <= 8)"); return S4G_ERR_STATE;
  }
  for (int r = 0; r < world; ++r) m->peer_boxes[r] = mailboxes[r];
  m->peer_rank = rank; m->peer_world = world;
  return 0;
}
int s4g_model_set_allreduce(S4GModel* m, s4g_allreduce_fn fn, void* user) {
  m->allreduce = fn; m->allreduce_user = user;
  return 0;
}
int s4g_model_set_bucket_events(S4GModel* m, void** events, int n) {
  m->events.clear();
  for (int i = 0; i < n; ++i) m->events.push_back(reinterpret_cast<cudaEvent_t>(events[i]));
  return 0;
}
static int bucket_of_block(int block, int n_buckets) {
  // bucket 0 = decoder; encoder blocks 11..0 spread over buckets 1..n-1 in backward order
  const int ne = n_buckets - 1;
  if (ne <= 0) return 0;
  const int rev = 11 - block;  // 0 = first finished
  return 1 + rev * ne / 12;
}
const char* s4g_model_bucket_params(S4GModel* m, int bucket, int n_buckets) {
  m->bucket_str.clear();
  for (auto& kv : m->params) {
    const std::string& name = kv.first;
    if (!kv.second.g) continue;
    int bkt;
    if (name.compare(0, 7, "decoder") == 0 || name.compare(0, 8, "lin_head") == 0 || name.compare(0, 5, "norm.") == 0 ||
        name == "mask_token") bkt = 0;   // decoder. (DPT) | decoder_blocks. / decoder_embed / decoder_norm / decoder_pred (MAE)
    else if (name.compare(0, 7, "blocks.") == 0) bkt = bucket_of_block(atoi(name.c_str() + 7), n_buckets);
    else bkt = n_buckets - 1;
    if (bkt == bucket) { m->bucket_str += name; m->bucket_str += '\n'; }
  }
  return m->bucket_str.c_str();
}
int s4g_model_last_launches(S4GModel* m) { return m->launches; }

int s4g_model_forward(S4GModel* m, const float* images, float* out, int training, uint64_t seed,
                      const uint8_t* dropout_mask, void* stream) {
  S4G_TRY(ensure_built(m));
  if (!m->ws) { s4g_set_error("model_forward: workspace not set"); return S4G_ERR_STATE; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  m->images = images; m->out = out; m->training = training; m->seed = seed; m->dropout_mask = dropout_mask;
  m->launches = 0;
  for (auto& op : m->pack) S4G_TRY(op(st));
  for (auto& op : m->fwd) S4G_TRY(op(st));
  if (m->d.dense == S4G_DENSE_MAE) {
    const size_t n = (size_t)m->d.batch * 196;
    if (m->mae_mask_out) S4G_CUDA(cudaMemcpyAsync(m->mae_mask_out, m->at<float>(m->mae_mask_off), sizeof(float) * n, cudaMemcpyDeviceToDevice, st));
    if (m->mae_ids_restore_out) S4G_CUDA(cudaMemcpyAsync(m->mae_ids_restore_out, m->at<int>(m->mae_ids_restore_off), sizeof(int) * n, cudaMemcpyDeviceToDevice, st));
  }
  return 0;
}
int s4g_model_set_seed_device(S4GModel* m, const unsigned long long* seed_dev) {
  if (!m) { s4g_set_error("model_set_seed_device: null model"); return S4G_ERR_STATE; }
  m->seed_dev = seed_dev;
  return 0;
}
int s4g_model_set_mae_io(S4GModel* m, const float* noise, float* mask_out, int* ids_restore_out) {
  if (!m || m->d.dense != S4G_DENSE_MAE) { s4g_set_error("model_set_mae_io: not an S4G_DENSE_MAE model"); return S4G_ERR_STATE; }
  m->mae_noise = noise; m->mae_mask_out = mask_out; m->mae_ids_restore_out = ids_restore_out;
  return 0;
}

int s4g_model_backward(S4GModel* m, const float* dout, void* stream) {
  if (!m->built || !m->ws) { s4g_set_error("model_backward: forward has not run"); return S4G_ERR_STATE; }
  if (!m->training) { s4g_set_error("model_backward: last forward was not in training mode"); return S4G_ERR_STATE; }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  m->dout = dout;
  m->launches = 0;
  const int nb = (int)m->events.size();
  m->pending.n = 0; m->pending.total_blocks = 0;
  for (size_t i = m->bwd_dec.size(); i-- > 0;) S4G_TRY(m->bwd_dec[i](st));
  // one bucket: it holds EVERY parameter (bucket_of_block() == 0), so its event is the end of backward, below
  if (nb > 1) { S4G_TRY(m->flush_pending(st)); S4G_CUDA(cudaEventRecord(m->events[0], st)); }
  // bwd_enc[0] = patch-embed, bwd_enc[1 + i] = block i
  for (size_t i = m->bwd_enc.size(); i-- > 0;) {
    S4G_TRY(m->bwd_enc[i](st));
    if (nb > 1 && i >= 1) {
      const int block = (int)i - 1;
      const int bkt = bucket_of_block(block, nb);
      const bool last_of_bucket = block == 0 ? false : bucket_of_block(block - 1, nb) != bkt;
      if (last_of_bucket && bkt < nb - 1) { S4G_TRY(m->flush_pending(st)); S4G_CUDA(cudaEventRecord(m->events[bkt], st)); }
    }
  }
  S4G_TRY(m->flush_pending(st));
  if (nb > 0) S4G_CUDA(cudaEventRecord(m->events[nb - 1], st));
  return 0;
}

}  // extern "C"
