// Training step of the PFPN engine (BASELINE.json config 4): what one batch of
// `training_model.fit_generator(...)` executes for the model compiled in PyraPose/bin/train.py:95-102 —
//   forward (engine.cu plan, raw '3Dbox' + sigmoid 'cls' / 'mask' outputs, PyraPose/models/retinanet.py:260-299)
//   losses  focal (cls, mask) + orthogonal_l1 (3Dbox) + the l2(0.001) kernel regulariser of the regression sub-model
//           (PyraPose/losses.py:22-68, 321-408; retinanet.py:101-131)
//   backward through every trainable convolution: data gradients are the SAME tcgen05 implicit-GEMM kernel run with
//           transposed weights and negated tap offsets (ReLU mask + skip-connection add in its epilogue), weight
//           gradients are conv_wgrad.cu; frozen parts (BatchNorm everywhere, conv1 + stage 2: the non-BN layers with
//           index < 40, PyraPose/models/resnet.py:100-103) get no gradient
//   update  keras.optimizers.Adam(lr, clipnorm) (Keras 2.3.1: global-norm clipping)
// Gradients live in ONE flat fp32 buffer so a data-parallel caller all-reduces them with a single NCCL call between
// ppb_engine_train_forward_backward and ppb_engine_train_apply.
#include "conv_wgrad.h"
#include "engine.h"
#include "train_kernels.h"

#include <algorithm>

namespace ppb {

enum TOpKind { kTConv, kTWgrad, kTBias, kTCombine, kTCount, kTFocal, kTOrth };

struct TOp {
  TOpKind kind;
  CUtensorMap ta, tb, to, tr;  // kTConv
  ConvParams p;
  CUtensorMap tg, tx;          // kTWgrad
  WgradParams wp;
  const void* in[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  void* out = nullptr;
  long long n[2] = {0, 0};
  int a[12] = {0};
  float f[4] = {0.f, 0.f, 0.f, 0.f};
  HeadGeom geom;
  char name[56] = {0};
  double flops = 0.0;
};

struct TrainLayer {
  bool trainable = false;
  bool need_dgrad = false;
  long long off_kernel = -1, off_bias = -1;
  float* scale = nullptr;            // device BatchNorm scale [cout] (null without BN)
  __nv_bfloat16* w_dgrad = nullptr;  // [cin][n_taps * g_ld]
  int g_ld = 0;
  int n_taps = 0;
  int tap_pos[9] = {0, 1, 2, 3, 4, 5, 6, 7, 8};
  int plane_tap_begin[5] = {0, 0, 0, 0, 0};  // stride-2 3x3 layers: taps grouped by input parity plane
  int last_grad_op = -1;  // index in TrainState::bwd of the last op that writes this layer's gradient (wgrad / bgrad)
};

// One CUDA graph of the step and the slice of the flat gradient buffer that is final when it has run. The step is cut
// after the op that completes each gradient bucket, so a data-parallel caller can start the NCCL all-reduce of bucket
// k on its communication stream while segment k + 1 (the rest of the backward pass) runs.
struct Segment {
  int op_begin = 0, op_end = 0;  // range of TrainState::bwd ops (segment 0 also holds the memsets + forward pass)
  long long g_off = 0, g_cnt = 0;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int kernels = 0;
};

struct TrainState {
  std::vector<TrainLayer> tl;
  long long n_params = 0;
  float* grads = nullptr;
  float* master = nullptr;
  float* m = nullptr;
  float* v = nullptr;
  double* loss = nullptr;  // [8] LossSlot
  int* counts = nullptr;   // [4]
  std::vector<ParamDesc> descs;
  ParamDesc* descs_dev = nullptr;
  int* chunk_desc = nullptr;
  int* chunk_index = nullptr;
  int n_chunks = 0;
  std::vector<TOp> bwd;
  std::vector<TOp> count_ops;  // positives per output: depend on the targets only, launched ahead of the graphs
  std::vector<Segment> segs;
  long long bucket_bytes = 0;  // 0: one segment = the whole step
  bool counts_ready = false;   // ppb_engine_train_count ran for this step (the caller may have all-reduced them)
  int bound_batch = 0;
  long long step = 0;
  int graph_kernels = 0;
  void drop_graphs() {
    for (Segment& sg : segs) {
      if (sg.exec) cudaGraphExecDestroy(sg.exec);
      if (sg.graph) cudaGraphDestroy(sg.graph);
      sg.exec = nullptr;
      sg.graph = nullptr;
    }
  }
  ~TrainState() { drop_graphs(); }
};

namespace {

#define PPB_TRY(expr)             \
  do {                            \
    int rc__ = (expr);            \
    if (rc__ != kOk) return rc__; \
  } while (0)

bool layer_trainable(const std::string& name) {
  // resnet.py:100-103: layers with index < 40 that are not BatchNorm are frozen = conv1 and every conv of stage 2
  // (keras-resnet 0.1.0 layer order: input, padding_conv1, conv1, bn_conv1, conv1_relu, pool1, then 13 + 11 + 11
  // layers for res2a/b/c; index 39 is the res2c Add). BatchNorm is frozen everywhere (freeze_bn=True, resnet.py:87).
  return !(name == "conv1" || name.compare(0, 4, "res2") == 0);
}

struct Builder {
  Engine& e;
  TrainState& t;
  int batch;

  Tensor* fwd(const std::string& name) { return e.tensor(name); }

  int gtensor(Tensor** out, const std::string& name, int h, int w, int c, int planes = 1, long long rows_override = 0) {
    *out = e.tensor("g:" + name);
    if (*out != nullptr) return kOk;
    return e.new_tensor(out, "g:" + name, h, w, c, 1, planes, rows_override);
  }

  const LayerW& L(const std::string& name) { return e.layers[e.layer_index.at(name)]; }
  TrainLayer& T(const std::string& name) { return t.tl[e.layer_index.at(name)]; }

  // ---- data gradient: conv_igemm_kernel over g with the transposed weights
  int finish_dgrad(TOp& op, const std::string& layer, const __nv_bfloat16* g, long long g_rows, int n_taps_launch,
                   int b_col_base, const __nv_bfloat16* residual, const __nv_bfloat16* mask, __nv_bfloat16* out,
                   long long out_rows, double px) {
    const LayerW& l = L(layer);
    const TrainLayer& tl = T(layer);
    if (tl.w_dgrad == nullptr) return fail(kErrState, "layer '%s' has no dgrad weights", layer.c_str());
    ConvParams& p = op.p;
    p.n_taps = n_taps_launch;
    p.cin_blocks = tl.g_ld / kConvBlockK;
    p.n_valid = l.cin;
    p.block_n = conv_pick_block_n(l.cin, p.n_taps * p.cin_blocks, p.n_mtiles, true);
    p.n_ntiles = (l.cin + p.block_n - 1) / p.block_n;
    p.epi = kEpiBf16;
    p.relu = 0;
    p.bias = nullptr;
    p.residual = residual;
    p.res_ld = l.cin;
    p.mask = mask;
    p.mask_ld = l.cin;
    p.b_col_base = b_col_base;
    p.out = out;
    p.out_ld = l.cin;
    p.pair = conv_pair_ok(p) ? 1 : 0;
    p.staged = (l.cin % 64 == 0 && p.block_n % 64 == 0) ? 1 : 0;
    p.slab3 = conv_slab3_mode(p);
    int rc = conv_encode_tmaps(&op.ta, &op.tb, g, g_rows, tl.g_ld, tl.g_ld, tl.w_dgrad, l.cin, tl.n_taps * tl.g_ld,
                               p.pair ? p.block_n / 2 : p.block_n, p.slab3 ? kConvSlabRows : kConvBlockM);
    if (rc != 0) return fail(kErrCuda, "cuTensorMapEncodeTiled failed for dgrad of '%s' (%d)", layer.c_str(), rc);
    memset(&op.to, 0, sizeof(op.to));
    memset(&op.tr, 0, sizeof(op.tr));
    p.staged = (l.cin % 64 == 0 && p.block_n % 64 == 0) ? 1 : 0;
    if (p.staged) {
      rc = conv_encode_rows_tmap(&op.to, out, out_rows, l.cin, l.cin);
      if (rc == 0 && residual != nullptr) rc = conv_encode_rows_tmap(&op.tr, residual, out_rows, l.cin, l.cin);
      if (rc != 0) return fail(kErrCuda, "cuTensorMapEncodeTiled (output) failed for dgrad of '%s' (%d)", layer.c_str(), rc);
    }
    if (conv_plan_smem(p) < 0) return fail(kErrInvalid, "dgrad of '%s' does not fit the shared-memory plan", layer.c_str());
    op.kind = kTConv;
    snprintf(op.name, sizeof(op.name), "dgrad:%s", layer.c_str());
    op.flops = 2.0 * px * l.cout * static_cast<double>(n_taps_launch) * l.cin;
    t.bwd.push_back(op);
    return kOk;
  }

  // stride-1 layer (or a 1x1/2 layer at its output resolution) on one padded-flat plane set
  int dgrad_plane(const std::string& layer, const __nv_bfloat16* g, int H, int W, const __nv_bfloat16* residual,
                  const __nv_bfloat16* mask, __nv_bfloat16* out) {
    const LayerW& l = L(layer);
    TOp op;
    memset(&op.p, 0, sizeof(op.p));
    conv_fill_single_seg(op.p, batch, H, W, l.ksize);
    for (int i = 0; i < op.p.n_taps; ++i) op.p.seg[0].tap_off[i] = -op.p.seg[0].tap_off[i];
    const long long rows = op.p.seg[0].rows;
    return finish_dgrad(op, layer, g, rows, op.p.n_taps, 0, residual, mask, out, rows,
                        static_cast<double>(batch) * H * W);
  }

  int dgrad_head(const std::string& layer, const __nv_bfloat16* g, int n_levels, const __nv_bfloat16* residual,
                 const __nv_bfloat16* mask, __nv_bfloat16* out) {
    TOp op;
    memset(&op.p, 0, sizeof(op.p));
    e.fill_head_segs(op.p, batch, n_levels);
    double px = 0;
    for (int i = 0; i < n_levels; ++i) {
      for (int k = 0; k < 9; ++k) op.p.seg[i].tap_off[k] = -op.p.seg[i].tap_off[k];
      px += static_cast<double>(batch) * e.lvl_h[i] * e.lvl_w[i];
    }
    const long long last_row = op.p.seg[n_levels - 1].row_begin + op.p.seg[n_levels - 1].rows;
    return finish_dgrad(op, layer, g, last_row, 9, 0, residual, mask, out, last_row, px);
  }

  // 3x3/2 'same' layer: the gradient of each of the four input parity planes is its own convolution over g
  int dgrad_s2planes(const std::string& layer, const __nv_bfloat16* g, int H, int W, int src_h, int src_w,
                     __nv_bfloat16* planes_out) {
    const LayerW& l = L(layer);
    const TrainLayer& tl = T(layer);
    const int pt = same_pad_before(src_h, 3, 2), pl = same_pad_before(src_w, 3, 2);
    for (int plane = 0; plane < 4; ++plane) {
      TOp op;
      memset(&op.p, 0, sizeof(op.p));
      conv_fill_single_seg(op.p, batch, H, W, 1);
      ConvSeg& s = op.p.seg[0];
      int nt = 0;
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
          const int u = r - pt, v = c - pl;
          const int ph = ((u % 2) + 2) % 2, pw = ((v % 2) + 2) % 2;
          if (ph * 2 + pw != plane) continue;
          const int dy = (u - ph) / 2, dx = (v - pw) / 2;
          if (tl.tap_pos[r * 3 + c] != tl.plane_tap_begin[plane] + nt)
            return fail(kErrState, "internal: dgrad tap order of '%s'", layer.c_str());
          s.tap_off[nt++] = -(dy * s.wp + dx);
        }
      const long long rows = s.rows;
      __nv_bfloat16* out = planes_out + static_cast<size_t>(plane) * rows * l.cin;
      if (nt == 0) return fail(kErrState, "internal: empty parity plane in '%s'", layer.c_str());
      PPB_TRY(finish_dgrad(op, layer, g, rows, nt, tl.plane_tap_begin[plane] * tl.g_ld, nullptr, nullptr, out, rows,
                           static_cast<double>(batch) * H * W));
    }
    return kOk;
  }

  // ---- weight (+ bias) gradient
  int finish_wgrad(TOp& op, const std::string& layer, const __nv_bfloat16* g, long long g_rows, int g_ld,
                   const __nv_bfloat16* x, long long x_rows, double px) {
    const LayerW& l = L(layer);
    const TrainLayer& tl = T(layer);
    if (!tl.trainable) return kOk;
    WgradParams& p = op.wp;
    p.n_taps = l.ksize * l.ksize;
    p.cout = l.cout;
    p.cin = l.cin;
    p.dw = t.grads + tl.off_kernel;
    p.dw_ld = static_cast<long long>(p.n_taps) * l.cin;
    wgrad_plan(p);
    const int rc = wgrad_encode_tmaps(&op.tg, &op.tx, g, g_rows, g_ld, g_ld, x, x_rows, l.cin, l.cin);
    if (rc != 0) return fail(kErrCuda, "cuTensorMapEncodeTiled failed for wgrad of '%s' (%d)", layer.c_str(), rc);
    op.kind = kTWgrad;
    snprintf(op.name, sizeof(op.name), "wgrad:%s", layer.c_str());
    op.flops = 2.0 * px * l.cout * static_cast<double>(p.n_taps) * l.cin;
    t.bwd.push_back(op);
    if (l.has_bias) {
      TOp b;
      b.kind = kTBias;
      b.in[0] = g;
      b.out = t.grads + tl.off_bias;
      b.n[0] = g_rows;
      b.a[0] = g_ld;
      b.a[1] = l.cout;
      snprintf(b.name, sizeof(b.name), "bgrad:%s", layer.c_str());
      t.bwd.push_back(b);
    }
    T(layer).last_grad_op = static_cast<int>(t.bwd.size()) - 1;
    return kOk;
  }

  int wgrad_plane(const std::string& layer, const __nv_bfloat16* g, const __nv_bfloat16* x, int H, int W) {
    const LayerW& l = L(layer);
    ConvParams cp;
    memset(&cp, 0, sizeof(cp));
    conv_fill_single_seg(cp, batch, H, W, l.ksize);
    TOp op;
    memset(&op.wp, 0, sizeof(op.wp));
    op.wp.n_segs = 1;
    op.wp.seg[0].row_begin = 0;
    op.wp.seg[0].k_iters = (cp.seg[0].rows + 63) / 64;
    for (int i = 0; i < cp.n_taps; ++i) op.wp.seg[0].tap_off[i] = cp.seg[0].tap_off[i];
    op.wp.total_k_iters = op.wp.seg[0].k_iters;
    return finish_wgrad(op, layer, g, cp.seg[0].rows, T(layer).g_ld, x, cp.seg[0].rows,
                        static_cast<double>(batch) * H * W);
  }

  int wgrad_head(const std::string& layer, const __nv_bfloat16* g, const __nv_bfloat16* x, int n_levels) {
    ConvParams cp;
    memset(&cp, 0, sizeof(cp));
    e.fill_head_segs(cp, batch, n_levels);
    TOp op;
    memset(&op.wp, 0, sizeof(op.wp));
    op.wp.n_segs = n_levels;
    double px = 0;
    for (int i = 0; i < n_levels; ++i) {
      op.wp.seg[i].row_begin = cp.seg[i].row_begin;
      op.wp.seg[i].k_iters = (cp.seg[i].rows + 63) / 64;
      for (int k = 0; k < 9; ++k) op.wp.seg[i].tap_off[k] = cp.seg[i].tap_off[k];
      op.wp.total_k_iters += op.wp.seg[i].k_iters;
      px += static_cast<double>(batch) * e.lvl_h[i] * e.lvl_w[i];
    }
    const long long last_row = cp.seg[n_levels - 1].row_begin + cp.seg[n_levels - 1].rows;
    return finish_wgrad(op, layer, g, last_row, T(layer).g_ld, x, last_row, px);
  }

  // 3x3/2 layer: x is the 4-plane tensor, taps carry the plane offsets exactly as in the forward plan (engine.cu)
  int wgrad_s2planes(const std::string& layer, const __nv_bfloat16* g, const __nv_bfloat16* x_planes, int H, int W,
                     int src_h, int src_w) {
    ConvParams cp;
    memset(&cp, 0, sizeof(cp));
    conv_fill_single_seg(cp, batch, H, W, 3);
    const ConvSeg& s = cp.seg[0];
    const long long plane_rows = s.rows;
    const int pt = same_pad_before(src_h, 3, 2), pl = same_pad_before(src_w, 3, 2);
    TOp op;
    memset(&op.wp, 0, sizeof(op.wp));
    op.wp.n_segs = 1;
    op.wp.seg[0].row_begin = 0;
    op.wp.seg[0].k_iters = (s.rows + 63) / 64;
    op.wp.total_k_iters = op.wp.seg[0].k_iters;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        const int u = r - pt, v = c - pl;
        const int ph = ((u % 2) + 2) % 2, pw = ((v % 2) + 2) % 2;
        const int dy = (u - ph) / 2, dx = (v - pw) / 2;
        op.wp.seg[0].tap_off[r * 3 + c] = static_cast<int>((ph * 2 + pw) * plane_rows) + dy * s.wp + dx;
      }
    return finish_wgrad(op, layer, g, s.rows, T(layer).g_ld, x_planes, plane_rows * 4,
                        static_cast<double>(batch) * H * W);
  }

  void combine(const char* name, const __nv_bfloat16* a, const __nv_bfloat16* b, const __nv_bfloat16* c,
               const __nv_bfloat16* d, int dH, int dW, const __nv_bfloat16* p, int pH, int pW, int n_planes,
               const __nv_bfloat16* mask, __nv_bfloat16* out, int H, int W, int C) {
    TOp op;
    op.kind = kTCombine;
    op.in[0] = a;
    op.in[1] = b;
    op.in[2] = c;
    op.in[3] = d;
    op.in[4] = p;
    op.in[5] = mask;
    op.out = out;
    const int args[9] = {dH, dW, pH, pW, n_planes, batch, H, W, C};
    for (int i = 0; i < 9; ++i) op.a[i] = args[i];
    snprintf(op.name, sizeof(op.name), "combine:%s", name);
    t.bwd.push_back(op);
  }
};

int round64(int v) { return (v + 63) / 64 * 64; }

// Master weights (+ BatchNorm scale, + the transposed bf16 data-gradient matrix) of one trainable layer from the host
// copies Engine::set_layer keeps. Also runs when weights are loaded into an engine that has already trained
// (model.load_weights / set_weights after compile): the Adam moments stay, the master copy follows the new weights.
int seed_layer(Engine& e, TrainState& t, int i) {
  const LayerW& l = e.layers[i];
  TrainLayer& tl = t.tl[i];
  if (!tl.trainable) return kOk;
  PPB_CUDA(cudaMemcpy(t.master + tl.off_kernel, l.raw_kernel.data(), l.raw_kernel.size() * sizeof(float),
                      cudaMemcpyHostToDevice));
  if (l.has_bias)
    PPB_CUDA(cudaMemcpy(t.master + tl.off_bias, l.raw_bias.data(), l.raw_bias.size() * sizeof(float),
                        cudaMemcpyHostToDevice));
  if (tl.scale != nullptr)
    PPB_CUDA(cudaMemcpy(tl.scale, l.bn_scale.data(), l.cout * sizeof(float), cudaMemcpyHostToDevice));
  if (tl.w_dgrad != nullptr) {
    const size_t cols = static_cast<size_t>(tl.n_taps) * tl.g_ld;
    std::vector<__nv_bfloat16> wt(static_cast<size_t>(l.cin) * cols, __float2bfloat16(0.0f));
    for (int co = 0; co < l.cout; ++co) {
      const float sc = l.bn_scale.empty() ? 1.0f : l.bn_scale[co];
      for (int tap = 0; tap < tl.n_taps; ++tap)
        for (int ci = 0; ci < l.cin; ++ci)
          wt[static_cast<size_t>(ci) * cols + static_cast<size_t>(tl.tap_pos[tap]) * tl.g_ld + co] =
              __float2bfloat16(l.raw_kernel[static_cast<size_t>(co) * l.k_total + tap * l.cin + ci] * sc);
    }
    PPB_CUDA(cudaMemcpy(tl.w_dgrad, wt.data(), wt.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  }
  return kOk;
}

// Allocates the optimizer state and the dgrad weights, seeds them from the host copies kept by set_layer.
int create_state(Engine& e) {
  std::shared_ptr<TrainState> st(new TrainState);
  TrainState& t = *st;
  t.tl.resize(e.layers.size());
  long long off = 0;
  for (size_t i = 0; i < e.layers.size(); ++i) {
    const LayerW& l = e.layers[i];
    TrainLayer& tl = t.tl[i];
    if (!l.loaded) return fail(kErrState, "layer '%s' has no weights (call ppb_engine_set_layer first)", l.name.c_str());
    tl.trainable = layer_trainable(l.name);
    tl.g_ld = round64(l.cout);
    tl.n_taps = l.ksize * l.ksize;
    if (!tl.trainable) continue;
    tl.need_dgrad = !(l.name == "res3a_branch2a" || l.name == "res3a_branch1");  // nothing trainable below them
    tl.off_kernel = off;
    off += static_cast<long long>(l.cout) * l.k_total;
    off = (off + 3) / 4 * 4;  // keep every tensor 16-byte aligned (vector atomics in the wgrad epilogue)
    if (l.has_bias) {
      tl.off_bias = off;
      off += (l.cout + 3) / 4 * 4;
    }
    if (l.stride == 2 && l.ksize == 3) {  // group taps by input parity plane (see Builder::dgrad_s2planes)
      const int src_h = l.name == "P3_down" ? e.lvl_h[0] : e.lvl_h[1];
      const int src_w = l.name == "P3_down" ? e.lvl_w[0] : e.lvl_w[1];
      const int pt = same_pad_before(src_h, 3, 2), pl = same_pad_before(src_w, 3, 2);
      int pos = 0;
      for (int plane = 0; plane < 4; ++plane) {
        tl.plane_tap_begin[plane] = pos;
        for (int r = 0; r < 3; ++r)
          for (int c = 0; c < 3; ++c) {
            const int u = r - pt, v = c - pl;
            if ((((u % 2) + 2) % 2) * 2 + (((v % 2) + 2) % 2) == plane) tl.tap_pos[r * 3 + c] = pos++;
          }
      }
      tl.plane_tap_begin[4] = pos;
    }
  }
  t.n_params = off;
  void* p = nullptr;
  PPB_TRY(e.device_alloc(&p, off * sizeof(float)));
  t.grads = static_cast<float*>(p);
  PPB_TRY(e.device_alloc(&p, off * sizeof(float)));
  t.master = static_cast<float*>(p);
  PPB_TRY(e.device_alloc(&p, off * sizeof(float)));
  t.m = static_cast<float*>(p);
  PPB_TRY(e.device_alloc(&p, off * sizeof(float)));
  t.v = static_cast<float*>(p);
  PPB_TRY(e.device_alloc(&p, 8 * sizeof(double)));
  t.loss = static_cast<double*>(p);
  PPB_TRY(e.device_alloc(&p, 4 * sizeof(int)));
  t.counts = static_cast<int*>(p);

  std::vector<int> chunk_desc, chunk_index;
  PPB_CUDA(cudaMemset(t.master, 0, off * sizeof(float)));
  for (size_t i = 0; i < e.layers.size(); ++i) {
    const LayerW& l = e.layers[i];
    TrainLayer& tl = t.tl[i];
    if (!tl.trainable) continue;
    if (!l.bn_scale.empty()) {
      PPB_TRY(e.device_alloc(&p, l.cout * sizeof(float)));
      tl.scale = static_cast<float*>(p);
    }
    if (tl.need_dgrad) {
      const size_t cols = static_cast<size_t>(tl.n_taps) * tl.g_ld;
      PPB_TRY(e.device_alloc(&p, static_cast<size_t>(l.cin) * cols * sizeof(__nv_bfloat16)));
      tl.w_dgrad = static_cast<__nv_bfloat16*>(p);
    }
    PPB_TRY(seed_layer(e, t, static_cast<int>(i)));
    ParamDesc d;
    memset(&d, 0, sizeof(d));
    d.offset = tl.off_kernel;
    d.count = static_cast<long long>(l.cout) * l.k_total;
    d.kind = 0;
    d.cout = l.cout;
    d.cin = l.cin;
    d.n_taps = tl.n_taps;
    d.k_total = l.k_total;
    d.g_ld = tl.g_ld;
    d.scale = tl.scale;
    d.l2 = l.name.compare(0, 20, "pyramid_regression3D") == 0 ? 0.001f : 0.0f;  // retinanet.py:108
    d.w_fwd = l.w;
    d.w_dgrad = tl.w_dgrad;
    for (int k = 0; k < 9; ++k) d.tap_pos[k] = tl.tap_pos[k];
    t.descs.push_back(d);
    if (l.has_bias) {
      ParamDesc b;
      memset(&b, 0, sizeof(b));
      b.offset = tl.off_bias;
      b.count = l.cout;
      b.kind = 1;
      b.cout = l.cout;
      b.k_total = 1;
      b.bias_fwd = l.bias;
      t.descs.push_back(b);
    }
  }
  for (size_t di = 0; di < t.descs.size(); ++di) {
    // kernels: one chunk per 64 x 64 (cout, k) tile so that adam_update_kernel can write the transposed dgrad copy in
    // coalesced runs (never fewer chunks than the contiguous split grad_prepare_kernel uses; its extra ones are empty)
    const ParamDesc& pd = t.descs[di];
    const long long chunks = pd.kind == 0 ? static_cast<long long>((pd.cout + 63) / 64) * ((pd.k_total + 63) / 64)
                                          : (pd.count + kOptChunk - 1) / kOptChunk;
    for (long long c = 0; c < chunks; ++c) {
      chunk_desc.push_back(static_cast<int>(di));
      chunk_index.push_back(static_cast<int>(c));
    }
  }
  t.n_chunks = static_cast<int>(chunk_desc.size());
  PPB_TRY(e.device_alloc(&p, t.descs.size() * sizeof(ParamDesc)));
  t.descs_dev = static_cast<ParamDesc*>(p);
  PPB_CUDA(cudaMemcpy(t.descs_dev, t.descs.data(), t.descs.size() * sizeof(ParamDesc), cudaMemcpyHostToDevice));
  PPB_TRY(e.device_alloc(&p, chunk_desc.size() * sizeof(int)));
  t.chunk_desc = static_cast<int*>(p);
  PPB_CUDA(cudaMemcpy(t.chunk_desc, chunk_desc.data(), chunk_desc.size() * sizeof(int), cudaMemcpyHostToDevice));
  PPB_TRY(e.device_alloc(&p, chunk_index.size() * sizeof(int)));
  t.chunk_index = static_cast<int*>(p);
  PPB_CUDA(cudaMemcpy(t.chunk_index, chunk_index.data(), chunk_index.size() * sizeof(int), cudaMemcpyHostToDevice));
  e.train = st;
  return kOk;
}

int build_backward(Engine& e, int batch, const float* y_reg, const float* y_cls, const float* y_mask,
                   const float* out_reg, const float* out_cls, const float* out_mask) {
  TrainState& t = *e.train;
  t.bwd.clear();
  t.count_ops.clear();
  for (TrainLayer& tl : t.tl) tl.last_grad_op = -1;
  Builder B{e, t, batch};
  const int A = e.num_anchors, C = e.num_classes;
  const long long featrows = e.feat_rows_total;
  const long long p3rows = (static_cast<long long>(e.max_batch) * (e.lvl_h[0] + 2) * (e.lvl_w[0] + 2) + 127) / 128 * 128;
  HeadGeom g3;
  g3.n_levels = 3;
  g3.locs = 0;
  for (int i = 0; i < 3; ++i) {
    g3.h[i] = e.lvl_h[i];
    g3.w[i] = e.lvl_w[i];
    g3.row_begin[i] = e.seg_row_begin[i];
    g3.loc_base[i] = g3.locs;
    g3.locs += e.lvl_h[i] * e.lvl_w[i];
  }
  HeadGeom g1 = g3;
  g1.n_levels = 1;
  g1.locs = e.lvl_h[0] * e.lvl_w[0];
  const long long n_anchors = static_cast<long long>(batch) * g3.locs * A;

  // ---- losses: positives per output, then loss value + dLoss/d(head conv output) in the padded-flat layout
  Tensor *g_reg, *g_cls, *g_msk;
  PPB_TRY(B.gtensor(&g_reg, "reg_out", 0, 0, round64(A * 16), 1, featrows));
  PPB_TRY(B.gtensor(&g_cls, "cls_out", 0, 0, round64(A * C), 1, featrows));
  PPB_TRY(B.gtensor(&g_msk, "mask_out", 0, 0, round64(C), 1, p3rows));
  {
    const float* ys[3] = {y_reg, y_cls, y_mask};
    const long long rows[3] = {n_anchors, n_anchors, static_cast<long long>(batch) * g1.locs};
    const int width[3] = {17, C + 1, C + 1};
    for (int i = 0; i < 3; ++i) {
      TOp op;
      op.kind = kTCount;
      op.in[0] = ys[i];
      op.n[0] = rows[i];
      op.a[0] = width[i];
      op.a[1] = i;
      snprintf(op.name, sizeof(op.name), "count_positive:%d", i);
      t.count_ops.push_back(op);
    }
    TOp r;
    r.kind = kTOrth;
    r.in[0] = out_reg;
    r.in[1] = y_reg;
    r.out = g_reg->ptr;
    r.a[0] = g_reg->C;
    r.a[1] = A;
    r.a[3] = kLossReg;
    r.f[0] = 0.125f;  // orthogonal_l1(weight=0.125, sigma=3.0), train.py:97 with the defaults of losses.py:321
    r.f[1] = 3.0f;
    r.geom = g3;
    snprintf(r.name, sizeof(r.name), "orthogonal_l1");
    t.bwd.push_back(r);
    TOp c;
    c.kind = kTFocal;
    c.in[0] = out_cls;
    c.in[1] = y_cls;
    c.out = g_cls->ptr;
    c.a[0] = g_cls->C;
    c.a[1] = A;
    c.a[2] = C;
    c.a[3] = kLossCls;
    c.f[0] = 0.25f;  // focal(alpha=0.25, gamma=2.0), losses.py:22
    c.f[1] = 2.0f;
    c.geom = g3;
    snprintf(c.name, sizeof(c.name), "focal:cls");
    t.bwd.push_back(c);
    TOp m = c;
    m.in[0] = out_mask;
    m.in[1] = y_mask;
    m.out = g_msk->ptr;
    m.a[0] = g_msk->C;
    m.a[1] = 1;
    m.a[3] = kLossMask;
    m.geom = g1;
    snprintf(m.name, sizeof(m.name), "focal:mask");
    t.bwd.push_back(m);
  }

  // ---- heads
  Tensor* feat = B.fwd("features");
  Tensor *g_feat, *gr[2], *gc[2];
  PPB_TRY(B.gtensor(&g_feat, "features", 0, 0, 256, 1, featrows));
  PPB_TRY(B.gtensor(&gr[0], "reg_tower_a", 0, 0, 512, 1, featrows));
  PPB_TRY(B.gtensor(&gr[1], "reg_tower_b", 0, 0, 512, 1, featrows));
  PPB_TRY(B.gtensor(&gc[0], "cls_tower_a", 0, 0, 256, 1, featrows));
  PPB_TRY(B.gtensor(&gc[1], "cls_tower_b", 0, 0, 256, 1, featrows));
  struct Tower {
    const char* out_layer;
    const char* prefix;
    const char* act;
    Tensor* g_out;
    Tensor** scratch;
    int n_levels;
  };
  const Tower towers[3] = {{"pyramid_regression3D", "pyramid_regression3D_", "reg_tower_t", g_reg, gr, 3},
                           {"pyramid_classification", "pyramid_classification_", "cls_tower_t", g_cls, gc, 3},
                           {"mask_out", "mask_", "mask_tower_t", g_msk, gc, 1}};
  for (int ti = 0; ti < 3; ++ti) {
    const Tower& tw = towers[ti];
    Tensor* act[4];
    for (int i = 0; i < 4; ++i) {
      act[i] = B.fwd(std::string(tw.act) + std::to_string(i));
      if (act[i] == nullptr) return fail(kErrState, "internal: tower activation %s%d missing", tw.act, i);
    }
    const __nv_bfloat16* g = tw.g_out->ptr;
    PPB_TRY(B.wgrad_head(tw.out_layer, g, act[3]->ptr, tw.n_levels));
    PPB_TRY(B.dgrad_head(tw.out_layer, g, tw.n_levels, nullptr, act[3]->ptr, tw.scratch[0]->ptr));
    int cur = 0;
    for (int i = 3; i >= 1; --i) {
      const std::string layer = std::string(tw.prefix) + std::to_string(i);
      PPB_TRY(B.wgrad_head(layer, tw.scratch[cur]->ptr, act[i - 1]->ptr, tw.n_levels));
      PPB_TRY(B.dgrad_head(layer, tw.scratch[cur]->ptr, tw.n_levels, nullptr, act[i - 1]->ptr, tw.scratch[cur ^ 1]->ptr));
      cur ^= 1;
    }
    const std::string first = std::string(tw.prefix) + "0";
    PPB_TRY(B.wgrad_head(first, tw.scratch[cur]->ptr, feat->ptr, tw.n_levels));
    // the three towers all read `features`: the first writes its gradient, the others accumulate into it in place
    PPB_TRY(B.dgrad_head(first, tw.scratch[cur]->ptr, tw.n_levels, ti == 0 ? nullptr : g_feat->ptr, nullptr,
                         g_feat->ptr));
  }

  // ---- __create_sparceFPN backward (forward: engine.cu build_plan / retinanet.py:180-214)
  const int h3 = e.lvl_h[0], w3 = e.lvl_w[0], h4 = e.lvl_h[1], w4 = e.lvl_w[1], h5 = e.lvl_h[2], w5 = e.lvl_w[2];
  const __nv_bfloat16* gf[3];
  for (int i = 0; i < 3; ++i) gf[i] = g_feat->ptr + static_cast<size_t>(e.seg_row_begin[i]) * 256;
  Tensor *gt5f, *gt4f, *gt3f, *gs4, *gs3, *gP4m, *gP3m, *gt3, *gt4, *gP3l, *gP4l, *gP5l;
  PPB_TRY(B.gtensor(&gt5f, "P5_fin", h5, w5, 256));
  PPB_TRY(B.gtensor(&gt4f, "P4_fin", h4, w4, 256));
  PPB_TRY(B.gtensor(&gt3f, "P3_fin", h3, w3, 256));
  PPB_TRY(B.gtensor(&gs4, "P4_mid_planes", h5, w5, 256, 4));
  PPB_TRY(B.gtensor(&gs3, "P3_mid_planes", h4, w4, 256, 4));
  PPB_TRY(B.gtensor(&gP4m, "P4_mid", h4, w4, 256));
  PPB_TRY(B.gtensor(&gP3m, "P3_mid", h3, w3, 256));
  PPB_TRY(B.gtensor(&gt3, "P3_mid_in", h3, w3, 256));
  PPB_TRY(B.gtensor(&gt4, "P4_mid_in", h4, w4, 256));
  PPB_TRY(B.gtensor(&gP3l, "P3_lateral", h3, w3, 256));
  PPB_TRY(B.gtensor(&gP4l, "P4_lateral", h4, w4, 256));
  PPB_TRY(B.gtensor(&gP5l, "P5_lateral", h5, w5, 256));
  PPB_TRY(B.wgrad_plane("P5", gf[2], B.fwd("P5_fin")->ptr, h5, w5));
  PPB_TRY(B.dgrad_plane("P5", gf[2], h5, w5, nullptr, nullptr, gt5f->ptr));
  PPB_TRY(B.wgrad_plane("P4", gf[1], B.fwd("P4_fin")->ptr, h4, w4));
  PPB_TRY(B.dgrad_plane("P4", gf[1], h4, w4, nullptr, nullptr, gt4f->ptr));
  PPB_TRY(B.wgrad_plane("P3", gf[0], B.fwd("P3_fin")->ptr, h3, w3));
  PPB_TRY(B.dgrad_plane("P3", gf[0], h3, w3, nullptr, nullptr, gt3f->ptr));
  PPB_TRY(B.wgrad_s2planes("P4_down", gt5f->ptr, B.fwd("P4_mid_planes")->ptr, h5, w5, h4, w4));
  PPB_TRY(B.dgrad_s2planes("P4_down", gt5f->ptr, h5, w5, h4, w4, gs4->ptr));
  PPB_TRY(B.wgrad_s2planes("P3_down", gt4f->ptr, B.fwd("P3_mid_planes")->ptr, h4, w4, h3, w3));
  PPB_TRY(B.dgrad_s2planes("P3_down", gt4f->ptr, h4, w4, h3, w3, gs3->ptr));
  B.combine("P4_mid", gt4f->ptr, nullptr, nullptr, nullptr, 0, 0, gs4->ptr, h5, w5, 4, nullptr, gP4m->ptr, h4, w4, 256);
  B.combine("P3_mid", gt3f->ptr, nullptr, nullptr, nullptr, 0, 0, gs3->ptr, h4, w4, 4, nullptr, gP3m->ptr, h3, w3, 256);
  PPB_TRY(B.wgrad_plane("P4_mid", gP4m->ptr, B.fwd("P4_mid_in")->ptr, h4, w4));
  PPB_TRY(B.dgrad_plane("P4_mid", gP4m->ptr, h4, w4, nullptr, nullptr, gt4->ptr));
  PPB_TRY(B.wgrad_plane("P3_mid", gP3m->ptr, B.fwd("P3_mid_in")->ptr, h3, w3));
  PPB_TRY(B.dgrad_plane("P3_mid", gP3m->ptr, h3, w3, nullptr, nullptr, gt3->ptr));
  B.combine("P3_lateral", gt3f->ptr, gt3->ptr, nullptr, nullptr, 0, 0, nullptr, 0, 0, 0, nullptr, gP3l->ptr, h3, w3, 256);
  B.combine("P4_lateral", gt4f->ptr, gt4->ptr, nullptr, gt3->ptr, h3, w3, nullptr, 0, 0, 0, nullptr, gP4l->ptr, h4, w4, 256);
  B.combine("P5_lateral", gt5f->ptr, nullptr, nullptr, gt4->ptr, h4, w4, nullptr, 0, 0, 0, nullptr, gP5l->ptr, h5, w5, 256);

  // ---- backbone stages 5, 4, 3 (stage 2 and the stem are frozen)
  const int blocks[4] = {3, 4, 6, 3};
  const char* lateral[4] = {nullptr, "P3_lateral", "P4_lateral", "P5_lateral"};
  Tensor* g_lat[4] = {nullptr, gP3l, gP4l, gP5l};
  Tensor* upstream = nullptr;  // gradient scattered down from the stage above (full resolution of this stage)
  for (int stage = 3; stage >= 1; --stage) {
    const int f = 64 << stage;
    const int sh = e.lvl_h[stage - 1], sw = e.lvl_w[stage - 1];
    const std::string sname = std::to_string(stage + 2);
    Tensor *GA, *GB, *G1, *G2;
    PPB_TRY(B.gtensor(&GA, "res" + sname + "_out_a", sh, sw, 4 * f));
    PPB_TRY(B.gtensor(&GB, "res" + sname + "_out_b", sh, sw, 4 * f));
    PPB_TRY(B.gtensor(&G1, "res" + sname + "_mid_a", sh, sw, f));
    PPB_TRY(B.gtensor(&G2, "res" + sname + "_mid_b", sh, sw, f));
    char last[8];
    snprintf(last, sizeof(last), "%d%c", stage + 2, 'a' + blocks[stage] - 1);
    Tensor* c_out = B.fwd(std::string("res") + last + "_relu");
    // stage output: lateral conv + (for stages 3, 4) the stage above
    PPB_TRY(B.wgrad_plane(lateral[stage], g_lat[stage]->ptr, c_out->ptr, sh, sw));
    PPB_TRY(B.dgrad_plane(lateral[stage], g_lat[stage]->ptr, sh, sw, upstream ? upstream->ptr : nullptr, c_out->ptr,
                          GA->ptr));
    Tensor* cur = GA;
    for (int b = blocks[stage] - 1; b >= 0; --b) {
      char tagc[8];
      snprintf(tagc, sizeof(tagc), "%d%c", stage + 2, 'a' + b);
      const std::string tag = std::string("res") + tagc;
      Tensor* y1 = B.fwd(tag + "_branch2a_relu");
      Tensor* y2 = B.fwd(tag + "_branch2b_relu");
      Tensor* xin;
      if (b > 0) {
        char prev[8];
        snprintf(prev, sizeof(prev), "%d%c", stage + 2, 'a' + b - 1);
        xin = B.fwd(std::string("res") + prev + "_relu");
      } else {
        xin = B.fwd(tag + "_sub");
      }
      if (!y1 || !y2 || !xin) return fail(kErrState, "internal: forward activations of %s missing", tag.c_str());
      PPB_TRY(B.wgrad_plane(tag + "_branch2c", cur->ptr, y2->ptr, sh, sw));
      PPB_TRY(B.dgrad_plane(tag + "_branch2c", cur->ptr, sh, sw, nullptr, y2->ptr, G2->ptr));
      PPB_TRY(B.wgrad_plane(tag + "_branch2b", G2->ptr, y1->ptr, sh, sw));
      PPB_TRY(B.dgrad_plane(tag + "_branch2b", G2->ptr, sh, sw, nullptr, y1->ptr, G1->ptr));
      PPB_TRY(B.wgrad_plane(tag + "_branch2a", G1->ptr, xin->ptr, sh, sw));
      if (b > 0) {
        Tensor* other = cur == GA ? GB : GA;
        PPB_TRY(B.dgrad_plane(tag + "_branch2a", G1->ptr, sh, sw, cur->ptr, xin->ptr, other->ptr));  // + identity skip
        cur = other;
      } else {
        PPB_TRY(B.wgrad_plane(tag + "_branch1", cur->ptr, xin->ptr, sh, sw));
        if (stage == 1) break;  // res3a reads stage 2, which is frozen: no data gradient needed
        Tensor *gsub, *up;
        PPB_TRY(B.gtensor(&gsub, tag + "_sub", sh, sw, 2 * f));
        PPB_TRY(B.gtensor(&up, tag + "_in", e.lvl_h[stage - 2], e.lvl_w[stage - 2], 2 * f));
        PPB_TRY(B.dgrad_plane(tag + "_branch2a", G1->ptr, sh, sw, nullptr, nullptr, gsub->ptr));
        PPB_TRY(B.dgrad_plane(tag + "_branch1", cur->ptr, sh, sw, gsub->ptr, nullptr, gsub->ptr));
        // 1x1 stride-2 sampling (out(i,j) = in(2i,2j)): scatter back, zeros elsewhere; the ReLU mask of the stage
        // below is applied when its lateral gradient is added (next iteration)
        B.combine((tag + "_in").c_str(), nullptr, nullptr, nullptr, nullptr, 0, 0, gsub->ptr, sh, sw, 1, nullptr,
                  up->ptr, e.lvl_h[stage - 2], e.lvl_w[stage - 2], 2 * f);
        upstream = up;
      }
    }
  }
  t.bound_batch = batch;
  return kOk;
}

int launch_op(Engine& e, const TOp& op, cudaStream_t st) {
  TrainState& t = *e.train;
  {
    switch (op.kind) {
      case kTConv:
        PPB_CUDA(conv_launch(op.ta, op.tb, op.to, op.tr, op.p, 0, st));
        break;
      case kTWgrad:
        PPB_CUDA(wgrad_launch(op.tg, op.tx, op.wp, st));
        break;
      case kTBias:
        PPB_CUDA(launch_bias_grad(static_cast<const __nv_bfloat16*>(op.in[0]), op.n[0], op.a[0], op.a[1],
                                  static_cast<float*>(op.out), st));
        break;
      case kTCombine:
        PPB_CUDA(launch_grad_combine(
            static_cast<const __nv_bfloat16*>(op.in[0]), static_cast<const __nv_bfloat16*>(op.in[1]),
            static_cast<const __nv_bfloat16*>(op.in[2]), static_cast<const __nv_bfloat16*>(op.in[3]), op.a[0], op.a[1],
            static_cast<const __nv_bfloat16*>(op.in[4]), op.a[2], op.a[3], op.a[4],
            static_cast<const __nv_bfloat16*>(op.in[5]), static_cast<__nv_bfloat16*>(op.out), op.a[5], op.a[6], op.a[7],
            op.a[8], st));
        break;
      case kTCount:
        PPB_CUDA(launch_count_positive(static_cast<const float*>(op.in[0]), op.n[0], op.a[0], t.counts, op.a[1], st));
        break;
      case kTFocal:
        PPB_CUDA(launch_focal_loss_grad(static_cast<const float*>(op.in[0]), static_cast<const float*>(op.in[1]),
                                        static_cast<__nv_bfloat16*>(op.out), op.a[0], op.geom, t.bound_batch, op.a[1],
                                        op.a[2], op.f[0], op.f[1], t.counts, op.a[3], t.loss, st));
        break;
      case kTOrth:
        PPB_CUDA(launch_orth_l1_loss_grad(static_cast<const float*>(op.in[0]), static_cast<const float*>(op.in[1]),
                                          static_cast<__nv_bfloat16*>(op.out), op.a[0], op.geom, t.bound_batch, op.a[1],
                                          op.f[0], op.f[1], t.counts, op.a[3], t.loss, st));
        break;
    }
  }
  return kOk;
}

int launch_backward(Engine& e, int op_begin, int op_end, cudaStream_t st) {
  TrainState& t = *e.train;
  for (int i = op_begin; i < op_end; ++i) PPB_TRY(launch_op(e, t.bwd[i], st));
  return kOk;
}

// Positives of the three outputs (the loss normalisers, losses.py:61-66 / 395-397). They depend on the targets only,
// so they run ahead of the forward pass and a data-parallel caller can sum them over the replicas first — the
// normaliser of keras.utils.multi_gpu_model, which computes each loss once over the concatenated outputs
// (PyraPose/bin/train.py:82-86).
int launch_counts(Engine& e, cudaStream_t st) {
  TrainState& t = *e.train;
  PPB_CUDA(cudaMemsetAsync(t.counts, 0, 4 * sizeof(int), st));
  for (const TOp& op : t.count_ops) PPB_TRY(launch_op(e, op, st));
  t.counts_ready = true;
  return kOk;
}

// Segment `si` as plain launches: segment 0 zeroes the accumulators and runs the forward pass first.
int launch_segment(Engine& e, int si, cudaStream_t st) {
  TrainState& t = *e.train;
  const Segment& sg = t.segs[si];
  if (si == 0) {
    PPB_CUDA(cudaMemsetAsync(t.grads, 0, t.n_params * sizeof(float), st));
    PPB_CUDA(cudaMemsetAsync(t.loss, 0, 8 * sizeof(double), st));
    PPB_TRY(e.launch_all(st));
  }
  return launch_backward(e, sg.op_begin, sg.op_end, st);
}

// Cuts the backward op list after the op that completes each gradient bucket. Buckets are contiguous slices of the
// flat buffer of about `bucket_bytes`, cut at layer boundaries; the backward pass runs heads -> PFPN -> res5 -> res3,
// i.e. roughly from the END of the buffer (layer order) to its beginning.
void plan_segments(Engine& e) {
  TrainState& t = *e.train;
  t.drop_graphs();
  t.segs.clear();
  const int n_ops = static_cast<int>(t.bwd.size());
  struct Bucket {
    long long lo, hi;
    int ready;
  };
  std::vector<Bucket> buckets;
  if (t.bucket_bytes > 0) {
    std::vector<int> order;  // trainable layers by descending offset
    for (size_t i = 0; i < t.tl.size(); ++i)
      if (t.tl[i].trainable && t.tl[i].last_grad_op >= 0) order.push_back(static_cast<int>(i));
    std::sort(order.begin(), order.end(), [&](int a, int b) { return t.tl[a].off_kernel > t.tl[b].off_kernel; });
    // The LAST bucket's all-reduce cannot hide behind any backward work, so it is kept small: the layers at the bottom
    // of the buffer (res3*, finished last) up to a quarter of bucket_bytes form their own bucket; everything above is
    // cut top-down into buckets of bucket_bytes. (Measured r02 at 2 GPUs with an up-to-25 MB tail: 0.29 ms exposed.)
    size_t n_tail = 0;
    long long tail_hi = 0;
    for (size_t k = order.size(); k-- > 0;) {
      const TrainLayer& tl = t.tl[order[k]];
      const long long next_hi = k > 0 ? t.tl[order[k - 1]].off_kernel : t.n_params;
      if (next_hi * static_cast<long long>(sizeof(float)) > t.bucket_bytes / 4 && n_tail > 0) break;
      (void)tl;
      tail_hi = next_hi;
      ++n_tail;
    }
    if (n_tail >= order.size()) {  // everything fits the tail: one bucket
      n_tail = 0;
      tail_hi = 0;
    }
    const size_t n_top = order.size() - n_tail;
    long long hi = t.n_params;
    Bucket cur{hi, hi, -1};
    for (size_t k = 0; k < n_top; ++k) {
      const TrainLayer& tl = t.tl[order[k]];
      cur.lo = tl.off_kernel;
      cur.ready = std::max(cur.ready, tl.last_grad_op);
      if ((cur.hi - cur.lo) * static_cast<long long>(sizeof(float)) >= t.bucket_bytes || k + 1 == n_top) {
        if (k + 1 == n_top) cur.lo = tail_hi;
        buckets.push_back(cur);
        cur = Bucket{cur.lo, cur.lo, -1};
      }
    }
    if (n_tail > 0) {
      Bucket tail{0, tail_hi, -1};
      for (size_t k = n_top; k < order.size(); ++k) tail.ready = std::max(tail.ready, t.tl[order[k]].last_grad_op);
      buckets.push_back(tail);
    }
    std::stable_sort(buckets.begin(), buckets.end(), [](const Bucket& a, const Bucket& b) { return a.ready < b.ready; });
  }
  int begin = 0;
  if (!buckets.empty()) {
    // Segment 0 = memsets + forward pass only (no backward op, no gradient slice): the data-parallel caller's
    // all-reduce of the loss normalisers (ppb_engine_train_counts, a few bytes but a full NCCL latency) runs while the
    // forward pass computes and is awaited only before segment 1, whose first launches are the loss kernels.
    t.segs.push_back(Segment());
  }
  for (const Bucket& b : buckets) {
    Segment sg;
    sg.op_begin = begin;
    sg.op_end = std::max(begin, b.ready + 1);
    sg.g_off = b.lo;
    sg.g_cnt = b.hi - b.lo;
    begin = sg.op_end;
    t.segs.push_back(sg);
  }
  if (begin < n_ops || t.segs.empty()) {  // the rest (data gradients after the last weight gradient), or the whole step
    Segment sg;
    sg.op_begin = begin;
    sg.op_end = n_ops;
    if (buckets.empty()) {
      sg.g_off = 0;
      sg.g_cnt = t.n_params;
    }
    t.segs.push_back(sg);
  }
}

int capture_segments(Engine& e, cudaStream_t st) {
  TrainState& t = *e.train;
  t.graph_kernels = static_cast<int>(e.ops.size() + t.bwd.size());
  for (size_t si = 0; si < t.segs.size(); ++si) {
    Segment& sg = t.segs[si];
    sg.kernels = sg.op_end - sg.op_begin + (si == 0 ? static_cast<int>(e.ops.size()) : 0);
    if (sg.kernels == 0) continue;  // a bucket that was already complete: nothing to launch
    PPB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const int rc = launch_segment(e, static_cast<int>(si), st);
    cudaGraph_t g = nullptr;
    const cudaError_t end = cudaStreamEndCapture(st, &g);
    if (rc != kOk) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    PPB_CUDA(end);
    sg.graph = g;
    PPB_CUDA(cudaGraphInstantiate(&sg.exec, sg.graph, 0));
  }
  return kOk;
}

}  // namespace

int train_refresh_layer(Engine& e, int layer) {
  if (!e.train) return kOk;
  return seed_layer(e, *e.train, layer);
}

}  // namespace ppb

using namespace ppb;

extern "C" {

int ppb_engine_train_bind(ppb_engine* h, int batch, const float* images, const float* y_reg, const float* y_cls,
                          const float* y_mask, float* out_reg, float* out_cls, float* out_mask, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_train_bind: null engine");
  Engine& e = h->e;
  PPB_REQUIRE(batch > 0 && batch <= e.max_batch, "ppb_engine_train_bind: batch %d outside [1, %d]", batch, e.max_batch);
  PPB_REQUIRE(images && y_reg && y_cls && y_mask && out_reg && out_cls && out_mask, "ppb_engine_train_bind: null buffer");
  PPB_REQUIRE((reinterpret_cast<uintptr_t>(out_reg) & 15) == 0, "ppb_engine_train_bind: outputs must be 16-byte aligned");
  if (e.split()) return fail(kErrState, "ppb_engine_train_bind: the fp32 path (precision 1) is inference only");
  if (e.alias_arena) return fail(kErrState, "ppb_engine_train_bind: the production arena (ppb_engine_set_arena_alias) is inference only");
  PPB_CUDA(cudaSetDevice(e.device));
  e.training_plan = true;
  int rc = e.build_plan(batch, images, out_reg, out_cls, out_mask, nullptr, 0);
  if (rc != kOk) return rc;
  if (!e.train) {
    rc = create_state(e);
    if (rc != kOk) return rc;
  }
  e.train->bucket_bytes = e.train_bucket_bytes;
  rc = build_backward(e, batch, y_reg, y_cls, y_mask, out_reg, out_cls, out_mask);
  if (rc != kOk) {
    e.train->bound_batch = 0;
    return rc;
  }
  TrainState& t = *e.train;
  plan_segments(e);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaStream_t own = nullptr;
  if (st == nullptr) {
    PPB_CUDA(cudaStreamCreateWithFlags(&own, cudaStreamNonBlocking));
    st = own;
  }
  rc = capture_segments(e, st);
  if (own) cudaStreamDestroy(own);
  if (rc != kOk) {
    t.bound_batch = 0;
    return rc;
  }
  t.counts_ready = false;
  e.bound_batch = 0;  // the inference graph (if any) no longer matches the plan
  return kOk;
}

int ppb_engine_train_set_bucket_bytes(ppb_engine* h, long long bucket_bytes) {
  PPB_REQUIRE(h != nullptr && bucket_bytes >= 0, "ppb_engine_train_set_bucket_bytes: bad argument");
  Engine& e = h->e;
  e.train_bucket_bytes = bucket_bytes;
  if (e.train && e.train->bucket_bytes != bucket_bytes) {
    e.train->bucket_bytes = bucket_bytes;
    e.train->bound_batch = 0;  // the segment graphs no longer match: force a re-bind
  }
  return kOk;
}

int ppb_engine_train_count(ppb_engine* h, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_train_count: null engine");
  Engine& e = h->e;
  if (!e.train || e.train->bound_batch == 0) return fail(kErrState, "ppb_engine_train_count: call ppb_engine_train_bind first");
  PPB_CUDA(cudaSetDevice(e.device));
  const int rc = launch_counts(e, static_cast<cudaStream_t>(stream));
  if (rc == kOk) count_launch(static_cast<int>(e.train->count_ops.size()));
  return rc;
}

int ppb_engine_train_counts(ppb_engine* h, int** counts_dev) {
  PPB_REQUIRE(h != nullptr && h->e.train && counts_dev != nullptr, "ppb_engine_train_counts: training state not created");
  *counts_dev = h->e.train->counts;
  return kOk;
}

int ppb_engine_train_num_segments(const ppb_engine* h) {
  return (h && h->e.train && h->e.train->bound_batch != 0) ? static_cast<int>(h->e.train->segs.size()) : 0;
}

int ppb_engine_train_segment_info(const ppb_engine* h, int segment, long long* grad_offset, long long* grad_count,
                                  int* n_launches) {
  PPB_REQUIRE(h != nullptr && h->e.train && segment >= 0 && segment < static_cast<int>(h->e.train->segs.size()),
              "ppb_engine_train_segment_info: bad segment %d", segment);
  const Segment& sg = h->e.train->segs[segment];
  if (grad_offset) *grad_offset = sg.g_off;
  if (grad_count) *grad_count = sg.g_cnt;
  if (n_launches) *n_launches = sg.kernels;
  return kOk;
}

int ppb_engine_train_run_segment(ppb_engine* h, int segment, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_train_run_segment: null engine");
  Engine& e = h->e;
  if (!e.train || e.train->bound_batch == 0)
    return fail(kErrState, "ppb_engine_train_run_segment: call ppb_engine_train_bind first");
  TrainState& t = *e.train;
  PPB_REQUIRE(segment >= 0 && segment < static_cast<int>(t.segs.size()), "ppb_engine_train_run_segment: bad segment %d",
              segment);
  PPB_CUDA(cudaSetDevice(e.device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (segment == 0 && !t.counts_ready) {
    const int rc = launch_counts(e, st);
    if (rc != kOk) return rc;
  }
  const Segment& sg = t.segs[segment];
  if (sg.exec != nullptr) PPB_CUDA(cudaGraphLaunch(sg.exec, st));
  count_launch(sg.kernels);
  if (segment + 1 == static_cast<int>(t.segs.size())) t.counts_ready = false;
  return kOk;
}

int ppb_engine_train_forward_backward(ppb_engine* h, int eager, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_train_forward_backward: null engine");
  Engine& e = h->e;
  if (!e.train || e.train->bound_batch == 0)
    return fail(kErrState, "ppb_engine_train_forward_backward: call ppb_engine_train_bind first");
  PPB_CUDA(cudaSetDevice(e.device));
  TrainState& t = *e.train;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!t.counts_ready) {
    const int rc = launch_counts(e, st);
    if (rc != kOk) return rc;
  }
  for (size_t si = 0; si < t.segs.size(); ++si) {
    if (eager) {
      const int rc = launch_segment(e, static_cast<int>(si), st);
      if (rc != kOk) return rc;
    } else if (t.segs[si].exec != nullptr) {
      PPB_CUDA(cudaGraphLaunch(t.segs[si].exec, st));
    }
  }
  t.counts_ready = false;
  count_launch(t.graph_kernels);
  return kOk;
}

int ppb_engine_train_grads(ppb_engine* h, float** grads, long long* count) {
  PPB_REQUIRE(h != nullptr && h->e.train, "ppb_engine_train_grads: training state not created");
  if (grads) *grads = h->e.train->grads;
  if (count) *count = h->e.train->n_params;
  return kOk;
}

int ppb_engine_train_apply(ppb_engine* h, float lr, float beta1, float beta2, float epsilon, float clipnorm,
                           float grad_scale, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_train_apply: null engine");
  Engine& e = h->e;
  if (!e.train || e.train->bound_batch == 0) return fail(kErrState, "ppb_engine_train_apply: call ppb_engine_train_bind first");
  PPB_CUDA(cudaSetDevice(e.device));
  TrainState& t = *e.train;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PPB_CUDA(launch_grad_prepare(t.descs_dev, static_cast<int>(t.descs.size()), t.chunk_desc, t.chunk_index, t.n_chunks,
                               t.grads, t.master, grad_scale, t.loss, st));
  t.step += 1;
  // Keras 2.3.1 Adam.get_updates: lr_t = lr * sqrt(1 - beta_2^t) / (1 - beta_1^t)
  const double tt = static_cast<double>(t.step);
  const float lr_t = static_cast<float>(lr * std::sqrt(1.0 - std::pow(static_cast<double>(beta2), tt)) /
                                        (1.0 - std::pow(static_cast<double>(beta1), tt)));
  PPB_CUDA(launch_adam_update(t.descs_dev, static_cast<int>(t.descs.size()), t.chunk_desc, t.chunk_index, t.n_chunks,
                              t.grads, t.master, t.m, t.v, lr_t, beta1, beta2, epsilon, clipnorm, t.loss, st));
  count_launch(2);
  return kOk;
}

int ppb_engine_train_losses(ppb_engine* h, double* out5_host, int* counts3_host, void* stream) {
  PPB_REQUIRE(h != nullptr && h->e.train, "ppb_engine_train_losses: training state not created");
  Engine& e = h->e;
  PPB_CUDA(cudaSetDevice(e.device));
  PPB_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  if (out5_host) PPB_CUDA(cudaMemcpy(out5_host, e.train->loss, 5 * sizeof(double), cudaMemcpyDeviceToHost));
  if (counts3_host) PPB_CUDA(cudaMemcpy(counts3_host, e.train->counts, 3 * sizeof(int), cudaMemcpyDeviceToHost));
  return kOk;
}

int ppb_engine_train_losses_async(ppb_engine* h, double* out5_host, int* counts3_host, void* stream) {
  PPB_REQUIRE(h != nullptr && h->e.train, "ppb_engine_train_losses_async: training state not created");
  PPB_REQUIRE(out5_host != nullptr, "ppb_engine_train_losses_async: null output");
  Engine& e = h->e;
  PPB_CUDA(cudaSetDevice(e.device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PPB_CUDA(cudaMemcpyAsync(out5_host, e.train->loss, 5 * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (counts3_host)
    PPB_CUDA(cudaMemcpyAsync(counts3_host, e.train->counts, 3 * sizeof(int), cudaMemcpyDeviceToHost, st));
  return kOk;
}

int ppb_engine_train_param_info(const ppb_engine* h, const char* layer, long long* kernel_offset,
                                long long* bias_offset, int* trainable) {
  PPB_REQUIRE(h != nullptr && h->e.train && layer != nullptr, "ppb_engine_train_param_info: training state not created");
  const Engine& e = h->e;
  auto it = e.layer_index.find(layer);
  PPB_REQUIRE(it != e.layer_index.end(), "ppb_engine_train_param_info: unknown layer '%s'", layer);
  const TrainLayer& tl = e.train->tl[it->second];
  if (kernel_offset) *kernel_offset = tl.off_kernel;
  if (bias_offset) *bias_offset = tl.off_bias;
  if (trainable) *trainable = tl.trainable ? 1 : 0;
  return kOk;
}

int ppb_engine_train_read(ppb_engine* h, int which, long long offset, long long count, float* out_host) {
  PPB_REQUIRE(h != nullptr && h->e.train && out_host != nullptr, "ppb_engine_train_read: training state not created");
  Engine& e = h->e;
  TrainState& t = *e.train;
  PPB_REQUIRE(which >= 0 && which <= 3 && offset >= 0 && count >= 0 && offset + count <= t.n_params,
              "ppb_engine_train_read: bad range");
  const float* src[4] = {t.grads, t.master, t.m, t.v};
  PPB_CUDA(cudaSetDevice(e.device));
  PPB_CUDA(cudaDeviceSynchronize());
  PPB_CUDA(cudaMemcpy(out_host, src[which] + offset, count * sizeof(float), cudaMemcpyDeviceToHost));
  return kOk;
}

int ppb_engine_train_num_ops(const ppb_engine* h) {
  return (h && h->e.train) ? static_cast<int>(h->e.ops.size() + h->e.train->bwd.size()) : 0;
}

int ppb_engine_train_op_info(const ppb_engine* h, int index, char* name, int name_cap, double* flops) {
  PPB_REQUIRE(h != nullptr && h->e.train, "ppb_engine_train_op_info: training state not created");
  const Engine& e = h->e;
  const int nf = static_cast<int>(e.ops.size());
  PPB_REQUIRE(index >= 0 && index < nf + static_cast<int>(e.train->bwd.size()), "ppb_engine_train_op_info: bad index");
  const char* nm = index < nf ? e.ops[index].name : e.train->bwd[index - nf].name;
  const double fl = index < nf ? e.ops[index].flops : e.train->bwd[index - nf].flops;
  if (name != nullptr && name_cap > 0) snprintf(name, name_cap, "%s", nm);
  if (flops) *flops = fl;
  return kOk;
}

/* Per-launch device times of one eager training step (forward ops first, then losses + backward), CUDA events on
 * `stream`; fills min(cap, ppb_engine_train_num_ops) entries. Synchronises the stream. */
int ppb_engine_train_profile(ppb_engine* h, void* stream, float* ms_out_host, int cap) {
  PPB_REQUIRE(h != nullptr && ms_out_host != nullptr, "ppb_engine_train_profile: null argument");
  Engine& e = h->e;
  if (!e.train || e.train->bound_batch == 0) return fail(kErrState, "ppb_engine_train_profile: call ppb_engine_train_bind first");
  PPB_CUDA(cudaSetDevice(e.device));
  TrainState& t = *e.train;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nf = static_cast<int>(e.ops.size()), nb = static_cast<int>(t.bwd.size());
  std::vector<cudaEvent_t> ev(nf + nb + 1);
  for (auto& x : ev) PPB_CUDA(cudaEventCreate(&x));
  PPB_CUDA(cudaMemsetAsync(t.grads, 0, t.n_params * sizeof(float), st));
  PPB_CUDA(cudaMemsetAsync(t.loss, 0, 8 * sizeof(double), st));
  PPB_TRY(launch_counts(e, st));
  t.counts_ready = false;
  std::vector<Op> all_f;
  all_f.swap(e.ops);
  int rc = kOk;
  cudaEventRecord(ev[0], st);
  for (int i = 0; i < nf && rc == kOk; ++i) {
    e.ops.assign(1, all_f[i]);
    rc = e.launch_all(st);
    cudaEventRecord(ev[i + 1], st);
  }
  e.ops.swap(all_f);
  for (int i = 0; i < nb && rc == kOk; ++i) {
    rc = launch_backward(e, i, i + 1, st);
    cudaEventRecord(ev[nf + i + 1], st);
  }
  const cudaError_t sync = cudaStreamSynchronize(st);
  if (rc == kOk && sync == cudaSuccess) {
    for (int i = 0; i < nf + nb && i < cap; ++i) cudaEventElapsedTime(&ms_out_host[i], ev[i], ev[i + 1]);
    count_launch(nf + nb);
  }
  for (auto& x : ev) cudaEventDestroy(x);
  if (rc != kOk) return rc;
  PPB_CUDA(sync);
  return kOk;
}

}  // extern "C"
