// PFPN engine internals shared by engine.cu (inference plan) and train.cu (training step).
#pragma once
#include "../../include/pyrapose_b200.h"
#include "conv_igemm.h"
#include "elementwise.h"
#include "errors.h"
#include "launch_count.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace ppb {

// fp32 path (Engine::precision 1, ConvParams::split): parts per value and the (activation part, weight part) of the
// six K segments, largest terms first.
constexpr int kSplitParts = 3;
constexpr int kSplitSegs = 6;
constexpr int kSplitAPart[kSplitSegs] = {0, 1, 0, 2, 1, 0};
constexpr int kSplitWPart[kSplitSegs] = {0, 0, 1, 0, 1, 2};

void conv_fill_single_seg(ConvParams& p, int batch, int h, int w, int ksize);

struct LayerW {
  std::string name;
  int ksize, cin, cout, stride;
  bool has_bn, has_bias;
  int k_total;   // columns of the device weight matrix (ksize*ksize*cin, stem: 192)
  int cout_pad;  // rows, multiple of 16
  int k_dev = 0;  // columns of the device matrix: k_total, or 6 * k_total on the fp32 path (six part blocks per tap)
  __nv_bfloat16* w = nullptr;
  float* w_f32 = nullptr;  // fp32 path, stem only: [7][7][3][64] with the BatchNorm scale folded in (launch_stem_conv_split)
  float* bias = nullptr;
  bool loaded = false;
  // host copies kept for the training step (master weights start from them): raw Keras kernel as [cout][k_total]
  // (un-folded), raw bias [cout], frozen-BatchNorm scale [cout] (empty without BN)
  std::vector<float> raw_kernel, raw_bias, bn_scale;
};

struct Tensor {  // bf16 activation on a pixel grid with pad_lo border pixels before and pad_hi after the interior
  std::string name;
  __nv_bfloat16* ptr = nullptr;
  int H = 0, W = 0, C = 0, pad_lo = 1, pad_hi = 1;  // padded-flat: 1/1; stem grid: 0/3
  int ld = 0;  // elements per row: C, or 3 * C on the fp32 path (three bf16 parts per value, ConvParams::split)
  int hp() const { return H + pad_lo + pad_hi; }
  int wp() const { return W + pad_lo + pad_hi; }
  long long rows(int B) const { return static_cast<long long>(B) * hp() * wp(); }
};

enum OpKind { kOpConv, kOpStemS2d, kOpMaxpool, kOpPlanes, kOpAdd, kOpStemSplit };

struct Op {
  OpKind kind;
  // conv
  CUtensorMap ta, tb, to, tr;
  ConvParams p;
  // elementwise
  const void* in0 = nullptr;
  const void* in1 = nullptr;
  const void* in2 = nullptr;
  const void* in3 = nullptr;
  void* out = nullptr;
  int a[8] = {0};
  char name[40] = {0};
  double flops = 0.0;  // conv: 2 * interior pixels * Cout * K (algorithmic)
};

inline int round_up(int v, int m) { return (v + m - 1) / m * m; }

inline int same_pad_before(int in, int k, int stride) {
  const int out = (in + stride - 1) / stride;
  const int total = std::max((out - 1) * stride + k - in, 0);
  return total / 2;
}

struct Engine {
  int device = 0, max_batch = 0, H = 0, W = 0, num_classes = 0, num_anchors = 0;
  std::vector<LayerW> layers;
  std::map<std::string, int> layer_index;
  std::vector<std::unique_ptr<Tensor>> tensors;
  std::map<std::string, Tensor*> tensor_index;
  std::vector<void*> allocations;
  // anchors / decode parameters (host copies; baked into the head conv's params)
  float base_anchor[kConvMaxSegs][9][4];
  float stride[kConvMaxSegs];
  float dec_mean[16], dec_std[16];
  bool anchors_set = false;
  // level geometry
  int lvl_h[3], lvl_w[3];
  int seg_row_begin[3], seg_tile_begin[3];
  long long feat_rows_total = 0;
  int feat_tiles_total = 0;
  // plan
  std::vector<Op> ops;
  int bound_batch = 0;  // batch of the captured inference graph (0 = none)
  int plan_batch = 0;   // batch of the last build_plan (inference or training)
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t graph_exec = nullptr;
  int graph_kernels = 0;
  // The same plan cut into three graphs at the points where an output tensor is complete — after the classification
  // head ('cls'), after the regression head ('boxes3D' / '3Dbox'), after the mask head — so that predict_on_batch can
  // start the device->host copy of each output while the remaining heads still compute (ppb_engine_run_part).
  // Captured lazily on the first ppb_engine_run_part; the single graph above stays the path of ppb_engine_run.
  int part_end[3] = {0, 0, 0};
  cudaGraph_t part_graph[3] = {nullptr, nullptr, nullptr};
  cudaGraphExec_t part_exec[3] = {nullptr, nullptr, nullptr};
  void drop_parts() {
    for (int i = 0; i < 3; ++i) {
      if (part_exec[i]) cudaGraphExecDestroy(part_exec[i]);
      if (part_graph[i]) cudaGraphDestroy(part_graph[i]);
      part_exec[i] = nullptr;
      part_graph[i] = nullptr;
    }
  }
  size_t arena_bytes = 0;
  // TMA-store epilogue through the shared-memory chunk ring (residual chunks prefetched by TMA): 0 never; 1 only convs
  // with a residual input; 2 every backbone / PFPN conv (measured best, profiles/r01_per_op_staged_variants.txt);
  // 3 head towers too. Override with PPB_STAGED_EPILOGUE for experiments.
  int staged_epilogue = 2;
  bool stem16 = true;  // stem convolution in pixel-slab mode (ConvParams::stem16); PPB_STEM16=0 turns it off
  // training step (train.cu): every head-tower activation keeps its own buffer so the backward pass can read it
  // input staging: images bound as raw uint8 BGR with the 'caffe' mean subtraction fused into the stem loader
  bool input_u8 = false;
  float input_mean[3] = {103.939f, 116.779f, 123.68f};
  // 0: bf16 activations and weights, fp32 accumulation (the production path, tolerance 1e-2).
  // 1: the fp32 path of BASELINE.json north_star (tolerance 1e-4): activations and weights as three bf16 parts each
  //    (24 significant bits), six tcgen05 MMAs per product (ConvParams::split), stem / Add / MaxPool in fp32
  //    (split_ops.cu). Inference only.
  int precision = 0;
  bool split() const { return precision == 1; }
  // Production arena (ppb_engine_set_arena_alias, inference plans only): the bottleneck intermediates of the backbone
  // share a handful of buffers instead of one per layer — all branch2a / branch2b outputs one buffer each, block
  // outputs ping-pong between two, one projection-shortcut and one subsample buffer — sized by their first (largest,
  // stage 2) use; C3 / C4 / C5, the pyramid and the head towers keep their own. ppb_engine_activation then only
  // returns tensors that still hold their value at the end of the pass; the parity tests run without it.
  bool alias_arena = false;
  std::map<std::string, __nv_bfloat16*> alias_pool;
  bool training_plan = false;
  long long train_bucket_bytes = 0;  // gradient bucket size of the segmented training step (0 = one graph)
  std::shared_ptr<struct TrainState> train;

  ~Engine() {
    drop_parts();
    if (graph_exec) cudaGraphExecDestroy(graph_exec);
    if (graph) cudaGraphDestroy(graph);
    for (void* p : allocations) cudaFree(p);
  }

  int add_layer(const std::string& name, int ksize, int cin, int cout, int stride, bool bn, bool bias) {
    LayerW l;
    l.name = name;
    l.ksize = ksize;
    l.cin = cin;
    l.cout = cout;
    l.stride = stride;
    l.has_bn = bn;
    l.has_bias = bias;
    l.k_total = (ksize == 7) ? 256 : ksize * ksize * cin;  // stem: 4 row taps x (4 px x 16 ch), see launch_stem_s2d
    l.cout_pad = round_up(cout, 16);
    layer_index[name] = static_cast<int>(layers.size());
    layers.push_back(l);
    return static_cast<int>(layers.size()) - 1;
  }

  int device_alloc(void** p, size_t bytes) {
    PPB_CUDA(cudaMalloc(p, bytes));
    PPB_CUDA(cudaMemset(*p, 0, bytes));
    allocations.push_back(*p);
    arena_bytes += bytes;
    return kOk;
  }

  Tensor* tensor(const std::string& name) {
    auto it = tensor_index.find(name);
    return it == tensor_index.end() ? nullptr : it->second;
  }

  // A tensor that lives in the shared buffer `pool` (allocated by its first user, which is the largest).
  int alias_tensor(Tensor** out, const std::string& name, const std::string& pool, int h, int w, int c) {
    auto it = alias_pool.find(pool);
    if (it == alias_pool.end()) {
      Tensor* first = nullptr;
      const int rc = new_tensor(&first, name, h, w, c, 1);
      if (rc != kOk) return rc;
      alias_pool[pool] = first->ptr;
      *out = first;
      return kOk;
    }
    std::unique_ptr<Tensor> t(new Tensor);
    t->name = name;
    t->H = h;
    t->W = w;
    t->C = c;
    t->ld = split() ? kSplitParts * c : c;
    t->ptr = it->second;
    *out = t.get();
    tensor_index[name] = t.get();
    tensors.push_back(std::move(t));
    return kOk;
  }

  int new_tensor(Tensor** out, const std::string& name, int h, int w, int c, int pad, int planes = 1,
                 long long rows_override = 0, int pad_hi = -1) {
    std::unique_ptr<Tensor> t(new Tensor);
    t->name = name;
    t->H = h;
    t->W = w;
    t->C = c;
    t->pad_lo = pad;
    t->pad_hi = pad_hi >= 0 ? pad_hi : pad;
    t->ld = split() ? kSplitParts * c : c;
    long long rows = rows_override ? rows_override : t->rows(max_batch) * planes;
    rows = (rows + kConvBlockM - 1) / kConvBlockM * kConvBlockM + kConvBlockM;  // slack: TMA rows may overlap the end
    void* p = nullptr;
    int rc = device_alloc(&p, static_cast<size_t>(rows) * t->ld * sizeof(__nv_bfloat16));
    if (rc != kOk) return rc;
    t->ptr = static_cast<__nv_bfloat16*>(p);
    *out = t.get();
    tensor_index[name] = t.get();
    tensors.push_back(std::move(t));
    return kOk;
  }

  void define_layers();
  int allocate_weights();
  int set_layer(const char* name, const float* kernel, const float* bias, const float* gamma, const float* beta,
                const float* mean, const float* var, float eps);
  int build_plan(int batch, const float* images, float* out_reg, float* out_cls, float* out_mask, float* out_hull,
                 int decode);
  // out2 / res2: the Add (+ UpsampleLike when res2_up) that follows the layer, fused into its epilogue
  // (ConvParams::out2): out2 = y + res2, or y + nearest-resize(res2) to y's size.
  int conv_op(const Tensor* x, int x_planes, const std::string& layer, const Tensor* res, Tensor* y, int relu, int batch,
              int src_h = 0, int src_w = 0, Tensor* out2 = nullptr, const Tensor* res2 = nullptr, bool res2_up = false);
  // PPB_FUSE_ADD=1: the five PFPN Add / UpsampleLike ops leave with the convolution that produces their first operand
  // instead of as elementwise launches. Off by default — measured neutral at B=32 (r02, same box, graph replay 7.40-7.49
  // ms unfused vs 7.46-7.48 fused): the direct epilogue's per-thread 64-byte row accesses cost the convolution about
  // what the coalesced add kernel costs (P3_lateral 0.079 -> 0.138 ms for a 0.048 ms add, P3_mid 0.142 -> 0.185 for
  // 0.046), so only bytes are saved, not time.
  bool fuse_adds = false;
  int head_conv_op(const __nv_bfloat16* x, int cin, const std::string& layer, __nv_bfloat16* y, int relu, int batch,
                   int n_levels, int epi, float* out_f32, float* out_hull);
  void conv_set_split(ConvParams& p, int cin, int out_c, int res_c) const;
  int capture(cudaStream_t st);
  int capture_parts(cudaStream_t st);
  int launch_all(cudaStream_t st) { return launch_range(0, static_cast<int>(ops.size()), st); }
  int launch_range(int begin, int end, cudaStream_t st);
  void fill_head_segs(ConvParams& p, int batch, int n_levels) const;
};

// train.cu: re-seeds the master weights / dgrad matrix of one layer after Engine::set_layer when a training state
// exists (no-op otherwise).
int train_refresh_layer(Engine& e, int layer);

}  // namespace ppb

// The opaque C handle (include/pyrapose_b200.h).
struct ppb_engine {
  ppb::Engine e;
};
