// PFPN inference engine: owns the bf16 weights, the activation arena and the launch plan of the whole
// network (ResNet-50 -> __create_sparceFPN -> heads -> decode) and replays it as one CUDA graph.
//
// Replaces what `model.predict_on_batch` executes for the graph built by
//   PyraPose/models/resnet.py:79-110        (resnet_retinanet -> keras_resnet.models.ResNet50)
//   PyraPose/models/retinanet.py:180-214    (__create_sparceFPN)
//   PyraPose/models/retinanet.py:9-131      (classification / mask / 3D-regression sub-models)
//   PyraPose/models/retinanet.py:260-335    (retinanet, retinanet_bbox: Anchors + RegressBoxes3D)
// Every convolution is one launch of the tcgen05 implicit-GEMM kernel (conv_igemm.cu); BatchNorm is folded
// into the weights, ReLU / residual add / sigmoid / Reshape+Concatenate / anchor decode live in its epilogue.
// The three pyramid levels share each head convolution as ONE launch over a 3-segment row space.
#include "engine.h"

namespace ppb {

void Engine::define_layers() {
  // keras-resnet 0.1.0 ResNet50 (Caffe-style bottlenecks, stride on the first 1x1 conv)
  add_layer("conv1", 7, 3, 64, 2, true, false);
  int cin = 64, features = 64;
  const int blocks[4] = {3, 4, 6, 3};
  for (int stage = 0; stage < 4; ++stage) {
    for (int b = 0; b < blocks[stage]; ++b) {
      char tag[8];
      snprintf(tag, sizeof(tag), "%d%c", stage + 2, 'a' + b);
      const int st = (b == 0 && stage > 0) ? 2 : 1;
      add_layer(std::string("res") + tag + "_branch2a", 1, cin, features, st, true, false);
      add_layer(std::string("res") + tag + "_branch2b", 3, features, features, 1, true, false);
      add_layer(std::string("res") + tag + "_branch2c", 1, features, features * 4, 1, true, false);
      if (b == 0) add_layer(std::string("res") + tag + "_branch1", 1, cin, features * 4, st, true, false);
      cin = features * 4;
    }
    features *= 2;
  }
  // __create_sparceFPN (retinanet.py:180-214)
  add_layer("P3_lateral", 1, 512, 256, 1, false, true);
  add_layer("P4_lateral", 1, 1024, 256, 1, false, true);
  add_layer("P5_lateral", 1, 2048, 256, 1, false, true);
  add_layer("P4_mid", 3, 256, 256, 1, false, true);
  add_layer("P3_mid", 3, 256, 256, 1, false, true);
  add_layer("P3_down", 3, 256, 256, 2, false, true);
  add_layer("P3", 3, 256, 256, 1, false, true);
  add_layer("P4_down", 3, 256, 256, 2, false, true);
  add_layer("P4", 3, 256, 256, 1, false, true);
  add_layer("P5", 3, 256, 256, 1, false, true);
  // heads (retinanet.py:9-131)
  for (int i = 0; i < 4; ++i)
    add_layer("pyramid_regression3D_" + std::to_string(i), 3, i == 0 ? 256 : 512, 512, 1, false, true);
  add_layer("pyramid_regression3D", 3, 512, num_anchors * 16, 1, false, true);
  for (int i = 0; i < 4; ++i) add_layer("pyramid_classification_" + std::to_string(i), 3, 256, 256, 1, false, true);
  add_layer("pyramid_classification", 3, 256, num_classes * num_anchors, 1, false, true);
  for (int i = 0; i < 4; ++i) add_layer("mask_" + std::to_string(i), 3, 256, 256, 1, false, true);
  add_layer("mask_out", 3, 256, num_classes, 1, false, true);
}

int Engine::allocate_weights() {
  for (LayerW& l : layers) {
    void* p = nullptr;
    l.k_dev = split() ? kSplitSegs * l.k_total : l.k_total;
    int rc = kOk;
    if (split() && l.ksize == 7) {  // fp32 path: the stem is a plain fp32 convolution (launch_stem_conv_split)
      rc = device_alloc(&p, static_cast<size_t>(7 * 7 * 3) * l.cout * sizeof(float));
      if (rc != kOk) return rc;
      l.w_f32 = static_cast<float*>(p);
    }
    rc = device_alloc(&p, static_cast<size_t>(l.cout_pad) * l.k_dev * sizeof(__nv_bfloat16));
    if (rc != kOk) return rc;
    l.w = static_cast<__nv_bfloat16*>(p);
    rc = device_alloc(&p, static_cast<size_t>(l.cout_pad) * sizeof(float));
    if (rc != kOk) return rc;
    l.bias = static_cast<float*>(p);
  }
  return kOk;
}

// Keras HWIO fp32 kernel (+ optional frozen BatchNorm) -> bf16 [cout_pad][K] with K = (r*k + s)*cin + c.
int Engine::set_layer(const char* name, const float* kernel, const float* bias, const float* gamma, const float* beta,
                      const float* mean, const float* var, float eps) {
  auto it = layer_index.find(name);
  PPB_REQUIRE(it != layer_index.end(), "ppb_engine_set_layer: unknown layer '%s'", name);
  LayerW& l = layers[it->second];
  PPB_REQUIRE(kernel != nullptr, "ppb_engine_set_layer: null kernel for '%s'", name);
  PPB_REQUIRE(!l.has_bn || (gamma && beta && mean && var), "ppb_engine_set_layer: '%s' needs BatchNorm parameters", name);
  const int k = l.ksize, cin = l.cin, cout = l.cout;
  std::vector<__nv_bfloat16> w(static_cast<size_t>(l.cout_pad) * l.k_dev, __float2bfloat16(0.0f));
  std::vector<float> w32;  // fp32 path, stem: [7][7][3][cout]
  if (split() && k == 7) w32.assign(static_cast<size_t>(7 * 7 * 3) * cout, 0.0f);
  std::vector<float> b(l.cout_pad, 0.0f);
  l.raw_kernel.assign(static_cast<size_t>(cout) * l.k_total, 0.0f);
  l.raw_bias.assign(cout, 0.0f);
  l.bn_scale.clear();
  if (l.has_bn) l.bn_scale.assign(cout, 1.0f);
  for (int o = 0; o < cout; ++o) {
    float scale = 1.0f, shift = 0.0f;
    if (l.has_bn) {  // y = gamma * (x - mean) / sqrt(var + eps) + beta
      scale = gamma[o] / std::sqrt(var[o] + eps);
      shift = beta[o] - mean[o] * scale;
    }
    const float b0 = (l.has_bias && bias != nullptr) ? bias[o] : 0.0f;
    b[o] = b0 * scale + shift;
    l.raw_bias[o] = b0;
    if (l.has_bn) l.bn_scale[o] = scale;
    for (int r = 0; r < k; ++r)
      for (int s = 0; s < k; ++s)
        for (int c = 0; c < cin; ++c) {
          const float v = kernel[((static_cast<size_t>(r) * k + s) * cin + c) * cout + o] * scale;
          size_t col = static_cast<size_t>(r * k + s) * cin + c;
          if (k == 7) {
            // stem over the space-to-depth grid: padded filter row r+1 = 2*dy + py, column s+1 = 2*dx + px;
            // K index = dy*64 + dx*16 + (py*2+px)*3 + c
            const int ry = r + 1, rx = s + 1;
            col = static_cast<size_t>(ry >> 1) * 64 + (rx >> 1) * 16 + ((ry & 1) * 2 + (rx & 1)) * 3 + c;
          }
          if (!split()) {
            w[static_cast<size_t>(o) * l.k_total + col] = __float2bfloat16(v);
          } else if (k == 7) {
            w32[(static_cast<size_t>(r * k + s) * cin + c) * cout + o] = v;
          } else {  // per tap the column blocks [w0 | w0 | w1 | w0 | w1 | w2] (ConvParams::split)
            __nv_bfloat16 part[3];
            float rem = v;
            for (int i = 0; i < 3; ++i) {
              part[i] = __float2bfloat16(rem);
              rem -= __bfloat162float(part[i]);
            }
            __nv_bfloat16* row = w.data() + static_cast<size_t>(o) * l.k_dev + static_cast<size_t>(r * k + s) * kSplitSegs * cin;
            for (int sg = 0; sg < kSplitSegs; ++sg) row[sg * cin + c] = part[kSplitWPart[sg]];
          }
          l.raw_kernel[static_cast<size_t>(o) * l.k_total + col] = kernel[((static_cast<size_t>(r) * k + s) * cin + c) * cout + o];
        }
  }
  PPB_CUDA(cudaMemcpy(l.w, w.data(), w.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  if (!w32.empty()) PPB_CUDA(cudaMemcpy(l.w_f32, w32.data(), w32.size() * sizeof(float), cudaMemcpyHostToDevice));
  PPB_CUDA(cudaMemcpy(l.bias, b.data(), b.size() * sizeof(float), cudaMemcpyHostToDevice));
  l.loaded = true;
  return kOk;
}

}  // namespace ppb

namespace ppb {

// One convolution over a single padded-flat plane set. `x` has the OUTPUT geometry (for stride-2 layers it is the
// parity-plane tensor produced by launch_parity_planes from a src_h x src_w input).
int Engine::conv_op(const Tensor* x, int x_planes, const std::string& layer, const Tensor* res, Tensor* y, int relu,
                    int batch, int src_h, int src_w, Tensor* out2, const Tensor* res2, bool res2_up) {
  const LayerW& l = layers[layer_index.at(layer)];
  if (!l.loaded) return fail(kErrState, "layer '%s' has no weights (call ppb_engine_set_layer first)", l.name.c_str());
  Op op;
  op.kind = kOpConv;
  ConvParams& p = op.p;
  memset(&p, 0, sizeof(p));
  int a_cols = x->ld, a_ld = x->ld;
  if (l.ksize == 7) {
    if (split()) return fail(kErrState, "internal: the fp32 path runs the stem as launch_stem_conv_split");  // stem over the space-to-depth grid: 4 row taps, A rows overlap (stride 16 < extent 64)
    p.n_segs = 1;
    ConvSeg& s = p.seg[0];
    s.row_begin = 0;
    s.hp = y->hp();
    s.wp = y->wp();
    s.pad_lo = y->pad_lo;
    s.pad_hi = y->pad_hi;
    s.rows = static_cast<int>(y->rows(batch));
    s.tile_begin = 0;
    for (int dy = 0; dy < 4; ++dy) s.tap_off[dy] = dy * s.wp;
    p.n_taps = 4;
    p.n_mtiles = (s.rows + kConvBlockM - 1) / kConvBlockM;
    a_cols = 64;
    a_ld = 16;
  } else {
    conv_fill_single_seg(p, batch, y->H, y->W, l.ksize);
    if (l.stride == 2 && l.ksize == 3) {
      // 3x3/2 'same' over four parity planes: input row 2i + r - pt lives in plane ((r-pt) mod 2) at row
      // i + floor((r-pt)/2); TF puts the odd padding pixel at the bottom/right (pt = total/2).
      ConvSeg& s = p.seg[0];
      const long long plane_rows = static_cast<long long>(batch) * s.hp * s.wp;
      const int pt = same_pad_before(src_h, 3, 2), pl = same_pad_before(src_w, 3, 2);
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
          const int u = r - pt, v = c - pl;
          const int ph = ((u % 2) + 2) % 2, pw = ((v % 2) + 2) % 2;
          const int dy = (u - ph) / 2, dx = (v - pw) / 2;
          s.tap_off[r * 3 + c] = static_cast<int>((ph * 2 + pw) * plane_rows) + dy * s.wp + dx;
        }
    }
  }
  p.cin_blocks = a_cols / kConvBlockK;
  if (split()) conv_set_split(p, x->C, y->C, res ? res->C : 0);
  p.n_valid = l.cout;
  p.block_n = conv_pick_block_n(l.cout_pad, p.n_taps * p.cin_blocks, p.n_mtiles, true, split() ? 128 : kConvMaxBlockN);
  p.n_ntiles = (l.cout_pad + p.block_n - 1) / p.block_n;
  p.epi = kEpiBf16;
  p.relu = relu;
  p.bias = l.bias;
  p.residual = res ? res->ptr : nullptr;
  p.res_ld = res ? res->ld : 0;
  p.out = y->ptr;
  p.out_ld = y->ld;
  const long long a_rows = x->rows(batch) * (x_planes > 0 ? x_planes : 1);
  if (out2 != nullptr) {  // fused Add: direct epilogue only
    if (split() || res != nullptr) return fail(kErrState, "internal: fused Add needs the bf16 path and no residual");
    p.out2 = out2->ptr;
    p.out2_ld = out2->ld;
    p.res2 = res2->ptr;
    p.res2_ld = res2->ld;
    p.res2_uh = res2_up ? res2->H : 0;
    p.res2_uw = res2_up ? res2->W : 0;
  }
  p.pair = conv_pair_ok(p) ? 1 : 0;
  p.staged = (!split() && out2 == nullptr && l.cout % 64 == 0 && p.block_n % 64 == 0 &&
              (staged_epilogue >= 2 || (staged_epilogue == 1 && res != nullptr))) ? 1 : 0;
  p.slab3 = conv_slab3_mode(p);
  if (l.ksize == 7 && stem16 && l.k_total == 256 && (p.block_n & 7) == 0) {
    p.stem16 = 1;
    p.pair = 0;
    p.slab3 = 0;
  }
  bool force_direct = out2 != nullptr;
  if (p.slab3 == 0 && p.staged && res == nullptr) {
    // a 256-wide pair tile in slab mode leaves no room for the epilogue ring: without a residual to prefetch the
    // direct epilogue costs nothing, so prefer the slab (P3, P3_mid)
    p.staged = 0;
    p.slab3 = conv_slab3_mode(p);
    force_direct = p.slab3 != 0;
    if (!force_direct) p.staged = 1;
  }
  // Without a residual the TMA-store ring only pays for shallow K loops (stem, stage-2 1x1, the 3x3 of stages 2-3).
  // From 8 pipeline iterations per tile on, its shared memory is worth more as pipeline stages: direct epilogue
  // (measured per layer at B=32: res4*/res5*_branch2b, res3b-d/res4*_branch2a, P3/P5_lateral, P4_mid, P4 4-10 % faster
  // direct; conv1, res2a_branch2a, the projection shortcuts and res3*_branch2b 5-40 % faster staged).
  static const bool direct_deep = [] {
    const char* e = getenv("PPB_DIRECT_DEEP");
    return e == nullptr || atoi(e) != 0;
  }();
  if (direct_deep && staged_epilogue == 2 && p.staged && res == nullptr && !p.stem16 &&
      (p.slab3 != 0 ? 3 : p.n_taps) * p.cin_blocks >= 8)
    force_direct = true;
  int rc = p.stem16 ? conv_encode_stem_tmaps(&op.ta, &op.tb, x->ptr, a_rows, l.w, l.cout_pad, l.k_total, p.block_n)
                    : conv_encode_tmaps(&op.ta, &op.tb, x->ptr, a_rows, a_cols, a_ld, l.w, l.cout_pad, l.k_dev,
                                        p.pair ? p.block_n / 2 : p.block_n, p.slab3 ? kConvSlabRows : kConvBlockM);
  if (rc != 0) return fail(kErrCuda, "cuTensorMapEncodeTiled failed for layer '%s' (%d)", l.name.c_str(), rc);
  memset(&op.to, 0, sizeof(op.to));
  memset(&op.tr, 0, sizeof(op.tr));
  p.staged = (!force_direct && !split() && l.cout % 64 == 0 && p.block_n % 64 == 0 &&
              (staged_epilogue >= 2 || (staged_epilogue == 1 && res != nullptr))) ? 1 : 0;
  if (p.staged) {
    rc = conv_encode_rows_tmap(&op.to, y->ptr, y->rows(batch), y->C, y->ld);
    if (rc == 0 && res != nullptr) rc = conv_encode_rows_tmap(&op.tr, res->ptr, res->rows(batch), res->C, res->ld);
    if (rc != 0) return fail(kErrCuda, "cuTensorMapEncodeTiled (output) failed for layer '%s' (%d)", l.name.c_str(), rc);
  }
  {  // experiments (measured in r02): tile walk direction alternating by layer, L2 eviction hints
    static const int snake = [] { const char* e = getenv("PPB_SNAKE"); return e ? atoi(e) : 0; }();
    static const int hint = [] { const char* e = getenv("PPB_L2HINT"); return e ? atoi(e) : 0; }();
    int n_conv = 0;
    for (const Op& o : ops) n_conv += o.kind == kOpConv;
    p.reverse = (snake != 0 && (n_conv & 1)) ? 1 : 0;
    p.l2_hint = (hint != 0 && !p.pair && !p.stem16) ? 1 : 0;
  }
  if (conv_plan_smem(p) < 0) return fail(kErrInvalid, "layer '%s' does not fit the shared-memory plan", l.name.c_str());
  snprintf(op.name, sizeof(op.name), "%s", l.name.c_str());
  op.flops = 2.0 * batch * y->H * y->W * static_cast<double>(l.cout) * l.ksize * l.ksize * l.cin;
  ops.push_back(op);
  return kOk;
}

// fp32 path: six channel-block passes per tap (activation part x weight part, every term >= 2^-16 of the product),
// three-part output (ConvParams::split).
void Engine::conv_set_split(ConvParams& p, int cin, int out_c, int res_c) const {
  p.split = kSplitSegs;
  p.parts = kSplitParts;
  p.a_cb = cin / kConvBlockK;
  for (int sg = 0; sg < kSplitSegs; ++sg) p.a_seg[sg] = kSplitAPart[sg];
  p.cin_blocks = kSplitSegs * p.a_cb;
  p.out_part_off = out_c;
  p.res_part_off = res_c;
}

// Row segments of the multi-level feature matrix (one per pyramid level) with the 3x3 'same' tap offsets.
void Engine::fill_head_segs(ConvParams& p, int batch, int n_levels) const {
  p.n_segs = n_levels;
  int tiles = 0, loc = 0;
  for (int i = 0; i < n_levels; ++i) {
    ConvSeg& s = p.seg[i];
    s.hp = lvl_h[i] + 2;
    s.wp = lvl_w[i] + 2;
    s.pad_lo = 1;
    s.pad_hi = 1;
    s.rows = batch * s.hp * s.wp;
    s.row_begin = seg_row_begin[i];  // laid out for max_batch so the buffers never move
    s.tile_begin = tiles;
    s.loc_base = loc;
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) s.tap_off[r * 3 + c] = (r - 1) * s.wp + (c - 1);
    tiles += (s.rows + kConvBlockM - 1) / kConvBlockM;
    loc += lvl_h[i] * lvl_w[i];
  }
  p.n_taps = 9;
  p.n_mtiles = tiles;
}

// A head convolution shared by `n_levels` pyramid levels: one launch over the multi-segment feature matrix.
int Engine::head_conv_op(const __nv_bfloat16* x, int cin, const std::string& layer, __nv_bfloat16* y, int relu,
                         int batch, int n_levels, int epi, float* out_f32, float* out_hull) {
  const LayerW& l = layers[layer_index.at(layer)];
  if (!l.loaded) return fail(kErrState, "layer '%s' has no weights (call ppb_engine_set_layer first)", l.name.c_str());
  Op op;
  op.kind = kOpConv;
  ConvParams& p = op.p;
  memset(&p, 0, sizeof(p));
  fill_head_segs(p, batch, n_levels);
  const long long last_row = p.seg[n_levels - 1].row_begin + p.seg[n_levels - 1].rows;
  p.cin_blocks = cin / kConvBlockK;
  if (split()) conv_set_split(p, cin, l.cout, 0);
  p.n_valid = l.cout;
  p.block_n = conv_pick_block_n(l.cout_pad, p.n_taps * p.cin_blocks, p.n_mtiles, epi == kEpiBf16,
                                split() ? 128 : kConvMaxBlockN);
  p.n_ntiles = (l.cout_pad + p.block_n - 1) / p.block_n;
  p.epi = epi;
  p.relu = relu;
  p.bias = l.bias;
  p.out = y;
  p.out_ld = split() ? kSplitParts * l.cout : l.cout;
  p.out_f32 = out_f32;
  p.out_hull = out_hull;
  int locs = 0;
  for (int i = 0; i < 3; ++i) locs += lvl_h[i] * lvl_w[i];
  p.locs_per_image = (n_levels == 1) ? lvl_h[0] * lvl_w[0] : locs;
  p.vals_per_loc = l.cout;
  memcpy(p.anchor, base_anchor, sizeof(base_anchor));
  memcpy(p.stride, stride, sizeof(stride));
  memcpy(p.dec_mean, dec_mean, sizeof(dec_mean));
  memcpy(p.dec_std, dec_std, sizeof(dec_std));
  p.pair = conv_pair_ok(p) ? 1 : 0;
  p.staged = (!split() && epi == kEpiBf16 && l.cout % 64 == 0 && p.block_n % 64 == 0 && staged_epilogue >= 3) ? 1 : 0;
  p.slab3 = conv_slab3_mode(p);
  const int a_ld = split() ? kSplitParts * cin : cin;
  int rc = conv_encode_tmaps(&op.ta, &op.tb, x, last_row, a_ld, a_ld, l.w, l.cout_pad, l.k_dev,
                             p.pair ? p.block_n / 2 : p.block_n, p.slab3 ? kConvSlabRows : kConvBlockM);
  if (rc != 0) return fail(kErrCuda, "cuTensorMapEncodeTiled failed for layer '%s' (%d)", l.name.c_str(), rc);
  memset(&op.to, 0, sizeof(op.to));
  memset(&op.tr, 0, sizeof(op.tr));
  p.staged = (!split() && epi == kEpiBf16 && l.cout % 64 == 0 && p.block_n % 64 == 0 && staged_epilogue >= 3) ? 1 : 0;
  if (p.staged) {
    rc = conv_encode_rows_tmap(&op.to, y, last_row, l.cout, l.cout);
    if (rc != 0) return fail(kErrCuda, "cuTensorMapEncodeTiled (output) failed for layer '%s' (%d)", l.name.c_str(), rc);
  }
  if (conv_plan_smem(p) < 0) return fail(kErrInvalid, "layer '%s' does not fit the shared-memory plan", l.name.c_str());
  snprintf(op.name, sizeof(op.name), "%s", l.name.c_str());
  {
    double px = 0;
    for (int i = 0; i < n_levels; ++i) px += static_cast<double>(lvl_h[i]) * lvl_w[i];
    op.flops = 2.0 * batch * px * l.cout * 9.0 * cin;
  }
  ops.push_back(op);
  return kOk;
}

}  // namespace ppb

namespace ppb {

int Engine::build_plan(int batch, const float* images, float* out_reg, float* out_cls, float* out_mask,
                       float* out_hull, int decode) {
  ops.clear();
  auto get = [&](Tensor** t, const std::string& name, int h, int w, int c, int pad, int planes = 1) -> int {
    *t = tensor(name);
    if (*t != nullptr) return kOk;
    return new_tensor(t, name, h, w, c, pad, planes);
  };
  // backbone intermediates: their own buffer, or (production arena) a view of a shared one
  const bool alias = alias_arena && !training_plan;
  auto get_bb = [&](Tensor** t, const std::string& name, const char* pool, int h, int w, int c) -> int {
    *t = tensor(name);
    if (*t != nullptr) return kOk;
    return alias ? alias_tensor(t, name, pool, h, w, c) : new_tensor(t, name, h, w, c, 1);
  };
  auto elementwise = [&](OpKind kind, const void* i0, const void* i1, const void* i2, const void* i3, void* out,
                         std::initializer_list<int> args) {
    Op op;
    op.kind = kind;
    op.in0 = i0;
    op.in1 = i1;
    op.in2 = i2;
    op.in3 = i3;
    op.out = out;
    int k = 0;
    for (int v : args) op.a[k++] = v;
    static const char* kNames[] = {"conv", "stem_s2d", "maxpool", "parity_planes", "add", "stem_conv_fp32"};
    snprintf(op.name, sizeof(op.name), "%s", kNames[kind]);
    ops.push_back(op);
  };
#define PPB_TRY(expr)          \
  do {                         \
    int rc__ = (expr);         \
    if (rc__ != kOk) return rc__; \
  } while (0)

  // ---- stem: ZeroPad(3) + 7x7/2 VALID + BN + ReLU, then MaxPool 3x3/2 'same'
  const int h1 = (H + 1) / 2, w1 = (W + 1) / 2;
  const int h2 = (h1 + 1) / 2, w2 = (w1 + 1) / 2;
  Tensor *s2d, *stem, *x;
  stem = tensor("conv1_relu");
  if (stem == nullptr) PPB_TRY(new_tensor(&stem, "conv1_relu", h1, w1, 64, 0, 1, 0, 3));
  PPB_TRY(get(&x, "pool1", h2, w2, 64, 1));
  if (split()) {  // fp32 path: the stem is one fp32 CUDA-core convolution straight from the image
    const LayerW& l1 = layers[layer_index.at("conv1")];
    if (!l1.loaded) return fail(kErrState, "layer 'conv1' has no weights (call ppb_engine_set_layer first)");
    elementwise(kOpStemSplit, images, l1.w_f32, l1.bias, nullptr, stem->ptr, {batch, H, W, h1, w1, input_u8 ? 1 : 0});
  } else {
    s2d = tensor("conv1_s2d");
    if (s2d == nullptr) PPB_TRY(new_tensor(&s2d, "conv1_s2d", h1, w1, 16, 0, 1, 0, 3));
    elementwise(kOpStemS2d, images, nullptr, nullptr, nullptr, s2d->ptr, {batch, H, W, h1, w1, input_u8 ? 1 : 0});
    PPB_TRY(conv_op(s2d, 1, "conv1", nullptr, stem, 1, batch));
  }
  elementwise(kOpMaxpool, stem->ptr, nullptr, nullptr, nullptr, x->ptr, {batch, h1, w1, stem->hp(), stem->wp(), 64, h2, w2});

  // ---- bottleneck stages
  const int blocks[4] = {3, 4, 6, 3};
  int features = 64, ch = h2, cw = w2, n_block = 0;
  Tensor* c_out[4] = {nullptr, nullptr, nullptr, nullptr};
  for (int stage = 0; stage < 4; ++stage) {
    for (int b = 0; b < blocks[stage]; ++b) {
      char tagc[8];
      snprintf(tagc, sizeof(tagc), "%d%c", stage + 2, 'a' + b);
      const std::string tag(tagc);
      const bool down = (b == 0 && stage > 0);
      Tensor* xin = x;
      if (down) {  // 1x1 stride-2 VALID: out(i,j) = in(2i,2j)
        const int nh = (ch + 1) / 2, nw = (cw + 1) / 2;
        Tensor* sub;
        PPB_TRY(get_bb(&sub, "res" + tag + "_sub", "bb_sub", nh, nw, x->C));
        elementwise(kOpPlanes, x->ptr, nullptr, nullptr, nullptr, sub->ptr, {batch, ch, cw, x->ld, nh, nw, 1});
        xin = sub;
        ch = nh;
        cw = nw;
      }
      Tensor *y1, *y2, *y3, *sc = x;
      PPB_TRY(get_bb(&y1, "res" + tag + "_branch2a_relu", "bb_y1", ch, cw, features));
      PPB_TRY(get_bb(&y2, "res" + tag + "_branch2b_relu", "bb_y2", ch, cw, features));
      // block outputs ping-pong (a block reads its input as the residual while it writes its output); C3, C4, C5 —
      // the last block of stages 3-5 — feed the pyramid much later and keep their own buffer
      if (stage > 0 && b == blocks[stage] - 1) PPB_TRY(get(&y3, "res" + tag + "_relu", ch, cw, features * 4, 1));
      else PPB_TRY(get_bb(&y3, "res" + tag + "_relu", (n_block & 1) ? "bb_x1" : "bb_x0", ch, cw, features * 4));
      ++n_block;
      PPB_TRY(conv_op(xin, 1, "res" + tag + "_branch2a", nullptr, y1, 1, batch));
      PPB_TRY(conv_op(y1, 1, "res" + tag + "_branch2b", nullptr, y2, 1, batch));
      if (b == 0) {
        PPB_TRY(get_bb(&sc, "res" + tag + "_branch1_bn", "bb_sc", ch, cw, features * 4));
        PPB_TRY(conv_op(xin, 1, "res" + tag + "_branch1", nullptr, sc, 0, batch));
      }
      PPB_TRY(conv_op(y2, 1, "res" + tag + "_branch2c", sc, y3, 1, batch));
      x = y3;
    }
    c_out[stage] = x;
    features *= 2;
  }
  Tensor *C3 = c_out[1], *C4 = c_out[2], *C5 = c_out[3];
  PPB_REQUIRE(C3->H == lvl_h[0] && C3->W == lvl_w[0] && C5->H == lvl_h[2] && C5->W == lvl_w[2],
              "internal: pyramid level shapes do not match the backbone outputs");

  // ---- __create_sparceFPN (retinanet.py:180-214)
  Tensor *P3l, *P4l, *P5l, *t4, *t3, *P4m, *P3m, *s3, *P3d, *t3f, *t4f, *s4, *P4d, *t5f;
  PPB_TRY(get(&P3l, "P3_lateral", C3->H, C3->W, 256, 1));
  PPB_TRY(get(&P4l, "P4_lateral", C4->H, C4->W, 256, 1));
  PPB_TRY(get(&P5l, "P5_lateral", C5->H, C5->W, 256, 1));
  PPB_TRY(get(&t4, "P4_mid_in", C4->H, C4->W, 256, 1));
  PPB_TRY(get(&t3, "P3_mid_in", C3->H, C3->W, 256, 1));
  PPB_TRY(get(&P4m, "P4_mid", C4->H, C4->W, 256, 1));
  PPB_TRY(get(&P3m, "P3_mid", C3->H, C3->W, 256, 1));
  PPB_TRY(get(&s3, "P3_mid_planes", C4->H, C4->W, 256, 1, 4));
  PPB_TRY(get(&P3d, "P3_down", C4->H, C4->W, 256, 1));
  PPB_TRY(get(&t3f, "P3_fin", C3->H, C3->W, 256, 1));
  PPB_TRY(get(&t4f, "P4_fin", C4->H, C4->W, 256, 1));
  PPB_TRY(get(&s4, "P4_mid_planes", C5->H, C5->W, 256, 1, 4));
  PPB_TRY(get(&P4d, "P4_down", C5->H, C5->W, 256, 1));
  PPB_TRY(get(&t5f, "P5_fin", C5->H, C5->W, 256, 1));
  if (fuse_adds && !split()) {
    // Every Add / UpsampleLike of the pyramid leaves with the convolution that produces its first operand
    // (ConvParams::out2): no elementwise launch, and the operand is never re-read from HBM.
    Tensor* P4ml;  // P4_mid + P4_lateral, the part of P4_fin (:206, 208) that does not wait for P3_down
    PPB_TRY(get(&P4ml, "P4_mid_plus_lateral", C4->H, C4->W, 256, 1));
    PPB_TRY(conv_op(C5, 1, "P5_lateral", nullptr, P5l, 0, batch));                                   // :194
    PPB_TRY(conv_op(C4, 1, "P4_lateral", nullptr, P4l, 0, batch, 0, 0, t4, P5l, true));              // :193 + :196, 198
    PPB_TRY(conv_op(C3, 1, "P3_lateral", nullptr, P3l, 0, batch, 0, 0, t3, P4l, true));              // :192 + :197, 200
    PPB_TRY(conv_op(t4, 1, "P4_mid", nullptr, P4m, 0, batch, 0, 0, P4ml, P4l, false));               // :199 (+ P4l of :208)
    PPB_TRY(conv_op(t3, 1, "P3_mid", nullptr, P3m, 0, batch, 0, 0, t3f, P3l, false));                // :201 + :203
    elementwise(kOpPlanes, P3m->ptr, nullptr, nullptr, nullptr, s3->ptr, {batch, C3->H, C3->W, P3m->ld, C4->H, C4->W, 4});
    PPB_TRY(conv_op(s3, 4, "P3_down", nullptr, P3d, 0, batch, C3->H, C3->W, t4f, P4ml, false));      // :202 + :206, 208
    elementwise(kOpPlanes, P4m->ptr, nullptr, nullptr, nullptr, s4->ptr, {batch, C4->H, C4->W, P4m->ld, C5->H, C5->W, 4});
    PPB_TRY(conv_op(s4, 4, "P4_down", nullptr, P4d, 0, batch, C4->H, C4->W, t5f, P5l, false));       // :207 + :211
  } else {
    PPB_TRY(conv_op(C3, 1, "P3_lateral", nullptr, P3l, 0, batch));                                   // :192
    PPB_TRY(conv_op(C4, 1, "P4_lateral", nullptr, P4l, 0, batch));                                   // :193
    PPB_TRY(conv_op(C5, 1, "P5_lateral", nullptr, P5l, 0, batch));                                   // :194
    elementwise(kOpAdd, P4l->ptr, nullptr, nullptr, P5l->ptr, t4->ptr, {batch, C4->H, C4->W, 256, C5->H, C5->W});  // :196,198
    PPB_TRY(conv_op(t4, 1, "P4_mid", nullptr, P4m, 0, batch));                                       // :199
    elementwise(kOpAdd, P3l->ptr, nullptr, nullptr, P4l->ptr, t3->ptr, {batch, C3->H, C3->W, 256, C4->H, C4->W});  // :197,200
    PPB_TRY(conv_op(t3, 1, "P3_mid", nullptr, P3m, 0, batch));                                       // :201
    elementwise(kOpPlanes, P3m->ptr, nullptr, nullptr, nullptr, s3->ptr, {batch, C3->H, C3->W, P3m->ld, C4->H, C4->W, 4});
    PPB_TRY(conv_op(s3, 4, "P3_down", nullptr, P3d, 0, batch, C3->H, C3->W));                        // :202
    elementwise(kOpAdd, P3m->ptr, P3l->ptr, nullptr, nullptr, t3f->ptr, {batch, C3->H, C3->W, 256, 0, 0});  // :203
    elementwise(kOpAdd, P3d->ptr, P4m->ptr, P4l->ptr, nullptr, t4f->ptr, {batch, C4->H, C4->W, 256, 0, 0});  // :206,208
    elementwise(kOpPlanes, P4m->ptr, nullptr, nullptr, nullptr, s4->ptr, {batch, C4->H, C4->W, P4m->ld, C5->H, C5->W, 4});
    PPB_TRY(conv_op(s4, 4, "P4_down", nullptr, P4d, 0, batch, C4->H, C4->W));                        // :207
    elementwise(kOpAdd, P4d->ptr, P5l->ptr, nullptr, nullptr, t5f->ptr, {batch, C5->H, C5->W, 256, 0, 0});  // :211
  }

  // ---- pyramid outputs P3, P4, P5 written into ONE multi-segment feature matrix
  Tensor* feat = tensor("features");
  if (feat == nullptr) PPB_TRY(new_tensor(&feat, "features", 0, 0, 256, 1, 1, feat_rows_total));
  Tensor fview[3];
  Tensor* fin[3] = {t3f, t4f, t5f};
  const char* pname[3] = {"P3", "P4", "P5"};
  for (int i = 0; i < 3; ++i) {
    fview[i].ptr = feat->ptr + static_cast<size_t>(seg_row_begin[i]) * feat->ld;
    fview[i].H = lvl_h[i];
    fview[i].W = lvl_w[i];
    fview[i].C = 256;
    fview[i].ld = feat->ld;
    PPB_TRY(conv_op(fin[i], 1, pname[i], nullptr, &fview[i], 0, batch));                           // :204,209,212
  }

  // ---- heads: every tower conv is one launch over the three levels (retinanet.py:224-225)
  auto tower_buf = [&](Tensor** t, const std::string& name, int c) -> int {
    *t = tensor(name);
    if (*t != nullptr) return kOk;
    return new_tensor(t, name, 0, 0, c, 1, 1, feat_rows_total);
  };
  // inference ping-pongs two buffers per tower; the training plan keeps all four activations (ReLU masks / wgrad inputs)
  Tensor *rt[4], *ct[4], *mt[4];
  for (int i = 0; i < 4; ++i) {
    const std::string sfx = training_plan ? "_t" + std::to_string(i) : ((i & 1) ? "_b" : "_a");
    PPB_TRY(tower_buf(&rt[i], "reg_tower" + sfx, 512));
    PPB_TRY(tower_buf(&ct[i], "cls_tower" + sfx, 256));
    PPB_TRY(tower_buf(&mt[i], "mask_tower" + sfx, 256));
  }
  const int kRaw = decode ? kEpiHeadDecode : kEpiHeadRaw;
  // Head order: classification, regression, mask. The two large outputs are complete as early as possible, so a
  // caller that wants them on the host (predict_on_batch: 234 MB per batch of 32) can copy 'cls' while the regression
  // head computes and 'boxes3D' while the mask head does (part_end / ppb_engine_run_part).
  PPB_TRY(head_conv_op(feat->ptr, 256, "pyramid_classification_0", ct[0]->ptr, 1, batch, 3, kEpiBf16, nullptr, nullptr));
  for (int i = 1; i < 4; ++i)
    PPB_TRY(head_conv_op(ct[i - 1]->ptr, 256, "pyramid_classification_" + std::to_string(i), ct[i]->ptr, 1, batch, 3,
                         kEpiBf16, nullptr, nullptr));
  PPB_TRY(head_conv_op(ct[3]->ptr, 256, "pyramid_classification", nullptr, 0, batch, 3, kEpiHeadSigmoid, out_cls, nullptr));
  part_end[0] = static_cast<int>(ops.size());
  PPB_TRY(head_conv_op(feat->ptr, 256, "pyramid_regression3D_0", rt[0]->ptr, 1, batch, 3, kEpiBf16, nullptr, nullptr));
  for (int i = 1; i < 4; ++i)
    PPB_TRY(head_conv_op(rt[i - 1]->ptr, 512, "pyramid_regression3D_" + std::to_string(i), rt[i]->ptr, 1, batch, 3,
                         kEpiBf16, nullptr, nullptr));
  PPB_TRY(head_conv_op(rt[3]->ptr, 512, "pyramid_regression3D", nullptr, 0, batch, 3, kRaw, out_reg,
                       decode ? out_hull : nullptr));
  part_end[1] = static_cast<int>(ops.size());
  if (out_mask != nullptr) {  // mask head on P3 only (retinanet.py:296-297)
    PPB_TRY(head_conv_op(feat->ptr, 256, "mask_0", mt[0]->ptr, 1, batch, 1, kEpiBf16, nullptr, nullptr));
    for (int i = 1; i < 4; ++i)
      PPB_TRY(head_conv_op(mt[i - 1]->ptr, 256, "mask_" + std::to_string(i), mt[i]->ptr, 1, batch, 1, kEpiBf16, nullptr,
                           nullptr));
    PPB_TRY(head_conv_op(mt[3]->ptr, 256, "mask_out", nullptr, 0, batch, 1, kEpiHeadSigmoid, out_mask, nullptr));
  }
  part_end[2] = static_cast<int>(ops.size());
  drop_parts();
#undef PPB_TRY
  bound_batch = batch;
  plan_batch = batch;
  return kOk;
}

int Engine::launch_range(int begin, int end, cudaStream_t st) {
  for (int oi = begin; oi < end; ++oi) {
    const Op& op = ops[oi];
    switch (op.kind) {
      case kOpConv:
        PPB_CUDA(conv_launch(op.ta, op.tb, op.to, op.tr, op.p, 0, st));
        break;
      case kOpStemS2d:
        if (op.a[5])
          PPB_CUDA(launch_stem_s2d_u8(static_cast<const unsigned char*>(op.in0), input_mean,
                                      static_cast<__nv_bfloat16*>(op.out), op.a[0], op.a[1], op.a[2], op.a[3], op.a[4],
                                      st));
        else
          PPB_CUDA(launch_stem_s2d(static_cast<const float*>(op.in0), static_cast<__nv_bfloat16*>(op.out), op.a[0],
                                   op.a[1], op.a[2], op.a[3], op.a[4], st));
        break;
      case kOpStemSplit:
        PPB_CUDA(launch_stem_conv_split(op.in0, op.a[5], input_mean, static_cast<const float*>(op.in1),
                                        static_cast<const float*>(op.in2), static_cast<__nv_bfloat16*>(op.out), op.a[0],
                                        op.a[1], op.a[2], op.a[3], op.a[4], st));
        break;
      case kOpMaxpool:
        if (split()) {
          PPB_CUDA(launch_maxpool3x3s2_split(static_cast<const __nv_bfloat16*>(op.in0), static_cast<__nv_bfloat16*>(op.out),
                                             op.a[0], op.a[1], op.a[2], op.a[3], op.a[4], op.a[5], op.a[6], op.a[7], st));
          break;
        }
        PPB_CUDA(launch_maxpool3x3s2(static_cast<const __nv_bfloat16*>(op.in0), static_cast<__nv_bfloat16*>(op.out),
                                     op.a[0], op.a[1], op.a[2], op.a[3], op.a[4], op.a[5], op.a[6], op.a[7], st));
        break;
      case kOpPlanes:
        PPB_CUDA(launch_parity_planes(static_cast<const __nv_bfloat16*>(op.in0), static_cast<__nv_bfloat16*>(op.out),
                                      op.a[0], op.a[1], op.a[2], op.a[3], op.a[4], op.a[5], op.a[6], st));
        break;
      case kOpAdd:
        if (split()) {
          PPB_CUDA(launch_add_split(static_cast<const __nv_bfloat16*>(op.in0), static_cast<const __nv_bfloat16*>(op.in1),
                                    static_cast<const __nv_bfloat16*>(op.in2), static_cast<const __nv_bfloat16*>(op.in3),
                                    op.a[4], op.a[5], static_cast<__nv_bfloat16*>(op.out), op.a[0], op.a[1], op.a[2],
                                    op.a[3], st));
          break;
        }
        PPB_CUDA(launch_add(static_cast<const __nv_bfloat16*>(op.in0), static_cast<const __nv_bfloat16*>(op.in1),
                            static_cast<const __nv_bfloat16*>(op.in2), static_cast<const __nv_bfloat16*>(op.in3),
                            op.a[4], op.a[5], static_cast<__nv_bfloat16*>(op.out), op.a[0], op.a[1], op.a[2], op.a[3],
                            st));
        break;
    }
  }
  return kOk;
}

int Engine::capture(cudaStream_t st) {
  if (graph_exec) {
    cudaGraphExecDestroy(graph_exec);
    graph_exec = nullptr;
  }
  if (graph) {
    cudaGraphDestroy(graph);
    graph = nullptr;
  }
  PPB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
  const int rc = launch_all(st);
  cudaGraph_t g = nullptr;
  const cudaError_t end = cudaStreamEndCapture(st, &g);
  if (rc != kOk) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  PPB_CUDA(end);
  graph = g;
  PPB_CUDA(cudaGraphInstantiate(&graph_exec, graph, 0));
  graph_kernels = static_cast<int>(ops.size());
  return kOk;
}

int Engine::capture_parts(cudaStream_t st) {
  drop_parts();
  int begin = 0;
  for (int i = 0; i < 3; ++i) {
    PPB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const int rc = launch_range(begin, part_end[i], st);
    cudaGraph_t g = nullptr;
    const cudaError_t end = cudaStreamEndCapture(st, &g);
    if (rc != kOk) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    PPB_CUDA(end);
    part_graph[i] = g;
    PPB_CUDA(cudaGraphInstantiate(&part_exec[i], g, 0));
    begin = part_end[i];
  }
  return kOk;
}

}  // namespace ppb

using namespace ppb;

extern "C" {

int ppb_engine_create(ppb_engine** out, int device, int max_batch, int height, int width, int num_classes,
                      int num_anchors) {
  return ppb_engine_create_ex(out, device, max_batch, height, width, num_classes, num_anchors, 0);
}

int ppb_engine_create_ex(ppb_engine** out, int device, int max_batch, int height, int width, int num_classes,
                         int num_anchors, int precision) {
  PPB_REQUIRE(out != nullptr, "ppb_engine_create: null out");
  PPB_REQUIRE(precision == 0 || precision == 1, "ppb_engine_create_ex: precision must be 0 (bf16) or 1 (fp32 path), got %d",
              precision);
  PPB_REQUIRE(max_batch > 0 && height >= 32 && width >= 32, "ppb_engine_create: bad shape batch=%d h=%d w=%d",
              max_batch, height, width);
  PPB_REQUIRE(num_classes > 0 && num_classes * num_anchors <= 2000 && num_anchors > 0 && num_anchors <= 9,
              "ppb_engine_create: unsupported head size classes=%d anchors=%d (classes*anchors <= 2000, anchors <= 9)",
              num_classes, num_anchors);
  PPB_CUDA(cudaSetDevice(device));
  std::unique_ptr<ppb_engine> h(new ppb_engine);
  Engine& e = h->e;
  e.device = device;
  e.max_batch = max_batch;
  e.H = height;
  e.W = width;
  e.num_classes = num_classes;
  e.num_anchors = num_anchors;
  e.precision = precision;
  int rows = 0, tiles = 0;
  for (int i = 0; i < 3; ++i) {  // guess_shapes (PyraPose/utils/anchors.py:357-369): (dim + 2^l - 1) // 2^l
    const int s = 8 << i;
    e.lvl_h[i] = (height + s - 1) / s;
    e.lvl_w[i] = (width + s - 1) / s;
    e.seg_row_begin[i] = rows;
    e.seg_tile_begin[i] = tiles;
    const int r = max_batch * (e.lvl_h[i] + 2) * (e.lvl_w[i] + 2);
    const int t = (r + kConvBlockM - 1) / kConvBlockM;
    rows += t * kConvBlockM;
    tiles += t;
  }
  e.feat_rows_total = rows;
  e.feat_tiles_total = tiles;
  for (int i = 0; i < 16; ++i) {
    e.dec_mean[i] = 0.0f;
    e.dec_std[i] = 0.2f;
  }
  memset(e.base_anchor, 0, sizeof(e.base_anchor));
  memset(e.stride, 0, sizeof(e.stride));
  if (const char* env = getenv("PPB_STAGED_EPILOGUE")) e.staged_epilogue = atoi(env);
  if (const char* env = getenv("PPB_STEM16")) e.stem16 = atoi(env) != 0;
  if (const char* env = getenv("PPB_FUSE_ADD")) e.fuse_adds = atoi(env) != 0;
  e.define_layers();
  int rc = e.allocate_weights();
  if (rc != kOk) return rc;
  *out = h.release();
  return kOk;
}

void ppb_engine_destroy(ppb_engine* h) {
  if (h == nullptr) return;
  cudaSetDevice(h->e.device);
  delete h;
}

int ppb_engine_num_layers(const ppb_engine* h) { return h ? static_cast<int>(h->e.layers.size()) : 0; }

int ppb_engine_layer_info(const ppb_engine* h, int index, char* name, int name_cap, int* ksize, int* cin, int* cout,
                          int* stride, int* has_bn, int* has_bias) {
  PPB_REQUIRE(h != nullptr && index >= 0 && index < static_cast<int>(h->e.layers.size()),
              "ppb_engine_layer_info: bad index %d", index);
  const LayerW& l = h->e.layers[index];
  if (name != nullptr && name_cap > 0) snprintf(name, name_cap, "%s", l.name.c_str());
  if (ksize) *ksize = l.ksize;
  if (cin) *cin = l.cin;
  if (cout) *cout = l.cout;
  if (stride) *stride = l.stride;
  if (has_bn) *has_bn = l.has_bn;
  if (has_bias) *has_bias = l.has_bias;
  return kOk;
}

int ppb_engine_set_layer(ppb_engine* h, const char* name, const float* kernel_hwio_host, const float* bias_host,
                         const float* bn_gamma_host, const float* bn_beta_host, const float* bn_mean_host,
                         const float* bn_var_host, float bn_eps) {
  PPB_REQUIRE(h != nullptr && name != nullptr, "ppb_engine_set_layer: null argument");
  PPB_CUDA(cudaSetDevice(h->e.device));
  const int rc = h->e.set_layer(name, kernel_hwio_host, bias_host, bn_gamma_host, bn_beta_host, bn_mean_host,
                                bn_var_host, bn_eps);
  if (rc != kOk) return rc;
  return train_refresh_layer(h->e, h->e.layer_index.at(name));  // an engine that has trained keeps its master copy in step
}

int ppb_engine_set_anchors(ppb_engine* h, const float* base_anchors_host, const float* strides_host,
                           const float* mean16_host, const float* std16_host) {
  PPB_REQUIRE(h != nullptr && base_anchors_host != nullptr && strides_host != nullptr,
              "ppb_engine_set_anchors: null argument");
  Engine& e = h->e;
  for (int l = 0; l < 3; ++l) {
    for (int a = 0; a < e.num_anchors; ++a)
      for (int j = 0; j < 4; ++j) e.base_anchor[l][a][j] = base_anchors_host[(l * e.num_anchors + a) * 4 + j];
    e.stride[l] = strides_host[l];
  }
  if (mean16_host) memcpy(e.dec_mean, mean16_host, sizeof(e.dec_mean));
  if (std16_host) memcpy(e.dec_std, std16_host, sizeof(e.dec_std));
  e.anchors_set = true;
  e.bound_batch = 0;  // parameters are baked into the plan: force a re-bind
  return kOk;
}

int ppb_engine_set_arena_alias(ppb_engine* h, int enable) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_set_arena_alias: null engine");
  Engine& e = h->e;
  if (!e.tensors.empty() && (enable != 0) != e.alias_arena)
    return fail(kErrState, "ppb_engine_set_arena_alias: call before the first ppb_engine_bind (the arena is already laid out)");
  e.alias_arena = enable != 0;
  return kOk;
}

int ppb_engine_set_input_u8(ppb_engine* h, int enable, const float* mean3_host) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_set_input_u8: null engine");
  Engine& e = h->e;
  e.input_u8 = enable != 0;
  if (mean3_host != nullptr)
    for (int i = 0; i < 3; ++i) e.input_mean[i] = mean3_host[i];
  e.bound_batch = 0;  // the input type is baked into the plan: force a re-bind
  return kOk;
}

int ppb_engine_bind(ppb_engine* h, int batch, const float* images, float* out_box3d, float* out_cls, float* out_mask,
                    float* out_hull, int decode, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_bind: null engine");
  Engine& e = h->e;
  PPB_REQUIRE(batch > 0 && batch <= e.max_batch, "ppb_engine_bind: batch %d outside [1, %d]", batch, e.max_batch);
  PPB_REQUIRE(images != nullptr && out_box3d != nullptr && out_cls != nullptr, "ppb_engine_bind: null buffer");
  PPB_REQUIRE((reinterpret_cast<uintptr_t>(out_box3d) & 15) == 0 && (reinterpret_cast<uintptr_t>(out_hull) & 15) == 0,
              "ppb_engine_bind: outputs must be 16-byte aligned");
  if (decode && !e.anchors_set) return fail(kErrState, "ppb_engine_bind: decode requested before ppb_engine_set_anchors");
  PPB_CUDA(cudaSetDevice(e.device));
  int rc = e.build_plan(batch, images, out_box3d, out_cls, out_mask, out_hull, decode);
  if (rc != kOk) {
    e.bound_batch = 0;
    return rc;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaStream_t own = nullptr;
  if (st == nullptr) {  // the legacy default stream cannot be captured
    PPB_CUDA(cudaStreamCreateWithFlags(&own, cudaStreamNonBlocking));
    st = own;
  }
  rc = e.capture(st);
  if (own) cudaStreamDestroy(own);
  if (rc != kOk) e.bound_batch = 0;
  return rc;
}

int ppb_engine_run(ppb_engine* h, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_run: null engine");
  Engine& e = h->e;
  if (e.bound_batch == 0 || e.graph_exec == nullptr) return fail(kErrState, "ppb_engine_run: call ppb_engine_bind first");
  PPB_CUDA(cudaGraphLaunch(e.graph_exec, static_cast<cudaStream_t>(stream)));
  count_launch(e.graph_kernels);
  return kOk;
}

/* One of the three parts of the forward pass (0: everything up to and including the classification head — 'cls' is
 * complete afterwards; 1: regression head — 'boxes3D' / '3Dbox' complete; 2: mask head). Running parts 0, 1, 2 in order
 * on one stream equals ppb_engine_run; between them the caller may record events and start copying finished outputs. */
int ppb_engine_run_part(ppb_engine* h, int part, void* stream) {
  PPB_REQUIRE(h != nullptr && part >= 0 && part < 3, "ppb_engine_run_part: bad argument (part %d)", part);
  Engine& e = h->e;
  if (e.bound_batch == 0 || e.graph_exec == nullptr) return fail(kErrState, "ppb_engine_run_part: call ppb_engine_bind first");
  PPB_CUDA(cudaSetDevice(e.device));
  if (e.part_exec[0] == nullptr) {
    cudaStream_t own = nullptr;
    PPB_CUDA(cudaStreamCreateWithFlags(&own, cudaStreamNonBlocking));
    const int rc = e.capture_parts(own);
    cudaStreamDestroy(own);
    if (rc != kOk) return rc;
  }
  PPB_CUDA(cudaGraphLaunch(e.part_exec[part], static_cast<cudaStream_t>(stream)));
  count_launch((part == 0 ? e.part_end[0] : e.part_end[part] - e.part_end[part - 1]));
  return kOk;
}

/* Same work as ppb_engine_run but as individual launches (no graph): for profiling with ncu. */
int ppb_engine_run_eager(ppb_engine* h, void* stream) {
  PPB_REQUIRE(h != nullptr, "ppb_engine_run_eager: null engine");
  Engine& e = h->e;
  if (e.bound_batch == 0) return fail(kErrState, "ppb_engine_run_eager: call ppb_engine_bind first");
  PPB_CUDA(cudaSetDevice(e.device));
  const int rc = e.launch_all(static_cast<cudaStream_t>(stream));
  if (rc == kOk) count_launch(static_cast<int>(e.ops.size()));
  return rc;
}

int ppb_engine_kernels_per_run(const ppb_engine* h) { return h ? static_cast<int>(h->e.ops.size()) : 0; }

size_t ppb_engine_device_bytes(const ppb_engine* h) { return h ? h->e.arena_bytes : 0; }

/* Debug/parity: copies a named intermediate activation to the host as fp32 NHWC [batch,H,W,C] (border removed).
 * Synchronises the device. shape4_host receives {batch,H,W,C}; pass out_host = NULL to query the shape only. */
int ppb_engine_activation(ppb_engine* h, const char* name, float* out_host, int* shape4_host) {
  PPB_REQUIRE(h != nullptr && name != nullptr, "ppb_engine_activation: null argument");
  Engine& e = h->e;
  PPB_REQUIRE(e.plan_batch > 0, "ppb_engine_activation: nothing bound");
  Tensor view;
  Tensor* t = e.tensor(name);
  const std::string nm(name);
  const size_t at = nm.find('@');
  if (t == nullptr && at != std::string::npos) {  // "<multi-level tensor>@<level>": one pyramid level of a head buffer
    Tensor* base = e.tensor(nm.substr(0, at));
    const int i = atoi(nm.c_str() + at + 1);
    PPB_REQUIRE(base != nullptr && i >= 0 && i < 3, "ppb_engine_activation: unknown tensor '%s'", name);
    view.ptr = base->ptr + static_cast<size_t>(e.seg_row_begin[i]) * base->ld;
    view.H = e.lvl_h[i];
    view.W = e.lvl_w[i];
    view.C = base->C;
    view.ld = base->ld;
    t = &view;
  }
  if (t == nullptr && (nm == "P3" || nm == "P4" || nm == "P5")) {
    Tensor* feat = e.tensor("features");
    PPB_REQUIRE(feat != nullptr, "ppb_engine_activation: features not allocated");
    const int i = nm[1] - '3';
    view.ptr = feat->ptr + static_cast<size_t>(e.seg_row_begin[i]) * feat->ld;
    view.H = e.lvl_h[i];
    view.W = e.lvl_w[i];
    view.C = 256;
    view.ld = feat->ld;
    t = &view;
  }
  PPB_REQUIRE(t != nullptr && t->H > 0, "ppb_engine_activation: unknown tensor '%s'", name);
  const int B = e.plan_batch;
  if (shape4_host) {
    shape4_host[0] = B;
    shape4_host[1] = t->H;
    shape4_host[2] = t->W;
    shape4_host[3] = t->C;
  }
  if (out_host == nullptr) return kOk;
  PPB_CUDA(cudaSetDevice(e.device));
  PPB_CUDA(cudaDeviceSynchronize());
  const size_t n = static_cast<size_t>(t->rows(B)) * t->ld;
  std::vector<__nv_bfloat16> host(n);
  PPB_CUDA(cudaMemcpy(host.data(), t->ptr, n * sizeof(__nv_bfloat16), cudaMemcpyDeviceToHost));
  const int hp = t->hp(), wp = t->wp();
  for (int b = 0; b < B; ++b)
    for (int y = 0; y < t->H; ++y)
      for (int x = 0; x < t->W; ++x) {
        const __nv_bfloat16* src = host.data() + ((static_cast<size_t>(b) * hp + y + t->pad_lo) * wp + x + t->pad_lo) * t->ld;
        float* dst = out_host + ((static_cast<size_t>(b) * t->H + y) * t->W + x) * t->C;
        for (int c = 0; c < t->C; ++c) {  // fp32 path: value = sum of the parts, smallest first
          float v = 0.0f;
          for (int part = t->ld / t->C - 1; part >= 0; --part) v += __bfloat162float(src[part * t->C + c]);
          dst[c] = v;
        }
      }
  return kOk;
}

/* Times every launch of one forward pass individually (CUDA events on `stream`, eager launches).
 * ms_out[i] / kind_out[i] (0 = tcgen05 conv, 1 = stem space-to-depth, 2 = max-pool, 3 = parity planes, 4 = add) for
 * i < min(cap, kernels_per_run). Synchronises the stream. Used by bench.py for the roofline numbers. */
int ppb_engine_profile(ppb_engine* h, void* stream, float* ms_out_host, int* kind_out_host, int cap) {
  PPB_REQUIRE(h != nullptr && ms_out_host != nullptr, "ppb_engine_profile: null argument");
  Engine& e = h->e;
  if (e.bound_batch == 0) return fail(kErrState, "ppb_engine_profile: call ppb_engine_bind first");
  PPB_CUDA(cudaSetDevice(e.device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int n = static_cast<int>(e.ops.size());
  std::vector<cudaEvent_t> ev(n + 1);
  for (auto& x : ev) PPB_CUDA(cudaEventCreate(&x));
  std::vector<Op> all;
  all.swap(e.ops);
  int rc = kOk;
  PPB_CUDA(cudaEventRecord(ev[0], st));
  for (int i = 0; i < n && rc == kOk; ++i) {
    e.ops.assign(1, all[i]);
    rc = e.launch_all(st);
    cudaEventRecord(ev[i + 1], st);
  }
  e.ops.swap(all);
  cudaError_t sync = cudaStreamSynchronize(st);
  if (rc == kOk && sync == cudaSuccess) {
    for (int i = 0; i < n && i < cap; ++i) {
      cudaEventElapsedTime(&ms_out_host[i], ev[i], ev[i + 1]);
      if (kind_out_host) kind_out_host[i] = static_cast<int>(e.ops[i].kind);
    }
    count_launch(n);
  }
  for (auto& x : ev) cudaEventDestroy(x);
  if (rc != kOk) return rc;
  PPB_CUDA(sync);
  return kOk;
}

/* Cycle-stamped trace of one CTA of one convolution launch (ConvParams::trace): runs op `index` of the bound plan
 * eagerly with the tracing instantiation of the kernel. trace_dev: device buffer of 8 * cap uint2 entries (zeroed by
 * the call); role r (0 producer, 1 MMA issuer, 2 residual producer, 3.. epilogue groups) owns entries [r*cap, (r+1)*cap):
 * entry 0 = {event count, end clock}, then {tag, clock64 low word} pairs. grid_out receives the CTA count. */
int ppb_engine_trace_op(ppb_engine* h, int index, int cta, void* trace_dev, int cap, int* info_out_host, void* stream) {
  PPB_REQUIRE(h != nullptr && trace_dev != nullptr && cap >= 16, "ppb_engine_trace_op: bad argument");
  Engine& e = h->e;
  PPB_REQUIRE(e.bound_batch != 0 || e.plan_batch != 0, "ppb_engine_trace_op: nothing bound");
  PPB_REQUIRE(index >= 0 && index < static_cast<int>(e.ops.size()) && e.ops[index].kind == kOpConv,
              "ppb_engine_trace_op: op %d is not a convolution", index);
  PPB_CUDA(cudaSetDevice(e.device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PPB_CUDA(cudaMemsetAsync(trace_dev, 0, static_cast<size_t>(8) * cap * sizeof(uint2), st));
  Op op = e.ops[index];
  op.p.trace = static_cast<uint2*>(trace_dev);
  op.p.trace_cap = cap;
  op.p.trace_cta = cta;
  PPB_CUDA(conv_launch(op.ta, op.tb, op.to, op.tr, op.p, 0, st));
  count_launch();
  if (info_out_host != nullptr) {  // the launch shape, for the tool's header
    const ConvParams& p = op.p;
    const int v[12] = {p.n_mtiles, p.n_ntiles, p.block_n, p.n_taps, p.cin_blocks, p.n_stages, p.ring_slots, p.pair,
                       p.slab3, p.bres, p.staged, p.epi_groups};
    memcpy(info_out_host, v, sizeof(v));
  }
  return kOk;
}

int ppb_engine_op_info(const ppb_engine* h, int index, char* name, int name_cap, double* flops) {
  PPB_REQUIRE(h != nullptr && index >= 0 && index < static_cast<int>(h->e.ops.size()), "ppb_engine_op_info: bad index");
  if (name != nullptr && name_cap > 0) snprintf(name, name_cap, "%s", h->e.ops[index].name);
  if (flops) *flops = h->e.ops[index].flops;
  return kOk;
}

}  // extern "C"
