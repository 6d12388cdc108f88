// C-ABI entry points that are not tied to the PFPN engine handle: library bookkeeping and the
// stand-alone convolution call (used by the parity tests and by layer-level callers).
#include "../../include/pyrapose_b200.h"
#include "conv_igemm.h"
#include "conv_wgrad.h"
#include "errors.h"
#include "launch_count.h"

#include <atomic>
#include <cstring>

namespace ppb {

char* last_error_buffer() {
  static thread_local char buf[kLastErrorSize] = {0};
  return buf;
}

std::atomic<long long> g_launches{0};

// Fills the segment/tap tables for a stride-1 ksize x ksize 'same' convolution over one padded-flat plane set.
void conv_fill_single_seg(ConvParams& p, int batch, int h, int w, int ksize) {
  p.n_segs = 1;
  ConvSeg& s = p.seg[0];
  s.row_begin = 0;
  s.hp = h + 2;
  s.wp = w + 2;
  s.pad_lo = 1;
  s.pad_hi = 1;
  s.rows = batch * s.hp * s.wp;
  s.tile_begin = 0;
  s.loc_base = 0;
  p.n_taps = ksize * ksize;
  const int half = ksize / 2;
  for (int r = 0; r < ksize; ++r)
    for (int c = 0; c < ksize; ++c) s.tap_off[r * ksize + c] = (r - half) * s.wp + (c - half);
  p.n_mtiles = (s.rows + kConvBlockM - 1) / kConvBlockM;
}

// UMMA N of one tile. Head output convolutions (fp32 epilogues) keep the whole padded Cout in one tile. bf16 layers
// choose among 256 / 128 / 64 by a static model: wave quantisation over the SMs times the relative efficiency of
// the tile shape -- K-deep (tensor-bound) layers prefer N = 256 (N = 64 MMAs are shared-memory-bandwidth bound),
// shallow-K (HBM / epilogue-bound) layers prefer N = 128 so that the epilogue ring holds several tiles of
// residual / output chunks in flight.
int conv_pick_block_n(int cout_pad, int k_iters, long m_tiles, bool bf16_out, int max_n) {
  if (max_n < kConvMaxBlockN) {  // fp32 path: two accumulators share a 256-column buffer (conv_igemm.cu kSideCol)
    if (bf16_out && cout_pad % 64 == 0) return cout_pad % max_n == 0 ? max_n : 64;
    const int nt = (cout_pad + max_n - 1) / max_n;
    return ((cout_pad + nt - 1) / nt + 15) / 16 * 16;
  }
  if (!bf16_out && cout_pad > kConvMaxBlockN) {  // wide head output (e.g. 30 classes x 9 anchors): equal N tiles
    const int nt = (cout_pad + kConvMaxBlockN - 1) / kConvMaxBlockN;
    return ((cout_pad + nt - 1) / nt + 15) / 16 * 16;
  }
  if (!bf16_out || cout_pad % 64 != 0) {
    if (cout_pad <= kConvMaxBlockN) return cout_pad;
    if (cout_pad % 256 == 0) return 256;
    if (cout_pad % 128 == 0) return 128;
    return 64;
  }
  const int sms = conv_num_sms() > 0 ? conv_num_sms() : 148;
  const bool deep = k_iters > 8;
  const int cand[3] = {256, 128, 64};
  const double shape_eff_deep[3] = {1.0, 0.93, 0.67};
  const double shape_eff_shallow[3] = {0.8, 1.0, 0.9};
  int best = 64;
  double best_score = -1.0;
  for (int i = 0; i < 3; ++i) {
    if (cout_pad % cand[i] != 0) continue;
    const long tiles = m_tiles * (cout_pad / cand[i]);
    const long waves = (tiles + sms - 1) / sms;
    const double score = static_cast<double>(tiles) / static_cast<double>(waves * sms) *
                         (deep ? shape_eff_deep[i] : shape_eff_shallow[i]);
    if (score > best_score + 1e-9) {
      best_score = score;
      best = cand[i];
    }
  }
  return best;
}

}  // namespace ppb

using namespace ppb;

extern "C" {

int ppb_version(void) { return 100; }

const char* ppb_last_error(void) { return last_error_buffer(); }

long long ppb_launch_count(void) { return g_launches.load(); }

int ppb_conv2d_bf16(const void* x, const void* weight, const float* bias, const void* residual, void* y, int batch,
                    int h, int w, int cin, int cout, int ksize, int relu, void* stream) {
  PPB_REQUIRE(x != nullptr && weight != nullptr && y != nullptr, "ppb_conv2d_bf16: null buffer");
  PPB_REQUIRE(batch > 0 && h > 0 && w > 0, "ppb_conv2d_bf16: bad shape batch=%d h=%d w=%d", batch, h, w);
  PPB_REQUIRE(ksize == 1 || ksize == 3, "ppb_conv2d_bf16: ksize must be 1 or 3, got %d", ksize);
  PPB_REQUIRE(cin > 0 && cin % kConvBlockK == 0, "ppb_conv2d_bf16: cin must be a multiple of 64, got %d", cin);
  PPB_REQUIRE(cout > 0 && cout % 16 == 0, "ppb_conv2d_bf16: cout must be a multiple of 16, got %d", cout);
  ConvParams p = {};
  conv_fill_single_seg(p, batch, h, w, ksize);
  p.cin_blocks = cin / kConvBlockK;
  p.n_valid = cout;
  p.block_n = conv_pick_block_n(cout, p.n_taps * p.cin_blocks, p.n_mtiles, true);
  p.n_ntiles = (cout + p.block_n - 1) / p.block_n;
  p.epi = kEpiBf16;
  p.relu = relu;
  p.bias = bias;
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.res_ld = cout;
  p.out = static_cast<__nv_bfloat16*>(y);
  p.out_ld = cout;
  CUtensorMap ta, tb, to, tr;
  p.pair = conv_pair_ok(p) ? 1 : 0;
  p.staged = (cout % 64 == 0 && p.block_n % 64 == 0) ? 1 : 0;
  p.slab3 = conv_slab3_mode(p);
  int rc = conv_encode_tmaps(&ta, &tb, x, p.seg[0].rows, cin, cin, weight, cout, ksize * ksize * cin,
                             p.pair ? p.block_n / 2 : p.block_n, p.slab3 ? kConvSlabRows : kConvBlockM);
  if (rc != 0) return fail(kErrCuda, "ppb_conv2d_bf16: cuTensorMapEncodeTiled failed (%d)", rc);
  memset(&to, 0, sizeof(to));
  memset(&tr, 0, sizeof(tr));
  p.staged = (cout % 64 == 0 && p.block_n % 64 == 0) ? 1 : 0;
  if (p.staged) {
    rc = conv_encode_rows_tmap(&to, y, p.seg[0].rows, cout, cout);
    if (rc == 0 && residual != nullptr) rc = conv_encode_rows_tmap(&tr, residual, p.seg[0].rows, cout, cout);
    if (rc != 0) return fail(kErrCuda, "ppb_conv2d_bf16: cuTensorMapEncodeTiled (output) failed (%d)", rc);
  }
  PPB_REQUIRE(conv_plan_smem(p) >= 0, "ppb_conv2d_bf16: cout=%d does not fit the shared-memory plan", cout);
  PPB_CUDA(conv_launch(ta, tb, to, tr, p, 0, static_cast<cudaStream_t>(stream)));
  count_launch();
  return kOk;
}

int ppb_conv2d_wgrad_bf16(const void* x, const void* g, float* dw, int batch, int h, int w, int cin, int cout, int g_ld,
                          int ksize, void* stream) {
  PPB_REQUIRE(x != nullptr && g != nullptr && dw != nullptr, "ppb_conv2d_wgrad_bf16: null buffer");
  PPB_REQUIRE(batch > 0 && h > 0 && w > 0, "ppb_conv2d_wgrad_bf16: bad shape batch=%d h=%d w=%d", batch, h, w);
  PPB_REQUIRE(ksize == 1 || ksize == 3, "ppb_conv2d_wgrad_bf16: ksize must be 1 or 3, got %d", ksize);
  PPB_REQUIRE(cin > 0 && cin % 64 == 0, "ppb_conv2d_wgrad_bf16: cin must be a multiple of 64, got %d", cin);
  PPB_REQUIRE(cout > 0 && g_ld >= cout && g_ld % 8 == 0, "ppb_conv2d_wgrad_bf16: bad cout=%d g_ld=%d", cout, g_ld);
  ConvParams cp = {};
  conv_fill_single_seg(cp, batch, h, w, ksize);
  WgradParams p = {};
  p.n_segs = 1;
  p.seg[0].row_begin = 0;
  p.seg[0].k_iters = (cp.seg[0].rows + 63) / 64;
  for (int t = 0; t < cp.n_taps; ++t) p.seg[0].tap_off[t] = cp.seg[0].tap_off[t];
  p.total_k_iters = p.seg[0].k_iters;
  p.n_taps = cp.n_taps;
  p.cout = cout;
  p.cin = cin;
  p.dw = dw;
  p.dw_ld = static_cast<long long>(cp.n_taps) * cin;
  wgrad_plan(p);
  CUtensorMap tg, tx;
  const int rc = wgrad_encode_tmaps(&tg, &tx, g, cp.seg[0].rows, g_ld, g_ld, x, cp.seg[0].rows, cin, cin);
  if (rc != 0) return fail(kErrCuda, "ppb_conv2d_wgrad_bf16: cuTensorMapEncodeTiled failed (%d)", rc);
  PPB_CUDA(wgrad_launch(tg, tx, p, static_cast<cudaStream_t>(stream)));
  count_launch();
  return kOk;
}

int ppb_conv2d_dgrad_bf16(const void* g, const void* weight_t, const void* residual, const void* relu_mask, void* dx,
                          int batch, int h, int w, int cin, int cout, int g_ld, int ksize, void* stream) {
  PPB_REQUIRE(g != nullptr && weight_t != nullptr && dx != nullptr, "ppb_conv2d_dgrad_bf16: null buffer");
  PPB_REQUIRE(batch > 0 && h > 0 && w > 0, "ppb_conv2d_dgrad_bf16: bad shape batch=%d h=%d w=%d", batch, h, w);
  PPB_REQUIRE(ksize == 1 || ksize == 3, "ppb_conv2d_dgrad_bf16: ksize must be 1 or 3, got %d", ksize);
  PPB_REQUIRE(cin > 0 && cin % 64 == 0, "ppb_conv2d_dgrad_bf16: cin must be a multiple of 64, got %d", cin);
  PPB_REQUIRE(cout > 0 && g_ld >= cout && g_ld % 64 == 0, "ppb_conv2d_dgrad_bf16: bad cout=%d g_ld=%d (multiple of 64)",
              cout, g_ld);
  ConvParams p = {};
  conv_fill_single_seg(p, batch, h, w, ksize);
  for (int t = 0; t < p.n_taps; ++t) p.seg[0].tap_off[t] = -p.seg[0].tap_off[t];  // transposed convolution
  p.cin_blocks = g_ld / kConvBlockK;
  p.n_valid = cin;
  p.block_n = conv_pick_block_n(cin, p.n_taps * p.cin_blocks, p.n_mtiles, true);
  p.n_ntiles = (cin + p.block_n - 1) / p.block_n;
  p.epi = kEpiBf16;
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.res_ld = cin;
  p.mask = static_cast<const __nv_bfloat16*>(relu_mask);
  p.mask_ld = cin;
  p.out = static_cast<__nv_bfloat16*>(dx);
  p.out_ld = cin;
  CUtensorMap ta, tb, to, tr;
  p.pair = conv_pair_ok(p) ? 1 : 0;
  p.staged = (p.block_n % 64 == 0) ? 1 : 0;
  p.slab3 = conv_slab3_mode(p);
  int rc = conv_encode_tmaps(&ta, &tb, g, p.seg[0].rows, g_ld, g_ld, weight_t, cin, ksize * ksize * g_ld,
                             p.pair ? p.block_n / 2 : p.block_n, p.slab3 ? kConvSlabRows : kConvBlockM);
  if (rc != 0) return fail(kErrCuda, "ppb_conv2d_dgrad_bf16: cuTensorMapEncodeTiled failed (%d)", rc);
  memset(&to, 0, sizeof(to));
  memset(&tr, 0, sizeof(tr));
  p.staged = (p.block_n % 64 == 0) ? 1 : 0;
  if (p.staged) {
    rc = conv_encode_rows_tmap(&to, dx, p.seg[0].rows, cin, cin);
    if (rc == 0 && residual != nullptr) rc = conv_encode_rows_tmap(&tr, residual, p.seg[0].rows, cin, cin);
    if (rc != 0) return fail(kErrCuda, "ppb_conv2d_dgrad_bf16: cuTensorMapEncodeTiled (output) failed (%d)", rc);
  }
  PPB_REQUIRE(conv_plan_smem(p) >= 0, "ppb_conv2d_dgrad_bf16: cin=%d does not fit the shared-memory plan", cin);
  PPB_CUDA(conv_launch(ta, tb, to, tr, p, 0, static_cast<cudaStream_t>(stream)));
  count_launch();
  return kOk;
}

}  // extern "C"
