// Weight-gradient GEMM of the implicit-GEMM convolution (training step):
//   dW[co][tap][ci] += sum over rows m of  g[m][co] * x[m + tap_off[tap]][ci]
// with g the (ReLU-masked) output gradient and x the layer input, both bf16 padded-flat matrices (conv_igemm.h).
// M = Cout (128 per tile), N = Cin (64..256 per tile), K = rows: both operands are "MN-major" for the tensor core
// (their contiguous dimension is the channel), which tcgen05 reads directly from [rows][64 channels] TMA boxes.
#pragma once
#include "conv_igemm.h"

namespace ppb {

struct WgradSeg {
  int row_begin;  // first row of the segment in g (and in x before the tap offset); multiple of 128
  int k_iters;    // ceil(rows / 64)
  int tap_off[kConvMaxTaps];
};

struct WgradParams {
  int n_segs;
  WgradSeg seg[kConvMaxSegs];
  int total_k_iters;  // over all segments
  int n_taps;
  int cout, cin;      // valid extents of dW
  int block_n;        // Cin columns per tile: 64, 128 or 256
  int n_mtiles;       // ceil(cout / 128)
  int n_ntiles;       // cin / block_n
  int k_splits;       // tasks sharing one (m, tap, n) tile along K (chosen by wgrad_plan so that tasks fill whole waves)
  int n_stages;
  int pair;           // 1: CTA pairs over two Cout tiles (tcgen05 cta_group::2), set by wgrad_plan() when cout % 256 == 0,
                      // cin tile >= 128 and PPB_PAIR != 0
  float* dw;          // fp32 [cout][n_taps * cin], accumulated with atomics (caller zeroes it)
  long long dw_ld;    // n_taps * cin
};

// g: [g_rows, g_ld] bf16 (g_ld >= cout, multiple of 8); x: [x_rows, x_ld] bf16.
int wgrad_encode_tmaps(CUtensorMap* tmap_g, CUtensorMap* tmap_x, const void* g, long g_rows, int g_cols, int g_ld,
                       const void* x, long x_rows, int x_cols, int x_ld);
void wgrad_plan(WgradParams& p);  // fills block_n, n_mtiles, n_ntiles, k_splits, n_stages from the shape fields
cudaError_t wgrad_launch(const CUtensorMap& tmap_g, const CUtensorMap& tmap_x, const WgradParams& p,
                         cudaStream_t stream);

}  // namespace ppb
