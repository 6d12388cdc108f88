// Internal interface of the tcgen05 implicit-GEMM convolution (shared by api.cu and engine.cu).
//
// Data layout ("padded-flat"): an NHWC bf16 activation at resolution HxW is stored as a 2-D matrix
// [B*(H+2)*(W+2), C] whose one-pixel border is kept at zero. A 3x3 stride-1 'same' convolution then
// is nine shifted GEMMs over the SAME matrix: tap (r,s) reads rows m + (r-1)*(W+2) + (s-1). Every A
// tile is therefore one plain 2-D TMA box {64 channels x 128 rows}; rows that fall outside the
// matrix are zero-filled by TMA. Outputs computed at border positions are written as zeros, which
// re-establishes the invariant for the next layer. Stride-2 convolutions run on four parity planes
// (space-to-depth) stored back to back in one matrix, so their taps are row offsets as well.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ppb {

constexpr int kConvBlockM = 128;
constexpr int kConvBlockK = 64;
constexpr int kConvMaxBlockN = 256;
constexpr int kConvMaxTaps = 9;
constexpr int kConvMaxSegs = 3;

enum ConvEpilogue : int {
  kEpiBf16 = 0,     // bias (+residual) (+ReLU) -> bf16, same row geometry as the input matrix
  kEpiHeadRaw = 1,  // bias -> fp32 [B, locs, vals]            (training-model '3Dbox' output)
  kEpiHeadSigmoid = 2,  // bias, sigmoid -> fp32 [B, locs, vals]   ('cls' and 'mask' outputs)
  kEpiHeadDecode = 3,   // bias, box3D_transform_inv against in-kernel anchors -> fp32 [B, locs*A, 16]
};

// One pyramid level (or the only plane) of the row space the M tiles walk over.
struct ConvSeg {
  int row_begin;   // first row in the A / output matrix; multiple of 128
  int rows;        // valid rows = B*hp*wp
  int hp, wp;      // plane geometry including the border
  int pad_lo;      // border rows/cols before the interior (1 for padded-flat, 0 for the stem's pixel grid)
  int pad_hi;      // border rows/cols after the interior (1 for padded-flat, 3 for the stem's pixel grid)
  int tile_begin;  // index of this segment's first M tile
  int loc_base;    // head epilogues: first location index of this level inside one image
  int tap_off[kConvMaxTaps];  // signed row offset of each filter tap
  // exact unsigned division by hp*wp and by wp as multiply-high + shifts (filled by conv_launch, see conv_fastdiv)
  unsigned plane_mul, plane_sh, wp_mul, wp_sh;
};

// Division of a 32-bit unsigned n by a fixed d >= 1 without the ~25-instruction emulated divide (Granlund-Montgomery
// round-up method, exact for every 32-bit n): t = umulhi(mul, n); q = (t + ((n - t) >> (sh & 255))) >> (sh >> 8).
inline void conv_fastdiv(unsigned d, unsigned* mul, unsigned* sh) {
  unsigned l = 0;
  while ((1ull << l) < d) ++l;  // ceil(log2 d)
  *mul = static_cast<unsigned>(((1ull << 32) * ((1ull << l) - d)) / d + 1);
  const unsigned sh1 = l < 1 ? l : 1;
  *sh = sh1 | ((l - sh1) << 8);
}

struct ConvParams {
  int n_segs;
  ConvSeg seg[kConvMaxSegs];
  int n_taps;
  int cin_blocks;   // Cin / 64
  int n_valid;      // Cout (columns beyond it are padding)
  int block_n;      // UMMA N of one tile: multiple of 16, <= 256
  int n_ntiles;
  int n_mtiles;     // over all segments
  int epi;          // ConvEpilogue
  int relu;
  int staged;       // kEpiBf16 only: epilogue through the shared-memory chunk ring + TMA store (needs tmap_out / tmap_res)
  int pair;         // 1: CTA pairs (cluster of 2, tcgen05 cta_group::2, M = 256 per MMA); needs tmap_b encoded with a
                    // block_n / 2 row box (pass block_n / 2 to conv_encode_tmaps) — see conv_pair_ok()
  int slab3;        // 3x3 stride-1 layers: one pipeline stage = one filter ROW of one 64-channel block. The A operand is
                    // ONE slab of 136 rows (tmap_a encoded with a 136-row box) and the three taps of the row are three
                    // MMAs whose descriptors start 0 / 1 / 2 rows (128 B) into it — the swizzle is a function of the
                    // shared-memory address, so a row-shifted start reads the right data (tools/probe_shift.py). Cuts
                    // the A bytes an SM pulls from L2 by 3x. +1 / -1: the taps of a row ascend / descend (dgrad) in
                    // row offset; 0: off. Set by conv_slab3_mode().
  int stem16;       // stem (7x7/2 over the space-to-depth grid, 16 channels per pixel): a stage = one filter row dy; the A
                    // operand is ONE slab of 136 pixels x 16 channels (32-byte rows, 32B swizzle) and the four column
                    // taps dx are four K = 16 MMAs whose descriptors start dx pixels (dx x 32 B) into it, instead of four
                    // overlapping 128-byte rows per pixel (4x the L2 traffic). Maps from conv_encode_stem_tmaps().
  int bres;         // weights resident in shared memory (filled by conv_plan_smem, see conv_bres_ok in conv_igemm.cu)
  int split;        // fp32-path layers (Engine precision 1): number of K segments per tap, 0 = off. Every
                    // activation value v is stored as `parts` bf16 numbers p0 = bf16(v), p1 = bf16(v - p0), p2 = bf16(v
                    // - p0 - p1) in columns [0, C), [C, 2C), [2C, 3C) of its row (24 significant bits), weights likewise
                    // as w0, w1, w2. A product is the six MMAs whose terms are >= 2^-16 of it, accumulated in the fp32
                    // TMEM accumulator: (a0,w0) (a1,w0) (a0,w1) (a2,w0) (a1,w1) (a0,w2). The weight matrix holds per tap
                    // the column blocks [w0 | w0 | w1 | w0 | w1 | w2]; K-loop channel block kb belongs to segment
                    // s = kb / a_cb and reads A columns (a_seg[s] * a_cb + kb - s * a_cb) * 64. The bf16 epilogue (direct
                    // only) writes part i at column n + i * out_part_off; the residual is the sum of its parts.
  int parts;        // 3 on the fp32 path
  int a_cb;         // channel blocks per segment (C / 64)
  int a_seg[6];     // which activation part each segment reads
  int out_part_off, res_part_off;
  int reverse;      // walk the tiles from the last to the first (engine: every other layer, PPB_SNAKE=1), so a layer
                    // starts on the rows its producer wrote last, which are the ones still in L2
  int l2_hint;      // activation / residual loads carry the evict-first L2 policy (read once, dead afterwards), weight
                    // loads evict-last (PPB_L2HINT=1)
  unsigned nt_mul, nt_sh;  // division by n_ntiles (conv_fastdiv), filled by conv_launch
  int epi_groups;   // epilogue warp groups (2 or 3)     }
  int early_free;   // staged + residual: a ring slot is handed back as soon as ITS store has read it, not when the  }
                    // group's next store is committed (more residual chunks in flight)                            }
  int n_stages;     // TMA->MMA pipeline depth (2..8)   } filled by conv_plan_smem()
  int ring_slots;   // 16 KiB epilogue ring slots        }
  // Cycle-stamped trace of ONE CTA (tools/trace_conv.py; the kTrace instantiations of the kernel): when non-null, CTA
  // trace_cta writes (tag, clock64 low word) pairs at every barrier hand-off of its producer / MMA / residual /
  // epilogue-group threads into trace[role * trace_cap + 1 ...]; entry 0 of a role's region holds its event count.
  // tag = event << 24 | (tile sequence number & 0xffff) << 8 | (k-iteration or chunk & 0xff); events in conv_igemm.cu.
  uint2* trace;
  int trace_cap, trace_cta;
  const float* bias;              // [n_valid] fp32, BatchNorm folded in; may be null
  const __nv_bfloat16* residual;  // kEpiBf16: added before ReLU; row stride res_ld; may be null
  int res_ld;
  const __nv_bfloat16* mask;      // kEpiBf16: ReLU backward — the result is zeroed where mask <= 0 (after the residual add)
  int mask_ld;
  int b_col_base;                 // first column of this launch's taps inside the weight matrix (multiple of 64)
  __nv_bfloat16* out;             // kEpiBf16
  int out_ld;
  // Second output of the DIRECT bf16 epilogue: the Keras Add (+ UpsampleLike) that follows the convolution in
  // __create_sparceFPN (PyraPose/models/retinanet.py:196-211) fused into it:
  //   out2[m] = bf16(float(out[m]) + res2[src(m)])    (out already rounded to bf16: bit-identical to a separate add)
  // src(m) = m when res2_uh == 0, else the row of the nearest-resized pixel of a [B, res2_uh + 2, res2_uw + 2] plane
  // (tf.image.resize NEAREST, half-pixel centres: layers/_misc.py:96-115). Border rows of out2 are written as zeros.
  __nv_bfloat16* out2;
  int out2_ld;
  const __nv_bfloat16* res2;
  int res2_ld, res2_uh, res2_uw;
  float* out_f32;                 // head epilogues
  float* out_hull;                // kEpiHeadDecode: optional [B, locs*A, 4] axis-aligned hull of the 8 corners
  int locs_per_image;             // sum over levels of H*W
  int vals_per_loc;               // A*16, A*C or C
  float anchor[kConvMaxSegs][9][4];  // kEpiHeadDecode: fp32 base anchors per level (Anchors layer)
  float stride[kConvMaxSegs];
  float dec_mean[16], dec_std[16];
};

// Encodes the two TMA descriptors for one convolution. a_rows x a_cols is the activation matrix
// (bf16, row stride a_ld elements); w is [w_rows, w_cols] bf16 row-major (Cout_pad x taps*Cin).
int conv_encode_tmaps(CUtensorMap* tmap_a, CUtensorMap* tmap_b, const void* a, long a_rows, int a_cols, int a_ld,
                      const void* w, int w_rows, int w_cols, int block_n, int a_box_rows = kConvBlockM);

// Stem maps: A = the space-to-depth pixel matrix [a_rows][16] (row stride 16 elements), box {16 x 136}; B = the
// [cout_pad][256] weight matrix, box {16 x block_n}; both 32B-swizzled.
int conv_encode_stem_tmaps(CUtensorMap* tmap_a, CUtensorMap* tmap_b, const void* a, long a_rows, const void* w,
                           int w_rows, int w_cols, int block_n);

constexpr int kConvSlabRows = 136;  // 128 + 2 rows of halo, rounded up to whole 8-row swizzle atoms

// +1 / -1 when the layer can run in slab mode (see ConvParams::slab3; needs n_taps == 9 with unit-spaced taps per
// filter row in every segment, p.pair and p.block_n already chosen), else 0. PPB_SLAB3=0 disables it.
int conv_slab3_mode(const ConvParams& p);

// Tensor map over a row-major bf16 matrix with a {64 cols x 128 rows} box (output store / residual load).
int conv_encode_rows_tmap(CUtensorMap* tmap, const void* base, long rows, int cols, int ld);

// Same with an explicit box height ({64 cols x box_rows rows}, 128B swizzle).
int conv_encode_box_tmap(CUtensorMap* tmap, const void* base, long rows, int cols, int ld, int box_rows);

// Chooses pipeline depth and epilogue-ring size for the shared-memory budget (needs block_n, n_taps, cin_blocks,
// epi, staged, vals_per_loc, n_ntiles). Returns the dynamic shared-memory bytes, or -1 if the shape does not fit.
int conv_plan_smem(ConvParams& p);

// Launches the persistent kernel on `stream` with at most `max_ctas` CTAs (0 = one per SM). tmap_out / tmap_res are
// only dereferenced when p.staged is set (tmap_res only when p.residual != nullptr).
cudaError_t conv_launch(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_out,
                        const CUtensorMap& tmap_res, const ConvParams& p, int max_ctas, cudaStream_t stream);

// Reads (and clears) the fault record a kernel leaves before trapping on a barrier timeout.
void conv_read_fault(unsigned int out[4]);

int conv_num_sms();

// UMMA N of one tile (api.cu). max_n < 256: the fp32 path's limit (two accumulators per buffer).
int conv_pick_block_n(int cout_pad, int k_iters, long m_tiles, bool bf16_out, int max_n = kConvMaxBlockN);

// CTA-pair mode is usable for this shape (N tile a multiple of 16: each CTA stages N/2 rows = whole 8-row swizzle atoms, K loop
// deep enough to be operand-feed bound) and enabled (PPB_PAIR: 0 off, 1 default, 2 every eligible shape).
bool conv_pair_ok(const ConvParams& p);

}  // namespace ppb
