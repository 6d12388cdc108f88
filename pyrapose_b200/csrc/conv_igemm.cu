// Implicit-GEMM convolution for sm_100a: TMA -> shared memory (128B swizzle) -> tcgen05.mma with
// fp32 accumulators in TMEM -> fused epilogue. Replaces the Conv2D / BatchNorm / ReLU / Add / Reshape /
// sigmoid / RegressBoxes3D nodes that the reference builds in PyraPose/models/retinanet.py:9-131,
// 180-214 and (via keras-resnet) PyraPose/models/resnet.py:87, and that TensorFlow executes.
//
// One persistent CTA per SM, 8 warps:
//   warp 0   : TMA producer (one elected lane)      global -> smem ring, 4 stages of {A 128x64, B Nx64}
//   warp 1   : MMA issuer (one elected lane)        4 x tcgen05.mma (K=16) per stage, commit -> empty[stage]
//   warp 2   : TMEM allocator (512 columns = two 256-column accumulators, double-buffered)
//   warps 4-7: epilogue, one TMEM lane (= output row) per thread
// M tile = 128 consecutive rows of the padded-flat activation matrix (see conv_igemm.h); the K loop
// runs over filter taps x 64-channel blocks, each tap being a row-shifted 2-D TMA box.
#include "conv_igemm.h"
#include "ptx.cuh"

#include <mutex>

namespace ppb {

namespace {

constexpr int kMaxStages = 4;
constexpr int kABytes = kConvBlockM * kConvBlockK * 2;        // 16 KiB
constexpr int kBBytesMax = kConvMaxBlockN * kConvBlockK * 2;  // 32 KiB
constexpr int kStageBytes = kABytes + kBBytesMax;
constexpr int kTmemCols = 512;
constexpr int kAccCols = 256;
constexpr int kThreads = 256;
// Shared memory map (after 1024-B alignment): direct epilogue = 4 pipeline stages; staged epilogue = 3 stages +
// a 64 KiB output/residual tile R (four 128x64 bf16 chunks in the 128B-swizzled TMA layout). Barriers follow.
constexpr int kRChunkBytes = kConvBlockM * 64 * 2;  // 16 KiB: 128 rows x 64 bf16
constexpr int kRBytes = 4 * kRChunkBytes;
constexpr int kBarOffset = 3 * kStageBytes + kRBytes;  // 208 KiB >= 4 * kStageBytes
constexpr int kSmemBytes = kBarOffset + 1024 /*alignment slack*/ + 256 /*barriers*/;
static_assert(kBarOffset >= kMaxStages * kStageBytes, "barriers must not overlap the 4-stage ring");

enum Fault : uint32_t {
  kFaultEmpty = 1, kFaultFull = 2, kFaultTmemEmpty = 3, kFaultTmemFull = 4, kFaultResEmpty = 5, kFaultResFull = 6
};

struct TileCoord {
  int seg;
  int row0;
  int n0;
};

__device__ __forceinline__ TileCoord tile_coord(const ConvParams& p, int tile) {
  const int m_tile = tile / p.n_ntiles;
  const int n_tile = tile - m_tile * p.n_ntiles;
  int s = 0;
  if (p.n_segs > 1 && m_tile >= p.seg[1].tile_begin) s = 1;
  if (p.n_segs > 2 && m_tile >= p.seg[2].tile_begin) s = 2;
  TileCoord t;
  t.seg = s;
  t.row0 = p.seg[s].row_begin + (m_tile - p.seg[s].tile_begin) * kConvBlockM;
  t.n0 = n_tile * p.block_n;
  return t;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

// Row bookkeeping of one epilogue thread for one tile.
struct RowInfo {
  bool in_buf;    // the row exists in the output matrix
  bool interior;  // the row is a real output pixel (not border)
  int b, y, x;    // image index and unpadded pixel coordinates (valid when interior)
};

__device__ __forceinline__ RowInfo row_info(const ConvSeg& sg, int m) {
  RowInfo r;
  const int local = m - sg.row_begin;
  r.in_buf = local < sg.rows;
  r.interior = r.in_buf;
  r.b = 0;
  r.y = 0;
  r.x = 0;
  const int plane = sg.hp * sg.wp;
  r.b = local / plane;
  const int rem = local - r.b * plane;
  const int hy = rem / sg.wp;
  const int wx = rem - hy * sg.wp;
  r.y = hy - sg.pad_lo;
  r.x = wx - sg.pad_lo;
  r.interior = r.in_buf && hy >= sg.pad_lo && hy < sg.hp - sg.pad_hi && wx >= sg.pad_lo && wx < sg.wp - sg.pad_hi;
  return r;
}

// bf16 epilogue for NC (16 or 32) accumulator columns of one row.
template <int NC>
__device__ __forceinline__ void epi_bf16(const ConvParams& p, const RowInfo& ri, int m, int n0, uint32_t (&v)[NC]) {
  if (!ri.in_buf || n0 >= p.n_valid) return;
  __nv_bfloat16* dst = p.out + static_cast<size_t>(m) * p.out_ld + n0;
  if (!ri.interior) {
    const uint4 z = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int i = 0; i < NC / 8; ++i) reinterpret_cast<uint4*>(dst)[i] = z;
    return;
  }
  float f[NC];
#pragma unroll
  for (int j = 0; j < NC; ++j) f[j] = __uint_as_float(v[j]);
  if (p.bias != nullptr) {
#pragma unroll
    for (int j = 0; j < NC; j += 4) {
      const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + j));
      f[j] += b4.x;
      f[j + 1] += b4.y;
      f[j + 2] += b4.z;
      f[j + 3] += b4.w;
    }
  }
  if (p.residual != nullptr) {
    const uint4* rp = reinterpret_cast<const uint4*>(p.residual + static_cast<size_t>(m) * p.res_ld + n0);
#pragma unroll
    for (int i = 0; i < NC / 8; ++i) {
      const uint4 r = __ldg(rp + i);
      f[i * 8 + 0] += bf16_lo(r.x);
      f[i * 8 + 1] += bf16_hi(r.x);
      f[i * 8 + 2] += bf16_lo(r.y);
      f[i * 8 + 3] += bf16_hi(r.y);
      f[i * 8 + 4] += bf16_lo(r.z);
      f[i * 8 + 5] += bf16_hi(r.z);
      f[i * 8 + 6] += bf16_lo(r.w);
      f[i * 8 + 7] += bf16_hi(r.w);
    }
  }
  if (p.relu) {
#pragma unroll
    for (int j = 0; j < NC; ++j) f[j] = fmaxf(f[j], 0.0f);
  }
#pragma unroll
  for (int i = 0; i < NC / 8; ++i) {
    uint4 o;
    o.x = pack_bf16x2(f[i * 8 + 0], f[i * 8 + 1]);
    o.y = pack_bf16x2(f[i * 8 + 2], f[i * 8 + 3]);
    o.z = pack_bf16x2(f[i * 8 + 4], f[i * 8 + 5]);
    o.w = pack_bf16x2(f[i * 8 + 6], f[i * 8 + 7]);
    reinterpret_cast<uint4*>(dst)[i] = o;
  }
}

// fp32 head epilogue for 16 accumulator columns (= the 16 regression values of one anchor when
// decoding). Output element (b, loc, n) lives at ((b*locs + loc)*vals + n): Keras' Reshape((-1, 16)) /
// Reshape((-1, C)) plus Concatenate(axis=1) over pyramid levels (retinanet.py:51, 129, 224-225).
__device__ __forceinline__ void epi_head16(const ConvParams& p, const ConvSeg& sg, int seg_idx, const RowInfo& ri,
                                           int n0, const uint32_t* v) {
  if (!ri.interior || n0 >= p.n_valid) return;
  const int w = sg.wp - sg.pad_lo - sg.pad_hi;
  const size_t loc = static_cast<size_t>(ri.b) * p.locs_per_image + sg.loc_base + ri.y * w + ri.x;
  float* dst = p.out_f32 + loc * p.vals_per_loc + n0;
  float f[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const int n = n0 + j;
    const float bias = (p.bias != nullptr && n < p.n_valid) ? __ldg(p.bias + n) : 0.0f;
    f[j] = __uint_as_float(v[j]) + bias;
  }
  if (p.epi == kEpiHeadSigmoid) {
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = 1.0f / (1.0f + expf(-f[j]));
  } else if (p.epi == kEpiHeadDecode) {
    // box3D_transform_inv (PyraPose/backend/common.py:25-56) against the Anchors layer's boxes
    // (PyraPose/layers/_misc.py:52-69, backend/common.py:93-116): fp32 base anchor + fp32 shift.
    const int a = n0 >> 4;
    const float st = p.stride[seg_idx];
    const float sx = __fmul_rn(static_cast<float>(ri.x) + 0.5f, st);
    const float sy = __fmul_rn(static_cast<float>(ri.y) + 0.5f, st);
    const float x1 = __fadd_rn(p.anchor[seg_idx][a][0], sx);
    const float y1 = __fadd_rn(p.anchor[seg_idx][a][1], sy);
    const float x2 = __fadd_rn(p.anchor[seg_idx][a][2], sx);
    const float y2 = __fadd_rn(p.anchor[seg_idx][a][3], sy);
    const float bw = __fsub_rn(x2, x1);
    const float bh = __fsub_rn(y2, y1);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float base = (j & 1) ? ((j & 2) ? y2 : y1) : ((j & 2) ? x2 : x1);
      const float size = (j & 1) ? bh : bw;
      const float d = __fadd_rn(__fmul_rn(f[j], p.dec_std[j]), p.dec_mean[j]);
      f[j] = __fadd_rn(base, __fmul_rn(d, size));
    }
    if (p.out_hull != nullptr) {
      float xmin = f[0], xmax = f[0], ymin = f[1], ymax = f[1];
#pragma unroll
      for (int j = 2; j < 16; j += 2) {
        xmin = fminf(xmin, f[j]);
        xmax = fmaxf(xmax, f[j]);
        ymin = fminf(ymin, f[j + 1]);
        ymax = fmaxf(ymax, f[j + 1]);
      }
      const size_t anchor_idx = loc * (p.vals_per_loc >> 4) + a;
      reinterpret_cast<float4*>(p.out_hull)[anchor_idx] = make_float4(xmin, ymin, xmax, ymax);
    }
  }
  if ((p.vals_per_loc & 3) == 0 && n0 + 16 <= p.n_valid) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      reinterpret_cast<float4*>(dst)[i] = make_float4(f[i * 4], f[i * 4 + 1], f[i * 4 + 2], f[i * 4 + 3]);
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (n0 + j < p.n_valid) dst[j] = f[j];
  }
}

__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
                  const __grid_constant__ ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kBarOffset);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kMaxStages;
  uint64_t* tmem_full_bar = bars + 2 * kMaxStages;
  uint64_t* tmem_empty_bar = bars + 2 * kMaxStages + 2;
  uint64_t* res_full_bar = bars + 2 * kMaxStages + 4;
  uint64_t* res_empty_bar = bars + 2 * kMaxStages + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kMaxStages + 6);
  const bool staged = p.staged != 0;
  const int n_stages = staged ? 3 : kMaxStages;
  uint8_t* rbuf = smem + 3 * kStageBytes;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&tmap_a);
      tma_prefetch_desc(&tmap_b);
      if (staged) {
        tma_prefetch_desc(&tmap_out);
        if (p.residual != nullptr) tma_prefetch_desc(&tmap_res);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      for (int i = 0; i < kMaxStages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tmem_full_bar[i], 1);
        mbar_init(&tmem_empty_bar[i], 128);
      }
      mbar_init(res_full_bar, 1);
      mbar_init(res_empty_bar, 1);
      fence_mbar_init();
    }
  } else if (warp == 2) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_tiles = p.n_mtiles * p.n_ntiles;
  const int k_iters = p.n_taps * p.cin_blocks;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx_bytes = kABytes + p.block_n * (kConvBlockK * 2);
      uint32_t res_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const TileCoord tc = tile_coord(p, tile);
        const ConvSeg& sg = p.seg[tc.seg];
        for (int tap = 0; tap < p.n_taps; ++tap) {
          const int a_row = tc.row0 + sg.tap_off[tap];
          for (int kb = 0; kb < p.cin_blocks; ++kb) {
            mbar_wait(&empty_bar[stage], phase ^ 1, kFaultEmpty);
            mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
            uint8_t* sa = smem + stage * kStageBytes;
            tma_load_2d(sa, &tmap_a, kb * kConvBlockK, a_row, &full_bar[stage]);
            tma_load_2d(sa + kABytes, &tmap_b, (tap * p.cin_blocks + kb) * kConvBlockK, tc.n0, &full_bar[stage]);
            if (++stage == n_stages) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        if (staged) {
          // The R tile becomes free once the previous tile's TMA store has read it; then either fetch this
          // tile's residual into it or just hand it to the epilogue warps.
          mbar_wait(res_empty_bar, res_phase ^ 1, kFaultResEmpty);
          if (p.residual != nullptr) {
            mbar_arrive_expect_tx(res_full_bar, static_cast<uint32_t>(kConvBlockM * 2) * p.block_n);
            for (int c = 0; c < p.block_n; c += 64)
              tma_load_2d(rbuf + (c >> 6) * kRChunkBytes, &tmap_res, tc.n0 + c, tc.row0, res_full_bar);
          } else {
            mbar_arrive(res_full_bar);
          }
          res_phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint32_t idesc = umma_idesc_bf16(kConvBlockM, p.block_n);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1, kFaultTmemEmpty);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kAccCols;
        for (int it = 0; it < k_iters; ++it) {
          mbar_wait(&full_bar[stage], phase, kFaultFull);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * kStageBytes);
          const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
          for (int k = 0; k < kConvBlockK / 16; ++k) {
            tc_mma_bf16(d_tmem, umma_desc_k128(a_addr + k * 32), umma_desc_k128(b_addr + k * 32), idesc,
                        (it | k) != 0 ? 1u : 0u);
          }
          tc_commit(&empty_bar[stage]);
          if (++stage == n_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        tc_commit(&tmem_full_bar[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue
    const int q = warp - 4;  // TMEM lane quadrant this warp may access
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t res_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const TileCoord tc = tile_coord(p, tile);
      const ConvSeg& sg = p.seg[tc.seg];
      const int m = tc.row0 + q * 32 + lane;
      const RowInfo ri = row_info(sg, m);
      mbar_wait(&tmem_full_bar[acc], acc_phase, kFaultTmemFull);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * kAccCols;
      if (staged) {
        // ---- staged bf16 epilogue: accumulators (+bias, +residual, ReLU) -> R tile in smem -> one TMA store
        mbar_wait(res_full_bar, res_phase, kFaultResFull);
        res_phase ^= 1;
        const int row = q * 32 + lane;
        const bool has_res = p.residual != nullptr;
        for (int c0 = 0; c0 < p.block_n; c0 += 64) {
          uint32_t v[64];
          tmem_ld_32x32(taddr + c0, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          tmem_ld_32x32(taddr + c0 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
          tmem_ld_wait();
          uint8_t* rowp = rbuf + (c0 >> 6) * kRChunkBytes + row * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4* sp = reinterpret_cast<uint4*>(rowp + ((j ^ (row & 7)) << 4));  // 128B swizzle: 16-B piece j of row r
            uint4 o = make_uint4(0u, 0u, 0u, 0u);
            if (ri.interior) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[j * 8 + e]);
              if (p.bias != nullptr) {
                const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + tc.n0 + c0 + j * 8));
                const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + tc.n0 + c0 + j * 8 + 4));
                f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
                f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
              }
              if (has_res) {
                const uint4 r = *sp;
                f[0] += bf16_lo(r.x); f[1] += bf16_hi(r.x); f[2] += bf16_lo(r.y); f[3] += bf16_hi(r.y);
                f[4] += bf16_lo(r.z); f[5] += bf16_hi(r.z); f[6] += bf16_lo(r.w); f[7] += bf16_hi(r.w);
              }
              if (p.relu) {
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.0f);
              }
              o.x = pack_bf16x2(f[0], f[1]);
              o.y = pack_bf16x2(f[2], f[3]);
              o.z = pack_bf16x2(f[4], f[5]);
              o.w = pack_bf16x2(f[6], f[7]);
            }
            *sp = o;  // border rows and rows past the segment end are stored as zeros
          }
        }
        tc_fence_before();
        mbar_arrive(&tmem_empty_bar[acc]);  // the accumulator is drained: the MMA warp may reuse it
        fence_proxy_async_smem();           // make the generic-proxy writes to R visible to the TMA store
        named_bar_sync(1, 128);
        if (warp == 4 && lane == 0) {  // always the same thread: bulk-async groups are per thread
          for (int c = 0; c < p.block_n; c += 64)
            tma_store_2d(&tmap_out, tc.n0 + c, tc.row0, rbuf + (c >> 6) * kRChunkBytes);
          tma_store_commit();
          tma_store_wait_read();
          mbar_arrive(res_empty_bar);
        }
      } else {
      for (int c0 = 0; c0 < p.block_n; c0 += 32) {
        if (p.block_n - c0 >= 32) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c0, v);
          tmem_ld_wait();
          if (p.epi == kEpiBf16) {
            epi_bf16<32>(p, ri, m, tc.n0 + c0, v);
          } else {
            epi_head16(p, sg, tc.seg, ri, tc.n0 + c0, v);
            epi_head16(p, sg, tc.seg, ri, tc.n0 + c0 + 16, v + 16);
          }
        } else {
          uint32_t v[16];
          tmem_ld_32x16(taddr + c0, v);
          tmem_ld_wait();
          if (p.epi == kEpiBf16) {
            epi_bf16<16>(p, ri, m, tc.n0 + c0, v);
          } else {
            epi_head16(p, sg, tc.seg, ri, tc.n0 + c0, v);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty_bar[acc]);
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  if (staged && warp == 4 && lane == 0) tma_store_wait_all();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  // Resolved through the runtime so the library has no link-time dependency on libcuda (it must
  // load on machines without a driver).
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

int encode_2d(CUtensorMap* map, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems, uint32_t box_rows) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) return -1;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kConvBlockK), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -static_cast<int>(r) - 1000;
}

}  // namespace

int conv_encode_rows_tmap(CUtensorMap* tmap, const void* base, long rows, int cols, int ld) {
  return encode_2d(tmap, base, static_cast<uint64_t>(rows), cols, ld, kConvBlockM);
}

int conv_encode_tmaps(CUtensorMap* tmap_a, CUtensorMap* tmap_b, const void* a, long a_rows, int a_cols, int a_ld,
                      const void* w, int w_rows, int w_cols, int block_n) {
  int rc = encode_2d(tmap_a, a, static_cast<uint64_t>(a_rows), a_cols, a_ld, kConvBlockM);
  if (rc != 0) return rc;
  return encode_2d(tmap_b, w, w_rows, w_cols, w_cols, block_n);
}

int conv_num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  return sms;
}

cudaError_t conv_launch(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_out,
                        const CUtensorMap& tmap_res, const ConvParams& p, int max_ctas, cudaStream_t stream) {
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(conv_igemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  });
  if (attr_err != cudaSuccess) return attr_err;
  const int total = p.n_mtiles * p.n_ntiles;
  if (total <= 0) return cudaSuccess;
  int ctas = max_ctas > 0 ? max_ctas : conv_num_sms();
  if (ctas > total) ctas = total;
  conv_igemm_kernel<<<ctas, kThreads, kSmemBytes, stream>>>(tmap_a, tmap_b, tmap_out, tmap_res, p);
  return cudaGetLastError();
}

void conv_read_fault(unsigned int out[4]) {
  unsigned int zero[4] = {0, 0, 0, 0};
  if (cudaMemcpyFromSymbol(out, g_kernel_fault, sizeof(zero)) != cudaSuccess) {
    out[0] = out[1] = out[2] = out[3] = 0xFFFFFFFFu;
    return;
  }
  cudaMemcpyToSymbol(g_kernel_fault, zero, sizeof(zero));
}

}  // namespace ppb
