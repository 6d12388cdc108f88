// Weight-gradient kernel of the convolution for sm_100a: see conv_wgrad.h. Same skeleton as conv_igemm.cu (persistent
// CTAs, TMA producer warp, single-thread tcgen05.mma issuer, double-buffered TMEM accumulators, epilogue warps), but
// the reduction dimension is the pixel row, so both operands are MN-major [64 rows][64 channels] TMA boxes, and the
// epilogue accumulates its fp32 tile into dW with 128-bit atomics (split-K across CTAs).
// Replaces the filter-gradient ops TensorFlow runs for keras.layers.Conv2D during model.fit_generator
// (PyraPose/bin/train.py:381-390).
#include "conv_wgrad.h"
#include "launch_count.h"
#include "ptx.cuh"

#include <cstdlib>
#include <mutex>

namespace ppb {
namespace {

constexpr int kWgThreads = 256;
constexpr int kWgBoxBytes = 64 * 64 * 2;    // one [64 rows][64 channels] bf16 box
constexpr int kWgABytes = 2 * kWgBoxBytes;  // 128 output channels
constexpr int kWgMaxStages = 6;
constexpr int kWgTmemCols = 512;
constexpr int kWgAccCols = 256;
constexpr int kWgBarBytes = 256;
constexpr int kWgSmemLimit = 232448;

enum WgFault : uint32_t { kWgFaultEmpty = 11, kWgFaultFull = 12, kWgFaultTmemEmpty = 13, kWgFaultTmemFull = 14 };

struct Task {
  int m_tile, tap, n_tile, k_begin, k_end;
};

// Work distribution: tasks = K splits x output tiles, task t = split * tiles + tile (tile = (m, tap, n), n fastest), worker
// w of G (a CTA, or a CTA pair) runs tasks w, w + G, ... . Concurrently running workers therefore sit in the SAME few
// row ranges of g and x (all tiles of one split move in lockstep), so every 64-row block comes from HBM once and is an
// L2 hit for the other tiles. (A stream-K walk that gives every worker an equal contiguous share of tiles x k-iterations
// was measured first: perfectly balanced, but the workers' row positions spread over the whole matrix, nothing is
// shared in L2 and the 512 -> 512 head layers went from 0.37 to 0.45 ms.) The balance comes from the choice of the
// split count instead: wgrad_plan() picks it so that tasks fill whole waves of G.
struct TaskWalk {
  int t, t_step, t_end, tiles;
  __device__ __forceinline__ bool next(const WgradParams& p, bool pair, uint32_t rank, Task& k) {
    if (t >= t_end) return false;
    const int split = t / tiles;
    int tile = t - split * tiles;
    t += t_step;
    k.k_begin = static_cast<int>(static_cast<long long>(p.total_k_iters) * split / p.k_splits);
    k.k_end = static_cast<int>(static_cast<long long>(p.total_k_iters) * (split + 1) / p.k_splits);
    k.n_tile = tile % p.n_ntiles;
    tile /= p.n_ntiles;
    k.tap = tile % p.n_taps;
    k.m_tile = tile / p.n_taps;
    if (pair) k.m_tile = (k.m_tile << 1) + static_cast<int>(rank);
    return true;
  }
};

__device__ __forceinline__ TaskWalk make_walk(const WgradParams& p, bool pair) {
  TaskWalk t;
  t.tiles = (pair ? p.n_mtiles >> 1 : p.n_mtiles) * p.n_taps * p.n_ntiles;
  t.t = pair ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  t.t_step = pair ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  t.t_end = t.tiles * p.k_splits;
  return t;
}

template <bool kPair>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_g, const __grid_constant__ CUtensorMap tmap_x,
                  const __grid_constant__ WgradParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // by offset: pointers stay shared-space
  // CTA pair (see conv_igemm.cu): the pair owns two consecutive Cout tiles (M = 256); each CTA stages its own 128
  // gradient columns and HALF of the input-channel boxes, rank 0 issues the cta_group::2 MMAs.
  constexpr bool pair = kPair;
  const uint32_t rank = pair ? cluster_ctarank() : 0u;
  const int n_stages = p.n_stages;
  const int n_boxes_b = pair ? (p.block_n >> 7) : (p.block_n >> 6);
  const int stage_bytes = kWgABytes + n_boxes_b * kWgBoxBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + n_stages * stage_bytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kWgMaxStages;
  uint64_t* tmem_full_bar = bars + 2 * kWgMaxStages;
  uint64_t* tmem_empty_bar = bars + 2 * kWgMaxStages + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kWgMaxStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&tmap_g);
      tma_prefetch_desc(&tmap_x);
    }
  } else if (warp == 1) {
    if (elect_one()) {
      for (int i = 0; i < kWgMaxStages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tmem_full_bar[i], 1);
        mbar_init(&tmem_empty_bar[i], pair ? 256 : 128);
      }
      fence_mbar_init();
    }
  } else if (warp == 2) {
    if (pair) {
      tmem_alloc2(tmem_slot, kWgTmemCols);
      tmem_relinquish2();
    } else {
      tmem_alloc(tmem_slot, kWgTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (pair) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // pair: tiles enumerate (Cout-tile PAIR, tap, n tile); this CTA's Cout tile is 2 * pair + rank

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      TaskWalk walk = make_walk(p, pair);
      Task k;
      while (walk.next(p, pair, rank, k)) {
        // locate k_begin inside the segment list
        int seg = 0, local = k.k_begin;
        while (seg < p.n_segs - 1 && local >= p.seg[seg].k_iters) {
          local -= p.seg[seg].k_iters;
          ++seg;
        }
        for (int it = k.k_begin; it < k.k_end; ++it) {
          const int row = p.seg[seg].row_begin + local * 64;
          const int xrow = row + p.seg[seg].tap_off[k.tap];
          mbar_wait(&empty_bar[stage], phase ^ 1, kWgFaultEmpty);
          uint8_t* sa = smem + stage * stage_bytes;
          if (pair) {
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], static_cast<uint32_t>(2 * stage_bytes));
            tma_load_2d_pair(sa, &tmap_g, k.m_tile * 128, row, &full_bar[stage]);
            tma_load_2d_pair(sa + kWgBoxBytes, &tmap_g, k.m_tile * 128 + 64, row, &full_bar[stage]);
            const int x_col = k.n_tile * p.block_n + static_cast<int>(rank) * (p.block_n >> 1);
            for (int j = 0; j < n_boxes_b; ++j)
              tma_load_2d_pair(sa + kWgABytes + j * kWgBoxBytes, &tmap_x, x_col + j * 64, xrow, &full_bar[stage]);
          } else {
            mbar_arrive_expect_tx(&full_bar[stage], static_cast<uint32_t>(stage_bytes));
            tma_load_2d(sa, &tmap_g, k.m_tile * 128, row, &full_bar[stage]);
            tma_load_2d(sa + kWgBoxBytes, &tmap_g, k.m_tile * 128 + 64, row, &full_bar[stage]);
            for (int j = 0; j < n_boxes_b; ++j)
              tma_load_2d(sa + kWgABytes + j * kWgBoxBytes, &tmap_x, k.n_tile * p.block_n + j * 64, xrow,
                          &full_bar[stage]);
          }
          if (++stage == n_stages) {
            stage = 0;
            phase ^= 1;
          }
          if (++local == p.seg[seg].k_iters && seg < p.n_segs - 1) {
            local = 0;
            ++seg;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    if (rank == 0 && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint32_t idesc = umma_idesc_bf16_mn(pair ? 256 : 128, p.block_n);
      TaskWalk walk = make_walk(p, pair);
      Task k;
      while (walk.next(p, pair, rank, k)) {
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1, kWgFaultTmemEmpty);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kWgAccCols;
        for (int it = k.k_begin; it < k.k_end; ++it) {
          mbar_wait(&full_bar[stage], phase, kWgFaultFull);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * stage_bytes);
          const uint32_t b_addr = a_addr + kWgABytes;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {  // 16 rows (K) per MMA = two 8-row groups of 1024 B
            const uint64_t da = umma_desc_mn128(a_addr + kk * 2048, kWgBoxBytes);
            const uint64_t db = umma_desc_mn128(b_addr + kk * 2048, kWgBoxBytes);
            const uint32_t accum = (it > k.k_begin || kk > 0) ? 1u : 0u;
            if (pair) tc_mma_bf16_2cta(d_tmem, da, db, idesc, accum);
            else tc_mma_bf16(d_tmem, da, db, idesc, accum);
          }
          if (pair) tc_commit2(&empty_bar[stage]);
          else tc_commit(&empty_bar[stage]);
          if (++stage == n_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (pair) tc_commit2(&tmem_full_bar[acc]);
        else tc_commit(&tmem_full_bar[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------ epilogue: fp32 tile -> atomics into dW
    const int q = warp - 4;
    int acc = 0;
    uint32_t acc_phase = 0;
    TaskWalk walk = make_walk(p, pair);
    Task k;
    while (walk.next(p, pair, rank, k)) {
      mbar_wait(&tmem_full_bar[acc], acc_phase, kWgFaultTmemFull);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * kWgAccCols;
      const int co = k.m_tile * 128 + q * 32 + lane;
      float* dst = p.dw + static_cast<long long>(co) * p.dw_ld + static_cast<long long>(k.tap) * p.cin +
                   k.n_tile * p.block_n;
      for (int c0 = 0; c0 < p.block_n; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(taddr + c0, v);
        tmem_ld_wait();
        if (co < p.cout) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            atomicAdd(reinterpret_cast<float4*>(dst + c0 + j),
                      make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                  __uint_as_float(v[j + 3])));
          }
        }
      }
      tc_fence_before();
      if (pair) mbar_arrive_rank0(&tmem_empty_bar[acc]);
      else mbar_arrive(&tmem_empty_bar[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (pair) cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    if (pair) tmem_dealloc2(tmem_base, kWgTmemCols);
    else tmem_dealloc(tmem_base, kWgTmemCols);
  }
}

}  // namespace

int wgrad_encode_tmaps(CUtensorMap* tmap_g, CUtensorMap* tmap_x, const void* g, long g_rows, int g_cols, int g_ld,
                       const void* x, long x_rows, int x_cols, int x_ld) {
  int rc = conv_encode_box_tmap(tmap_g, g, g_rows, g_cols, g_ld, 64);
  if (rc != 0) return rc;
  return conv_encode_box_tmap(tmap_x, x, x_rows, x_cols, x_ld, 64);
}

// Tile shape and split count, by a small cost model over (pair, block_n, splits):
//   cycles ~ MMA cycles / (shape efficiency x wave efficiency) + flushed bytes / atomic throughput.
// Every task ends in a 128 x block_n fp32 flush into dW with 128-bit atomics, which the L2 retires at ~1 TB/s
// (measured r02: 25-38 MB of flushes <-> 35-40 us on the stage-4/5 and PFPN layers, whose MMAs need ~5 us), so small
// layers want narrow tiles and few splits while large ones want the 256 x 256 CTA-pair tile (92 % tensor-pipe activity
// in the head layers) and a split count whose tasks fill whole waves (tiles x splits a multiple of the workers: the old
// fixed rule ran 36 x 5 = 180 tasks on 74 pairs, 3 waves for 2.43 waves of work).
// PPB_WGRAD_BN / PPB_WGRAD_PAIR / PPB_WGRAD_SPLITS force a choice for A/B runs.
void wgrad_plan(WgradParams& p) {
  static const int pair_mode = [] {
    const char* e = getenv("PPB_PAIR");
    return e != nullptr ? atoi(e) : 1;
  }();
  static const int force_bn = [] {
    const char* e = getenv("PPB_WGRAD_BN");
    return e != nullptr ? atoi(e) : 0;
  }();
  static const int force_pair = [] {
    const char* e = getenv("PPB_WGRAD_PAIR");
    return e != nullptr ? atoi(e) : -1;
  }();
  static const int force_splits = [] {
    const char* e = getenv("PPB_WGRAD_SPLITS");
    return e != nullptr ? atoi(e) : 0;
  }();
  const int sms = conv_num_sms() > 0 ? conv_num_sms() : 148;
  p.n_mtiles = (p.cout + 127) / 128;
  // per-SM MMA cycles of the whole layer at full rate: a k-iteration (64 rows) of a 128 x 64 tile is 4 MMAs x 32 clk
  const double mma_clk = static_cast<double>(p.n_mtiles) * p.n_taps * (p.cin / 64.0) * p.total_k_iters * 128.0 / sms;
  const int max_splits = p.total_k_iters / 8 > 0 ? p.total_k_iters / 8 : 1;
  double best_cost = 0.0;
  int best_bn = 0, best_pair = 0, best_splits = 1;
  for (int pr = 1; pr >= 0; --pr)
    for (int bn = 256; bn >= 64; bn >>= 1) {
      if (p.cin % bn != 0) continue;
      if (pr && (pair_mode == 0 || p.cout % 256 != 0 || bn < 128)) continue;
      if (force_bn != 0 && bn != force_bn && p.cin % force_bn == 0) continue;
      if (force_pair >= 0 && pr != force_pair && !(force_pair == 1 && (p.cout % 256 != 0 || bn < 128))) continue;
      // relative MMA efficiency: narrow tiles are bound by the shared-memory reads of the A operand, single-CTA wide
      // tiles by the L2 -> SM operand feed
      const double shape_eff = bn == 64 ? 0.67 : (bn == 128 ? (pr ? 0.95 : 0.9) : (pr ? 1.0 : 0.87));
      const int tiles = (pr ? p.n_mtiles / 2 : p.n_mtiles) * p.n_taps * (p.cin / bn);
      const int workers = pr ? sms / 2 : sms;
      const int s_hi = (4 * workers + tiles - 1) / tiles < max_splits ? (4 * workers + tiles - 1) / tiles : max_splits;
      for (int sp = 1; sp <= s_hi; ++sp) {
        if (force_splits != 0 && sp != (force_splits < s_hi ? force_splits : s_hi)) continue;
        const int tasks = tiles * sp;
        const int waves = (tasks + workers - 1) / workers;
        const double wave_eff = static_cast<double>(tasks) / (static_cast<double>(waves) * workers);
        const double flush_bytes = static_cast<double>(tasks) * (pr ? 2 : 1) * 128.0 * bn * 4.0;
        // the flushes of all but the last wave overlap the next task's MMAs (double-buffered accumulators)
        const double exposed = flush_bytes / waves;
        const double cost = mma_clk / (shape_eff * wave_eff) + exposed / 550.0 + 0.1 * (flush_bytes - exposed) / 550.0;
        if (best_bn == 0 || cost < best_cost) {
          best_cost = cost;
          best_bn = bn;
          best_pair = pr;
          best_splits = sp;
        }
      }
    }
  p.block_n = best_bn;
  p.pair = best_pair;
  p.k_splits = best_splits;
  p.n_ntiles = p.cin / p.block_n;
  const int stage_bytes = kWgABytes + (p.pair ? p.block_n >> 7 : p.block_n >> 6) * kWgBoxBytes;
  int stages = (kWgSmemLimit - 1024 - kWgBarBytes) / stage_bytes;
  if (stages > kWgMaxStages) stages = kWgMaxStages;
  p.n_stages = stages;
}

cudaError_t wgrad_launch(const CUtensorMap& tmap_g, const CUtensorMap& tmap_x, const WgradParams& p,
                         cudaStream_t stream) {
  static PerDeviceOnce once;  // per device, see launch_count.h
  if (once.first_on_current_device()) {
    cudaError_t attr_err = cudaFuncSetAttribute(conv_wgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmemLimit);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(conv_wgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmemLimit);
    if (attr_err != cudaSuccess) return attr_err;
  }
  const int total = (p.pair ? p.n_mtiles / 2 : p.n_mtiles) * p.n_taps * p.n_ntiles * p.k_splits;
  if (total <= 0 || p.total_k_iters <= 0) return cudaSuccess;
  const int stage_bytes = kWgABytes + (p.pair ? p.block_n >> 7 : p.block_n >> 6) * kWgBoxBytes;
  const int smem_bytes = p.n_stages * stage_bytes + kWgBarBytes + 1024;
  int ctas = conv_num_sms();
  if (!p.pair) {
    if (ctas > total) ctas = total;
    conv_wgrad_kernel<false><<<ctas, kWgThreads, smem_bytes, stream>>>(tmap_g, tmap_x, p);
    return cudaGetLastError();
  }
  ctas = (ctas / 2 < total ? ctas / 2 : total) * 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(kWgThreads);
  cfg.dynamicSmemBytes = static_cast<size_t>(smem_bytes);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, conv_wgrad_kernel<true>, tmap_g, tmap_x, p);
}

}  // namespace ppb
