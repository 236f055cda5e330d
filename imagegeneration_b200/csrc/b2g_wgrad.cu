// b2g_wgrad.cu — weight-gradient implicit GEMM on tcgen05 with MN-major (pixel-reduction) operands.
//
//   dW[tap][a][b] = sum over grid pixels  S[pixel shifted by tap, a] * U[pixel, b]        (fp32, atomically
//   accumulated into the caller's zero-initialised gradient buffer; split-K over pixel ranges)
//
//   S = the tensor the taps shift over (conv: the layer input x; convT: the output gradient), read through the
//       same TMA tap plan as the forward conv (K3S1: 9 shifts; K4S2: 16 taps over 4 parity sub-grids);
//   U = the tensor living on the reduction grid itself (conv: output gradient; convT: the layer input).
// Both tiles arrive as [64 pixels][64 channels] 128B-swizzled TMA boxes, i.e. the reduction index (pixel) is the
// slow one: they are fed to tcgen05.mma as MN-major A (M = channels of one operand) and MN-major B.
// Whichever operand has a multiple of 128 channels becomes the M side; the epilogue transposes if needed.
//
// Replaces Conv2DBackpropFilter (tape.gradient at sndcgan/SNDCGAN.py:248,:256,:266) for the Conv2D layers
// :80-118 and the Conv2DTranspose layers :36-53.
#include <stdlib.h>

#include "b2g_conv.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) WgradKernelParams {
  CUtensorMap smap[4];  // shifted operand (1 or 4 maps)
  CUtensorMap umap;     // unshifted operand
  ConvPlan plan;        // tap plan + 64-pixel tiling of the reduction grid
  int m_is_shifted;     // 1: M side = S (a), N side = U (b);  0: M side = U (b), N side = S (a) -> transposed store
  int Cm, Cn;           // channels on the M / N side
  int m_tiles, n_tiles, splits, ptiles;
  int Ca, Cb;           // output is [taps][Ca][Cb]
  int ca_valid;         // rows a < ca_valid are stored; row a == ca_valid goes to dbias (edge layers), else == Ca
  float* out;
  float* dbias;
};

// kTf32: fp32 operands (32 channels per 128-byte line), tcgen05.mma kind::tf32 with K = 8 pixels per instruction.
template <int BLOCK_N, bool kTf32 = false>
struct WgradCfg {
  static constexpr int kTileK = 64;                 // pixels per k-step
  static constexpr int kBoxCh = kTf32 ? 32 : 64;    // channels per 128-byte line of a box
  static constexpr int kBoxBytes = kTileK * 128;    // one [64 pix][kBoxCh ch] box
  static constexpr int kMBoxes = 128 / kBoxCh;      // 128 channels on the M side
  static constexpr int kNBoxes = BLOCK_N / kBoxCh;
  static constexpr int kMBytes = kMBoxes * kBoxBytes;
  static constexpr int kNBytes = kNBoxes * kBoxBytes;
  static constexpr int kStageBytes = kMBytes + kNBytes;
  static constexpr int kStages = kTf32 ? (BLOCK_N == 128 ? 3 : 4) : (BLOCK_N == 256 ? 4 : (BLOCK_N == 128 ? 6 : 8));
  static constexpr int kTmemCols = 2 * BLOCK_N;
  static constexpr int kSmemBytes = kStages * kStageBytes + 256 + 1024;
};

constexpr int kWgradThreads = 192;

struct WgradItem {
  int tap, m_tile, n_tile, p_begin, p_end;
};
__device__ __forceinline__ WgradItem decode_item(const WgradKernelParams& p, int item) {
  WgradItem w;
  const int taps = p.plan.taps;
  w.tap = item % taps;
  int r = item / taps;
  w.m_tile = r % p.m_tiles;
  r /= p.m_tiles;
  w.n_tile = r % p.n_tiles;
  const int split = r / p.n_tiles;
  w.p_begin = (int)((long)split * p.ptiles / p.splits);
  w.p_end = (int)((long)(split + 1) * p.ptiles / p.splits);
  return w;
}

template <int BLOCK_N, bool kTf32 = false>
__global__ void __launch_bounds__(kWgradThreads, 1) wgrad_kernel(const __grid_constant__ WgradKernelParams p) {
  using Cfg = WgradCfg<BLOCK_N, kTf32>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const ConvPlan& pl = p.plan;
  const int total_items = pl.taps * p.m_tiles * p.n_tiles * p.splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4);
    }
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < pl.n_amaps; ++i) tma_prefetch_desc(&p.smap[i]);
    tma_prefetch_desc(&p.umap);
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    {  // whole warp converged on uniform values; one elected lane issues the TMA loads
      const uint32_t leader = elect_one_sync();
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        const WgradItem wi = decode_item(p, item);
        const CUtensorMap* sm = &p.smap[pl.tap_map[wi.tap]];
        const int dy = pl.tap_dy[wi.tap], dx = pl.tap_dx[wi.tap];
        const int m_c0 = wi.m_tile * 128, n_c0 = wi.n_tile * BLOCK_N;
        for (int pt = wi.p_begin; pt < wi.p_end; ++pt) {
          const int tw = pt % pl.tiles_w;
          const int r2 = pt / pl.tiles_w;
          const int th = r2 % pl.tiles_h;
          const int tn = r2 / pl.tiles_h;
          const int w0 = tw * pl.bw, h0 = th * pl.bh, n0 = tn * pl.bn;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* m_dst = smem + stage * Cfg::kStageBytes;
          uint8_t* n_dst = m_dst + Cfg::kMBytes;
          if (leader) {
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          if (p.m_is_shifted) {
#pragma unroll
            for (int g = 0; g < Cfg::kMBoxes; ++g)
              tma_load_4d(m_dst + g * Cfg::kBoxBytes, sm, &full_bar[stage], m_c0 + g * Cfg::kBoxCh, w0 + dx, h0 + dy, n0);
#pragma unroll
            for (int g = 0; g < Cfg::kNBoxes; ++g)
              tma_load_4d(n_dst + g * Cfg::kBoxBytes, &p.umap, &full_bar[stage], n_c0 + g * Cfg::kBoxCh, w0, h0, n0);
          } else {
#pragma unroll
            for (int g = 0; g < Cfg::kMBoxes; ++g)
              tma_load_4d(m_dst + g * Cfg::kBoxBytes, &p.umap, &full_bar[stage], m_c0 + g * Cfg::kBoxCh, w0, h0, n0);
#pragma unroll
            for (int g = 0; g < Cfg::kNBoxes; ++g)
              tma_load_4d(n_dst + g * Cfg::kBoxBytes, sm, &full_bar[stage], n_c0 + g * Cfg::kBoxCh, w0 + dx, h0 + dy, n0);
          }
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // whole warp converged on uniform values; one elected lane issues (see elect_one_sync in b2g_ptx.cuh)
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*bf16 = 1, tf32 = 2*/ kTf32 ? 2 : 1, /*A MN-major*/ 1, /*B MN-major*/ 1, 128, BLOCK_N);
    const uint32_t smem_base = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++it) {
      const WgradItem wi = decode_item(p, item);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(&tempty_bar[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BLOCK_N;
      const int k_steps = wi.p_end - wi.p_begin;
      for (int ks = 0; ks < k_steps; ++ks) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t m_addr = smem_base + stage * Cfg::kStageBytes;
        const uint64_t mdesc = kTf32 ? umma_smem_desc_sw128_32b(m_addr, Cfg::kBoxBytes, 512)
                                     : umma_smem_desc(m_addr, Cfg::kBoxBytes, 1024);
        const uint64_t ndesc = kTf32 ? umma_smem_desc_sw128_32b(m_addr + Cfg::kMBytes, Cfg::kBoxBytes, 512)
                                     : umma_smem_desc(m_addr + Cfg::kMBytes, Cfg::kBoxBytes, 1024);
        if (leader) {
          if constexpr (kTf32) {
#pragma unroll
            for (int k = 0; k < 8; ++k)  // 8 pixels per MMA = two 4-line swizzle groups = 1024 bytes (>> 4 = 64)
              umma_tf32(d_tmem, mdesc + 64 * k, ndesc + 64 * k, idesc, (ks | k) != 0);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)  // 16 pixels per MMA = two 8-line swizzle groups = 2048 bytes (>> 4 = 128)
              umma_f16(d_tmem, mdesc + 128 * k, ndesc + 128 * k, idesc, (ks | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
      if (leader) umma_commit(&tfull_bar[as]);
      __syncwarp();
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;  // M index inside the tile
    int it = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++it) {
      const WgradItem wi = decode_item(p, item);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const bool nonempty = wi.p_end > wi.p_begin;
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const int m = wi.m_tile * 128 + row;
      float* obase = p.out + (size_t)wi.tap * p.ca_valid * p.Cb;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + c * 32, r);
        tmem_ld_wait();
        if (nonempty) {
          const int n = wi.n_tile * BLOCK_N + c * 32;
          if (p.m_is_shifted) {  // out[a = m][b = n..n+31]: contiguous per thread
            float* dst = obase + (size_t)m * p.Cb + n;
            bool store = true;
            if (m >= p.ca_valid) {  // padded rows: nothing, except the all-ones row = bias gradient
              store = (m == p.ca_valid) && (p.dbias != nullptr);
              dst = p.dbias + n;
            }
            if (store) {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j),
                             "f"(__uint_as_float(r[j])), "f"(__uint_as_float(r[j + 1])),
                             "f"(__uint_as_float(r[j + 2])), "f"(__uint_as_float(r[j + 3]))
                             : "memory");
              }
            }
          } else {  // out[a = n + j][b = m]: coalesced across the warp's lanes
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (n + j < p.ca_valid) atomicAdd(obase + (size_t)(n + j) * p.Cb + m, __uint_as_float(r[j]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int BLOCK_N, bool kTf32 = false>
static int launch_wgrad(const WgradKernelParams& kp, int total_items, cudaStream_t stream) {
  using Cfg = WgradCfg<BLOCK_N, kTf32>;
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(wgrad_kernel<BLOCK_N, kTf32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::kSmemBytes));
    attr_set = true;
  }
  int grid = total_items < num_sms() ? total_items : num_sms();
  wgrad_kernel<BLOCK_N, kTf32><<<grid, kWgradThreads, Cfg::kSmemBytes, stream>>>(kp);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}


// ------------------------------------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2) for the layers with >= 256 channels on both sides.  Round 2's ncu of wgrad_kernel<256>:
// tensor pipe 67 %, 1.8 GB per launch from L2 = 15 TB/s — the one-CTA tile (M = 128, N = 256) needs 48 KB of operands
// per 128 x 256 x 64 MMA group, 96 B/clk/SM, more than the L2 -> SM path sustains.  A pair of CTAs shares one
// M = 256 x N = 256 tile: each loads its own 128 M-side channels and HALF of the N-side channels (32 KB per k-step
// instead of 48), the leader issues one tcgen05.mma.cta_group::2 that reads both halves, each CTA's TMEM holds its own
// 128 accumulator rows and its four epilogue warps add them into dW (same vector reds as the one-CTA kernel).
// Same protocol as conv_pair_kernel (b2g_conv_pair.cu): the leader's full barrier counts the TMA bytes of both CTAs,
// tcgen05.commit is multicast to both, accumulator release = remote arrive on the leader's barrier.
// ------------------------------------------------------------------------------------------------------------
struct WgradPairCfg {
  static constexpr int kBlockN = 256;
  static constexpr int kBoxBytes = 64 * 128;            // one [64 pixels][64 channels] box
  static constexpr int kMBytes = 2 * kBoxBytes;         // this CTA's 128 M-side channels
  static constexpr int kNBytes = 2 * kBoxBytes;         // this CTA's half (128) of the 256 N-side channels
  static constexpr int kStageBytes = kMBytes + kNBytes;  // 32 KB
  static constexpr int kStages = 6;
  static constexpr int kTmemCols = 2 * kBlockN;
  static constexpr int kSmemBytes = kStages * kStageBytes + 256 + 1024;
};

__global__ void __launch_bounds__(kWgradThreads, 1) wgrad_pair_kernel(const __grid_constant__ WgradKernelParams p) {
  using Cfg = WgradPairCfg;
  constexpr int BLOCK_N = Cfg::kBlockN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = MMA leader
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const ConvPlan& pl = p.plan;
  const int total_items = pl.taps * p.m_tiles * p.n_tiles * p.splits;  // m_tiles counts 256-channel PAIRS of M tiles here

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);   // leader's producer: one arrive.expect_tx covering the bytes of both CTAs
      mbar_init(&empty_bar[s], 1);  // multicast tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);   // multicast tcgen05.commit
      mbar_init(&tempty_bar[a], 8);  // four epilogue warps of BOTH CTAs (the leader's copy is used)
    }
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < pl.n_amaps; ++i) tma_prefetch_desc(&p.smap[i]);
    tma_prefetch_desc(&p.umap);
  }
  if (warp == 1) tmem_alloc_pair(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's barriers are initialised and its TMEM is allocated before any remote signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer (both CTAs)
    const uint32_t leader = elect_one_sync();
    int stage = 0;
    uint32_t phase = 0;
    for (int item = cluster_id; item < total_items; item += n_clusters) {
      const WgradItem wi = decode_item(p, item);
      const CUtensorMap* sm = &p.smap[pl.tap_map[wi.tap]];
      const int dy = pl.tap_dy[wi.tap], dx = pl.tap_dx[wi.tap];
      const int m_c0 = (wi.m_tile * 2 + rank) * 128, n_c0 = wi.n_tile * BLOCK_N + rank * (BLOCK_N / 2);
      for (int pt = wi.p_begin; pt < wi.p_end; ++pt) {
        const int tw = pt % pl.tiles_w;
        const int r2 = pt / pl.tiles_w;
        const int th = r2 % pl.tiles_h;
        const int tn = r2 / pl.tiles_h;
        const int w0 = tw * pl.bw, h0 = th * pl.bh, n0 = tn * pl.bn;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (leader) {
          uint8_t* m_dst = smem + stage * Cfg::kStageBytes;
          uint8_t* n_dst = m_dst + Cfg::kMBytes;
          const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);  // the MMA leader's full barrier
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
          const CUtensorMap* mm = p.m_is_shifted ? sm : &p.umap;
          const CUtensorMap* nm = p.m_is_shifted ? &p.umap : sm;
          const int mdx = p.m_is_shifted ? dx : 0, mdy = p.m_is_shifted ? dy : 0;
          const int ndx = p.m_is_shifted ? 0 : dx, ndy = p.m_is_shifted ? 0 : dy;
#pragma unroll
          for (int g = 0; g < 2; ++g)
            tma_load_4d_pair(m_dst + g * Cfg::kBoxBytes, mm, bar, m_c0 + g * 64, w0 + mdx, h0 + mdy, n0);
#pragma unroll
          for (int g = 0; g < 2; ++g)
            tma_load_4d_pair(n_dst + g * Cfg::kBoxBytes, nm, bar, n_c0 + g * 64, w0 + ndx, h0 + ndy, n0);
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer (leader CTA only)
    if (rank == 0) {
      const uint32_t leader = elect_one_sync();
      constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, /*A MN-major*/ 1, /*B MN-major*/ 1, 256, BLOCK_N);
      const uint32_t smem_base = smem_u32(smem);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int item = cluster_id; item < total_items; item += n_clusters, ++it) {
        const WgradItem wi = decode_item(p, item);
        const int as = it & 1;
        mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BLOCK_N;
        const int k_steps = wi.p_end - wi.p_begin;
        for (int ks = 0; ks < k_steps; ++ks) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t m_addr = smem_base + stage * Cfg::kStageBytes;
          const uint64_t mdesc = umma_smem_desc(m_addr, Cfg::kBoxBytes, 1024);
          const uint64_t ndesc = umma_smem_desc(m_addr + Cfg::kMBytes, Cfg::kBoxBytes, 1024);
          if (leader) {
#pragma unroll
            for (int k = 0; k < 4; ++k)  // 16 pixels per MMA = two 8-line swizzle groups = 2048 bytes (>> 4 = 128)
              umma_f16_pair(d_tmem, mdesc + 128 * k, ndesc + 128 * k, idesc, (ks | k) != 0);
            umma_commit_pair(&empty_bar[stage], 3);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        if (leader) umma_commit_pair(&tfull_bar[as], 3);
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------ epilogue: this CTA's 128 accumulator rows (both CTAs)
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t tempty_leader = mapa_shared(smem_u32(&tempty_bar[0]), 0);
    int it = 0;
    for (int item = cluster_id; item < total_items; item += n_clusters, ++it) {
      const WgradItem wi = decode_item(p, item);
      const int as = it & 1;
      const bool nonempty = wi.p_end > wi.p_begin;
      mbar_wait(&tfull_bar[as], (it >> 1) & 1);
      tc_fence_after();
      const int m = (wi.m_tile * 2 + rank) * 128 + row;
      float* obase = p.out + (size_t)wi.tap * p.Ca * p.Cb;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + c * 32, r);
        tmem_ld_wait();
        if (nonempty) {
          const int n = wi.n_tile * BLOCK_N + c * 32;
          if (p.m_is_shifted) {  // out[a = m][b = n..n+31]: contiguous per thread
            float* dst = obase + (size_t)m * p.Cb + n;
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(__uint_as_float(r[j])),
                           "f"(__uint_as_float(r[j + 1])), "f"(__uint_as_float(r[j + 2])), "f"(__uint_as_float(r[j + 3]))
                           : "memory");
          } else {  // out[a = n + j][b = m]: coalesced across the warp's lanes
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(obase + (size_t)(n + j) * p.Cb + m, __uint_as_float(r[j]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tempty_leader + as * 8);
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // no CTA frees TMEM / exits while its peer may still signal or read it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
  }
}

static int launch_wgrad_pair(const WgradKernelParams& kp, int total_items, cudaStream_t stream) {
  using Cfg = WgradPairCfg;
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(wgrad_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  int clusters = num_sms() / 2;
  if (total_items < clusters) clusters = total_items;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * clusters, 1, 1);
  cfg.blockDim = dim3(kWgradThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B2G_CUDA_CHECK(cudaLaunchKernelEx(&cfg, wgrad_pair_kernel, kp));
  return B2G_OK;
}


// ------------------------------------------------------------------------------------------------------------
// Tap-grouped variant (the default for the 128+-channel layers): M side = 128 channels of the UNSHIFTED operand,
// N side = up to 256 columns made of `T` taps x `cc` channels of the shifted operand (cc = min(Ca, 256),
// T = 256 / cc).  One unshifted tile feeds T taps, every MMA is N=256 wide (MMA-bound instead of smem-bound for
// the 64/128-channel layers), and the accumulator is stored transposed: out[tap][a][b = M row].
// ------------------------------------------------------------------------------------------------------------
struct alignas(64) WgradTapsParams {
  CUtensorMap smap[4];
  CUtensorMap umap;
  ConvPlan plan;
  int Ca, Cb;
  int cc;          // shifted channels per tap inside one N tile
  int T;           // taps per group
  int groups;      // ceil(taps / T)
  int a_chunks;    // Ca / cc
  int m_tiles;     // Cb / 128
  int splits, ptiles;
  float* out;
};

struct WgradTapsCfg {
  static constexpr int kBoxBytes = 64 * 128;
  static constexpr int kMBytes = 2 * kBoxBytes;
  static constexpr int kNBytes = 4 * kBoxBytes;
  static constexpr int kStageBytes = kMBytes + kNBytes;  // 48 KB
  static constexpr int kStages = 4;
  static constexpr int kTmemCols = 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + 256 + 1024;
};

struct WgradTapsItem {
  int tg, ac, m_tile, nt, p_begin, p_end;
};
__device__ __forceinline__ WgradTapsItem decode_taps_item(const WgradTapsParams& p, int item) {
  WgradTapsItem w;
  w.tg = item % p.groups;
  int r = item / p.groups;
  w.ac = r % p.a_chunks;
  r /= p.a_chunks;
  w.m_tile = r % p.m_tiles;
  const int split = r / p.m_tiles;
  const int left = p.plan.taps - w.tg * p.T;
  w.nt = left < p.T ? left : p.T;
  w.p_begin = (int)((long)split * p.ptiles / p.splits);
  w.p_end = (int)((long)(split + 1) * p.ptiles / p.splits);
  return w;
}

__global__ void __launch_bounds__(kWgradThreads, 1) wgrad_taps_kernel(const __grid_constant__ WgradTapsParams p) {
  using Cfg = WgradTapsCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const ConvPlan& pl = p.plan;
  const int total_items = p.groups * p.a_chunks * p.m_tiles * p.splits;
  const int boxes_per_tap = p.cc / 64;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4);
    }
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < pl.n_amaps; ++i) tma_prefetch_desc(&p.smap[i]);
    tma_prefetch_desc(&p.umap);
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    {
      const uint32_t leader = elect_one_sync();
      int stage = 0;
      uint32_t phase = 0;
      for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        const WgradTapsItem wi = decode_taps_item(p, item);
        const int m_c0 = wi.m_tile * 128, a_c0 = wi.ac * p.cc;
        const uint32_t bytes = Cfg::kMBytes + wi.nt * boxes_per_tap * Cfg::kBoxBytes;
        for (int pt = wi.p_begin; pt < wi.p_end; ++pt) {
          const int tw = pt % pl.tiles_w;
          const int r2 = pt / pl.tiles_w;
          const int th = r2 % pl.tiles_h;
          const int tn = r2 / pl.tiles_h;
          const int w0 = tw * pl.bw, h0 = th * pl.bh, n0 = tn * pl.bn;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* m_dst = smem + stage * Cfg::kStageBytes;
          uint8_t* n_dst = m_dst + Cfg::kMBytes;
          if (leader) {
          mbar_expect_tx(&full_bar[stage], bytes);
          tma_load_4d(m_dst, &p.umap, &full_bar[stage], m_c0, w0, h0, n0);
          tma_load_4d(m_dst + Cfg::kBoxBytes, &p.umap, &full_bar[stage], m_c0 + 64, w0, h0, n0);
          for (int t = 0; t < wi.nt; ++t) {
            const int tap = wi.tg * p.T + t;
            const CUtensorMap* sm = &p.smap[pl.tap_map[tap]];
            const int dy = pl.tap_dy[tap], dx = pl.tap_dx[tap];
            for (int g = 0; g < boxes_per_tap; ++g)
              tma_load_4d(n_dst + (t * boxes_per_tap + g) * Cfg::kBoxBytes, sm, &full_bar[stage], a_c0 + g * 64,
                          w0 + dx, h0 + dy, n0);
          }
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t leader = elect_one_sync();
    const uint32_t smem_base = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++it) {
      const WgradTapsItem wi = decode_taps_item(p, item);
      const uint32_t idesc = umma_idesc(/*bf16*/ 1, /*A MN-major*/ 1, /*B MN-major*/ 1, 128, wi.nt * p.cc);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(&tempty_bar[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * 256;
      const int k_steps = wi.p_end - wi.p_begin;
      for (int ks = 0; ks < k_steps; ++ks) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t m_addr = smem_base + stage * Cfg::kStageBytes;
        const uint64_t mdesc = umma_smem_desc(m_addr, Cfg::kBoxBytes, 1024);
        const uint64_t ndesc = umma_smem_desc(m_addr + Cfg::kMBytes, Cfg::kBoxBytes, 1024);
        if (leader) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_f16(d_tmem, mdesc + 128 * k, ndesc + 128 * k, idesc, (ks | k) != 0);
          umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
      if (leader) umma_commit(&tfull_bar[as]);
      __syncwarp();
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    int it = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++it) {
      const WgradTapsItem wi = decode_taps_item(p, item);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const bool nonempty = wi.p_end > wi.p_begin;
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const int b = wi.m_tile * 128 + row;
      const int n_chunks = wi.nt * p.cc / 32;
#pragma unroll 1
      for (int c = 0; c < n_chunks; ++c) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * 256 + c * 32, r);
        tmem_ld_wait();
        if (nonempty) {
          const int n = c * 32;
          const int tap = wi.tg * p.T + n / p.cc;
          const int a0 = wi.ac * p.cc + n % p.cc;
          float* dst = p.out + ((size_t)tap * p.Ca + a0) * p.Cb + b;
#pragma unroll
          for (int j = 0; j < 32; ++j) atomicAdd(dst + (size_t)j * p.Cb, __uint_as_float(r[j]));
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

static int launch_wgrad_taps(const WgradTapsParams& kp, int total_items, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(wgrad_taps_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        WgradTapsCfg::kSmemBytes));
    attr_set = true;
  }
  int grid = total_items < num_sms() ? total_items : num_sms();
  wgrad_taps_kernel<<<grid, kWgradThreads, WgradTapsCfg::kSmemBytes, stream>>>(kp);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

// tensor maps of both operands ([64 pixel][64 channel] boxes) shared by the two kernel variants
static int make_wgrad_maps(const b2g_wgrad_args* a, const ConvPlan& pl, CUtensorMap* smap, CUtensorMap* umap) {
  const size_t es = 2;
  const uint32_t box[4] = {64u, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bn};
  const char* sbase = static_cast<const char*>(a->shifted);
  if (pl.n_amaps == 1) {
    const uint64_t dims[4] = {(uint64_t)a->Ca, (uint64_t)a->Ws, (uint64_t)a->Hs, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Ca * es, (uint64_t)a->Ws * a->Ca * es, (uint64_t)a->Hs * a->Ws * a->Ca * es};
    int rc = make_tmap(&smap[0], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, sbase, dims, str, box);
    if (rc) return rc;
  } else {
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        const uint64_t dims[4] = {(uint64_t)a->Ca, (uint64_t)a->Ws / 2, (uint64_t)a->Hs / 2, (uint64_t)a->N};
        const uint64_t str[3] = {(uint64_t)2 * a->Ca * es, (uint64_t)2 * a->Ws * a->Ca * es,
                                 (uint64_t)a->Hs * a->Ws * a->Ca * es};
        int rc = make_tmap(&smap[py * 2 + px], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                           sbase + ((size_t)py * a->Ws + px) * a->Ca * es, dims, str, box);
        if (rc) return rc;
      }
  }
  const uint64_t dims[4] = {(uint64_t)a->Cb, (uint64_t)pl.Wg, (uint64_t)pl.Hg, (uint64_t)a->N};
  const uint64_t str[3] = {(uint64_t)a->Cb * es, (uint64_t)pl.Wg * a->Cb * es, (uint64_t)pl.Hg * pl.Wg * a->Cb * es};
  return make_tmap(umap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a->unshifted, dims, str, box);
}

}  // namespace b2g

using namespace b2g;

extern "C" int b2g_conv_wgrad(const b2g_wgrad_args* a, b2g_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a || !a->shifted || !a->unshifted || !a->dw) return fail(B2G_E_INVALID_ARG, "b2g_conv_wgrad: null pointer");
  if (a->kind == B2G_CONV_K3S1_C3) return launch_wgrad_c3(a, stream_);  // 3-channel edge layers (b2g_wgrad_c3.cu)
  if (a->kind != B2G_CONV_K3S1 && a->kind != B2G_CONV_K4S2 && a->kind != B2G_CONV_K1S1)
    return fail(B2G_E_INVALID_ARG, "b2g_conv_wgrad: kind must be K3S1, K4S2 or K1S1 (tap plan of the shifted operand)");
  if (a->Ca % 64 || a->Cb % 64 || a->Ca <= 0 || a->Cb <= 0)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad: channels (%d,%d) must be multiples of 64", a->Ca, a->Cb);
  if (a->ca_valid < 0 || a->ca_valid > a->Ca || (a->dbias && (a->ca_valid <= 0 || a->ca_valid >= a->Ca)))
    return fail(B2G_E_INVALID_ARG, "b2g_conv_wgrad: bad ca_valid=%d for Ca=%d", a->ca_valid, a->Ca);

  static int force_taps = []() { const char* e = getenv("B2G_WGRAD_TAPS"); return e ? atoi(e) : -1; }();
  const int tf32 = a->tf32 ? 1 : 0;  // fp32 operands, kind::tf32 (the per-layer parity path)
  if (tf32 && a->kind == B2G_CONV_K1S1)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad: the tf32 path covers K3S1 / K4S2");
  const bool taps_ok = !tf32 && a->Cb % 128 == 0 && (a->ca_valid == 0 || a->ca_valid == a->Ca) && !a->dbias &&
                       (a->Ca <= 256 || a->Ca % 256 == 0);  // the tap-grouped kernels split Ca into chunks of min(Ca, 256)
  // Measured on B200 (profiles/r1_wgrad_variants.md): grouping taps wins when the other kernel's N tile would be
  // narrow (Cb == 128: 64->128 and 128->128 layers); for Cb >= 256 the per-tap kernel with vector reds is faster.
  const bool use_taps = taps_ok && (force_taps < 0 ? a->Cb == 128 : force_taps != 0);
  if (use_taps) {  // same layers, shifted operand read once per pixel row when the grid is >= 64 pixels wide
    bool handled = false;
    int rc = launch_wgrad_slab(a, stream_, &handled);
    if (rc || handled) return rc;
  }
  if (use_taps) {
    // tap-grouped kernel (M = unshifted operand, N = taps x shifted channels)
    WgradTapsParams tp;
    memset(&tp, 0, sizeof(tp));
    if (!build_conv_plan(a->kind, a->N, a->Hs, a->Ws, a->Ca, a->Cb, 64, &tp.plan))
      return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad: cannot tile N=%d Hs=%d Ws=%d", a->N, a->Hs, a->Ws);
    const ConvPlan& tpl = tp.plan;
    tp.Ca = a->Ca;
    tp.Cb = a->Cb;
    tp.cc = a->Ca < 256 ? a->Ca : 256;
    if (a->Ca % tp.cc) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad: Ca=%d unsupported", a->Ca);
    tp.T = 256 / tp.cc;
    tp.groups = (tpl.taps + tp.T - 1) / tp.T;
    tp.a_chunks = a->Ca / tp.cc;
    tp.m_tiles = a->Cb / 128;
    tp.ptiles = tpl.m_tiles;
    tp.out = a->dw;
    const int base = tp.groups * tp.a_chunks * tp.m_tiles;
    // equal-length work items: keep their count just below a whole number of rounds over the SMs (no ragged tail)
    static int rounds = []() { const char* e = getenv("B2G_WGRAD_ROUNDS"); int r = e ? atoi(e) : 2; return r > 0 ? r : 2; }();
    int sp = rounds * num_sms() / base;
    if (sp > tp.ptiles / 8) sp = tp.ptiles / 8;
    if (sp < 1) sp = 1;
    tp.splits = sp;
    int rc = make_wgrad_maps(a, tpl, tp.smap, &tp.umap);
    if (rc) return rc;
    return launch_wgrad_taps(tp, base * sp, stream);
  }

  WgradKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  // plan over the shifted operand's geometry; its grid (Hg, Wg) is the unshifted operand's spatial size
  if (!build_conv_plan(a->kind, a->N, a->Hs, a->Ws, a->Ca, a->Cb, 64, &kp.plan))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad: cannot tile N=%d Hs=%d Ws=%d", a->N, a->Hs, a->Ws);
  const ConvPlan& pl = kp.plan;
  // M side needs 128-channel tiles: prefer the operand with a multiple of 128 channels; when neither has one
  // (64 x 64 edge layers) the shifted operand goes on M and its upper 64 rows are TMA zero fill (box beyond C).
  kp.m_is_shifted = (a->Ca % 128 == 0 || a->Cb % 128 != 0) ? 1 : 0;
  kp.Cm = kp.m_is_shifted ? a->Ca : a->Cb;
  kp.Cn = kp.m_is_shifted ? a->Cb : a->Ca;
  const int block_n = (kp.Cn % 256 == 0 && !tf32) ? 256 : (kp.Cn % 128 == 0 ? 128 : 64);
  kp.m_tiles = (kp.Cm + 127) / 128;
  kp.ca_valid = a->ca_valid > 0 ? a->ca_valid : a->Ca;
  kp.dbias = a->dbias;
  if (kp.ca_valid != a->Ca && !kp.m_is_shifted)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad: ca_valid needs the shifted operand on the M side");
  kp.n_tiles = kp.Cn / block_n;
  kp.ptiles = pl.m_tiles;
  kp.Ca = a->Ca;
  kp.Cb = a->Cb;
  kp.out = a->dw;
  // CTA-pair kernel: both sides in whole 256-channel tiles, plain bf16 gradient (B2G_WGRAD_PAIR=0 turns it off)
  static int pair_env = []() { const char* e = getenv("B2G_WGRAD_PAIR"); return e ? atoi(e) : 1; }();
  const bool use_pair = pair_env && !tf32 && block_n == 256 && kp.Cm % 256 == 0 && kp.ca_valid == a->Ca && !a->dbias;
  if (use_pair) kp.m_tiles = kp.Cm / 256;  // the pair kernel counts PAIRS of 128-channel M tiles
  // split-K: equal-length work items, just under 4 whole rounds over the SMs (no ragged last round)
  const int base_items = pl.taps * kp.m_tiles * kp.n_tiles;
  static int rounds1 = []() { const char* e = getenv("B2G_WGRAD_ROUNDS1"); int r = e ? atoi(e) : 4; return r > 0 ? r : 4; }();
  int splits = rounds1 * (use_pair ? num_sms() / 2 : num_sms()) / base_items;
  if (splits > kp.ptiles / 8) splits = kp.ptiles / 8;
  if (splits < 1) splits = 1;
  kp.splits = splits;

  const size_t es = tf32 ? 4 : 2;
  const CUtensorMapDataType dt = tf32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  // MN-major 32-bit operands must use the 32-byte-atom flavour of the 128-byte swizzle (see umma_smem_desc_sw128_32b)
  const CUtensorMapSwizzle sw = tf32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
  const uint32_t box[4] = {tf32 ? 32u : 64u, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bn};
  const char* sbase = static_cast<const char*>(a->shifted);
  if (pl.n_amaps == 1) {
    const uint64_t dims[4] = {(uint64_t)a->Ca, (uint64_t)a->Ws, (uint64_t)a->Hs, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Ca * es, (uint64_t)a->Ws * a->Ca * es, (uint64_t)a->Hs * a->Ws * a->Ca * es};
    int rc = make_tmap(&kp.smap[0], dt, 4, sbase, dims, str, box, sw);
    if (rc) return rc;
  } else {
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        const uint64_t dims[4] = {(uint64_t)a->Ca, (uint64_t)a->Ws / 2, (uint64_t)a->Hs / 2, (uint64_t)a->N};
        const uint64_t str[3] = {(uint64_t)2 * a->Ca * es, (uint64_t)2 * a->Ws * a->Ca * es,
                                 (uint64_t)a->Hs * a->Ws * a->Ca * es};
        int rc = make_tmap(&kp.smap[py * 2 + px], dt, 4,
                           sbase + ((size_t)py * a->Ws + px) * a->Ca * es, dims, str, box, sw);
        if (rc) return rc;
      }
  }
  {
    const uint64_t dims[4] = {(uint64_t)a->Cb, (uint64_t)pl.Wg, (uint64_t)pl.Hg, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Cb * es, (uint64_t)pl.Wg * a->Cb * es, (uint64_t)pl.Hg * pl.Wg * a->Cb * es};
    int rc = make_tmap(&kp.umap, dt, 4, a->unshifted, dims, str, box, sw);
    if (rc) return rc;
  }
  const int total_items = base_items * kp.splits;
  if (use_pair) return launch_wgrad_pair(kp, total_items, stream);
  if (tf32) return block_n == 128 ? launch_wgrad<128, true>(kp, total_items, stream)
                                  : launch_wgrad<64, true>(kp, total_items, stream);
  switch (block_n) {
    case 256: return launch_wgrad<256>(kp, total_items, stream);
    case 128: return launch_wgrad<128>(kp, total_items, stream);
    default: return launch_wgrad<64>(kp, total_items, stream);
  }
}
