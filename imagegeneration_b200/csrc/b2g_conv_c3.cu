// b2g_conv_c3.cu — 3x3 stride-1 SAME convolution over a 3-channel fp32 tensor (3 -> 64 channels) on tcgen05, with
// the im2col operand built IN SHARED MEMORY instead of in HBM.
//
//   warp 0      : loads the packed [64][64] weight tile once (TMA; resident for the whole kernel), then streams the
//                 fp32 input: per 128-pixel tile ONE TMA box = the tile's pixels plus a one-pixel halo
//                 ((bw+8)*3 floats x (bh+2) rows x bn images; out-of-image parts arrive as zeros = SAME padding;
//                 the box starts 4 pixels left of the tile because TMA wants a 16-byte aligned start offset)
//                 through an 8-stage ring — the HBM/L2 latency is hidden by the ring, not by the builder threads
//   warp 1      : TMEM allocator + MMA issuer — one K=32 step (two K=16 MMAs, M=128, N=64) per 128-pixel tile
//   warps 2..3  : patch builders, two pixels per thread: 27 conflict-free LDS.32 from the staged box + a ones
//                 column -> 32 bf16 = 4 x 16 B written at the SW128-swizzled position of the pixel's row,
//                 fence.proxy.async, mbarrier arrive.  This is the im2col27 tensor of b2g_edge.cu, never leaving
//                 the SM.
//   warps 4..15 : epilogue (3 groups x 4 TMEM lane quarters, 4 accumulator buffers).  A warp takes BOTH 32-channel
//                 halves of its 32 pixels in one visit (two tcgen05.ld in flight, the accumulator is released as soon as
//                 they land): bias / LeakyReLU / dropout or ReLU' from the saved output (lean math of b2g_conv_epi.cuh)
//                 into a [32 pixels][64 channels] 128-byte-swizzled box, ONE TMA store of whole 128-byte lines per
//                 visit, two boxes per warp.  Round 2 measured the 16-warp / 32-channel-item form latency-bound: per
//                 item ~4000 cycles of one warp for ~250 instructions, a third of the stall samples in per-item
//                 bookkeeping (tile walk, coordinates, fence / elect / store) — twice the work per visit halves that
//                 and doubles the independent work a warp has in flight.
//
// HBM traffic per pixel: 12 B read + 128 B written (+128 B for the dact variant), against 12+128 (im2col27) +
// 128+128 (K1S1 GEMM) for the two-kernel route.  Replaces D.conv1 forward (Conv2D 3->64 k3 s1 + bias + LeakyReLU +
// Dropout, sndcgan/SNDCGAN.py:73-78) and the input gradient of G.convT4 (Conv2DTranspose 64->3 k3 s1, :57-61).
#include <string.h>

#include "b2g_conv.cuh"
#include "b2g_conv_epi.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) C3KernelParams {
  CUtensorMap bmap;
  CUtensorMap imap;  // fp32 input seen as 8-byte elements: dims {W*3/2, H, N}, box {(bw+8)*3/2, bh+2, bn}
  ConvPlan plan;     // K1S1 geometry: 128-pixel boxes (bw, bh, bn) over the N x H x W grid
  EpiParams e;
  EpiTmaMaps em;     // 32-channel maps of b2g_conv_epi.cuh (unused here: kept so fill_epi_tma decides e.tma / e.lean)
  CUtensorMap omap64, ymap64;  // TMA epilogue (used when e.tma): bf16 NHWC output / saved output, box {64 channels, qw,
                               // qh, qn} = 32 pixels x 128 bytes, 128-byte swizzle
  int in_box_bytes;  // bytes one input box delivers
};

struct C3Cfg {
  static constexpr int kBlockN = 64;
  static constexpr int kABytes = 128 * 128;  // 128 pixel rows of 128 B; only the first 64 B (K = 32) are used
  static constexpr int kStages = 4;
  static constexpr int kInStages = 6;           // staged fp32 input boxes
  static constexpr int kInStageBytes = 8192;    // upper bound of one box (128-px row tile: 136 * 12 * 3 = 4896)
  static constexpr int kBBytes = kBlockN * 128;
  static constexpr int kAccBufs = 4;            // 8 (all of TMEM) measured the same (round 2)
  static constexpr int kTmemCols = kAccBufs * kBlockN;
  static constexpr int kBuilderWarps = 2;       // 64 threads, two pixel rows each (16 warps per CTA = 4 per scheduler: 128 registers)
  static constexpr int kEpiGroups = 3;
  static constexpr int kThreads = 64 + 32 * kBuilderWarps + 128 * kEpiGroups;
  static constexpr int kMainBytes = kStages * kABytes + kBBytes + kInStages * kInStageBytes;  // multiple of 1 KiB
  static constexpr int kEpiWarps = 4 * kEpiGroups;
  static constexpr int kEpiBoxes = 2;           // boxes per epilogue warp
  static constexpr int kEpiBoxBytes = 4096;     // [32 pixels][64 channels] bf16
  static constexpr int kEpiBytes = kEpiWarps * kEpiBoxes * kEpiBoxBytes + 256;  // boxes, then their mbarriers
  static constexpr int kSmemBytes = kMainBytes + kEpiBytes + 512 + 1024;
};

__global__ void __launch_bounds__(C3Cfg::kThreads, 1) conv_c3_kernel(const __grid_constant__ C3KernelParams p) {
  using Cfg = C3Cfg;
  constexpr int BLOCK_N = Cfg::kBlockN;
  __shared__ float sbias[512];     // bias table, or 8 copies of the [64] sums of squares (EpiParams::cs_copies)
  __shared__ float s_colsum[512];  // 8 copies of the [64] column sums: one per pair of epilogue warps
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* bsm = smem + Cfg::kStages * Cfg::kABytes;
  uint8_t* insm = bsm + Cfg::kBBytes;
  uint8_t* epism = smem + Cfg::kMainBytes;  // per-warp boxes of the TMA epilogue
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epism + Cfg::kEpiBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + Cfg::kAccBufs;
  uint64_t* bfull_bar = tempty_bar + Cfg::kAccBufs;
  uint64_t* in_full = bfull_bar + 1;
  uint64_t* in_empty = in_full + Cfg::kInStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(in_empty + Cfg::kInStages);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const ConvPlan& pl = p.plan;
  const int total_tiles = pl.m_tiles;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], Cfg::kBuilderWarps);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < Cfg::kAccBufs; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4);  // the four lane-quarter warps of the group that owns the tile
    }
    mbar_init(bfull_bar, 1);
    for (int s = 0; s < Cfg::kInStages; ++s) {
      mbar_init(&in_full[s], 1);
      mbar_init(&in_empty[s], Cfg::kBuilderWarps);
    }
    if (p.e.tma)
      for (int w = 0; w < Cfg::kEpiWarps * Cfg::kEpiBoxes; ++w)
        mbar_init(reinterpret_cast<uint64_t*>(epism + Cfg::kEpiWarps * Cfg::kEpiBoxes * Cfg::kEpiBoxBytes) + w, 1);
    fence_barrier_init();
  }
  epi_load_bias(p.e, sbias, BLOCK_N);
  epi_colsum_zero(p.e, s_colsum, BLOCK_N, sbias);
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ weight tile: loaded once, resident; then the input stream
    const uint32_t leader = elect_one_sync();
    if (leader) {
      tma_prefetch_desc(&p.bmap);
      tma_prefetch_desc(&p.imap);
      mbar_expect_tx(bfull_bar, Cfg::kBBytes);
      tma_load_2d(bsm, &p.bmap, bfull_bar, 0, 0);
    }
    __syncwarp();
    int stage = 0;
    uint32_t phase = 0;
    TileWalk tc = tile_walk_begin(pl, blockIdx.x, gridDim.x);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, tile_walk_next(pl, tc)) {
      mbar_wait(&in_empty[stage], phase ^ 1);
      if (leader) {
        mbar_expect_tx(&in_full[stage], p.in_box_bytes);
        // box origin: pixel (w0 - 4, h0 - 1) of image n0; 3 floats per pixel, 2 floats per 8-byte element.
        // (Measured: a start offset that is not a multiple of 16 bytes is an illegal-instruction fault.)
        tma_load_3d(insm + stage * Cfg::kInStageBytes, &p.imap, &in_full[stage], (tc.tw * pl.bw - 4) * 3 / 2,
                    tc.th * pl.bh - 1, tc.tn * pl.bn);
      }
      __syncwarp();
      if (++stage == Cfg::kInStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer (whole warp converged; one elected lane issues)
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, 0, 0, 128, BLOCK_N);
    const uint32_t smem_base = smem_u32(smem);
    const uint64_t bdesc = umma_desc_kmajor(smem_u32(bsm));
    mbar_wait(bfull_bar, 0);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int as = it % Cfg::kAccBufs;
      mbar_wait(&tempty_bar[as], ((it / Cfg::kAccBufs) & 1) ^ 1);
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BLOCK_N;
      const uint64_t adesc = umma_desc_kmajor(smem_base + stage * Cfg::kABytes);
      if (leader) {
        umma_f16(d_tmem, adesc, bdesc, idesc, 0);          // K = 0..15
        umma_f16(d_tmem, adesc + 2, bdesc + 2, idesc, 1);  // K = 16..31 (27 taps*channels, ones column, zeros)
        umma_commit(&empty_bar[stage]);
        umma_commit(&tfull_bar[as]);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp < 2 + Cfg::kBuilderWarps) {
    // ------------------------------------------------ patch builders: thread t owns pixel rows t and t + 64 of the A tile
    const int t = threadIdx.x - 64;
    const int pitch = (pl.bw + 8) * 3;
    uint32_t row_off[2], swz[2];
    int win_off[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int m = t + 64 * r;
      const int wl = m % pl.bw, hl = (m / pl.bw) % pl.bh, nl = m / (pl.bw * pl.bh);
      row_off[r] = m * 128;
      swz[r] = m & 7;
      // the row's window inside a staged box [bn][bh + 2][(bw + 8) * 3 floats]: rows hl..hl+2, nine consecutive floats
      // from pixel wl + 3 (= image pixel x - 1).  Consecutive lanes are 3 words apart: conflict-free LDS.32.
      win_off[r] = ((nl * (pl.bh + 2) + hl) * (pl.bw + 8) + wl + 3) * 3;
    }
    int stage = 0, istage = 0;
    uint32_t phase = 0, iphase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      float v[2][28];
      // the builders are on the critical path of every tile: plain try_wait polling, no back-off sleep (a parked wait
      // woke them up to 256 ns late, half of the per-tile budget)
      mbar_wait(&in_full[istage], iphase);
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const float* src = reinterpret_cast<const float*>(insm + istage * Cfg::kInStageBytes) + win_off[r];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
          for (int j = 0; j < 9; ++j) v[r][ky * 9 + j] = src[ky * pitch + j];  // j = kx * 3 + c
        v[r][27] = 1.f;
      }
      mbar_wait(&empty_bar[stage], phase ^ 1);
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        uint8_t* dst = smem + stage * Cfg::kABytes + row_off[r];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 o;
          o.x = pack_bf16x2(v[r][j * 8 + 0], v[r][j * 8 + 1]);
          o.y = pack_bf16x2(v[r][j * 8 + 2], v[r][j * 8 + 3]);
          if (j < 3) {
            o.z = pack_bf16x2(v[r][j * 8 + 4], v[r][j * 8 + 5]);
            o.w = pack_bf16x2(v[r][j * 8 + 6], v[r][j * 8 + 7]);
          } else {
            o.z = 0u;  // k = 28 .. 31: zero padding of the K = 32 step
            o.w = 0u;
          }
          *reinterpret_cast<uint4*>(dst + ((j ^ swz[r]) << 4)) = o;  // 128-byte swizzle: 16-byte chunk j of row m -> j ^ (m & 7)
        }
      }
      fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&in_empty[istage]);  // the box was consumed into registers before the stores above
        mbar_arrive(&full_bar[stage]);
      }
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      if (++istage == Cfg::kInStages) { istage = 0; iphase ^= 1; }
    }
  } else {
    // ------------------------------------------------ epilogue (warp w owns TMEM lanes 32*(w%4)..+31)
    const EpiParams& e = p.e;                 // stays in the constant bank
    const EpiSeed sd = epi_resolve_seed(p.e);  // dropout seed possibly read from device memory
    const int q = warp & 3;
    const int ew = warp - 2 - Cfg::kBuilderWarps;
    const int grp = ew >> 2;
    const int row = q * 32 + lane;
    const int wl = row % pl.bw, hl = (row / pl.bw) % pl.bh, nl = row / (pl.bw * pl.bh);
    // origin of this warp's 32-pixel box inside a tile (TMA epilogue)
    const int wl0 = (q * 32) % pl.bw, hl0 = ((q * 32) / pl.bw) % pl.bh, nl0 = (q * 32) / (pl.bw * pl.bh);
    constexpr int NB = Cfg::kEpiBoxes;
    uint8_t* const boxes = epism + ew * NB * Cfg::kEpiBoxBytes;
    uint64_t* const ybar = reinterpret_cast<uint64_t*>(epism + Cfg::kEpiWarps * NB * Cfg::kEpiBoxBytes) + ew * NB;
    const bool use_tma = e.tma != 0;
    const bool bwd = e.dact != B2G_ACT_NONE;
    const int lean = e.lean;
    const int cs_off = (warp % e.cs_copies) * BLOCK_N;
    // this thread's 128-byte row of a box: 16-byte group j lives at j ^ (lane & 7) (128-byte swizzle); the two 32-channel
    // halves are groups 0..3 and 4..7 = two 64-byte half rows with the 64-byte-style swizzle (lane & 3) inside
    const uint32_t s7 = lane & 7, sw = s7 & 3;
    const uint32_t half0 = (s7 & 4) << 4, half1 = half0 ^ 64;
    // saved-output load of the warp's visit number k into box k % NB (the store that last used that box must have read it)
    uint32_t k_begin = 0, k_done = 0;
    auto begin = [&](int bx, int by, int bn) {
      const uint32_t b = k_begin % NB;
      ++k_begin;
      if (elect_one_sync()) {
        tma_store_wait_read_n<0>();
        mbar_expect_tx(&ybar[b], Cfg::kEpiBoxBytes);
        tma_load_4d(boxes + b * Cfg::kEpiBoxBytes, &p.ymap64, &ybar[b], 0, bx, by, bn);
      }
      __syncwarp();
    };
    // group g owns every third tile of the CTA, starting with tile number g: a division-free walk per warp
    const int stride = Cfg::kEpiGroups * gridDim.x;
    int it = grp;
    int tile = blockIdx.x + it * gridDim.x;
    TileWalk tc = tile_walk_begin(pl, tile, stride);
    for (; tile < total_tiles; tile += stride, it += Cfg::kEpiGroups, tile_walk_next(pl, tc)) {
      const int as = it % Cfg::kAccBufs;
      const int bx = tc.tw * pl.bw + wl0, by = tc.th * pl.bh + hl0, bn = tc.tn * pl.bn + nl0;
      const int gw = tc.tw * pl.bw + wl, gh = tc.th * pl.bh + hl, gn = tc.tn * pl.bn + nl;
      const size_t pix = (static_cast<size_t>(gn) * pl.Hout + gh) * pl.Wout + gw;
      const bool dbg = (e.dbg & 1) != 0;
      if (use_tma && bwd && !dbg && k_begin == k_done) begin(bx, by, bn);  // the very first visit; later ones are begun one ahead
      mbar_wait_parked(&tfull_bar[as], (it / Cfg::kAccBufs) & 1);
      tc_fence_after();
      uint32_t r0[32], r1[32];
      __syncwarp();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N;
      tmem_ld_32x32(taddr, r0);
      tmem_ld_32x32(taddr + 32, r1);
      if (use_tma && bwd && !dbg && tile + stride < total_tiles) {
        // the warp's next visit is begun into the other box now: its load has until the next accumulator to arrive
        TileWalk nx = tc;
        tile_walk_next(pl, nx);
        begin(nx.tw * pl.bw + wl0, nx.th * pl.bh + hl0, nx.tn * pl.bn + nl0);
      }
      tmem_ld_wait();
      tc_fence_before();  // both halves are in registers: the accumulator goes back to the MMA warp before the math
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
      if (dbg) continue;
      const bool valid = (gh < pl.Hg) && (gn < pl.N);
      if (!use_tma) {
        conv_epilogue_chunk(e, sbias, r0, pix, 0, valid, s_colsum, sd);
        conv_epilogue_chunk(e, sbias, r1, pix, 32, valid, s_colsum, sd);
        continue;
      }
      if (e.colsum && (!bwd || e.dact_zero != 0.f) && !valid) {  // see conv_pair_kernel: pixels outside the tensor
#pragma unroll
        for (int j = 0; j < 32; ++j) { r0[j] = 0u; r1[j] = 0u; }
      }
      const uint32_t k = k_done++;
      const uint32_t b = k % NB;
      uint8_t* box = boxes + b * Cfg::kEpiBoxBytes;
      uint8_t* rowp = box + lane * 128;
      if (bwd) {
        mbar_wait(&ybar[b], (k / NB) & 1);
      } else {
        if (elect_one_sync()) tma_store_wait_read_n<NB - 1>();  // the store of visit k - NB has read this box
        __syncwarp();
      }
      const uint32_t grp32 = static_cast<uint32_t>((pix * BLOCK_N) >> 5);
      if (lean == 1) {
        epi_lean_fwd(e, sbias, r0, grp32, 0, rowp + half0, sw, sd);
        epi_lean_fwd(e, sbias, r1, grp32 + 1, 32, rowp + half1, sw, sd);
      } else if (lean == 2) {
        epi_lean_bwd<2>(e, r0, rowp + half0, sw);
        epi_lean_bwd<2>(e, r1, rowp + half1, sw);
      } else if (lean == 3) {
        epi_lean_bwd<3>(e, r0, rowp + half0, sw);
        epi_lean_bwd<3>(e, r1, rowp + half1, sw);
      } else if (lean == 4) {
        epi_lean_bwd<4>(e, r0, rowp + half0, sw);
        epi_lean_bwd<4>(e, r1, rowp + half1, sw);
      } else {
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          uint8_t* hp = rowp + (c ? half1 : half0);
          uint32_t rr[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) rr[j] = c ? r1[j] : r0[j];
          conv_epilogue_math(
              e, sbias, rr, pix, c * 32, true,
              [&](int g) { return *reinterpret_cast<const uint4*>(hp + ((g ^ sw) << 4)); },
              [&](int g, const uint4& o) { *reinterpret_cast<uint4*>(hp + ((g ^ sw) << 4)) = o; }, nullptr, false, nullptr,
              sd);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (elect_one_sync()) tma_store_4d(&p.omap64, box, 0, bx, by, bn);
      __syncwarp();
      if (e.colsum) {
        // Column sums straight from the box (the TMA store only READS it).  Lane = channel pair: word `lane` of every
        // pixel row, conflict free (a row's 32 words are a permutation of the banks); rows of pixels outside the tensor
        // hold exact zeros.  One mixed-precision add per element, no shuffles: each lane owns its two channels.
        const uint32_t gq = lane >> 2, wo = (lane & 3) * 4;
        float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
        if (e.colsumsq) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const uint32_t w = *reinterpret_cast<const uint32_t*>(box + j * 128 + ((gq ^ (j & 7)) << 4) + wo);
            bf2_acc(w, s0, s1);
            bf2_acc_sq(w, q0, q1);
          }
          atomicAdd(&sbias[cs_off + 2 * lane], q0);
          atomicAdd(&sbias[cs_off + 2 * lane + 1], q1);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            bf2_acc(*reinterpret_cast<const uint32_t*>(box + j * 128 + ((gq ^ (j & 7)) << 4) + wo), s0, s1);
        }
        atomicAdd(&s_colsum[cs_off + 2 * lane], s0);
        atomicAdd(&s_colsum[cs_off + 2 * lane + 1], s1);
      }
    }
    if (use_tma) epi_tma_end();
  }

  tc_fence_before();
  __syncthreads();
  epi_colsum_flush(p.e, s_colsum, BLOCK_N, sbias);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

int launch_conv_c3(const b2g_conv_args* a, void* stream_) {
  if (a->Cin != 3) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm(K3S1_C3): Cin=%d, expected 3", a->Cin);
  if (a->Cout != 64 || a->out_f32)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm(K3S1_C3): Cout=%d (bf16 output) expected 64", a->Cout);
  if (a->dact != B2G_ACT_NONE && !a->yprev) return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dact without yprev");
  if (a->dropout_thresh8 < 0 || a->dropout_thresh8 > 255)
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dropout_thresh8 out of range");
  if (a->colsum && (a->bias || (a->dact == B2G_ACT_NONE && !a->colsumsq)))
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: colsum needs a bf16 backward epilogue (dact, no bias)");
  if (a->colsumsq && (!a->colsum || a->dropout_thresh8))
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: colsumsq needs colsum, no bias and no dropout (BatchNorm statistics)");
  if (a->dropout_thresh8 > 0 && a->act == B2G_ACT_TANH)
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dropout needs a positively homogeneous activation (none/leaky/relu)");
  C3KernelParams kp;
  memset(&kp, 0, sizeof(kp));
  if (!build_conv_plan(a->kind, a->N, a->Hin, a->Win, 64, a->Cout, 128, &kp.plan))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm(K3S1_C3): cannot tile N=%d H=%d W=%d", a->N, a->Hin, a->Win);
  fill_epi_params(a, &kp.e);
  {
    int rc = fill_epi_tma(a, kp.plan, true, &kp.e, &kp.em);
    if (rc) return rc;
    if (kp.e.tma) {  // the same 32-pixel boxes, all 64 channels of a pixel per row
      const ConvPlan& pl = kp.plan;
      const uint64_t dims[4] = {64u, (uint64_t)pl.Wout, (uint64_t)pl.Hout, (uint64_t)pl.N};
      const uint64_t str[3] = {64u * 2u, (uint64_t)pl.Wout * 64u * 2u, (uint64_t)pl.Hout * pl.Wout * 64u * 2u};
      const uint32_t box[4] = {64u, (uint32_t)kp.e.qw, (uint32_t)kp.e.qh, (uint32_t)kp.e.qn};
      rc = make_tmap(&kp.omap64, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a->out, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) return rc;
      if (a->dact != B2G_ACT_NONE) {
        rc = make_tmap(&kp.ymap64, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a->yprev, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
      }
    }
  }
  {
    const ConvPlan& pl = kp.plan;
    kp.in_box_bytes = pl.bn * (pl.bh + 2) * (pl.bw + 8) * 12;
    if (pl.bw % 4 || a->Win % 4 || kp.in_box_bytes > C3Cfg::kInStageBytes)
      return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm(K3S1_C3): cannot stage N=%d H=%d W=%d (box %d bytes)", a->N,
                  a->Hin, a->Win, kp.in_box_bytes);
    // 8-byte elements (two floats): the 3-float pixels cannot be a TMA dimension (12 bytes), 2-float units can
    const uint64_t dims[3] = {(uint64_t)a->Win * 3 / 2, (uint64_t)a->Hin, (uint64_t)a->N};
    const uint64_t str[2] = {(uint64_t)a->Win * 12, (uint64_t)a->Hin * a->Win * 12};
    const uint32_t ibox[3] = {(uint32_t)(pl.bw + 8) * 3 / 2, (uint32_t)pl.bh + 2, (uint32_t)pl.bn};
    int rc = make_tmap(&kp.imap, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, a->in, dims, str, ibox, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {64u, (uint64_t)a->Cout};
    const uint64_t str[1] = {64u * 2u};
    const uint32_t bbox[2] = {64u, (uint32_t)C3Cfg::kBlockN};
    int rc = make_tmap(&kp.bmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a->wpacked, dims, str, bbox);
    if (rc) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(conv_c3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C3Cfg::kSmemBytes));
    attr_set = true;
  }
  const int total = kp.plan.m_tiles;
  const int grid = total < num_sms() ? total : num_sms();
  conv_c3_kernel<<<grid, C3Cfg::kThreads, C3Cfg::kSmemBytes, static_cast<cudaStream_t>(stream_)>>>(kp);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

}  // namespace b2g
