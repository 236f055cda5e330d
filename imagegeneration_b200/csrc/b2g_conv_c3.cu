// b2g_conv_c3.cu — 3x3 stride-1 SAME convolution over a 3-channel fp32 tensor (3 -> 64 channels) on tcgen05, with
// the im2col operand built IN SHARED MEMORY instead of in HBM.
//
//   warp 0      : loads the packed [64][64] weight tile once (TMA; resident for the whole kernel), then streams the
//                 fp32 input: per 128-pixel tile ONE TMA box = the tile's pixels plus a one-pixel halo
//                 ((bw+8)*3 floats x (bh+2) rows x bn images; out-of-image parts arrive as zeros = SAME padding;
//                 the box starts 4 pixels left of the tile because TMA wants a 16-byte aligned start offset)
//                 through an 8-stage ring — the HBM/L2 latency is hidden by the ring, not by the builder threads
//   warp 1      : TMEM allocator + MMA issuer — one K=32 step (two K=16 MMAs, M=128, N=64) per 128-pixel tile
//   warps 2..5  : patch builders, one pixel per thread: 27 conflict-free LDS.32 from the staged box + a ones
//                 column -> 32 bf16 = 4 x 16 B written at the SW128-swizzled position of the pixel's row,
//                 fence.proxy.async, mbarrier arrive.  This is the im2col27 tensor of b2g_edge.cu, never leaving
//                 the SM.
//   warps 6..21 : epilogue (4 groups x 4 TMEM lane quarters, 4 accumulator buffers): bias / LeakyReLU / dropout or
//                 ReLU' from the saved output, bf16 NHWC stores (b2g_conv_epi.cuh)
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
  EpiTmaMaps em;     // TMA epilogue (used when e.tma)
  int in_box_bytes;  // bytes one input box delivers
};

struct C3Cfg {
  static constexpr int kBlockN = 64;
  static constexpr int kABytes = 128 * 128;  // 128 pixel rows of 128 B; only the first 64 B (K = 32) are used
  static constexpr int kStages = 6;
  static constexpr int kInStages = 8;           // staged fp32 input boxes
  static constexpr int kInStageBytes = 8192;    // upper bound of one box (128-px row tile: 136 * 12 * 3 = 4896)
  static constexpr int kBBytes = kBlockN * 128;
  static constexpr int kAccBufs = 4;
  static constexpr int kTmemCols = kAccBufs * kBlockN;
  static constexpr int kBuilderWarps = 4;
  static constexpr int kEpiGroups = 4;
  static constexpr int kThreads = 64 + 32 * kBuilderWarps + 128 * kEpiGroups;
  static constexpr int kMainBytes = kStages * kABytes + kBBytes + kInStages * kInStageBytes;  // multiple of 1 KiB
  static constexpr int kEpiWarps = 4 * kEpiGroups;
  static constexpr int kEpiBoxes = 1;           // boxes per epilogue warp (b2g_conv_epi.cuh); a second one measured neutral
  static constexpr int kEpiBytes = epi_tma_smem_bytes(kEpiWarps, kEpiBoxes);
  static constexpr int kSmemBytes = kMainBytes + kEpiBytes + 512 + 1024;
};

__global__ void __launch_bounds__(C3Cfg::kThreads, 1) conv_c3_kernel(const __grid_constant__ C3KernelParams p) {
  using Cfg = C3Cfg;
  constexpr int BLOCK_N = Cfg::kBlockN;
  constexpr int kChunks = BLOCK_N / 32;
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
      mbar_init(&tempty_bar[a], 4 * kChunks);
    }
    mbar_init(bfull_bar, 1);
    for (int s = 0; s < Cfg::kInStages; ++s) {
      mbar_init(&in_full[s], 1);
      mbar_init(&in_empty[s], Cfg::kBuilderWarps);
    }
    if (p.e.tma) epi_tma_init(epism, Cfg::kEpiWarps, Cfg::kEpiBoxes);
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
    // ------------------------------------------------ patch builders: thread m owns pixel row m of the A tile
    const int m = threadIdx.x - 64;
    const int wl = m % pl.bw, hl = (m / pl.bw) % pl.bh, nl = m / (pl.bw * pl.bh);
    const uint32_t row_off = m * 128, sw = m & 7;
    // this thread's window inside a staged box [bn][bh + 2][(bw + 8) * 3 floats]: rows hl..hl+2, nine consecutive
    // floats from pixel wl + 3 (= image pixel x - 1).  Consecutive lanes are 3 words apart: conflict-free LDS.32.
    const int pitch = (pl.bw + 8) * 3;
    const int win_off = ((nl * (pl.bh + 2) + hl) * (pl.bw + 8) + wl + 3) * 3;
    int stage = 0, istage = 0;
    uint32_t phase = 0, iphase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      float v[32];
      mbar_wait_parked(&in_full[istage], iphase);
      const float* src = reinterpret_cast<const float*>(insm + istage * Cfg::kInStageBytes) + win_off;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int j = 0; j < 9; ++j) v[ky * 9 + j] = src[ky * pitch + j];  // j = kx * 3 + c
      v[27] = 1.f;
#pragma unroll
      for (int k = 28; k < 32; ++k) v[k] = 0.f;
      mbar_wait_parked(&empty_bar[stage], phase ^ 1);
      uint8_t* dst = smem + stage * Cfg::kABytes + row_off;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 o;
        o.x = pack_bf16x2(v[j * 8 + 0], v[j * 8 + 1]);
        o.y = pack_bf16x2(v[j * 8 + 2], v[j * 8 + 3]);
        o.z = pack_bf16x2(v[j * 8 + 4], v[j * 8 + 5]);
        o.w = pack_bf16x2(v[j * 8 + 6], v[j * 8 + 7]);
        *reinterpret_cast<uint4*>(dst + ((j ^ sw) << 4)) = o;  // 128-byte swizzle: 16-byte chunk j of row m -> j ^ (m & 7)
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
    const int grp = (warp - 2 - Cfg::kBuilderWarps) >> 2;
    const int row = q * 32 + lane;
    const int wl = row % pl.bw, hl = (row / pl.bw) % pl.bh, nl = row / (pl.bw * pl.bh);
    // origin of this warp's 32-pixel box inside a tile (TMA epilogue)
    const int wl0 = (q * 32) % pl.bw, hl0 = ((q * 32) / pl.bw) % pl.bh, nl0 = (q * 32) / (pl.bw * pl.bh);
    constexpr int NB = Cfg::kEpiBoxes;
    EpiWarpStageN<NB> st = epi_tma_stage<NB>(epism, Cfg::kEpiWarps, warp - 2 - Cfg::kBuilderWarps);
    const bool use_tma = p.e.tma != 0;
    // Chunk c of tile number i (counted per CTA) belongs to epilogue group (2 * i + c) % 4, i.e. group g owns chunk g & 1 of
    // every other tile, starting with tile number g >> 1: one division-free walk with stride 2 * gridDim.x per warp.
    static_assert(kChunks == 2 && Cfg::kEpiGroups == 4, "chunk-to-group rotation below assumes 2 chunks, 4 groups");
    const int c = grp & 1;
    const int stride = 2 * gridDim.x;
    int it = grp >> 1;
    int tile = blockIdx.x + it * gridDim.x;
    TileWalk tc = tile_walk_begin(pl, tile, stride);
    for (; tile < total_tiles; tile += stride, it += 2, tile_walk_next(pl, tc)) {
      const int as = it % Cfg::kAccBufs;
      const int gw = tc.tw * pl.bw + wl, gh = tc.th * pl.bh + hl, gn = tc.tn * pl.bn + nl;
      const bool valid = (gh < pl.Hg) && (gn < pl.N);
      const size_t pix = (static_cast<size_t>(gn) * pl.Hout + gh) * pl.Wout + gw;
      const int bx = tc.tw * pl.bw + wl0, by = tc.th * pl.bh + hl0, bn = tc.tn * pl.bn + nl0;
      // the saved-output load of this item (backward epilogue) was started right after the previous item (second box);
      // the very first one starts here, before the accumulator wait
      if (use_tma && st.k_begin == st.k_chunk) epi_tma_begin(e, &p.em.ymap[0], st, c * 32, bx, by, bn);
      mbar_wait_parked(&tfull_bar[as], (it / Cfg::kAccBufs) & 1);
      tc_fence_after();
      uint32_t r[32];
      __syncwarp();
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + c * 32, r);
      if (NB >= 2 && use_tma && tile + stride < total_tiles) {
        // (two boxes only) the warp's next item (two tiles on) is begun into the other box now: the store that last used that box is two
        // items old, and the load has until the next accumulator to arrive
        TileWalk nx = tc;
        tile_walk_next(pl, nx);
        epi_tma_begin(e, &p.em.ymap[0], st, c * 32, nx.tw * pl.bw + wl0, nx.th * pl.bh + hl0, nx.tn * pl.bn + nl0);
      }
      tmem_ld_wait();
      if (use_tma) epi_tma_chunk(e, &p.em.omap[0], st, sbias, r, pix, c * 32, lane, bx, by, bn, s_colsum, sd);
      else conv_epilogue_chunk(e, sbias, r, pix, c * 32, valid, s_colsum, sd);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
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
