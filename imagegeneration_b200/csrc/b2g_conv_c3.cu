// b2g_conv_c3.cu — 3x3 stride-1 SAME convolution over a 3-channel fp32 tensor (3 -> 64 channels) on tcgen05, with
// the im2col operand built IN SHARED MEMORY instead of in HBM.
//
//   warp 0      : loads the packed [64][64] weight tile once (TMA); it stays resident for the whole kernel
//   warp 1      : TMEM allocator + MMA issuer — one K=32 step (two K=16 MMAs, M=128, N=64) per 128-pixel tile
//   warps 2..5  : patch builders, one pixel per thread: 27 fp32 loads (zero outside the image) + a ones column ->
//                 32 bf16 = 4 x 16 B written at the SW128-swizzled position of the pixel's row, fence.proxy.async,
//                 mbarrier arrive.  This is the im2col27 tensor of b2g_edge.cu, never leaving the SM.
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
  ConvPlan plan;  // K1S1 geometry: 128-pixel boxes (bw, bh, bn) over the N x H x W grid
  EpiParams e;
  const float* in;  // fp32 [N, H, W, 3]
};

struct C3Cfg {
  static constexpr int kBlockN = 64;
  static constexpr int kABytes = 128 * 128;  // 128 pixel rows of 128 B; only the first 64 B (K = 32) are used
  static constexpr int kStages = 6;
  static constexpr int kBBytes = kBlockN * 128;
  static constexpr int kAccBufs = 4;
  static constexpr int kTmemCols = kAccBufs * kBlockN;
  static constexpr int kBuilderWarps = 4;
  static constexpr int kEpiGroups = 4;
  static constexpr int kThreads = 64 + 32 * kBuilderWarps + 128 * kEpiGroups;
  static constexpr int kSmemBytes = kStages * kABytes + kBBytes + 256 + 1024;
};

__global__ void __launch_bounds__(C3Cfg::kThreads, 1) conv_c3_kernel(const __grid_constant__ C3KernelParams p) {
  using Cfg = C3Cfg;
  constexpr int BLOCK_N = Cfg::kBlockN;
  constexpr int kChunks = BLOCK_N / 32;
  __shared__ float sbias[64];
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* bsm = smem + Cfg::kStages * Cfg::kABytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bsm + Cfg::kBBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + Cfg::kAccBufs;
  uint64_t* bfull_bar = tempty_bar + Cfg::kAccBufs;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bfull_bar + 1);

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
    fence_barrier_init();
  }
  if (p.e.bias && threadIdx.x < BLOCK_N) sbias[threadIdx.x] = p.e.bias[threadIdx.x];
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ weight tile: loaded once, resident
    if (elect_one_sync()) {
      tma_prefetch_desc(&p.bmap);
      mbar_expect_tx(bfull_bar, Cfg::kBBytes);
      tma_load_2d(bsm, &p.bmap, bfull_bar, 0, 0);
    }
    __syncwarp();
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
    const int H = pl.Hg, W = pl.Wg;
    const uint32_t row_off = m * 128, sw = m & 7;
    // gathers the 27 taps*channels of this thread's pixel of `tile` (zeros outside the image, ones column at 27).
    // The loads are volatile asm so that the compiler keeps them BEFORE the mbarrier wait / pack of the previous
    // tile (it otherwise sinks them behind the pack and the whole L2 latency is exposed once per tile).
    auto gather = [&](int tile, float (&v)[28]) {
      const TileCoord tc = decode_tile(pl, 1, tile);
      const int x = tc.tw * pl.bw + wl, y = tc.th * pl.bh + hl, n = tc.tn * pl.bn + nl;
#pragma unroll
      for (int k = 0; k < 28; ++k) v[k] = 0.f;
      if (y < H && n < pl.N) {
        const float* img = p.in + (size_t)n * H * W * 3;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const int yy = y + ky - 1;
          if (yy < 0 || yy >= H) continue;
          const float* rowp = img + ((size_t)yy * W + x) * 3;
#pragma unroll
          for (int j = 0; j < 9; ++j) {  // j = kx * 3 + c: nine consecutive floats starting at pixel x - 1
            const int xx = x + j / 3 - 1;
            if (xx >= 0 && xx < W)
              asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v[ky * 9 + j]) : "l"(rowp + j - 3));
          }
        }
        v[27] = 1.f;
      }
    };
    int stage = 0;
    uint32_t phase = 0;
    // packs one gathered pixel into its swizzled 64 bytes of the A stage and hands the stage to the MMA warp
    auto emit = [&](const float (&v)[28]) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t* dst = smem + stage * Cfg::kABytes + row_off;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint4 o;
        o.x = pack_bf16x2(v[j * 8 + 0], v[j * 8 + 1]);
        o.y = pack_bf16x2(v[j * 8 + 2], v[j * 8 + 3]);
        o.z = j < 3 ? pack_bf16x2(v[j < 3 ? j * 8 + 4 : 0], v[j < 3 ? j * 8 + 5 : 0]) : 0u;  // columns 28..31 = 0
        o.w = j < 3 ? pack_bf16x2(v[j < 3 ? j * 8 + 6 : 0], v[j < 3 ? j * 8 + 7 : 0]) : 0u;
        *reinterpret_cast<uint4*>(dst + ((j ^ sw) << 4)) = o;  // 128-byte swizzle: 16-byte chunk j of row m -> j ^ (m & 7)
      }
      fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_bar[stage]);
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    };
    // software pipeline, two register sets: the next tile's loads are in flight while this tile is packed
    float va[28], vb[28];
    const int step = gridDim.x;
    int tile = blockIdx.x;
    if (tile < total_tiles) gather(tile, va);
    while (tile < total_tiles) {
      if (tile + step < total_tiles) gather(tile + step, vb);
      emit(va);
      tile += step;
      if (tile >= total_tiles) break;
      if (tile + step < total_tiles) gather(tile + step, va);
      emit(vb);
      tile += step;
    }
  } else {
    // ------------------------------------------------ epilogue (warp w owns TMEM lanes 32*(w%4)..+31)
    const int q = warp & 3;
    const int grp = (warp - 2 - Cfg::kBuilderWarps) >> 2;
    const int row = q * 32 + lane;
    const int wl = row % pl.bw, hl = (row / pl.bw) % pl.bh, nl = row / (pl.bw * pl.bh);
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int c = (grp - it * kChunks) & (Cfg::kEpiGroups - 1);  // chunk c of tile `it` belongs to group (2*it + c) % 4
      if (c >= kChunks) continue;
      const int as = it % Cfg::kAccBufs;
      const TileCoord tc = decode_tile(pl, 1, tile);
      const int gw = tc.tw * pl.bw + wl, gh = tc.th * pl.bh + hl, gn = tc.tn * pl.bn + nl;
      const bool valid = (gh < pl.Hg) && (gn < pl.N);
      const size_t pix = (static_cast<size_t>(gn) * pl.Hout + gh) * pl.Wout + gw;
      mbar_wait(&tfull_bar[as], (it / Cfg::kAccBufs) & 1);
      tc_fence_after();
      uint32_t r[32];
      __syncwarp();
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + c * 32, r);
      tmem_ld_wait();
      conv_epilogue_chunk(p.e, sbias, r, pix, c * 32, valid);
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

int launch_conv_c3(const b2g_conv_args* a, void* stream_) {
  if (a->Cin != 3) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm(K3S1_C3): Cin=%d, expected 3", a->Cin);
  if (a->Cout != 64 || a->out_f32)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm(K3S1_C3): Cout=%d (bf16 output) expected 64", a->Cout);
  if (a->dact != B2G_ACT_NONE && !a->yprev) return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dact without yprev");
  if (a->dropout_thresh8 < 0 || a->dropout_thresh8 > 255)
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dropout_thresh8 out of range");
  C3KernelParams kp;
  memset(&kp, 0, sizeof(kp));
  if (!build_conv_plan(a->kind, a->N, a->Hin, a->Win, 64, a->Cout, 128, &kp.plan))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm(K3S1_C3): cannot tile N=%d H=%d W=%d", a->N, a->Hin, a->Win);
  fill_epi_params(a, &kp.e);
  kp.in = static_cast<const float*>(a->in);
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
