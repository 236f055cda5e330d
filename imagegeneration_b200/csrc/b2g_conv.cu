// b2g_conv.cu — persistent, warp-specialised implicit-GEMM convolution for sm_100a.
//
//   warp 0      : TMA producer  (one lane) — per k-step one 4-D box of the NHWC activation (128 pixels x 64 ch,
//                 shifted by the tap offset, zero-filled outside the image) + one 2-D box of packed weights
//   warp 1      : TMEM allocator + MMA issuer (one lane) — tcgen05.mma kind::f16, M=128, N=BLOCK_N, K=16,
//                 fp32 accumulators double-buffered in TMEM so the epilogue of tile i overlaps tile i+1
//   warps 2..5  : epilogue — tcgen05.ld, bias / activation / dropout / activation-derivative, bf16 NHWC stores
//
// Replaces the cuDNN convolutions TensorFlow dispatches for sndcgan/SNDCGAN.py:36-53 (Conv2DTranspose),
// :80-118 (Conv2D) and their gradients w.r.t. the input (tape.gradient at :248,:256,:266).
#include "b2g_conv.cuh"
#include "b2g_conv_epi.cuh"
#include "b2g_conv_rows.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) ConvKernelParams {
  CUtensorMap amap[4];
  CUtensorMap bmap;
  ConvPlan plan;
  int cin_chunks;  // Cin / 64
  int n_tiles;     // Cout / BLOCK_N
  EpiParams e;     // fused epilogue (b2g_conv_epi.cuh)
  int brows;       // rows of the packed weight matrix per phase
};

template <int BLOCK_N>
struct ConvCfg {
  static constexpr int kTileM = 128;
  static constexpr int kABytes = kTileM * 128;
  static constexpr int kBBytes = BLOCK_N * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = BLOCK_N == 256 ? 4 : (BLOCK_N == 128 ? 6 : 8);
  // accumulator buffers in TMEM: 2 for the wide tiles; 4 for N <= 64, whose tiles have only 1-2 column chunks, so that
  // the 4 epilogue groups work on 2-4 consecutive tiles at once (a K1S1 tile is one MMA step: epilogue-latency bound)
  static constexpr int kAccBufs = BLOCK_N <= 64 ? 4 : 2;
  static constexpr int kTmemCols = kAccBufs * BLOCK_N;  // 128 / 256 / 256 / 512: powers of two >= 32
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = kStages * kStageBytes + kBarBytes + 1024;  // +1024: manual alignment
};

// epilogue warp groups: each group is 4 warps (one per TMEM lane quarter); groups interleave over the 32-column
// chunks of a tile so that several chunks are in flight (the epilogue is latency-bound with one warp per SMSP)
template <int BLOCK_N>
constexpr int epi_groups() { return 4; }
template <int BLOCK_N>
constexpr int conv_threads() { return 64 + 128 * epi_groups<BLOCK_N>(); }

// kTf32: fp32 activations / weights (32 channels per 128-byte k-step row), tcgen05.mma kind::tf32 (K = 8 per
// instruction = the same 32 bytes as K = 16 bf16, so tiles, swizzle and descriptors are byte-identical).
template <int BLOCK_N, bool kTf32 = false>
__global__ void __launch_bounds__(conv_threads<BLOCK_N>(), 1)
    conv_igemm_kernel(const __grid_constant__ ConvKernelParams p) {
  constexpr int kChunk = kTf32 ? 32 : 64;  // channels per k-step
  using Cfg = ConvCfg<BLOCK_N>;
  constexpr int kEpiGroups = epi_groups<BLOCK_N>();
  constexpr int kChunks = BLOCK_N / 32;
  constexpr int kGroupsPerTile = kChunks < kEpiGroups ? kChunks : kEpiGroups;  // groups that work on one tile
  __shared__ float sbias[512];
  __shared__ float s_colsum[512];
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + Cfg::kAccBufs;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + Cfg::kAccBufs);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const ConvPlan& pl = p.plan;
  const int total_tiles = pl.phases * pl.m_tiles * p.n_tiles;
  const int k_steps = pl.taps * p.cin_chunks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < Cfg::kAccBufs; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4 * kGroupsPerTile);
    }
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < pl.n_amaps; ++i) tma_prefetch_desc(&p.amap[i]);
    tma_prefetch_desc(&p.bmap);
  }
  if (p.e.bias) {
    const int nb = p.e.out_f32 ? p.e.cs : pl.Cout;
    for (int i = threadIdx.x; i < nb && i < 512; i += blockDim.x) sbias[i] = p.e.bias[i] * p.e.pre_scale;
  }
  epi_colsum_zero(p.e, s_colsum, pl.Cout, sbias);
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer (whole warp converged; one elected lane issues)
    const uint32_t leader = elect_one_sync();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const TileCoord tc = decode_tile(pl, p.n_tiles, tile);
      const int w0 = tc.tw * pl.bw, h0 = tc.th * pl.bh, n0 = tc.tn * pl.bn;
      const int brow = tc.phase * p.brows + tc.n_tile * BLOCK_N;
      int tap = 0, chunk = 0;
      for (int ks = 0; ks < k_steps; ++ks) {
        const int e = tc.phase * pl.taps + tap;
        const int amap = pl.tap_map[e], dx = pl.tap_dx[e], dy = pl.tap_dy[e];
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (leader) {
          uint8_t* a_dst = smem + stage * Cfg::kStageBytes;
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          tma_load_4d(a_dst, &p.amap[amap], &full_bar[stage], chunk * kChunk, w0 + dx, h0 + dy, n0);
          tma_load_2d(a_dst + Cfg::kABytes, &p.bmap, &full_bar[stage], ks * kChunk, brow);
        }
        __syncwarp();
        if (++chunk == p.cin_chunks) { chunk = 0; ++tap; }
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer (whole warp converged; one elected lane issues)
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*bf16 = 1, tf32 = 2*/ kTf32 ? 2 : 1, 0, 0, 128, BLOCK_N);
    const uint32_t smem_base = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int as = it % Cfg::kAccBufs;
      const uint32_t aphase = (it / Cfg::kAccBufs) & 1;
      mbar_wait(&tempty_bar[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BLOCK_N;
      for (int ks = 0; ks < k_steps; ++ks) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a_addr = smem_base + stage * Cfg::kStageBytes;
        const uint64_t adesc = umma_desc_kmajor(a_addr);
        const uint64_t bdesc = umma_desc_kmajor(a_addr + Cfg::kABytes);
        if (leader) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if constexpr (kTf32) umma_tf32(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (ks | k) != 0);
            else umma_f16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (ks | k) != 0);
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
    // ------------------------------------------------ epilogue (warp w owns TMEM lanes 32*(w%4)..+31)
    const EpiParams& e = p.e;                 // stays in the constant bank
    const EpiSeed sd = epi_resolve_seed(p.e);  // dropout seed possibly read from device memory
    const int q = warp & 3;
    const int grp = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const int wl = row % pl.bw;
    const int hl = (row / pl.bw) % pl.bh;
    const int nl = row / (pl.bw * pl.bh);
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int as = it % Cfg::kAccBufs;
      const uint32_t aphase = (it / Cfg::kAccBufs) & 1;
      // wide tiles: every group takes the chunks c = grp (mod 4); narrow tiles (1-2 chunks): chunk c of tile `it`
      // belongs to group (it * kChunks + c) mod 4, so consecutive tiles are drained by different groups
      int c_first = grp;
      if (kChunks < kEpiGroups) {
        c_first = (grp - it * kChunks) & (kEpiGroups - 1);
        if (c_first >= kChunks) continue;
      }
      const TileCoord tc = decode_tile(pl, p.n_tiles, tile);
      const int gw = tc.tw * pl.bw + wl, gh = tc.th * pl.bh + hl, gn = tc.tn * pl.bn + nl;
      const bool valid = (gh < pl.Hg) && (gn < pl.N);
      const int oy = gh * pl.os + pl.ph_y[tc.phase], ox = gw * pl.os + pl.ph_x[tc.phase];
      const size_t pix = (static_cast<size_t>(gn) * pl.Hout + oy) * pl.Wout + ox;
      const int col0 = tc.n_tile * BLOCK_N;
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
#pragma unroll 1
      for (int c = c_first; c < kChunks; c += kEpiGroups) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + c * 32, r);
        tmem_ld_wait();
        conv_epilogue_chunk(e, sbias, r, pix, col0 + c * 32, valid, s_colsum, sd);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  epi_colsum_flush(p.e, s_colsum, pl.Cout, sbias);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------------------
template <int BLOCK_N, bool kTf32 = false>
static int launch_conv(const ConvKernelParams& kp, int total_tiles, cudaStream_t stream) {
  using Cfg = ConvCfg<BLOCK_N>;
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(conv_igemm_kernel<BLOCK_N, kTf32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::kSmemBytes));
    attr_set = true;
  }
  int grid = total_tiles < num_sms() ? total_tiles : num_sms();
  conv_igemm_kernel<BLOCK_N, kTf32><<<grid, conv_threads<BLOCK_N>(), Cfg::kSmemBytes, stream>>>(kp);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

}  // namespace b2g

using namespace b2g;

extern "C" int b2g_conv_igemm(const b2g_conv_args* a, b2g_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a || !a->in || !a->wpacked || !a->out) return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: null pointer");
  if (a->kind == B2G_CONV_K3S1_C3) return launch_conv_c3(a, stream_);
  const int tf32 = a->tf32 ? 1 : 0;
  const int kchunk = tf32 ? 32 : 64;  // channels per 128-byte k-step row
  if (a->Cin % kchunk || a->Cin <= 0)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: Cin=%d not a multiple of %d", a->Cin, kchunk);
  if (tf32 && (a->out_f32 || a->kind == B2G_CONV_K1S1))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: the tf32 path covers K3S1 / K4S2 / CONVT with full-width output");
  if (a->out_f32) {
    if (a->Cout != 32 || a->cout_store < 1 || a->cout_store > 8 || a->kind == B2G_CONVT_K4S2)
      return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: fp32 narrow output needs Cout=32, 1<=cout_store<=8, 1 phase");
    if (a->dropout_thresh8 || (a->act != B2G_ACT_NONE && a->act != B2G_ACT_TANH) ||
        (a->dact != B2G_ACT_NONE && a->dact != B2G_ACT_TANH))
      return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: fp32 narrow output supports bias/tanh/tanh' epilogues only");
  } else if (a->Cout % 64 || a->Cout <= 0) {
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: Cout=%d not a multiple of 64", a->Cout);
  }
  if (a->dact != B2G_ACT_NONE && !a->yprev) return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dact without yprev");
  if (a->Cout > 512) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: Cout=%d > 512", a->Cout);
  if (a->dropout_thresh8 < 0 || a->dropout_thresh8 > 255) return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dropout_thresh8 out of range");
  if (a->colsum && (a->out_f32 || a->tf32 || a->bias || (a->dact == B2G_ACT_NONE && !a->colsumsq)))
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: colsum needs a bf16 backward epilogue (dact, no bias)");
  if (a->colsumsq && (!a->colsum || a->dropout_thresh8))
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: colsumsq needs colsum, no bias and no dropout (BatchNorm statistics)");
  if (a->dropout_thresh8 > 0 && a->act == B2G_ACT_TANH)  // the dropout scale is applied before the activation
    return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: dropout needs a positively homogeneous activation (none/leaky/relu)");

  if (a->out_f32 && !tf32) {  // 64 -> 3 channels on 128/256-pixel rows: per-pixel GEMM + gather
    bool handled = false;
    const int rc = launch_conv_o3(a, stream_, &handled);
    if (rc != B2G_OK || handled) return rc;
  }
  ConvKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  if (!build_conv_plan(a->kind, a->N, a->Hin, a->Win, a->Cin, a->Cout, 128, &kp.plan))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: kind %d cannot tile N=%d H=%d W=%d", a->kind, a->N, a->Hin, a->Win);
  const ConvPlan& pl = kp.plan;
  const int block_n = a->out_f32 ? 32 : (a->Cout % 256 == 0 ? 256 : (a->Cout % 128 == 0 ? 128 : 64));
  if (!tf32) {  // wide layers of the model (>= 128 pixels per row, Cout <= 128): row-slab kernel, compile-time schedule
    bool handled = false;
    const int rc = launch_conv_rows(a, pl, block_n, stream_, &handled, /*static_only=*/true);
    if (rc != B2G_OK || handled) return rc;
  }
  if (!tf32) {  // CTA-pair tiles (cta_group::2) whenever the number of pixel tiles is even
    bool handled = false;
    const int rc = launch_conv_pair(a, pl, block_n, stream_, &handled);
    if (rc != B2G_OK || handled) return rc;
  }
  if (!tf32) {  // wide layers with Cout <= 128: row-slab kernel (b2g_conv_rows.cu) re-uses each input row across taps
    bool handled = false;
    const int rc = launch_conv_rows(a, pl, block_n, stream_, &handled, /*static_only=*/false);
    if (rc != B2G_OK || handled) return rc;
  }
  fill_epi_params(a, &kp.e);
  kp.brows = a->out_f32 ? a->cout_store : a->Cout;
  kp.cin_chunks = a->Cin / kchunk;
  kp.n_tiles = a->Cout / block_n;

  // A tensor maps (innermost first: C, W, H, N)
  const size_t es = tf32 ? 4 : 2;
  const CUtensorMapDataType dt = tf32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const char* base = static_cast<const char*>(a->in);
  const uint32_t box[4] = {(uint32_t)kchunk, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bn};
  if (pl.n_amaps == 1) {
    const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->Win, (uint64_t)a->Hin, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Cin * es, (uint64_t)a->Win * a->Cin * es, (uint64_t)a->Hin * a->Win * a->Cin * es};
    int rc = make_tmap(&kp.amap[0], dt, 4, base, dims, str, box);
    if (rc) return rc;
  } else {
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->Win / 2, (uint64_t)a->Hin / 2, (uint64_t)a->N};
        const uint64_t str[3] = {(uint64_t)2 * a->Cin * es, (uint64_t)2 * a->Win * a->Cin * es,
                                 (uint64_t)a->Hin * a->Win * a->Cin * es};
        const char* b = base + ((size_t)py * a->Win + px) * a->Cin * es;
        int rc = make_tmap(&kp.amap[py * 2 + px], dt, 4, b, dims, str, box);
        if (rc) return rc;
      }
  }
  {
    const uint64_t ktot = (uint64_t)pl.taps * a->Cin;
    const uint64_t dims[2] = {ktot, (uint64_t)pl.phases * kp.brows};
    const uint64_t str[1] = {ktot * es};
    const uint32_t bbox[2] = {(uint32_t)kchunk, (uint32_t)block_n};
    int rc = make_tmap(&kp.bmap, dt, 2, a->wpacked, dims, str, bbox);
    if (rc) return rc;
  }
  const int total_tiles = pl.phases * pl.m_tiles * kp.n_tiles;
  if (tf32) {
    switch (block_n) {
      case 256: return launch_conv<256, true>(kp, total_tiles, stream);
      case 128: return launch_conv<128, true>(kp, total_tiles, stream);
      default: return launch_conv<64, true>(kp, total_tiles, stream);
    }
  }
  switch (block_n) {
    case 256: return launch_conv<256>(kp, total_tiles, stream);
    case 128: return launch_conv<128>(kp, total_tiles, stream);
    case 32: return launch_conv<32>(kp, total_tiles, stream);
    default: return launch_conv<64>(kp, total_tiles, stream);
  }
}
