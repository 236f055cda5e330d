// b2g_conv_pair.cu — CTA-pair (tcgen05 cta_group::2) implicit-GEMM convolution for sm_100a.
//
// Two CTAs of a thread-block cluster (the two SMs of a TPC) compute ONE 256-pixel x BLOCK_N tile per step:
// CTA r owns pixel tile 2*pair + r (its 128 A rows and the 128 TMEM lanes of its accumulator) and HALF of the weight
// tile (BLOCK_N/2 rows); the leader CTA issues tcgen05.mma.cta_group::2 (M = 256) for both.  Against the one-CTA
// kernel (b2g_conv.cu) every SM writes and reads half as many weight bytes in shared memory per MMA — the 128 B/clk
// shared-memory port is what bounds the 128 x N tiles (N = 256: 192 B/clk needed at the nominal MMA rate, N = 128:
// 256 B/clk) — and one issuing thread feeds two tensor cores.
//
//   warp 0 (both CTAs)   : TMA producer — own A box + own half of the weight box per k-step; completion bytes of BOTH
//                          CTAs land on the LEADER's full barrier (cp.async.bulk.tensor.cta_group::2)
//   warp 1 (leader only) : MMA issuer; tcgen05.commit multicast releases the smem stage / publishes the accumulator
//                          in both CTAs
//   warps 2..17          : epilogue of the CTA's own 128 pixels (b2g_conv_epi.cuh); accumulator release = remote
//                          mbarrier arrive on the leader's tempty barrier
//
// Same roles / arguments as conv_igemm_kernel; replaces the cuDNN convolutions behind sndcgan/SNDCGAN.py:36-53,
// :80-118 and their input gradients.
#include <stdlib.h>
#include <string.h>

#include "b2g_conv.cuh"
#include "b2g_conv_epi.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) PairKernelParams {
  CUtensorMap amap[4];
  CUtensorMap bmap;  // box = 64 k x BLOCK_N/2 rows
  ConvPlan plan;
  EpiParams e;
  int cin_chunks;   // Cin / 64
  int n_tiles;      // Cout / BLOCK_N
  EpiTmaMaps em;    // TMA epilogue (used when e.tma)
  int brows;        // rows of the packed weight matrix per phase
  int arrive_release_cluster;  // measurement switch: 1 = the old .release.cluster remote arrive
  int m_pairs;      // plan.m_tiles / 2
  int total_pairs;  // phases * m_pairs * n_tiles
  FastDiv fd_nt, fd_tw, fd_th;  // tile decoding without divisions (n_tiles, tiles_w, tiles_h)
};

// BOXES: N = 256 only — give up one pipeline stage for the TMA epilogue's boxes.  Used for the BACKWARD roles of the
// >= 256-channel layers: their direct epilogue loads the saved output with one 64-byte global load per lane and pixel row
// (uncoalesced, a fifth of the stall samples on the first use of what it loaded) and sums columns with a 31-shuffle
// butterfly per chunk (ncu round 2: D.conv6 input gradient 134 us at 66 % tensor-pipe activity against 117 us / 88 % for a
// forward role of the same size).
template <int BLOCK_N, bool BOXES = false>
struct PairCfg {
  static constexpr int kABytes = 128 * 128;
  static constexpr int kBBytes = (BLOCK_N / 2) * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  // N <= 128 tiles: the TMA epilogue's per-warp boxes (b2g_conv_epi.cuh) take the room of one pipeline stage
  static constexpr bool kEpiTma = BLOCK_N <= 128 || BOXES;
  static constexpr int kStages = BLOCK_N == 256 ? (BOXES ? 5 : 6) : (BLOCK_N == 128 ? 7 : 8);
  // N = 128: four accumulators since round 2 (TMEM has the columns).  A same-box A/B against two buffers measured NO
  // difference (G.convT2 forward 152.6 us, D.conv4 input gradient 164.9 us both ways) — an apparent 10 % gain in an earlier
  // run was the difference between two boxes of the pool; kept because it costs nothing.
  static constexpr int kAccBufs = BLOCK_N <= 128 ? 4 : 2;  // N = 128: four accumulators since round 2 (two before; TMEM has the columns)
  static constexpr int kTmemCols = kAccBufs * BLOCK_N;
  // 16 epilogue warps; N <= 128 (TMA epilogue): one 2 KiB box per warp (b2g_conv_epi.cuh).  Measured in round 2: a second
  // box (paid for with one pipeline stage) changed nothing here (D.conv4 dgrad 181 -> 187 us; again neutral after the
  // four accumulator buffers and the cheap remote arrive: 162 -> 161 us), 8 warps with 3-box rings
  // were 15-45 % slower — the backward epilogue is bound by per-warp instruction latency, not by the box hand-over
  static constexpr int kEpiGroups = 4;
  static constexpr int kEpiBoxes = kEpiTma ? 1 : 0;
  static constexpr int kThreads = 64 + 128 * kEpiGroups;
  static constexpr int kEpiWarps = 4 * kEpiGroups;
  static constexpr int kEpiBytes = kEpiTma ? epi_tma_smem_bytes(kEpiWarps, kEpiBoxes) : 0;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 256 + 1024;
  static_assert(kSmemBytes + 4096 <= 227 * 1024, "shared memory budget (incl. the static bias / column-sum tables)");
};

struct PairCoord {
  int phase, n_tile, tn, th, tw;
};
__device__ __forceinline__ PairCoord decode_pair(const PairKernelParams& p, int pair, int rank) {
  PairCoord c;
  int rest;
  fast_divmod(pair, p.fd_nt, rest, c.n_tile);
  c.phase = 0;  // rest / m_pairs is the phase (0..3): comparisons, m_pairs is too large for the multiply-high bound
  while (rest >= p.m_pairs) { rest -= p.m_pairs; ++c.phase; }
  const int m = 2 * rest + rank;
  int r2;
  fast_divmod(m, p.fd_tw, r2, c.tw);
  fast_divmod(r2, p.fd_th, c.tn, c.th);
  return c;
}

template <int BLOCK_N, bool BOXES = false>
__global__ void __launch_bounds__(PairCfg<BLOCK_N, BOXES>::kThreads, 1)
    conv_pair_kernel(const __grid_constant__ PairKernelParams p) {
  using Cfg = PairCfg<BLOCK_N, BOXES>;
  constexpr int kChunks = BLOCK_N / 32;
  constexpr int kGroupsPerTile = kChunks < Cfg::kEpiGroups ? kChunks : Cfg::kEpiGroups;
  __shared__ float sbias[512];
  __shared__ float s_colsum[512];
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epism = smem + Cfg::kStages * Cfg::kStageBytes;  // stage sizes are multiples of 1 KiB
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epism + Cfg::kEpiBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + Cfg::kAccBufs;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + Cfg::kAccBufs);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = MMA leader
  const ConvPlan& pl = p.plan;
  const int k_steps = pl.taps * p.cin_chunks;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);   // leader's producer: one arrive.expect_tx covering the bytes of both CTAs
      mbar_init(&empty_bar[s], 1);  // multicast tcgen05.commit
    }
    for (int a = 0; a < Cfg::kAccBufs; ++a) {
      mbar_init(&tfull_bar[a], 1);                         // multicast tcgen05.commit
      mbar_init(&tempty_bar[a], 2 * 4 * kGroupsPerTile);   // epilogue warps of BOTH CTAs (leader's copy is used)
    }
    if (Cfg::kEpiTma && p.e.tma) epi_tma_init(epism, Cfg::kEpiWarps, Cfg::kEpiBoxes);
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
    for (int pair = cluster_id; pair < p.total_pairs; pair += n_clusters) {
      const PairCoord tc = decode_pair(p, pair, rank);
      const int w0 = tc.tw * pl.bw, h0 = tc.th * pl.bh, n0 = tc.tn * pl.bn;
      const int brow = tc.phase * p.brows + tc.n_tile * BLOCK_N + rank * (BLOCK_N / 2);
      int tap = 0, chunk = 0;
      for (int ks = 0; ks < k_steps; ++ks) {
        const int e = tc.phase * pl.taps + tap;
        const int amap = pl.tap_map[e], dx = pl.tap_dx[e], dy = pl.tap_dy[e];
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (leader) {
          uint8_t* a_dst = smem + stage * Cfg::kStageBytes;
          const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);  // the MMA leader's full barrier
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
          tma_load_4d_pair(a_dst, &p.amap[amap], bar, chunk * 64, w0 + dx, h0 + dy, n0);
          tma_load_2d_pair(a_dst + Cfg::kABytes, &p.bmap, bar, ks * 64, brow);
        }
        __syncwarp();
        if (++chunk == p.cin_chunks) { chunk = 0; ++tap; }
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer (leader CTA only)
    if (rank == 0) {
      const uint32_t leader = elect_one_sync();
      constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, 0, 0, 256, BLOCK_N);
      const uint32_t smem_base = smem_u32(smem);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int pair = cluster_id; pair < p.total_pairs; pair += n_clusters, ++it) {
        const int as = it % Cfg::kAccBufs;
        mbar_wait(&tempty_bar[as], ((it / Cfg::kAccBufs) & 1) ^ 1);
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
            for (int k = 0; k < 4; ++k) umma_f16_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (ks | k) != 0);
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
    // ------------------------------------------------ epilogue (warp w owns TMEM lanes 32*(w%4)..+31)
    const EpiParams& e = p.e;                 // stays in the constant bank
    const EpiSeed sd = epi_resolve_seed(p.e);  // dropout seed possibly read from device memory
    const int q = warp & 3;
    const int grp = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const int wl = row % pl.bw;
    const int hl = (row / pl.bw) % pl.bh;
    const int nl = row / (pl.bw * pl.bh);
    const uint32_t tempty_leader = mapa_shared(smem_u32(&tempty_bar[0]), 0);
    // origin of this warp's 32-pixel box inside a tile (TMA epilogue)
    const int wl0 = (q * 32) % pl.bw, hl0 = ((q * 32) / pl.bw) % pl.bh, nl0 = (q * 32) / (pl.bw * pl.bh);
    constexpr int NB = Cfg::kEpiBoxes > 0 ? Cfg::kEpiBoxes : 1;
    EpiWarpStageN<NB> st = epi_tma_stage<NB>(epism, Cfg::kEpiWarps, warp - 2);
    const bool use_tma = Cfg::kEpiTma && p.e.tma != 0;
    // Column sums are read back from the warp's box, which also holds the pixels of a ragged tile that lie OUTSIDE the
    // tensor (the TMA store clips them).  Their accumulators are not zero (the taps of a pixel just outside reach into the
    // image); with act'(0) = 0 the zero-filled saved output masks them, with act'(0) != 0 (no-dropout layers, WGAN) the
    // accumulators are cleared instead.
    const bool zero_oob = e.colsum != nullptr && (e.dact == B2G_ACT_NONE || e.dact_zero != 0.f);  // forward statistics too
    // first chunk of tile number `i` (counted per cluster) that belongs to this warp's group; >= kChunks: none
    auto first_chunk = [&](int i) {
      return kChunks < Cfg::kEpiGroups ? ((grp - i * kChunks) & (Cfg::kEpiGroups - 1)) : grp;
    };
    // begin pointer of the TMA epilogue (see b2g_conv_rows.cu): the next item (pair, chunk) that has not been begun
    int b_pair = cluster_id, b_it = 0, b_c = first_chunk(0);
    int c_pair = -1;   // the pair the main loop decoded last: begin() of the same pair reuses its coordinates (the
    PairCoord c_tc{};  // multiply-high divisions are a dependent chain the epilogue warps stall on, ncu round 2)
    auto ensure_begun = [&](uint32_t n) {
      if (!use_tma) return;
      while (st.k_begin - st.k_chunk < n) {
        while (b_pair < p.total_pairs && b_c >= kChunks) {
          b_pair += n_clusters;
          ++b_it;
          b_c = first_chunk(b_it);
        }
        if (b_pair >= p.total_pairs) return;
        const PairCoord bt = b_pair == c_pair ? c_tc : decode_pair(p, b_pair, rank);
        epi_tma_begin(e, &p.em.ymap[bt.phase], st, bt.n_tile * BLOCK_N + b_c * 32, bt.tw * pl.bw + wl0,
                      bt.th * pl.bh + hl0, bt.tn * pl.bn + nl0);
        b_c += Cfg::kEpiGroups;
      }
    };
    int it = 0;
    for (int pair = cluster_id; pair < p.total_pairs; pair += n_clusters, ++it) {
      const int as = it % Cfg::kAccBufs;
      const int c_first = first_chunk(it);
      if (c_first >= kChunks) continue;
      const PairCoord tc = decode_pair(p, pair, rank);
      c_pair = pair;
      c_tc = tc;
      const int gw = tc.tw * pl.bw + wl, gh = tc.th * pl.bh + hl, gn = tc.tn * pl.bn + nl;
      const bool valid = (gh < pl.Hg) && (gn < pl.N);
      const int oy = gh * pl.os + pl.ph_y[tc.phase], ox = gw * pl.os + pl.ph_x[tc.phase];
      const size_t pix = (static_cast<size_t>(gn) * pl.Hout + oy) * pl.Wout + ox;
      const int col0 = tc.n_tile * BLOCK_N;
      const int bx = tc.tw * pl.bw + wl0, by = tc.th * pl.bh + hl0, bn = tc.tn * pl.bn + nl0;
      ensure_begun(1);
      mbar_wait_parked(&tfull_bar[as], (it / Cfg::kAccBufs) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = c_first; c < kChunks; c += Cfg::kEpiGroups) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + c * 32, r);
        ensure_begun(NB);
        tmem_ld_wait();
        if (use_tma && zero_oob && !valid) {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0u;
        }
        if (use_tma) epi_tma_chunk(e, &p.em.omap[tc.phase], st, sbias, r, pix, col0 + c * 32, lane, bx, by, bn, s_colsum, sd);
        else conv_epilogue_chunk(e, sbias, r, pix, col0 + c * 32, valid, s_colsum, sd);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (p.arrive_release_cluster) mbar_arrive_cluster_release(tempty_leader + as * 8);
        else mbar_arrive_cluster(tempty_leader + as * 8);
      }
    }
    if (use_tma) epi_tma_end(e, st, s_colsum, lane);
  }

  tc_fence_before();
  __syncthreads();
  epi_colsum_flush(p.e, s_colsum, pl.Cout, sbias);
  cluster_sync_all();  // no CTA frees TMEM / exits while its peer may still signal or read it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
  }
}

template <int BLOCK_N, bool BOXES = false>
static int launch_pair(const PairKernelParams& kp, cudaStream_t stream) {
  using Cfg = PairCfg<BLOCK_N, BOXES>;
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(conv_pair_kernel<BLOCK_N, BOXES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::kSmemBytes));
    attr_set = true;
  }
  int clusters = num_sms() / 2;
  if (kp.total_pairs < clusters) clusters = kp.total_pairs;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(2 * clusters, 1, 1);
  cfg.blockDim = dim3(Cfg::kThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B2G_CUDA_CHECK(cudaLaunchKernelEx(&cfg, conv_pair_kernel<BLOCK_N, BOXES>, kp));
  return B2G_OK;
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

// Launches the CTA-pair kernel when the role fits it (even number of pixel tiles, BLOCK_N in {64, 128, 256});
// *handled tells the caller whether it did.  B2G_CONV_PAIR=0 disables it (tests compare both kernels).
int launch_conv_pair(const b2g_conv_args* a, const ConvPlan& pl, int block_n, void* stream_, bool* handled) {
  *handled = false;
  if (!env_int("B2G_CONV_PAIR", 1)) return B2G_OK;
  if (block_n != 64 && block_n != 128 && block_n != 256) return B2G_OK;
  if (a->out_f32 || (pl.m_tiles & 1)) return B2G_OK;
  PairKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  kp.plan = pl;
  // backward roles at N = 256 take the TMA epilogue (one stage fewer); B2G_PAIR_BOXES256=0 keeps the direct one
  const int boxes_env = env_int("B2G_PAIR_BOXES256", 1);  // 2: forward roles too — measured neutral (conv / wgrad launches
                                                          // of a step 7.76 vs 7.78 ms): their direct stores hide behind the MMAs
  const bool boxes256 = block_n == 256 && !a->tf32 && !a->out_f32 && !a->colsumsq &&
                        (boxes_env == 2 || (boxes_env == 1 && a->dact != B2G_ACT_NONE));
  fill_epi_params(a, &kp.e);
  {
    int rc = fill_epi_tma(a, pl, block_n <= 128 || boxes256, &kp.e, &kp.em);
    if (rc) return rc;
  }
  kp.arrive_release_cluster = env_int("B2G_PAIR_ARRIVE_RELEASE_CLUSTER", 0);
  kp.brows = a->Cout;
  kp.cin_chunks = a->Cin / 64;
  kp.n_tiles = a->Cout / block_n;
  kp.m_pairs = pl.m_tiles / 2;
  kp.total_pairs = pl.phases * kp.m_pairs * kp.n_tiles;
  kp.fd_nt = make_fastdiv(kp.n_tiles, kp.total_pairs);
  kp.fd_tw = make_fastdiv(pl.tiles_w, pl.m_tiles);
  kp.fd_th = make_fastdiv(pl.tiles_h, pl.m_tiles);

  const size_t es = 2;
  const char* base = static_cast<const char*>(a->in);
  const uint32_t box[4] = {64u, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bn};
  if (pl.n_amaps == 1) {
    const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->Win, (uint64_t)a->Hin, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Cin * es, (uint64_t)a->Win * a->Cin * es,
                             (uint64_t)a->Hin * a->Win * a->Cin * es};
    int rc = make_tmap(&kp.amap[0], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, dims, str, box);
    if (rc) return rc;
  } else {
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->Win / 2, (uint64_t)a->Hin / 2, (uint64_t)a->N};
        const uint64_t str[3] = {(uint64_t)2 * a->Cin * es, (uint64_t)2 * a->Win * a->Cin * es,
                                 (uint64_t)a->Hin * a->Win * a->Cin * es};
        const char* b = base + ((size_t)py * a->Win + px) * a->Cin * es;
        int rc = make_tmap(&kp.amap[py * 2 + px], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, b, dims, str, box);
        if (rc) return rc;
      }
  }
  {
    const uint64_t ktot = (uint64_t)pl.taps * a->Cin;
    const uint64_t dims[2] = {ktot, (uint64_t)pl.phases * kp.brows};
    const uint64_t str[1] = {ktot * es};
    const uint32_t bbox[2] = {64u, (uint32_t)block_n / 2};
    int rc = make_tmap(&kp.bmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a->wpacked, dims, str, bbox);
    if (rc) return rc;
  }
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int rc;
  switch (block_n) {
    case 256: rc = (boxes256 && kp.e.tma) ? launch_pair<256, true>(kp, stream) : launch_pair<256>(kp, stream); break;
    case 128: rc = launch_pair<128>(kp, stream); break;
    default: rc = launch_pair<64>(kp, stream); break;
  }
  if (rc == B2G_OK) *handled = true;
  return rc;
}

}  // namespace b2g
