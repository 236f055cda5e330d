// b2g_conv_rows.cu — row-slab implicit-GEMM convolution for sm_100a (layers >= 128 pixels wide, Cout <= 128).
//
//   warp 0      : slab producer (one lane) — per needed input row ONE TMA box of 130 pixels x 64 channels
//   warp 1      : TMEM allocator + MMA issuer (one lane) — walks the static schedule of b2g_conv_rows.cuh:
//                 per step one weight tile and R tcgen05.mma groups (one per row of the super-tile), the dx tap
//                 selected by offsetting the SW128 smem descriptor by dx*128 bytes inside the slab
//   warp 2      : weight-tile producer (one lane) — its own ring, so neither producer can starve the other
//   warps 3..18 : epilogue (4 groups x 4 TMEM lane quarters) over the R*P accumulators, double-buffered in TMEM
//
// Same roles, arguments and fused epilogues as conv_igemm_kernel (b2g_conv.cu); replaces the cuDNN convolutions
// behind sndcgan/SNDCGAN.py:50-61 (G.convT3, G.convT4), :80-90 (D.conv2, D.conv3) and their input gradients.
#include <stdlib.h>
#include <string.h>

#include <utility>

#include "b2g_conv_epi.cuh"
#include "b2g_conv_rows.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) RowsKernelParams {
  CUtensorMap amap[4];
  CUtensorMap bmap;
  ConvPlan plan;
  EpiParams e;
  RowsGeom geom;
  RowsSched sched[2];
  EpiTmaMaps em;     // TMA epilogue (used when e.tma)
  int n_tiles;       // Cout / BLOCK_N
  int brows;         // rows of the packed weight matrix per phase
  int total_tiles;   // PG * n_tiles * N * row_blocks * tiles_w
  int tiles_per_pg;  // n_tiles * N * row_blocks * tiles_w
  FastDiv fd_tw, fd_rb, fd_n, fd_nt;  // tile decoding without divisions (tiles_w, row_blocks, N, n_tiles)
};

template <int BLOCK_N, class SS = RowsDyn>
struct RowsCfg {
  static constexpr int kSlabBytes = 17 * 1024;                       // 130 px * 128 B = 16640, padded to 1 KiB
  static constexpr int kSlabTx = kRowsSlabPixels * 128;              // bytes one slab TMA delivers
  static constexpr int kSlabs = SS::kStatic ? SS::kS : (BLOCK_N == 128 ? 7 : 8);
  static constexpr int kBTile = BLOCK_N * 128;                       // one [BLOCK_N][64] weight tile
  static constexpr int kBBytes = (SS::kMerge ? 2 : 1) * kBTile;      // merged schedules stack two tiles per stage
  static constexpr int kBStages = SS::kStatic ? SS::kSB : (BLOCK_N == 128 ? 6 : 8);
  static constexpr int kBarBytes = 512;
  static constexpr int kMainBytes = kSlabs * kSlabBytes + kBStages * kBBytes;  // multiple of 1 KiB
  // Epilogue: 4 groups x 4 TMEM lane quarters = 16 warps, each with kEpiBoxes 2 KiB output boxes for the TMA epilogue
  // (b2g_conv_epi.cuh).  Measured in round 2 (ncu source page + tools/conv_probe.py): a backward epilogue item is ~600-1000
  // instructions of ONE warp's dependent work — 8 warps with run-ahead loads were 45 % slower than 16 warps; the
  // epilogue needs warps and fewer instructions, not deeper load pipelines.
  static constexpr int kEpiGroups = 4;
  static constexpr int kEpiWarps = 4 * kEpiGroups;
  static constexpr bool fits(int nb) {
    return kMainBytes + epi_tma_smem_bytes(kEpiWarps, nb) + kBarBytes + 1024 + 4096 <= 227 * 1024;
  }
  // Tried after the merged schedule (round 2) and NOT kept: the conv_c3-style "fat" epilogue for the N = 64 role — a warp
  // takes all 64 channels of one accumulator per visit (two tcgen05.ld in flight, early accumulator release, one 4 KiB
  // 128-byte-swizzled box).  With 19 warps the CTA is capped at 96 registers (64 of them accumulators) and one box per
  // warp: G.convT3 forward 167 -> 189 us, D.conv2 input gradient unchanged at 196 us, MMA pipeline alone 150 -> 157 us.
  static constexpr int kEpiBoxes = BLOCK_N < 64 ? 0 : (fits(2) ? 2 : (fits(1) ? 1 : 0));
  static constexpr bool kEpiTma = kEpiBoxes > 0;
  static constexpr int kEpiBytes = kEpiTma ? epi_tma_smem_bytes(kEpiWarps, kEpiBoxes) : 0;
  static constexpr int kSmemBytes = kMainBytes + kEpiBytes + kBarBytes + 1024;
  static_assert(2 * (kSlabs + kBStages) * 8 + 4 * 8 + 8 <= kBarBytes, "barrier area");
  static_assert(kSmemBytes + 4096 <= 227 * 1024, "shared memory budget (incl. the static bias / column-sum tables)");
  static constexpr int kThreads = 96 + 128 * kEpiGroups;
};

struct RowsTile {
  int pg, n_tile, n, y0, w0;
};
__device__ __forceinline__ RowsTile decode_rows_tile(const RowsKernelParams& p, int tile) {
  RowsTile t;
  int tw, yb;
  fast_divmod(tile, p.fd_tw, tile, tw);
  t.w0 = tw * 128;
  fast_divmod(tile, p.fd_rb, tile, yb);
  t.y0 = yb * p.geom.R;
  fast_divmod(tile, p.fd_n, tile, t.n);
  fast_divmod(tile, p.fd_nt, t.pg, t.n_tile);
  return t;
}

// Division-free tile iterator for the single-lane issue warps: integer division is a vector-ALU sequence, and
// anything derived from it leaves the uniform datapath (R2UR before every TMA / MMA operand).  The mixed-radix
// coordinate (tw, yb, n, n_tile, pg) of tile blockIdx.x is found by repeated subtraction once, then advanced by the
// digits of gridDim.x with carries.
struct RowsCoord {
  int tw, yb, n, nt, pg;
};
__device__ __forceinline__ void rows_coord_normalise(const RowsKernelParams& p, RowsCoord& c) {
  while (c.tw >= p.plan.tiles_w) { c.tw -= p.plan.tiles_w; ++c.yb; }
  while (c.yb >= p.geom.row_blocks) { c.yb -= p.geom.row_blocks; ++c.n; }
  while (c.n >= p.plan.N) { c.n -= p.plan.N; ++c.nt; }
  while (c.nt >= p.n_tiles) { c.nt -= p.n_tiles; ++c.pg; }
}
__device__ __forceinline__ void rows_coord_advance(const RowsKernelParams& p, RowsCoord& c, const RowsCoord& d) {
  c.tw += d.tw;
  int cy = 0;
  if (c.tw >= p.plan.tiles_w) { c.tw -= p.plan.tiles_w; cy = 1; }
  c.yb += d.yb + cy;
  cy = 0;
  if (c.yb >= p.geom.row_blocks) { c.yb -= p.geom.row_blocks; cy = 1; }
  c.n += d.n + cy;
  cy = 0;
  if (c.n >= p.plan.N) { c.n -= p.plan.N; cy = 1; }
  c.nt += d.nt + cy;
  cy = 0;
  if (c.nt >= p.n_tiles) { c.nt -= p.n_tiles; cy = 1; }
  c.pg += d.pg + cy;
}

// One step of a compile-time schedule: every stage, barrier, parity bit and accumulator is an immediate.
template <class SS, int BLOCK_N, int J>
__device__ __forceinline__ void rows_static_step(uint32_t leader, uint32_t slab_lo0, uint32_t b_lo0, uint32_t d_base,
                                                 uint64_t* sfull, uint64_t* sempty, uint64_t* bfull, uint64_t* bempty,
                                                 uint32_t par_s0, uint32_t par_b0) {
  using Cfg = RowsCfg<BLOCK_N, SS>;
  constexpr RowsStep st = SS::step(J);
  constexpr int S = SS::kS, SB = SS::kSB, R = SS::kR, P = SS::kP;
  constexpr int bst = J % SB;
  constexpr int need = st.slab0 + R - 1, prev = SS::need_before(J), rel0 = SS::released_before(J);
  constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, 0, 0, 128, st.wide ? 2 * BLOCK_N : BLOCK_N);
  static_assert(!st.wide || SS::kMerge, "wide step in a schedule without stacked weight stages");
  mbar_wait(&bfull[bst], par_b0 ^ ((J / SB) & 1));
#pragma unroll
  for (int i = prev + 1; i <= need; ++i) mbar_wait(&sfull[i % S], par_s0 ^ ((i / S) & 1));
  tc_fence_after();
  if (leader) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int acc = st.acc0 + r * P;
      static_assert(!st.wide || SS::written_before(J, st.acc0) == SS::written_before(J, st.acc0 + 1),
                    "a wide step has ONE accumulate flag for both accumulators");
      umma_f16_k64(d_base + acc * BLOCK_N,
                   slab_lo0 + ((st.slab0 + r) % S) * (Cfg::kSlabBytes >> 4) + st.dxi * 8,
                   b_lo0 + bst * (Cfg::kBBytes >> 4), idesc, SS::written_before(J, acc) ? 1u : 0u);
    }
    umma_commit(&bempty[bst]);
#pragma unroll
    for (int q = 0; q < st.rel; ++q) umma_commit(&sempty[(rel0 + q) % S]);
  }
  __syncwarp();
}
template <class SS, int BLOCK_N, int... Js>
__device__ __forceinline__ void rows_static_tile(std::integer_sequence<int, Js...>, uint32_t leader, uint32_t slab_lo0,
                                                 uint32_t b_lo0, uint32_t d_base, uint64_t* sfull, uint64_t* sempty,
                                                 uint64_t* bfull, uint64_t* bempty, uint32_t par_s0, uint32_t par_b0) {
  (rows_static_step<SS, BLOCK_N, Js>(leader, slab_lo0, b_lo0, d_base, sfull, sempty, bfull, bempty, par_s0, par_b0),
   ...);
}

template <int BLOCK_N, class SS>
__global__ void __launch_bounds__(RowsCfg<BLOCK_N, SS>::kThreads, 1)
    conv_rows_kernel(const __grid_constant__ RowsKernelParams p) {
  using Cfg = RowsCfg<BLOCK_N, SS>;
  __shared__ float sbias[512];
  __shared__ float s_colsum[512];
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* bsm = smem + Cfg::kSlabs * Cfg::kSlabBytes;
  uint8_t* epism = smem + Cfg::kMainBytes;
  uint64_t* sfull = reinterpret_cast<uint64_t*>(epism + Cfg::kEpiBytes);
  uint64_t* sempty = sfull + Cfg::kSlabs;
  uint64_t* bfull = sempty + Cfg::kSlabs;
  uint64_t* bempty = bfull + Cfg::kBStages;
  uint64_t* tfull = bempty + Cfg::kBStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const ConvPlan& pl = p.plan;
  // compile-time schedules know R and P: the per-item "acc / P", "acc % P" of the epilogue become shifts instead of
  // runtime integer divisions
  int R = p.geom.R, P = p.geom.P, n_acc = p.geom.n_acc;
  if constexpr (SS::kStatic) {
    R = SS::kR;
    P = SS::kP;
    n_acc = SS::kR * SS::kP;
  }
  const uint32_t tmem_cols = 2u * n_acc * BLOCK_N;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kSlabs; ++s) {
      mbar_init(&sfull[s], 1);
      mbar_init(&sempty[s], 1);
    }
    for (int s = 0; s < Cfg::kBStages; ++s) {
      mbar_init(&bfull[s], 1);
      mbar_init(&bempty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 4 * Cfg::kEpiGroups);
    }
    if (Cfg::kEpiTma && p.e.tma) epi_tma_init(epism, Cfg::kEpiWarps, Cfg::kEpiBoxes);
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < pl.n_amaps; ++i) tma_prefetch_desc(&p.amap[i]);
  }
  if (warp == 2 && lane == 0) tma_prefetch_desc(&p.bmap);
  if (p.e.bias) {
    const int nb = p.e.out_f32 ? p.e.cs : pl.Cout;
    for (int i = threadIdx.x; i < nb && i < 512; i += blockDim.x) sbias[i] = p.e.bias[i] * p.e.pre_scale;
  }
  epi_colsum_zero(p.e, s_colsum, pl.Cout, sbias);
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ slab producer (whole warp converged; one elected lane issues)
    const uint32_t leader = elect_one_sync();
    RowsCoord c = {static_cast<int>(blockIdx.x), 0, 0, 0, 0}, d = {static_cast<int>(gridDim.x), 0, 0, 0, 0};
    rows_coord_normalise(p, c);
    rows_coord_normalise(p, d);
    uint32_t stage = 0, par = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const RowsSched& sc = p.sched[c.pg];
      const int n_slabs = sc.n_slabs;
      const int w0 = c.tw * 128 - 1, y0 = c.yb * R;
      for (int i = 0; i < n_slabs; ++i) {
        const int amap = sc.slab_map[i], chunk = sc.slab_chunk[i], dy = sc.slab_dy[i];
        mbar_wait(&sempty[stage], par ^ 1);
        if (leader) {
          mbar_expect_tx(&sfull[stage], Cfg::kSlabTx);
          tma_load_4d(smem + stage * Cfg::kSlabBytes, &p.amap[amap], &sfull[stage], chunk * 64, w0, y0 + dy, c.n);
        }
        __syncwarp();
        if (++stage == Cfg::kSlabs) { stage = 0; par ^= 1; }
      }
      rows_coord_advance(p, c, d);
    }
  } else if (warp == 2) {
    // ------------------------------------------------ weight-tile producer
    const uint32_t leader = elect_one_sync();
    RowsCoord c = {static_cast<int>(blockIdx.x), 0, 0, 0, 0}, d = {static_cast<int>(gridDim.x), 0, 0, 0, 0};
    rows_coord_normalise(p, c);
    rows_coord_normalise(p, d);
    uint32_t stage = 0, par = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
      const RowsSched& sc = p.sched[c.pg];
      const int n_steps = sc.n_steps;
      const int brow0 = c.pg * P * p.brows + c.nt * BLOCK_N;
      for (int j = 0; j < n_steps; ++j) {
        const int bcol = sc.steps[j].bcol, phl = sc.steps[j].phase_l, wide = SS::kMerge ? sc.steps[j].wide : 0;
        mbar_wait(&bempty[stage], par ^ 1);
        if (leader) {
          mbar_expect_tx(&bfull[stage], wide ? 2 * Cfg::kBTile : Cfg::kBTile);
          tma_load_2d(bsm + stage * Cfg::kBBytes, &p.bmap, &bfull[stage], bcol * 64, brow0 + phl * p.brows);
          if (wide)  // the next local phase's tile for the same shift, stacked behind the first: rows BLOCK_N .. 2 BLOCK_N - 1
            tma_load_2d(bsm + stage * Cfg::kBBytes + Cfg::kBTile, &p.bmap, &bfull[stage], sc.steps[j].bcol2 * 64,
                        brow0 + (phl + 1) * p.brows);
        }
        __syncwarp();
        if (++stage == Cfg::kBStages) { stage = 0; par ^= 1; }
      }
      rows_coord_advance(p, c, d);
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer (whole warp converged; one elected lane issues)
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, 0, 0, 128, BLOCK_N);
    constexpr uint32_t kSlabD = Cfg::kSlabBytes >> 4, kBD = Cfg::kBBytes >> 4;  // stage strides in descriptor units
    const uint32_t slab_lo0 = (smem_u32(smem) & 0x3FFFFu) >> 4, b_lo0 = (smem_u32(bsm) & 0x3FFFFu) >> 4;
    if constexpr (SS::kStatic) {
      // compile-time schedule: the rings divide the slabs / weight tiles of a super-tile, so only one parity bit
      // per ring changes from tile to tile
      static_assert(SS::ok(), "schedule does not fit");
      static_assert(SS::n_slabs() % SS::kS == 0 && SS::n_steps() % SS::kSB == 0, "rings must divide the schedule");
      constexpr uint32_t us = SS::n_slabs() / SS::kS, ub = SS::n_steps() / SS::kSB;  // uses of a stage per tile
      uint32_t cs = 0, cb = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        mbar_wait(&tempty[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        rows_static_tile<SS, BLOCK_N>(std::make_integer_sequence<int, SS::n_steps()>{}, leader, slab_lo0, b_lo0,
                                      tmem_base + as * (n_acc * BLOCK_N), sfull, sempty, bfull, bempty, cs & 1, cb & 1);
        cs += us;
        cb += ub;
        if (leader) umma_commit(&tfull[as]);
        __syncwarp();
      }
    } else {
    uint32_t s_head = 0, s_head_par = 0;  // ring position / parity of slab 0 of the current tile
    uint32_t b_stage = 0, b_par = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const RowsSched& sc = p.sched[tile >= p.tiles_per_pg ? 1 : 0];
      const int as = it & 1;
      mbar_wait(&tempty[as], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_base = tmem_base + as * (n_acc * BLOCK_N);
      uint32_t written = 0;
      int waited = 0;
      uint32_t w_stage = s_head, w_par = s_head_par;  // next slab to wait for
      uint32_t r_stage = s_head;                      // next slab to release
      const int n_steps = sc.n_steps;
      for (int j = 0; j < n_steps; ++j) {
        const int slab0 = sc.steps[j].slab0, dxi = sc.steps[j].dxi, acc0 = sc.steps[j].acc0, rel = sc.steps[j].rel;
        mbar_wait(&bfull[b_stage], b_par);
        for (; waited < slab0 + R; ++waited) {
          mbar_wait(&sfull[w_stage], w_par);
          if (++w_stage == Cfg::kSlabs) { w_stage = 0; w_par ^= 1; }
        }
        tc_fence_after();
        const uint32_t b_lo = b_lo0 + b_stage * kBD;
        uint32_t sstage = s_head + slab0;  // slab0 < 3 * kSlabs (kRowsMaxSlabs = 24, ring >= 7)
        if (sstage >= 2 * Cfg::kSlabs) sstage -= 2 * Cfg::kSlabs;
        if (sstage >= Cfg::kSlabs) sstage -= Cfg::kSlabs;
        if (sstage >= Cfg::kSlabs) sstage -= Cfg::kSlabs;
#pragma unroll 1
        for (int r = 0; r < R; ++r) {
          const uint32_t a_lo = slab_lo0 + sstage * kSlabD + dxi * 8;
          const int acc = acc0 + r * P;
          const uint32_t d_tmem = d_base + acc * BLOCK_N;
          const uint32_t have = (written >> acc) & 1;
          if (leader) umma_f16_k64(d_tmem, a_lo, b_lo, idesc, have);
          __syncwarp();
          written |= 1u << acc;
          if (++sstage == Cfg::kSlabs) sstage = 0;
        }
        if (leader) {
          umma_commit(&bempty[b_stage]);
          for (int q = 0; q < rel; ++q) {
            uint32_t rs = r_stage + q;
            if (rs >= Cfg::kSlabs) rs -= Cfg::kSlabs;
            umma_commit(&sempty[rs]);
          }
        }
        __syncwarp();
        r_stage += rel;
        if (r_stage >= Cfg::kSlabs) r_stage -= Cfg::kSlabs;
        if (++b_stage == Cfg::kBStages) { b_stage = 0; b_par ^= 1; }
      }
      s_head = w_stage;  // every slab of the tile has been waited for by its last step
      s_head_par = w_par;
      if (leader) umma_commit(&tfull[as]);
      __syncwarp();
    }
    }
  } else {
    // ------------------------------------------------ epilogue (warp w owns TMEM lanes 32*(w%4)..+31)
    const EpiParams& e = p.e;                 // stays in the constant bank
    const EpiSeed sd = epi_resolve_seed(p.e);  // dropout seed possibly read from device memory
    const int q = warp & 3;
    const int grp = (warp - 3) >> 2;
    const int row = q * 32 + lane;
    constexpr int kChunks = BLOCK_N / 32;
    constexpr int NB = Cfg::kEpiBoxes > 0 ? Cfg::kEpiBoxes : 1;
    EpiWarpStageN<NB> st = epi_tma_stage<NB>(epism, Cfg::kEpiWarps, warp - 3);
    const bool use_tma = Cfg::kEpiTma && p.e.tma != 0;
    const bool zero_oob = e.colsum != nullptr && (e.dact == B2G_ACT_NONE || e.dact_zero != 0.f);
    const int n_items = n_acc * kChunks;  // items (accumulator, 32-channel chunk) of a tile; this warp: grp, grp + G, ...
    // Begin pointer of the TMA epilogue: the next item (tile, idx) of this warp that has not been begun (box freed,
    // saved-output load started).  ensure_begun(n) keeps n items begun beyond the ones already processed: 1 before the
    // accumulator wait (the first item's load overlaps the wait), NB inside the item loop (NB = 2: one item ahead).
    int b_tile = blockIdx.x, b_idx = grp;
    RowsTile bt = decode_rows_tile(p, b_tile);  // decoded once per tile, not per item (the divisions are a dependent
                                                // chain of ~40 instructions: a quarter of the epilogue warps' stall
                                                // samples sat in this bookkeeping, ncu round 2)
    auto ensure_begun = [&](uint32_t n) {
      if (!use_tma || grp >= n_items) return;
      while (st.k_begin - st.k_chunk < n) {
        if (b_idx >= n_items) {
          b_tile += gridDim.x;
          b_idx = grp;
          if (b_tile < p.total_tiles) bt = decode_rows_tile(p, b_tile);
        }
        if (b_tile >= p.total_tiles) return;
        const int acc = b_idx / kChunks, c = b_idx % kChunks;
        epi_tma_begin(e, &p.em.ymap[bt.pg * P + acc % P], st, bt.n_tile * BLOCK_N + c * 32, bt.w0 + q * 32,
                      bt.y0 + acc / P, bt.n);
        b_idx += Cfg::kEpiGroups;
      }
    };
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1;
      const RowsTile t = decode_rows_tile(p, tile);
      ensure_begun(1);
      // backward epilogues: the saved-output boxes of the warp's remaining items of this tile go to L2 now, while the
      // MMAs of the tile still run
      if (use_tma && e.dact != B2G_ACT_NONE) {
        for (int idx = grp + Cfg::kEpiGroups; idx < n_items; idx += Cfg::kEpiGroups) {
          const int acc = idx / kChunks, c = idx % kChunks;
          if (elect_one_sync())
            tma_prefetch_4d(&p.em.ymap[t.pg * P + acc % P], t.n_tile * BLOCK_N + c * 32, t.w0 + q * 32, t.y0 + acc / P, t.n);
        }
        __syncwarp();
      }
      mbar_wait_parked(&tfull[as], (it >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int idx = grp; idx < n_items; idx += Cfg::kEpiGroups) {
        const int acc = idx / kChunks, c = idx % kChunks;
        const int r = acc / P, phase = t.pg * P + acc % P;
        const int gy = t.y0 + r;
        const bool valid = gy < pl.Hg;
        const int oy = gy * pl.os + pl.ph_y[phase], ox = (t.w0 + row) * pl.os + pl.ph_x[phase];
        const size_t pix = (static_cast<size_t>(t.n) * pl.Hout + oy) * pl.Wout + ox;
        const int cb = t.n_tile * BLOCK_N + c * 32;
        uint32_t v[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * (n_acc * BLOCK_N) + acc * BLOCK_N + c * 32,
                      v);
        ensure_begun(NB);
        tmem_ld_wait();
        if (use_tma && zero_oob && !valid) {  // see conv_pair_kernel: pixels of a ragged tile outside the tensor
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if (use_tma) epi_tma_chunk(e, &p.em.omap[phase], st, sbias, v, pix, cb, lane, t.w0 + q * 32, gy, t.n, s_colsum, sd);
        else conv_epilogue_chunk(e, sbias, v, pix, cb, valid, s_colsum, sd);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
    }
    if (use_tma) epi_tma_end(e, st, s_colsum, lane);
  }

  tc_fence_before();
  __syncthreads();
  epi_colsum_flush(p.e, s_colsum, pl.Cout, sbias);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

template <int BLOCK_N, class SS = RowsDyn>
static int launch_rows(const RowsKernelParams& kp, cudaStream_t stream) {
  using Cfg = RowsCfg<BLOCK_N, SS>;
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(conv_rows_kernel<BLOCK_N, SS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::kSmemBytes));
    attr_set = true;
  }
  const int grid = kp.total_tiles < num_sms() ? kp.total_tiles : num_sms();
  conv_rows_kernel<BLOCK_N, SS><<<grid, Cfg::kThreads, Cfg::kSmemBytes, stream>>>(kp);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

static int rows_ring(int block_n) { return block_n == 128 ? RowsCfg<128>::kSlabs : RowsCfg<64>::kSlabs; }

// compile-time schedules of the model's four wide roles: <kind, Cin/64, R, P, N tile, slab ring, weight-tile ring>
using RowsConv3 = RowsStatic<B2G_CONV_K3S1, 2, 2, 1, 128, 4, 6>;    // D.conv3 fwd / dgrad (128 -> 128 @72x128); 3 weight
                                                                    // stages + a second epilogue box measured WORSE on the
                                                                    // same box (fwd 138 -> 144 us, dgrad 142 -> 151 us)
using RowsConv2 = RowsStatic<B2G_CONV_K4S2, 1, 2, 1, 128, 4, 8>;    // D.conv2 fwd (64 -> 128, 144x256 -> 72x128)
using RowsConv2b = RowsStatic<B2G_CONV_K4S2, 1, 2, 1, 128, 4, 4>;   // same role, 4 weight stages: room for 3 epilogue boxes
using RowsConvT3 = RowsStatic<B2G_CONVT_K4S2, 2, 2, 2, 64, 6, 3, true>;  // G.convT3 fwd / D.conv2 dgrad (128 -> 64): merged
                                                                         // dx = 0 steps (12 instead of 16 per super-tile),
                                                                         // 3 stacked weight stages of 16 KB
using RowsConvT4 = RowsStatic<B2G_CONV_K3S1, 1, 4, 1, 32, 6, 9>;    // G.convT4 fwd / D.conv1 dgrad (64 -> 3)

// Tries the compile-time kernel SS for this role: the host schedule is rebuilt with SS's slab ring and must equal
// the compile-time one step for step.
template <int BLOCK_N, class SS>
static bool try_static(const ConvPlan& pl, int block_n, RowsKernelParams* kp) {
  if (block_n != BLOCK_N || pl.kind != SS::kKind || pl.Cin != SS::kCh * 64) return false;
  RowsGeom g;
  RowsSched sc[2];
  if (!build_rows_sched(pl, block_n, SS::kS, &g, sc, SS::kMerge)) return false;
  if (!rows_static_matches<SS>(g, sc)) return false;
  kp->geom = g;
  kp->sched[0] = sc[0];
  kp->sched[1] = sc[1];
  return true;
}

int launch_conv_rows(const b2g_conv_args* a, const ConvPlan& pl, int block_n, void* stream_, bool* handled,
                     bool static_only) {
  *handled = false;
  const int enabled = env_int("B2G_CONV_ROWS", 1);  // read per call: tests switch kernels
  if (!enabled) return B2G_OK;
  if (block_n != 32 && block_n != 64 && block_n != 128) return B2G_OK;
  RowsKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  int which = 0;  // 0 = generic table-driven kernel, 1..4 = compile-time schedule
  if (env_int("B2G_ROWS_STATIC", 1)) {
    if (try_static<128, RowsConv3>(pl, block_n, &kp)) which = 1;
    else if (try_static<128, RowsConv2>(pl, block_n, &kp)) which = env_int("B2G_ROWS_CONV2_SB4", 1) ? 5 : 2;
    else if (try_static<64, RowsConvT3>(pl, block_n, &kp)) which = 3;
    else if (try_static<32, RowsConvT4>(pl, block_n, &kp)) which = 4;
  }
  if (!which && static_only) return B2G_OK;
  if (!which && !build_rows_sched(pl, block_n, rows_ring(block_n), &kp.geom, kp.sched)) return B2G_OK;
  kp.plan = pl;
  fill_epi_params(a, &kp.e);
  {
    const bool room = which == 1   ? RowsCfg<128, RowsConv3>::kEpiTma
                      : which == 2 ? RowsCfg<128, RowsConv2>::kEpiTma
                      : which == 3 ? RowsCfg<64, RowsConvT3>::kEpiTma
                      : which == 5 ? RowsCfg<128, RowsConv2b>::kEpiTma
                                   : false;
    int rc = fill_epi_tma(a, pl, room, &kp.e, &kp.em, /*row32=*/true);
    if (rc) return rc;
  }
  kp.brows = a->out_f32 ? a->cout_store : a->Cout;
  kp.n_tiles = a->Cout / block_n;
  kp.tiles_per_pg = kp.n_tiles * pl.N * kp.geom.row_blocks * pl.tiles_w;
  kp.total_tiles = kp.geom.PG * kp.tiles_per_pg;
  kp.fd_tw = make_fastdiv(pl.tiles_w, kp.total_tiles);
  kp.fd_rb = make_fastdiv(kp.geom.row_blocks, kp.total_tiles);
  kp.fd_n = make_fastdiv(pl.N, kp.total_tiles);
  kp.fd_nt = make_fastdiv(kp.n_tiles, kp.total_tiles);

  const size_t es = 2;
  const char* base = static_cast<const char*>(a->in);
  const uint32_t box[4] = {64u, (uint32_t)kRowsSlabPixels, 1u, 1u};
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
    const uint32_t bbox[2] = {64u, (uint32_t)block_n};
    int rc = make_tmap(&kp.bmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a->wpacked, dims, str, bbox);
    if (rc) return rc;
  }
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int rc;
  switch (which) {
    case 1: rc = launch_rows<128, RowsConv3>(kp, stream); break;
    case 2: rc = launch_rows<128, RowsConv2>(kp, stream); break;
    case 3: rc = launch_rows<64, RowsConvT3>(kp, stream); break;
    case 4: rc = launch_rows<32, RowsConvT4>(kp, stream); break;
    case 5: rc = launch_rows<128, RowsConv2b>(kp, stream); break;
    default:
      switch (block_n) {
        case 128: rc = launch_rows<128>(kp, stream); break;
        case 64: rc = launch_rows<64>(kp, stream); break;
        default: rc = launch_rows<32>(kp, stream); break;
      }
  }
  if (rc == B2G_OK) *handled = true;
  return rc;
}

}  // namespace b2g

// host-only introspection for the CPU test-suite: the row-slab schedule of a conv role.
//   out[0..7] = {handled, R, P, PG, n_acc, row_blocks, n_slabs(pg), n_steps(pg)};
//   out[8 + 3*i + {0,1,2}] = slab i {map, dy, chunk};  out[8 + 72 + 6*j + {0..5}] = step j {bcol, phase_l, slab0, dxi, acc0, rel}
extern "C" int b2g_conv_rows_plan_query(int kind, int N, int Hin, int Win, int Cin, int block_n, int pg, int* out,
                                        int out_len) {
  using namespace b2g;
  if (!out || out_len < 8 + 3 * kRowsMaxSlabs + 6 * kRowsMaxSteps) return fail(B2G_E_INVALID_ARG, "rows_plan_query: out too small");
  memset(out, 0, sizeof(int) * out_len);
  ConvPlan pl;
  if (!build_conv_plan(kind, N, Hin, Win, Cin, block_n, 128, &pl)) return B2G_OK;
  RowsGeom g;
  RowsSched sc[2];
  if (block_n != 32 && block_n != 64 && block_n != 128) return B2G_OK;
  if (!build_rows_sched(pl, block_n, rows_ring(block_n), &g, sc)) return B2G_OK;
  if (pg < 0 || pg >= g.PG) return fail(B2G_E_INVALID_ARG, "rows_plan_query: phase group %d out of range", pg);
  out[0] = 1; out[1] = g.R; out[2] = g.P; out[3] = g.PG; out[4] = g.n_acc; out[5] = g.row_blocks;
  out[6] = sc[pg].n_slabs; out[7] = sc[pg].n_steps;
  for (int i = 0; i < sc[pg].n_slabs; ++i) {
    out[8 + 3 * i] = sc[pg].slab_map[i]; out[9 + 3 * i] = sc[pg].slab_dy[i]; out[10 + 3 * i] = sc[pg].slab_chunk[i];
  }
  for (int j = 0; j < sc[pg].n_steps; ++j) {
    const RowsStep& st = sc[pg].steps[j];
    int* o = out + 8 + 3 * kRowsMaxSlabs + 6 * j;
    o[0] = st.bcol; o[1] = st.phase_l; o[2] = st.slab0; o[3] = st.dxi; o[4] = st.acc0; o[5] = st.rel;
  }
  return B2G_OK;
}

// The schedule with merged dx = 0 steps (transposed conv, two phases per group; RowsStep::wide): as above, 8 ints per
// step from out[8 + 72]: {bcol, phase_l, slab0, dxi, acc0, rel, wide, bcol2}; out needs 8 + 72 + 8 * 64 ints.
extern "C" int b2g_conv_rows_plan_query_merged(int kind, int N, int Hin, int Win, int Cin, int block_n, int pg, int ring,
                                               int* out, int out_len) {
  using namespace b2g;
  if (!out || out_len < 8 + 3 * kRowsMaxSlabs + 8 * kRowsMaxSteps) return fail(B2G_E_INVALID_ARG, "rows_plan_query_merged: out too small");
  memset(out, 0, sizeof(int) * out_len);
  ConvPlan pl;
  if (!build_conv_plan(kind, N, Hin, Win, Cin, block_n, 128, &pl)) return B2G_OK;
  RowsGeom g;
  RowsSched sc[2];
  if (block_n != 32 && block_n != 64 && block_n != 128) return B2G_OK;
  if (!build_rows_sched(pl, block_n, ring > 0 ? ring : rows_ring(block_n), &g, sc, true)) return B2G_OK;
  if (pg < 0 || pg >= g.PG) return fail(B2G_E_INVALID_ARG, "rows_plan_query_merged: phase group %d out of range", pg);
  out[0] = 1; out[1] = g.R; out[2] = g.P; out[3] = g.PG; out[4] = g.n_acc; out[5] = g.row_blocks;
  out[6] = sc[pg].n_slabs; out[7] = sc[pg].n_steps;
  for (int i = 0; i < sc[pg].n_slabs; ++i) {
    out[8 + 3 * i] = sc[pg].slab_map[i]; out[9 + 3 * i] = sc[pg].slab_dy[i]; out[10 + 3 * i] = sc[pg].slab_chunk[i];
  }
  for (int j = 0; j < sc[pg].n_steps; ++j) {
    const RowsStep& st = sc[pg].steps[j];
    int* o = out + 8 + 3 * kRowsMaxSlabs + 8 * j;
    o[0] = st.bcol; o[1] = st.phase_l; o[2] = st.slab0; o[3] = st.dxi; o[4] = st.acc0; o[5] = st.rel; o[6] = st.wide;
    o[7] = st.bcol2;
  }
  return B2G_OK;
}
