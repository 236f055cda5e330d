// b2g_wgrad_slab.cu — weight gradient of the 64/128-channel layers with the shifted operand read ONCE per pixel row.
//
// The tap-grouped kernel (b2g_wgrad.cu) loads one [64 px][64 ch] box of the shifted operand per tap and k-step:
// 48 KB per 64 pixels, which is the L2->SM ingest limit (measured 11.2 TB/s, tensor pipe 45 % busy, profiles/
// r1_v8_wgrad_taps.md).  Taps that differ only in dx are windows of the same pixel row, one pixel apart.  Here a
// k-step loads a SLAB = 66 (3x3) or 65 (4x4 stride 2) pixels of one row and parity map, and the MMA reads the
// taps out of it: with MN-major operands a pixel is one 128-byte line, so tap dx starts dx * 128 bytes further
// and "the next tap" is simply the next 64-column atom at a leading-dimension offset of 128 bytes.  One MMA of
// N = taps_per_slab * 64 covers all dx of a slab box.
//
//   D[b][col] += sum_px U[px][b] * S_slab[px + dx(col)][a(col)]      M = 128 channels of the unshifted operand
//   work item = (slab group, 128-channel chunk of the shifted operand, M tile, pixel split); <= 512 TMEM columns
//   3x3 s1 : slab = (dy), 3 taps, groups of 1 slab x 2 channel boxes (N = 384)
//   4x4 s2 : slab = (parity map, row offset), 2 taps, groups of 4 slabs x 1 channel box (N = 512) for 64 channels
//
// Replaces Conv2DBackpropFilter (tape.gradient at sndcgan/SNDCGAN.py:248,:256,:266) for D.conv2 / D.conv3.
#include <stdlib.h>
#include <string.h>

#include "b2g_conv.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

constexpr int kSlabMaxBoxes = 4;   // slab boxes ([<= 66 px][64 ch]) per k-step
constexpr int kSlabMaxGroups = 8;
constexpr int kSlabBoxBytes = 9 * 1024;  // 66 px * 128 B = 8448, padded to 1 KiB

struct SlabBox {
  int8_t map, dy, dx0, cbox;  // tensor map, row offset, first pixel offset, 64-channel box inside the item's chunk
  int8_t ntaps;               // taps (consecutive dx) read from this box
  int8_t tap[3];              // their tap indices (output blocks)
};
struct SlabGroup {
  int n_boxes;
  SlabBox box[kSlabMaxBoxes];
};

struct alignas(64) WgradSlabParams {
  CUtensorMap smap[4];  // shifted operand, box {64 ch, slab_px, 1, 1}
  CUtensorMap umap;     // unshifted operand, box {64 ch, 64 px, 1, 1}
  SlabGroup group[kSlabMaxGroups];
  int n_groups, a_chunks, m_tiles, splits, ptiles;
  int tiles_w, Hg;      // pixel tiles: 64 consecutive pixels of one grid row
  int slab_px;          // 66 or 65
  int chunk_ch;         // shifted channels per item (64 or 128)
  int Ca, Cb;
  float* out;           // [taps][Ca][Cb]
};

struct WgradSlabCfg {
  static constexpr int kUBytes = 2 * 64 * 128;  // 128 channels x 64 pixels
  static constexpr int kStageBytes = kUBytes + kSlabMaxBoxes * kSlabBoxBytes;  // 52 KB
  static constexpr int kStages = 4;
  static constexpr int kSmemBytes = kStages * kStageBytes + 256 + 1024;
};

constexpr int kWgradSlabThreads = 192;

struct SlabItem {
  int grp, ac, m_tile, p_begin, p_end;
};
__device__ __forceinline__ SlabItem decode_slab_item(const WgradSlabParams& p, int item) {
  SlabItem w;
  w.grp = item % p.n_groups;
  int r = item / p.n_groups;
  w.ac = r % p.a_chunks;
  r /= p.a_chunks;
  w.m_tile = r % p.m_tiles;
  const int split = r / p.m_tiles;
  w.p_begin = (int)((long)split * p.ptiles / p.splits);
  w.p_end = (int)((long)(split + 1) * p.ptiles / p.splits);
  return w;
}

__global__ void __launch_bounds__(kWgradSlabThreads, 1) wgrad_slab_kernel(const __grid_constant__ WgradSlabParams p) {
  using Cfg = WgradSlabCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_items = p.n_groups * p.a_chunks * p.m_tiles * p.splits;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tfull_bar, 1);
    mbar_init(tempty_bar, 4);
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < 4; ++i) tma_prefetch_desc(&p.smap[i]);
    tma_prefetch_desc(&p.umap);
  }
  if (warp == 1) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer
    const uint32_t leader = elect_one_sync();
    int stage = 0;
    uint32_t phase = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
      const SlabItem wi = decode_slab_item(p, item);
      const SlabGroup& g = p.group[wi.grp];
      const int m_c0 = wi.m_tile * 128, a_c0 = wi.ac * p.chunk_ch;
      const uint32_t bytes = Cfg::kUBytes + g.n_boxes * p.slab_px * 128;
      for (int pt = wi.p_begin; pt < wi.p_end; ++pt) {
        const int tw = pt % p.tiles_w;
        const int r2 = pt / p.tiles_w;
        const int y = r2 % p.Hg, n = r2 / p.Hg;
        const int x0 = tw * 64;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* u_dst = smem + stage * Cfg::kStageBytes;
        uint8_t* s_dst = u_dst + Cfg::kUBytes;
        if (leader) {
          mbar_expect_tx(&full_bar[stage], bytes);
          tma_load_4d(u_dst, &p.umap, &full_bar[stage], m_c0, x0, y, n);
          tma_load_4d(u_dst + 64 * 128, &p.umap, &full_bar[stage], m_c0 + 64, x0, y, n);
          for (int b = 0; b < g.n_boxes; ++b) {
            const SlabBox& sb = g.box[b];
            tma_load_4d(s_dst + b * kSlabBoxBytes, &p.smap[sb.map], &full_bar[stage], a_c0 + sb.cbox * 64, x0 + sb.dx0,
                        y + sb.dy, n);
          }
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer
    const uint32_t leader = elect_one_sync();
    const uint32_t smem_base = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++it) {
      const SlabItem wi = decode_slab_item(p, item);
      const SlabGroup& g = p.group[wi.grp];
      if (it > 0) {  // one accumulator set (up to 512 columns): the epilogue must have drained the previous item
        mbar_wait(tempty_bar, (it - 1) & 1);
        tc_fence_after();
      }
      const int k_steps = wi.p_end - wi.p_begin;
      for (int ks = 0; ks < k_steps; ++ks) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t u_addr = smem_base + stage * Cfg::kStageBytes;
        const uint64_t udesc = umma_smem_desc(u_addr, 64 * 128, 1024);
        if (leader) {
          uint32_t col = 0;
          for (int b = 0; b < g.n_boxes; ++b) {
            const int nt = g.box[b].ntaps;
            const uint32_t idesc = umma_idesc(/*bf16*/ 1, /*A MN-major*/ 1, /*B MN-major*/ 1, 128, nt * 64);
            // taps of this box = 64-column atoms one pixel line (128 bytes) apart
            const uint64_t sdesc = umma_smem_desc(u_addr + Cfg::kUBytes + b * kSlabBoxBytes, 128, 1024);
#pragma unroll
            for (int k = 0; k < 4; ++k)  // 16 pixels per MMA = 2048 bytes of both operands
              umma_f16(tmem_base + col, udesc + 128 * k, sdesc + 128 * k, idesc, (ks | k) != 0);
            col += nt * 64;
          }
          umma_commit(&empty_bar[stage]);
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
      if (leader) umma_commit(tfull_bar);
      __syncwarp();
    }
  } else {
    // ------------------------------------------------ epilogue: out[tap][a][b = M row], coalesced reds over b
    const int q = warp & 3;
    const int row = q * 32 + lane;
    int it = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x, ++it) {
      const SlabItem wi = decode_slab_item(p, item);
      const SlabGroup& g = p.group[wi.grp];
      const bool nonempty = wi.p_end > wi.p_begin;
      mbar_wait(tfull_bar, it & 1);
      tc_fence_after();
      const int b = wi.m_tile * 128 + row;
      uint32_t col = 0;
      for (int sbx = 0; sbx < g.n_boxes; ++sbx) {
        const SlabBox& sb = g.box[sbx];
        for (int t = 0; t < sb.ntaps; ++t) {
          const int a0 = wi.ac * p.chunk_ch + sb.cbox * 64;
          float* dst0 = p.out + ((size_t)sb.tap[t] * p.Ca + a0) * p.Cb + b;
#pragma unroll 1
          for (int h = 0; h < 2; ++h) {
            uint32_t r[32];
            __syncwarp();
            tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + col + h * 32, r);
            tmem_ld_wait();
            if (nonempty) {
              float* dst = dst0 + (size_t)(h * 32) * p.Cb;
#pragma unroll
              for (int j = 0; j < 32; ++j) atomicAdd(dst + (size_t)j * p.Cb, __uint_as_float(r[j]));
            }
          }
          col += 64;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// Host: groups the taps of the plan into slabs (same tensor map and row offset, consecutive dx) and the slabs into
// work-item groups of at most 512 accumulator columns / kSlabMaxBoxes boxes.  Returns false when the shape is not
// one this kernel is meant for.
static bool build_slab_groups(const ConvPlan& pl, int Ca, WgradSlabParams* kp) {
  struct Slab { int map, dy, dx0, n; int tap[3]; };
  Slab slabs[16];
  int ns = 0;
  for (int t = 0; t < pl.taps; ++t) {
    int s = -1;
    for (int i = 0; i < ns; ++i)
      if (slabs[i].map == pl.tap_map[t] && slabs[i].dy == pl.tap_dy[t]) s = i;
    if (s < 0) {
      s = ns++;
      slabs[s] = {pl.tap_map[t], pl.tap_dy[t], pl.tap_dx[t], 0, {0, 0, 0}};
    }
    Slab& sl = slabs[s];
    if (sl.n >= 3) return false;
    if (pl.tap_dx[t] < sl.dx0) return false;  // taps arrive in increasing dx
    if (pl.tap_dx[t] != sl.dx0 + sl.n) return false;
    sl.tap[sl.n++] = t;
  }
  const int ntaps = slabs[0].n;
  for (int i = 0; i < ns; ++i)
    if (slabs[i].n != ntaps) return false;
  kp->slab_px = 64 + ntaps - 1;
  kp->chunk_ch = Ca >= 128 ? 128 : 64;
  if (Ca % kp->chunk_ch) return false;
  kp->a_chunks = Ca / kp->chunk_ch;
  const int boxes_per_slab = kp->chunk_ch / 64;
  const int cols_per_slab = boxes_per_slab * ntaps * 64;
  int slabs_per_group = 512 / cols_per_slab;
  if (slabs_per_group * boxes_per_slab > kSlabMaxBoxes) slabs_per_group = kSlabMaxBoxes / boxes_per_slab;
  if (slabs_per_group < 1) return false;
  while (ns % slabs_per_group) --slabs_per_group;  // equal-sized groups
  kp->n_groups = ns / slabs_per_group;
  if (kp->n_groups > kSlabMaxGroups) return false;
  for (int g = 0; g < kp->n_groups; ++g) {
    SlabGroup& sg = kp->group[g];
    sg.n_boxes = 0;
    for (int i = 0; i < slabs_per_group; ++i) {
      const Slab& sl = slabs[g * slabs_per_group + i];
      for (int cb = 0; cb < boxes_per_slab; ++cb) {
        SlabBox& b = sg.box[sg.n_boxes++];
        b.map = (int8_t)sl.map; b.dy = (int8_t)sl.dy; b.dx0 = (int8_t)sl.dx0; b.cbox = (int8_t)cb;
        b.ntaps = (int8_t)sl.n;
        for (int j = 0; j < 3; ++j) b.tap[j] = (int8_t)sl.tap[j];
      }
    }
  }
  return true;
}

int launch_wgrad_slab(const b2g_wgrad_args* a, void* stream_, bool* handled) {
  *handled = false;
  static int enabled = []() { const char* e = getenv("B2G_WGRAD_SLAB"); return e ? atoi(e) : 1; }();
  if (!enabled) return B2G_OK;
  if (a->tf32 || a->dbias || (a->ca_valid != 0 && a->ca_valid != a->Ca)) return B2G_OK;
  if (a->kind != B2G_CONV_K3S1 && a->kind != B2G_CONV_K4S2) return B2G_OK;
  if (a->Cb % 128 || a->Ca % 64) return B2G_OK;
  WgradSlabParams kp;
  memset(&kp, 0, sizeof(kp));
  ConvPlan pl;
  memset(&pl, 0, sizeof(pl));
  if (!build_conv_plan(a->kind, a->N, a->Hs, a->Ws, a->Ca, a->Cb, 64, &pl)) return B2G_OK;
  if (pl.Wg % 64) return B2G_OK;  // pixel tiles are 64 consecutive pixels of one grid row
  if (!build_slab_groups(pl, a->Ca, &kp)) return B2G_OK;
  kp.tiles_w = pl.Wg / 64;
  kp.Hg = pl.Hg;
  kp.ptiles = a->N * pl.Hg * kp.tiles_w;
  kp.m_tiles = a->Cb / 128;
  kp.Ca = a->Ca;
  kp.Cb = a->Cb;
  kp.out = a->dw;
  const int base = kp.n_groups * kp.a_chunks * kp.m_tiles;
  // one accumulator set per CTA: one round of equal-length items over the SMs (the epilogue of an item is not
  // overlapped with the next item's MMAs)
  static int rounds = []() { const char* e = getenv("B2G_WGRAD_SLAB_ROUNDS"); int r = e ? atoi(e) : 1; return r > 0 ? r : 1; }();
  int sp = rounds * num_sms() / base;
  if (sp > kp.ptiles / 8) sp = kp.ptiles / 8;
  if (sp < 1) sp = 1;
  kp.splits = sp;

  const size_t es = 2;
  const char* sbase = static_cast<const char*>(a->shifted);
  const uint32_t sbox[4] = {64u, (uint32_t)kp.slab_px, 1u, 1u};
  if (pl.n_amaps == 1) {
    const uint64_t dims[4] = {(uint64_t)a->Ca, (uint64_t)a->Ws, (uint64_t)a->Hs, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Ca * es, (uint64_t)a->Ws * a->Ca * es, (uint64_t)a->Hs * a->Ws * a->Ca * es};
    for (int i = 0; i < 4; ++i) {
      int rc = make_tmap(&kp.smap[i], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, sbase, dims, str, sbox);
      if (rc) return rc;
    }
  } else {
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        const uint64_t dims[4] = {(uint64_t)a->Ca, (uint64_t)a->Ws / 2, (uint64_t)a->Hs / 2, (uint64_t)a->N};
        const uint64_t str[3] = {(uint64_t)2 * a->Ca * es, (uint64_t)2 * a->Ws * a->Ca * es,
                                 (uint64_t)a->Hs * a->Ws * a->Ca * es};
        int rc = make_tmap(&kp.smap[py * 2 + px], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                           sbase + ((size_t)py * a->Ws + px) * a->Ca * es, dims, str, sbox);
        if (rc) return rc;
      }
  }
  {
    const uint64_t dims[4] = {(uint64_t)a->Cb, (uint64_t)pl.Wg, (uint64_t)pl.Hg, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Cb * es, (uint64_t)pl.Wg * a->Cb * es, (uint64_t)pl.Hg * pl.Wg * a->Cb * es};
    const uint32_t ubox[4] = {64u, 64u, 1u, 1u};
    int rc = make_tmap(&kp.umap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a->unshifted, dims, str, ubox);
    if (rc) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(wgrad_slab_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        WgradSlabCfg::kSmemBytes));
    attr_set = true;
  }
  const int total_items = base * kp.splits;
  const int grid = total_items < num_sms() ? total_items : num_sms();
  wgrad_slab_kernel<<<grid, kWgradSlabThreads, WgradSlabCfg::kSmemBytes, static_cast<cudaStream_t>(stream_)>>>(kp);
  B2G_LAUNCH_CHECK();
  *handled = true;
  return B2G_OK;
}

}  // namespace b2g
