// b2g_wgrad_c3.cu — weight (and bias) gradient of the 3-channel 3x3 edge layers on tcgen05 with the im2col operand
// built in shared memory (the companion of b2g_conv_c3.cu for the backward pass).
//
//   dW[tap*3 + c][b] = sum_px in[px shifted by tap][c] * U[px][b]      dbias[b] = sum_px U[px][b]
//
//   warp 0      : TMA producer of U — per 64-pixel k-step one [64 px][64 ch] bf16 box (12-deep ring)
//   warp 2      : TMA producer of the fp32 halo boxes of the 3-channel tensor ((bw+8)*3 floats x (bh+2) rows x bn
//                 images, zero fill outside the image).  Two producer warps with division-free tile iterators: a
//                 k-step moves only 10.6 KB, so the per-iteration cost of ONE issuing warp (measured ~800 clk
//                 with the decode's integer divisions) was the bound of the whole kernel
//   warp 1      : MMA issuer — 4 MMAs (K = 16 pixels each, M = 128 with the upper 64 rows a constant zero box,
//                 N = 64), MN-major operands: a pixel is one 128-byte line of either tile
//   warps 3..6  : patch builders, two groups of 64 threads taking alternate k-steps (one k-step is a serial chain
//                 wait -> 27 LDS -> pack -> 8 STS -> proxy fence -> arrive; two chains in flight hide it), one pixel
//                 per thread: 27 LDS.32 from the halo box -> one 128-byte line [27 taps*channels, 1.0, zeros] at
//                 its SW128-swizzled place (= a row of the im2col27 tensor)
//   warps 7..10 : epilogue after the CTA's last k-step: rows 0..26 -> dW (fp32 reds), row 27 -> dbias
//
// HBM traffic per pixel: 12 B + 128 B, against 12 + 128 (im2col27) + 128 + 128 (K1S1 weight-gradient GEMM) for the
// two-kernel route.  Replaces Conv2DBackpropFilter / BiasAddGrad of D.conv1 (sndcgan/SNDCGAN.py:73-78 under
// tape.gradient :256,:266) and Conv2DBackpropFilter of G.convT4 (:57-61, :248).
#include <stdlib.h>
#include <string.h>

#include "b2g_conv.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) WgradC3Params {
  CUtensorMap umap;  // U bf16 [N,H,W,64], box {64 ch, bw, bh, bn}
  CUtensorMap imap;  // fp32 [N,H,W,3] as rows of floats: dims {W*3, H, N}, box {(bw+8)*3, bh+2, bn}
  ConvPlan plan;     // K1S1 geometry, 64-pixel boxes
  int in_box_bytes;
  int ptiles;
  float* dw;         // [27][64]
  float* dbias;      // [64] or null
};

struct WgradC3Cfg {
  static constexpr int kTile = 64 * 128;      // one [64 px][64 ch] bf16 tile
  static constexpr int kStages = 4;           // built tiles; ONE constant zero tile (M rows 64..127) follows them
  static constexpr int kUStages = 12;         // U tiles: 96 KB in flight per SM (the kernel is an HBM stream over U)
  static constexpr int kInStages = 16;        // halo boxes: small, slow-to-arrive TMA boxes (8-byte elements) -> deep ring
  static constexpr int kInStageBytes = 4096;  // 64-pixel row tile: 3 x 72 x 12 = 2592 bytes
  static constexpr int kBuilderGroups = 2;
  static constexpr int kThreads = 96 + 64 * kBuilderGroups + 128;
  static constexpr int kSmemBytes = (kStages + 1 + kUStages) * kTile + kInStages * kInStageBytes + 512 + 1024;
};

__global__ void __launch_bounds__(WgradC3Cfg::kThreads, 1) wgrad_c3_kernel(const __grid_constant__ WgradC3Params p) {
  using Cfg = WgradC3Cfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* zsm = smem + Cfg::kStages * Cfg::kTile;  // the zero tile
  uint8_t* usm = zsm + Cfg::kTile;
  uint8_t* insm = usm + Cfg::kUStages * Cfg::kTile;
  uint64_t* ufull = reinterpret_cast<uint64_t*>(insm + Cfg::kInStages * Cfg::kInStageBytes);
  uint64_t* uempty = ufull + Cfg::kUStages;
  uint64_t* sfull = uempty + Cfg::kUStages;
  uint64_t* empty = sfull + Cfg::kStages;
  uint64_t* in_full = empty + Cfg::kStages;
  uint64_t* in_empty = in_full + Cfg::kInStages;
  uint64_t* tfull = in_empty + Cfg::kInStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const ConvPlan& pl = p.plan;
  // this CTA's contiguous range of 64-pixel tiles
  const int p_begin = (int)((long)blockIdx.x * p.ptiles / gridDim.x);
  const int p_end = (int)((long)(blockIdx.x + 1) * p.ptiles / gridDim.x);

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&sfull[s], 2);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < Cfg::kUStages; ++s) {
      mbar_init(&ufull[s], 1);
      mbar_init(&uempty[s], 1);
    }
    for (int s = 0; s < Cfg::kInStages; ++s) {
      mbar_init(&in_full[s], 1);
      mbar_init(&in_empty[s], 2);
    }
    mbar_init(tfull, 1);
    fence_barrier_init();
  }
  // the zero tile (M rows 64..127 of the shifted operand), written once
  for (int i = threadIdx.x; i < Cfg::kTile / 16; i += Cfg::kThreads)
    *reinterpret_cast<uint4*>(zsm + i * 16) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 || warp == 2) {
    // ------------------------------------------------ TMA producers: warp 0 = U tiles, warp 2 = fp32 halo boxes
    const uint32_t leader = elect_one_sync();
    const bool is_u = warp == 0;
    if (leader) tma_prefetch_desc(is_u ? &p.umap : &p.imap);
    // division-free iterator over this CTA's consecutive tiles
    int tw = p_begin % pl.tiles_w;
    int r2 = p_begin / pl.tiles_w;
    int th = r2 % pl.tiles_h, tn = r2 / pl.tiles_h;
    const int n_stages = is_u ? Cfg::kUStages : Cfg::kInStages;
    uint64_t* fullb = is_u ? ufull : in_full;
    uint64_t* emptyb = is_u ? uempty : in_empty;
    int stage = 0;
    uint32_t phase = 0;
    for (int pt = p_begin; pt < p_end; ++pt) {
      const int w0 = tw * pl.bw, h0 = th * pl.bh, n0 = tn * pl.bn;
      mbar_wait(&emptyb[stage], phase ^ 1);
      if (leader) {
        if (is_u) {
          mbar_expect_tx(&fullb[stage], Cfg::kTile);
          tma_load_4d(usm + stage * Cfg::kTile, &p.umap, &fullb[stage], 0, w0, h0, n0);
        } else {
          mbar_expect_tx(&fullb[stage], p.in_box_bytes);
          tma_load_3d(insm + stage * Cfg::kInStageBytes, &p.imap, &fullb[stage], (w0 - 4) * 3, h0 - 1, n0);
        }
      }
      __syncwarp();
      if (++stage == n_stages) { stage = 0; phase ^= 1; }
      if (++tw == pl.tiles_w) {
        tw = 0;
        if (++th == pl.tiles_h) { th = 0; ++tn; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, /*A MN-major*/ 1, /*B MN-major*/ 1, 128, 64);
    const uint32_t smem_base = smem_u32(smem);
    int stage = 0, ustage = 0;
    uint32_t phase = 0, uphase = 0;
    for (int pt = p_begin; pt < p_end; ++pt) {
      mbar_wait(&ufull[ustage], uphase);
      mbar_wait(&sfull[stage], phase);
      tc_fence_after();
      // atoms of the M side: built tile of this stage, then the zero tile behind the ring
      const uint64_t sdesc = umma_smem_desc(smem_base + stage * Cfg::kTile, (Cfg::kStages - stage) * Cfg::kTile, 1024);
      const uint64_t udesc = umma_smem_desc(smem_u32(usm) + ustage * Cfg::kTile, Cfg::kTile, 1024);
      if (leader) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16(tmem_base, sdesc + 128 * k, udesc + 128 * k, idesc, (pt > p_begin) || k != 0);
        umma_commit(&empty[stage]);
        umma_commit(&uempty[ustage]);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      if (++ustage == Cfg::kUStages) { ustage = 0; uphase ^= 1; }
    }
    if (leader) umma_commit(tfull);
    __syncwarp();
  } else if (warp < 3 + 2 * Cfg::kBuilderGroups) {
    // ------------------------------------------------ patch builders: thread m owns pixel (k) line m of the tile
    const int grp = (warp - 3) >> 1;
    const int m = (threadIdx.x - 96) & 63;
    const int wl = m % pl.bw, hl = (m / pl.bw) % pl.bh, nl = m / (pl.bw * pl.bh);
    const uint32_t row_off = m * 128, sw = m & 7;
    const int pitch = (pl.bw + 8) * 3;
    const int win_off = ((nl * (pl.bh + 2) + hl) * (pl.bw + 8) + wl + 3) * 3;
    for (int i = grp; i < p_end - p_begin; i += Cfg::kBuilderGroups) {
      const int stage = i % Cfg::kStages, istage = i % Cfg::kInStages;
      const uint32_t phase = (i / Cfg::kStages) & 1, iphase = (i / Cfg::kInStages) & 1;
      float v[32];
      mbar_wait_parked(&in_full[istage], iphase);
      const float* src = reinterpret_cast<const float*>(insm + istage * Cfg::kInStageBytes) + win_off;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int j = 0; j < 9; ++j) v[ky * 9 + j] = src[ky * pitch + j];
      v[27] = 1.f;  // ones channel: its row of the product is the bias gradient
#pragma unroll
      for (int k = 28; k < 32; ++k) v[k] = 0.f;
      mbar_wait_parked(&empty[stage], phase ^ 1);
      uint8_t* dst = smem + stage * Cfg::kTile + row_off;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint4 o = make_uint4(0, 0, 0, 0);
        if (j < 4) {
          o.x = pack_bf16x2(v[j * 8 + 0], v[j * 8 + 1]);
          o.y = pack_bf16x2(v[j * 8 + 2], v[j * 8 + 3]);
          o.z = pack_bf16x2(v[j * 8 + 4], v[j * 8 + 5]);
          o.w = pack_bf16x2(v[j * 8 + 6], v[j * 8 + 7]);
        }
        *reinterpret_cast<uint4*>(dst + ((j ^ sw) << 4)) = o;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&in_empty[istage]);
        mbar_arrive(&sfull[stage]);
      }
    }
  } else {
    // ------------------------------------------------ epilogue: accumulator rows = im2col channel a, columns = b
    const int q = warp & 3;
    const int a = q * 32 + lane;
    if (p_end > p_begin) {
      mbar_wait_parked(tfull, 0);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
        tmem_ld_wait();
        float* dst = nullptr;
        if (a < 27) dst = p.dw + (size_t)a * 64 + c * 32;
        else if (a == 27 && p.dbias) dst = p.dbias + c * 32;
        if (dst) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(__uint_as_float(r[j])),
                         "f"(__uint_as_float(r[j + 1])), "f"(__uint_as_float(r[j + 2])), "f"(__uint_as_float(r[j + 3]))
                         : "memory");
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

// b2g_conv_wgrad(kind = B2G_CONV_K3S1_C3): shifted = fp32 [N,Hs,Ws,3], unshifted = bf16 [N,Hs,Ws,64], dw = [27][64]
int launch_wgrad_c3(const b2g_wgrad_args* a, void* stream_) {
  if (a->Ca != 3 || a->Cb != 64 || a->tf32)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad(K3S1_C3): needs Ca=3 (fp32), Cb=64 (bf16), got %d, %d", a->Ca, a->Cb);
  WgradC3Params kp;
  memset(&kp, 0, sizeof(kp));
  if (!build_conv_plan(B2G_CONV_K1S1, a->N, a->Hs, a->Ws, 64, 64, 64, &kp.plan))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad(K3S1_C3): cannot tile N=%d H=%d W=%d", a->N, a->Hs, a->Ws);
  const ConvPlan& pl = kp.plan;
  kp.in_box_bytes = pl.bn * (pl.bh + 2) * (pl.bw + 8) * 12;
  if (pl.bw % 4 || a->Ws % 4 || kp.in_box_bytes > WgradC3Cfg::kInStageBytes)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_wgrad(K3S1_C3): cannot stage N=%d H=%d W=%d (box %d bytes)", a->N,
                a->Hs, a->Ws, kp.in_box_bytes);
  {
    // rows of (bw + 8) * 3 <= 216 floats: the box fits the 256-element limit as plain fp32
    const uint64_t dims[3] = {(uint64_t)a->Ws * 3, (uint64_t)a->Hs, (uint64_t)a->N};
    const uint64_t str[2] = {(uint64_t)a->Ws * 12, (uint64_t)a->Hs * a->Ws * 12};
    const uint32_t ibox[3] = {(uint32_t)(pl.bw + 8) * 3, (uint32_t)pl.bh + 2, (uint32_t)pl.bn};
    int rc = make_tmap(&kp.imap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, a->shifted, dims, str, ibox, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {64u, (uint64_t)a->Ws, (uint64_t)a->Hs, (uint64_t)a->N};
    const uint64_t str[3] = {128u, (uint64_t)a->Ws * 128, (uint64_t)a->Hs * a->Ws * 128};
    const uint32_t box[4] = {64u, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bn};
    int rc = make_tmap(&kp.umap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a->unshifted, dims, str, box);
    if (rc) return rc;
  }
  kp.ptiles = pl.m_tiles;
  kp.dw = a->dw;
  kp.dbias = a->dbias;
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(wgrad_c3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WgradC3Cfg::kSmemBytes));
    attr_set = true;
  }
  const int grid = kp.ptiles < num_sms() ? kp.ptiles : num_sms();
  wgrad_c3_kernel<<<grid, WgradC3Cfg::kThreads, WgradC3Cfg::kSmemBytes, static_cast<cudaStream_t>(stream_)>>>(kp);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

}  // namespace b2g
