// b2g_conv_o3.cu — 3x3 stride-1 SAME convolution from 64 channels DOWN TO 3 (fp32 output) as ONE 64 -> 27 GEMM per
// pixel plus a 9-neighbour gather in shared memory ("col2im"), instead of 9 tap GEMMs whose N = 32 tile is 29/32
// padding:
//
//   Y[p][tap*3 + co] = sum_c x[p][c] * w[tap][c][co]            (tcgen05, M = 128 pixels, N = 32, K = 64)
//   out[y][x][co]    = sum_{ky,kx} Y[(y+ky-1, x+kx-1)][(ky*3+kx)*3 + co]        (zero outside the image)
//
// The tap-by-tap formulation reads the 128 x 16 A operand from shared memory 36 times per 128 pixels (32 clk each:
// the MMA is bound by its A read, not by N); here it is read 4 times.  A CTA owns a contiguous range of image rows
// and walks down: per row one TMA load of the whole row (W/128 boxes), 4 MMAs per box, the accumulators are written
// to a 4-row ring of Y in shared memory (27 floats per pixel: conflict-free strides), and the output row above is
// produced from the three rows around it — x is read from HBM once, Y never leaves the SM.
//
//   warp 0      : TMA producer (one row of x per step, 3-stage ring)
//   warp 1      : TMEM allocator + MMA issuer (accumulators double-buffered per row)
//   warps 2..9  : epilogue: TMEM -> Y ring, named barrier, gather + bias / tanh / tanh' -> coalesced fp32 rows
//
// Replaces G.convT4 forward (Conv2DTranspose 64 -> 3 k3 s1 + tanh, sndcgan/SNDCGAN.py:57-61) and the input gradient
// of D.conv1 (:73-78 under tape.gradient :248).
#include <stdlib.h>
#include <string.h>

#include "b2g_conv.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) O3Params {
  CUtensorMap amap;       // x bf16 [N,H,W,64], box {64, 128, 1, 1}
  const __nv_bfloat16* wpacked;  // [3][9*64]: row co, column tap*64 + c (the K3S1 packing of an out_f32 call)
  const float* bias;      // [3] or null
  const float* yprev;     // fp32 [N,H,W,3] or null (dact = tanh')
  float* out;             // fp32 [N,H,W,3]
  int N, H, W, tiles_w;   // tiles_w = W / 128 (1 or 2)
  int act, dact;
};

struct O3Cfg {
  static constexpr int kTile = 128 * 128;              // [128 px][64 ch] bf16
  static constexpr int kMaxTilesW = 2;
  static constexpr int kStageBytes = kMaxTilesW * kTile;
  static constexpr int kStages = 3;
  static constexpr int kBBytes = 32 * 128;             // [32 rows (27 used)][64 ch] bf16
  static constexpr int kYRowFloats = (kMaxTilesW * 128 + 2) * 27;  // one zero pixel left and right
  static constexpr int kYSlots = 4;
  static constexpr int kYBytes = ((kYSlots * kYRowFloats * 4 + 1023) / 1024) * 1024;
  static constexpr int kEpiWarps = 8;
  static constexpr int kThreads = 64 + 32 * kEpiWarps;
  static constexpr int kSmemBytes = kStages * kStageBytes + kBBytes + kYBytes + 256 + 1024;
};

__global__ void __launch_bounds__(O3Cfg::kThreads, 1) conv_o3_kernel(const __grid_constant__ O3Params p) {
  using Cfg = O3Cfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* bsm = smem + Cfg::kStages * Cfg::kStageBytes;
  float* yring = reinterpret_cast<float*>(bsm + Cfg::kBBytes);
  uint64_t* xfull = reinterpret_cast<uint64_t*>(bsm + Cfg::kBBytes + Cfg::kYBytes);
  uint64_t* xempty = xfull + Cfg::kStages;
  uint64_t* tfull = xempty + Cfg::kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rows_total = p.N * p.H;
  // output rows [r0, r1) of this CTA; Y is computed for rows [gs, ge] = one halo row on either side
  const int r0 = (int)((long)blockIdx.x * rows_total / gridDim.x);
  const int r1 = (int)((long)(blockIdx.x + 1) * rows_total / gridDim.x);
  const int gs = r0 > 0 ? r0 - 1 : 0;
  const int ge = r1 < rows_total ? r1 : rows_total - 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&xfull[s], 1);
      mbar_init(&xempty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], Cfg::kEpiWarps);
    }
    fence_barrier_init();
  }
  // B operand, built once: row r = tap*3 + co (rows 27..31 zero), 64 channels, K-major with the 128-byte swizzle
  for (int i = threadIdx.x; i < 32 * 8; i += Cfg::kThreads) {
    const int r = i >> 3, ch = i & 7;  // 16-byte chunk ch = channels 8 ch .. 8 ch + 7
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < 27) {
      const int tap = r / 3, co = r % 3;
      v = *reinterpret_cast<const uint4*>(p.wpacked + (size_t)co * 576 + tap * 64 + ch * 8);
    }
    *reinterpret_cast<uint4*>(bsm + r * 128 + ((ch ^ (r & 7)) << 4)) = v;
  }
  // Y ring: zero everything once (the halo pixels of every slot stay zero for the whole kernel)
  for (int i = threadIdx.x; i < Cfg::kYSlots * Cfg::kYRowFloats; i += Cfg::kThreads) yring[i] = 0.f;
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer: one image row per step
    const uint32_t leader = elect_one_sync();
    if (leader) tma_prefetch_desc(&p.amap);
    int n = gs / p.H, y = gs % p.H;
    int stage = 0;
    uint32_t phase = 0;
    for (int g = gs; g <= ge; ++g) {
      mbar_wait(&xempty[stage], phase ^ 1);
      if (leader) {
        mbar_expect_tx(&xfull[stage], p.tiles_w * Cfg::kTile);
        for (int t = 0; t < p.tiles_w; ++t)
          tma_load_4d(smem + stage * Cfg::kStageBytes + t * Cfg::kTile, &p.amap, &xfull[stage], 0, t * 128, y, n);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      if (++y == p.H) { y = 0; ++n; }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, 0, 0, 128, 32);
    const uint32_t smem_base = smem_u32(smem);
    const uint64_t bdesc = umma_desc_kmajor(smem_u32(bsm));
    int stage = 0, it = 0;
    uint32_t phase = 0;
    for (int g = gs; g <= ge; ++g, ++it) {
      const int as = it & 1;
      mbar_wait(&tempty[as], ((it >> 1) & 1) ^ 1);
      mbar_wait(&xfull[stage], phase);
      tc_fence_after();
      if (leader) {
        for (int t = 0; t < p.tiles_w; ++t) {
          const uint64_t adesc = umma_desc_kmajor(smem_base + stage * Cfg::kStageBytes + t * Cfg::kTile);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16(tmem_base + as * 64 + t * 32, adesc + 2 * k, bdesc + 2 * k, idesc, k != 0);
        }
        umma_commit(&xempty[stage]);
        umma_commit(&tfull[as]);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else {
    // ------------------------------------------------ epilogue
    const int ew = warp - 2;            // 0..7
    const int q = warp & 3;             // TMEM lane quarter this warp may read
    const int tile = ew >> 2;           // which 128-pixel box of the row this warp drains
    const int px_a = tile * 128 + q * 32 + lane;        // pixel whose accumulator row this thread holds
    const int x = ew * 32 + lane;                       // pixel whose outputs this thread produces
    const int W = p.W, H = p.H;
    float bias3[3] = {0.f, 0.f, 0.f};
    if (p.bias) { bias3[0] = p.bias[0]; bias3[1] = p.bias[1]; bias3[2] = p.bias[2]; }
    int it = 0;
    for (int g = gs; g <= r1; ++g) {
      if (g <= ge) {   // A: accumulators of row g -> ring slot g % 4
        const int as = it & 1;
        mbar_wait_parked(&tfull[as], (it >> 1) & 1);
        tc_fence_after();
        if (tile < p.tiles_w) {
          uint32_t r[32];
          __syncwarp();
          tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * 64 + tile * 32, r);
          tmem_ld_wait();
          float* dst = yring + (g & 3) * Cfg::kYRowFloats + (px_a + 1) * 27;
#pragma unroll
          for (int j = 0; j < 27; ++j) dst[j] = __uint_as_float(r[j]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty[as]);
        ++it;
      }
      asm volatile("bar.sync 1, %0;" ::"n"(32 * Cfg::kEpiWarps) : "memory");  // row g is in the ring for everyone
      const int go = g - 1;             // B: output row go from the Y rows go-1, go, go+1
      if (go >= r0 && x < W) {
        const int y = go % H;
        float acc[3] = {bias3[0], bias3[1], bias3[2]};
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
          const int yy = y + ky - 1;
          if (yy < 0 || yy >= H) continue;  // uniform per row
          const float* src = yring + ((go + ky - 1) & 3) * Cfg::kYRowFloats + x * 27 + ky * 9;
#pragma unroll
          for (int kx = 0; kx < 3; ++kx) {  // pixel x + kx - 1 is stored at index x + kx
            acc[0] += src[kx * 27 + kx * 3 + 0];
            acc[1] += src[kx * 27 + kx * 3 + 1];
            acc[2] += src[kx * 27 + kx * 3 + 2];
          }
        }
        const size_t o = ((size_t)go * W + x) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float v = acc[c];
          if (p.act == B2G_ACT_TANH) v = tanhf(v);
          if (p.dact == B2G_ACT_TANH) {
            const float yv = __ldg(p.yprev + o + c);
            v *= 1.f - yv * yv;
          }
          p.out[o + c] = v;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// Takes a K3S1 call with fp32 3-channel output when the image is 128 or 256 pixels wide and has 64 input channels
// (both uses in the model); every other shape stays with the tap kernels.  B2G_CONV_O3=0 switches it off (tests).
int launch_conv_o3(const b2g_conv_args* a, void* stream_, bool* handled) {
  *handled = false;
  const char* env = getenv("B2G_CONV_O3");
  if (env && atoi(env) == 0) return B2G_OK;
  if (a->kind != B2G_CONV_K3S1 || !a->out_f32 || a->cout_store != 3 || a->Cin != 64) return B2G_OK;
  if (a->Win != 128 && a->Win != 256) return B2G_OK;
  if ((long)a->N * a->Hin < 4) return B2G_OK;
  O3Params p;
  memset(&p, 0, sizeof(p));
  const uint64_t dims[4] = {64u, (uint64_t)a->Win, (uint64_t)a->Hin, (uint64_t)a->N};
  const uint64_t str[3] = {128u, (uint64_t)a->Win * 128u, (uint64_t)a->Hin * a->Win * 128u};
  const uint32_t box[4] = {64u, 128u, 1u, 1u};
  int rc = make_tmap(&p.amap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a->in, dims, str, box);
  if (rc) return rc;
  p.wpacked = static_cast<const __nv_bfloat16*>(a->wpacked);
  p.bias = a->bias;
  p.yprev = static_cast<const float*>(a->yprev);
  p.out = static_cast<float*>(a->out);
  p.N = a->N; p.H = a->Hin; p.W = a->Win; p.tiles_w = a->Win / 128;
  p.act = a->act; p.dact = a->dact;
  static bool attr_done = false;
  if (!attr_done) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(conv_o3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, O3Cfg::kSmemBytes));
    attr_done = true;
  }
  const long rows = (long)a->N * a->Hin;
  const int grid = rows / 2 < num_sms() ? (int)(rows / 2) : num_sms();  // at least two rows per CTA
  conv_o3_kernel<<<grid, O3Cfg::kThreads, O3Cfg::kSmemBytes, static_cast<cudaStream_t>(stream_)>>>(p);
  B2G_LAUNCH_CHECK();
  *handled = true;
  return B2G_OK;
}

}  // namespace b2g
