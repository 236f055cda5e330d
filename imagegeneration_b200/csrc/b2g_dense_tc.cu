// b2g_dense_tc.cu — the generator stem on tcgen05: Dense(128 -> F, no bias) + inference BatchNorm + ReLU and its
// backward (sndcgan/SNDCGAN.py:31-34).  F = 512 * (H/8) * (W/8) = 294 912 at 256x144, so the fp32 weight is 151 MB:
// both passes are HBM-bound streams over W (forward: W read once; backward: W read + dW written once), the tensor
// core only has to keep the 4.8 GFLOP out of the CUDA cores' way (the fp32 FMA version was compute-bound there).
//
// Orientation (both passes): GEMM M = 128 consecutive features j (TMEM lane = feature), so every global access of
// an epilogue warp is 32 consecutive features = one coalesced 64/128-byte line.
//   forward : D[j][b] = sum_k W[k][j] z[b][k]      kind::tf32; A = W tile straight from HBM by TMA (MN-major fp32:
//             128-byte swizzle with 32-byte atoms), B = z (K-major fp32), N = 64 samples, K = 128
//   backward: D[j][k] = sum_b dq[b][j] z[b][k]     kind::f16;  A = dq tile by TMA (MN-major bf16), B = z split into
//             bf16 hi + lo parts (error 2^-17) built in shared memory once, plus a ones column whose accumulator
//             column is dbeta[j]; K = 64 samples per launch, N = 128 + 16
//             epilogue per feature j: dW[k][j] = gscale[j] * D[j][k], dgamma[j] = inv_std * sum_k W[k][j] D[j][k]
#include <string.h>

#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

constexpr int kStemK = 128;  // z size (sndcgan/Trainer.py:8)
constexpr int kStemBC = 64;  // samples per pass
constexpr int kStemThreads = 64 + 256;  // TMA warp, MMA warp, 2 epilogue groups x 4 TMEM lane quarters

// ------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------
struct alignas(64) StemFwdParams {
  CUtensorMap wmap;  // W fp32 [128][F]: dims {F, 128}, box {32, 128}, 128-byte swizzle with 32-byte atoms
  CUtensorMap zmap;  // z fp32 [B][128]: dims {128, B}, box {32, 64}, 128-byte swizzle
  const float* gamma;
  const float* beta;
  float inv_std;
  float alpha;  // LeakyReLU slope of the output activation (0 = ReLU)
  __nv_bfloat16* out;
  float* out_f32;  // non-null: full-precision output (the fp32/tf32 parity path) instead of `out`
  int B, F, b0, tiles;
};

struct StemFwdCfg {
  static constexpr int kWBox = 128 * 128;        // [128 k][32 j] fp32
  static constexpr int kWStage = 4 * kWBox;      // 128 features
  static constexpr int kStages = 3;
  static constexpr int kZBytes = 4 * 64 * 128;   // 4 k-chunks of [64 b][32 k] fp32
  static constexpr int kSmemBytes = kStages * kWStage + kZBytes + 256 + 1024;
};

__global__ void __launch_bounds__(kStemThreads, 1) stem_fwd_kernel(const __grid_constant__ StemFwdParams p) {
  using Cfg = StemFwdCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* zsm = smem + Cfg::kStages * Cfg::kWStage;
  uint64_t* wfull = reinterpret_cast<uint64_t*>(zsm + Cfg::kZBytes);
  uint64_t* wempty = wfull + Cfg::kStages;
  uint64_t* tfull = wempty + Cfg::kStages;
  uint64_t* tempty = tfull + 2;
  uint64_t* zfull = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(zfull + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&wfull[s], 1);
      mbar_init(&wempty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 4);
    }
    mbar_init(zfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    const uint32_t leader = elect_one_sync();
    if (leader) {
      tma_prefetch_desc(&p.wmap);
      tma_prefetch_desc(&p.zmap);
      mbar_expect_tx(zfull, Cfg::kZBytes);
      for (int c = 0; c < 4; ++c) tma_load_2d(zsm + c * 64 * 128, &p.zmap, zfull, c * 32, p.b0);
    }
    __syncwarp();
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
      mbar_wait(&wempty[stage], phase ^ 1);
      if (leader) {
        mbar_expect_tx(&wfull[stage], Cfg::kWStage);
        for (int g = 0; g < 4; ++g)
          tma_load_2d(smem + stage * Cfg::kWStage + g * Cfg::kWBox, &p.wmap, &wfull[stage], tile * 128 + g * 32, 0);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*tf32*/ 2, /*A MN-major*/ 1, /*B K-major*/ 0, 128, kStemBC);
    const uint32_t zaddr = smem_u32(zsm);
    mbar_wait(zfull, 0);
    int stage = 0, it = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      const int as = it & 1;
      mbar_wait(&tempty[as], ((it >> 1) & 1) ^ 1);
      mbar_wait(&wfull[stage], phase);
      tc_fence_after();
      const uint64_t adesc = umma_smem_desc_sw128_32b(smem_u32(smem) + stage * Cfg::kWStage, Cfg::kWBox, 512);
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {  // K = 8 per MMA: 8 k-lines of A (1024 B), 32 bytes along a z row
          const uint64_t bdesc = umma_desc_kmajor(zaddr + (kk >> 2) * 64 * 128) + 2 * (kk & 3);
          umma_tf32(tmem_base + as * kStemBC, adesc + 64 * kk, bdesc, idesc, kk != 0);
        }
        umma_commit(&wempty[stage]);
        umma_commit(&tfull[as]);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else {
    const int q = warp & 3, grp = (warp - 2) >> 2;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      if ((it & 1) != grp) continue;  // group g drains accumulator buffer g
      const int as = it & 1;
      const int j = tile * 128 + q * 32 + lane;
      const float g = __ldg(p.gamma + j) * p.inv_std, be = __ldg(p.beta + j);
      mbar_wait_parked(&tfull[as], (it >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kStemBC + h * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const int b = p.b0 + h * 32 + c;
          if (b < p.B)  // 32 consecutive features per warp: one 64-byte line per sample
          {
            const float y = fmaf(__uint_as_float(r[c]), g, be);
            const float a = fmaxf(y, y * p.alpha);  // alpha in [0, 1)
            if (p.out_f32) p.out_f32[(size_t)b * p.F + j] = a;
            else p.out[(size_t)b * p.F + j] = __float2bfloat16_rn(a);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

// ------------------------------------------------------------------------------------------------------------
// backward (one launch per chunk of 64 samples; chunks after the first accumulate)
// ------------------------------------------------------------------------------------------------------------
struct alignas(64) StemBwdParams {
  CUtensorMap qmap;  // dq bf16 [B][F]: dims {F, B}, box {64, 64}, 128-byte swizzle
  const float* z;
  const float* W;
  const float* gamma;
  float inv_std;
  float* dW;
  float* dgamma;
  float* dbeta;
  int B, F, b0, tiles, accumulate;
};

struct StemBwdCfg {
  static constexpr int kQBox = 64 * 128;       // [64 b][64 j] bf16
  static constexpr int kQStage = 2 * kQBox;    // 128 features
  static constexpr int kStages = 4;
  static constexpr int kZBlock = 64 * 128;     // [64 b][64 k] bf16
  static constexpr int kZPart = 3 * kZBlock;   // k 0..63, k 64..127, ones block
  static constexpr int kN = kStemK + 16;       // 128 z columns + the ones column (N must be a multiple of 16)
  static constexpr int kAccCols = 256;         // TMEM columns per accumulator buffer
  static constexpr int kSmemBytes = kStages * kQStage + 2 * kZPart + 256 + 1024;
};

__global__ void __launch_bounds__(kStemThreads, 1) stem_bwd_kernel(const __grid_constant__ StemBwdParams p) {
  using Cfg = StemBwdCfg;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* zsm = smem + Cfg::kStages * Cfg::kQStage;  // hi part, then lo part
  uint64_t* qfull = reinterpret_cast<uint64_t*>(zsm + 2 * Cfg::kZPart);
  uint64_t* qempty = qfull + Cfg::kStages;
  uint64_t* tfull = qempty + Cfg::kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&qfull[s], 1);
      mbar_init(&qempty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 4);
    }
    fence_barrier_init();
  }
  // B operand, built once: z[b][k] -> bf16 hi + lo parts in the MN-major 128-byte-swizzled layout a TMA box
  // [64 b][64 k] would have; third block: column 128 = 1 (hi part only), the rest 0
  for (int i = threadIdx.x; i < 64 * 192 / 8; i += kStemThreads) {
    const int b = i / 24, c8 = i % 24;  // 8 consecutive k (one 16-byte chunk)
    const int kb = c8 / 8, ch = c8 % 8;
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float v[2] = {0.f, 0.f};
      if (kb < 2 && p.b0 + b < p.B) {
        const float2 z2 = *reinterpret_cast<const float2*>(p.z + (size_t)(p.b0 + b) * kStemK + kb * 64 + ch * 8 + e * 2);
        v[0] = z2.x;
        v[1] = z2.y;
      } else if (kb == 2 && ch == 0 && e == 0) {
        v[0] = 1.f;
      }
      const float h0 = __bfloat162float(__float2bfloat16_rn(v[0])), h1 = __bfloat162float(__float2bfloat16_rn(v[1]));
      hi[e] = pack_bf16x2(h0, h1);
      lo[e] = pack_bf16x2(v[0] - h0, v[1] - h1);
    }
    const uint32_t off = kb * Cfg::kZBlock + (b >> 3) * 1024 + (b & 7) * 128 + ((ch ^ (b & 7)) << 4);
    *reinterpret_cast<uint4*>(zsm + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(zsm + Cfg::kZPart + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc(tmem_slot, 2 * Cfg::kAccCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    const uint32_t leader = elect_one_sync();
    if (leader) tma_prefetch_desc(&p.qmap);
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
      mbar_wait(&qempty[stage], phase ^ 1);
      if (leader) {
        mbar_expect_tx(&qfull[stage], Cfg::kQStage);
        for (int g = 0; g < 2; ++g)
          tma_load_2d(smem + stage * Cfg::kQStage + g * Cfg::kQBox, &p.qmap, &qfull[stage], tile * 128 + g * 64, p.b0);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    const uint32_t leader = elect_one_sync();
    constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, /*A MN-major*/ 1, /*B MN-major*/ 1, 128, Cfg::kN);
    const uint64_t zhi = umma_smem_desc(smem_u32(zsm), Cfg::kZBlock, 1024);
    const uint64_t zlo = umma_smem_desc(smem_u32(zsm) + Cfg::kZPart, Cfg::kZBlock, 1024);
    int stage = 0, it = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      const int as = it & 1;
      mbar_wait(&tempty[as], ((it >> 1) & 1) ^ 1);
      mbar_wait(&qfull[stage], phase);
      tc_fence_after();
      const uint64_t adesc = umma_smem_desc(smem_u32(smem) + stage * Cfg::kQStage, Cfg::kQBox, 1024);
      const uint32_t d = tmem_base + as * Cfg::kAccCols;
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {  // 16 samples per MMA = 2048 bytes of both operands
          umma_f16(d, adesc + 128 * kk, zhi + 128 * kk, idesc, kk != 0);
          umma_f16(d, adesc + 128 * kk, zlo + 128 * kk, idesc, 1);
        }
        umma_commit(&qempty[stage]);
        umma_commit(&tfull[as]);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else {
    const int q = warp & 3, grp = (warp - 2) >> 2;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
      if ((it & 1) != grp) continue;
      const int as = it & 1;
      const int j = tile * 128 + q * 32 + lane;
      const float g = __ldg(p.gamma + j) * p.inv_std;
      const float* wcol = p.W + j;
      float* dwcol = p.dW + j;
      float sg = 0.f;
      mbar_wait_parked(&tfull[as], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t tbase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * Cfg::kAccCols;
#pragma unroll 1
      for (int h = 0; h < 4; ++h) {
        uint32_t r[32];
        float w[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) w[c] = __ldg(wcol + (size_t)(h * 32 + c) * p.F);  // coalesced across the warp
        __syncwarp();
        tmem_ld_32x32(tbase + h * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const float G = __uint_as_float(r[c]);
          sg = fmaf(w[c], G, sg);
          float o = G * g;
          float* dst = dwcol + (size_t)(h * 32 + c) * p.F;
          if (p.accumulate) o += *dst;
          *dst = o;
        }
      }
      {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tbase + 128, r);  // column 128 = sum_b dq[b][j]
        tmem_ld_wait();
        const float db = __uint_as_float(r[0]);
        if (p.accumulate) {
          p.dgamma[j] += sg * p.inv_std;
          p.dbeta[j] += db;
        } else {
          p.dgamma[j] = sg * p.inv_std;
          p.dbeta[j] = db;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[as]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * Cfg::kAccCols);
  }
}

}  // namespace b2g

using namespace b2g;

static int dense_g_fwd_any(const float* z, const float* W, const float* gamma, const float* beta, float inv_std,
                           float alpha, void* out, int B, int F, b2g_stream_t stream_, bool out_f32 = false) {
  if (alpha < 0.f || alpha >= 1.f) return fail(B2G_E_INVALID_ARG, "b2g_dense_g_fwd: alpha must be in [0, 1)");
  if (!z || !W || !gamma || !beta || !out) return fail(B2G_E_INVALID_ARG, "b2g_dense_g_fwd: null pointer");
  if (F % 128 || F <= 0 || B <= 0) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_dense_g_fwd: F=%d must be a multiple of 128", F);
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(stem_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, StemFwdCfg::kSmemBytes));
    attr_set = true;
  }
  StemFwdParams p;
  memset(&p, 0, sizeof(p));
  {
    const uint64_t dims[2] = {(uint64_t)F, (uint64_t)kStemK};
    const uint64_t str[1] = {(uint64_t)F * 4};
    const uint32_t box[2] = {32u, 128u};
    int rc = make_tmap(&p.wmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, W, dims, str, box, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)kStemK, (uint64_t)B};
    const uint64_t str[1] = {(uint64_t)kStemK * 4};
    const uint32_t box[2] = {32u, 64u};
    int rc = make_tmap(&p.zmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, z, dims, str, box);
    if (rc) return rc;
  }
  p.gamma = gamma; p.beta = beta; p.inv_std = inv_std; p.alpha = alpha;
  p.out = out_f32 ? nullptr : static_cast<__nv_bfloat16*>(out);
  p.out_f32 = out_f32 ? static_cast<float*>(out) : nullptr;
  p.B = B; p.F = F; p.tiles = F / 128;
  const int grid = p.tiles < num_sms() ? p.tiles : num_sms();
  for (int b0 = 0; b0 < B; b0 += kStemBC) {
    p.b0 = b0;
    stem_fwd_kernel<<<grid, kStemThreads, StemFwdCfg::kSmemBytes, static_cast<cudaStream_t>(stream_)>>>(p);
    B2G_LAUNCH_CHECK();
  }
  return B2G_OK;
}

extern "C" int b2g_dense_g_fwd(const float* z, const float* W, const float* gamma, const float* beta, float inv_std,
                               void* out, int B, int F, b2g_stream_t stream) {
  return dense_g_fwd_any(z, W, gamma, beta, inv_std, 0.f, out, B, F, stream);
}
extern "C" int b2g_dense_g_fwd_leaky(const float* z, const float* W, const float* gamma, const float* beta,
                                     float inv_std, float alpha, void* out, int B, int F, b2g_stream_t stream) {
  return dense_g_fwd_any(z, W, gamma, beta, inv_std, alpha, out, B, F, stream);
}

extern "C" int b2g_dense_g_fwd_f32(const float* z, const float* W, const float* gamma, const float* beta, float inv_std,
                                   float* out, int B, int F, b2g_stream_t stream) {
  return dense_g_fwd_any(z, W, gamma, beta, inv_std, 0.f, out, B, F, stream, true);
}

// `first` = false: accumulate onto dW / dgamma / dbeta from the first sample chunk on (second pass of the fp32 variant)
static int dense_g_bwd_any(const float* z, const float* W, const float* gamma, float inv_std, const void* dq,
                           float* dW, float* dgamma, float* dbeta, int B, int F, bool first, b2g_stream_t stream_) {
  if (!z || !W || !gamma || !dq || !dW || !dgamma || !dbeta) return fail(B2G_E_INVALID_ARG, "b2g_dense_g_bwd: null pointer");
  if (F % 128 || F <= 0 || B <= 0) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_dense_g_bwd: F=%d must be a multiple of 128", F);
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(stem_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, StemBwdCfg::kSmemBytes));
    attr_set = true;
  }
  StemBwdParams p;
  memset(&p, 0, sizeof(p));
  {
    const uint64_t dims[2] = {(uint64_t)F, (uint64_t)B};
    const uint64_t str[1] = {(uint64_t)F * 2};
    const uint32_t box[2] = {64u, 64u};
    int rc = make_tmap(&p.qmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dq, dims, str, box);
    if (rc) return rc;
  }
  p.z = z; p.W = W; p.gamma = gamma; p.inv_std = inv_std;
  p.dW = dW; p.dgamma = dgamma; p.dbeta = dbeta;
  p.B = B; p.F = F; p.tiles = F / 128;
  const int grid = p.tiles < num_sms() ? p.tiles : num_sms();
  for (int b0 = 0; b0 < B; b0 += kStemBC) {
    p.b0 = b0;
    p.accumulate = b0 > 0 || !first;
    stem_bwd_kernel<<<grid, kStemThreads, StemBwdCfg::kSmemBytes, static_cast<cudaStream_t>(stream_)>>>(p);
    B2G_LAUNCH_CHECK();
  }
  return B2G_OK;
}

extern "C" int b2g_dense_g_bwd(const float* z, const float* W, const float* gamma, float inv_std, const void* dq,
                               float* dW, float* dgamma, float* dbeta, int B, int F, b2g_stream_t stream) {
  return dense_g_bwd_any(z, W, gamma, inv_std, dq, dW, dgamma, dbeta, B, F, true, stream);
}

namespace b2g {
// fp32 -> bf16 hi + bf16 lo (x = hi + lo up to 2^-17 relative): lets the bf16 tensor-core backward consume fp32 gradients
__global__ void __launch_bounds__(256) split_hi_lo_kernel(const float4* __restrict__ x, uint2* __restrict__ hi,
                                                          uint2* __restrict__ lo, long n4) {
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n4; i += gridDim.x * 256L) {
    const float4 v = __ldg(x + i);
    const float f[4] = {v.x, v.y, v.z, v.w};
    float h[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __bfloat162float(__float2bfloat16_rn(f[k]));
    hi[i] = make_uint2(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]));
    lo[i] = make_uint2(pack_bf16x2(f[0] - h[0], f[1] - h[1]), pack_bf16x2(f[2] - h[2], f[3] - h[3]));
  }
}
}  // namespace b2g

// fp32 gradient dq (the fp32/tf32 parity path): split into bf16 hi + lo in `scratch` (2 * B * F bf16 elements), then two
// passes of the tensor-core backward, the second accumulating — products exact to 2^-17.
extern "C" int b2g_dense_g_bwd_f32(const float* z, const float* W, const float* gamma, float inv_std, const float* dq,
                                   void* scratch, float* dW, float* dgamma, float* dbeta, int B, int F,
                                   b2g_stream_t stream_) {
  if (!dq || !scratch) return fail(B2G_E_INVALID_ARG, "b2g_dense_g_bwd_f32: null pointer");
  if (F % 128 || F <= 0 || B <= 0) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_dense_g_bwd_f32: F=%d must be a multiple of 128", F);
  const long n4 = (long)B * F / 4;
  __nv_bfloat16* hi = static_cast<__nv_bfloat16*>(scratch);
  __nv_bfloat16* lo = hi + (size_t)B * F;
  long blocks = (n4 + 255) / 256;
  if (blocks > num_sms() * 16L) blocks = num_sms() * 16L;
  split_hi_lo_kernel<<<(int)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<const float4*>(dq), reinterpret_cast<uint2*>(hi), reinterpret_cast<uint2*>(lo), n4);
  B2G_LAUNCH_CHECK();
  int rc = dense_g_bwd_any(z, W, gamma, inv_std, hi, dW, dgamma, dbeta, B, F, true, stream_);
  if (rc) return rc;
  return dense_g_bwd_any(z, W, gamma, inv_std, lo, dW, dgamma, dbeta, B, F, false, stream_);
}
