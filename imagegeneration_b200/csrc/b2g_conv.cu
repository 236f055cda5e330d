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
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct alignas(64) ConvKernelParams {
  CUtensorMap amap[4];
  CUtensorMap bmap;
  ConvPlan plan;
  int cin_chunks;  // Cin / 64
  int n_tiles;     // Cout / BLOCK_N
  __nv_bfloat16* out;
  const float* bias;
  const __nv_bfloat16* yprev;
  int act;
  float alpha;
  int drop_thresh8;
  float drop_scale;
  uint32_t seed_lo, seed_hi;
  int dact;
  float dact_alpha, dact_scale, dact_zero;
  int out_f32;     // narrow fp32 output: `cs` real channels, out/yprev are float [pixels][cs]
  int cs;          // channel stride of the output tensor (== plan.Cout unless out_f32)
  int brows;       // rows of the packed weight matrix per phase
};

template <int BLOCK_N>
struct ConvCfg {
  static constexpr int kTileM = 128;
  static constexpr int kABytes = kTileM * 128;
  static constexpr int kBBytes = BLOCK_N * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = BLOCK_N == 256 ? 4 : (BLOCK_N == 128 ? 6 : 8);
  static constexpr int kTmemCols = 2 * BLOCK_N;  // 64 / 128 / 256 / 512: powers of two >= 32
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = kStages * kStageBytes + kBarBytes + 1024;  // +1024: manual alignment
};

// epilogue warp groups: each group is 4 warps (one per TMEM lane quarter); groups interleave over the 32-column
// chunks of a tile so that several chunks are in flight (the epilogue is latency-bound with one warp per SMSP)
template <int BLOCK_N>
constexpr int epi_groups() { return BLOCK_N >= 128 ? 4 : BLOCK_N / 32; }
template <int BLOCK_N>
constexpr int conv_threads() { return 64 + 128 * epi_groups<BLOCK_N>(); }

struct TileCoord {
  int phase, n_tile, tn, th, tw;
};
__device__ __forceinline__ TileCoord decode_tile(const ConvPlan& pl, int n_tiles, int tile) {
  TileCoord c;
  c.n_tile = tile % n_tiles;
  int rest = tile / n_tiles;
  int m = rest % pl.m_tiles;
  c.phase = rest / pl.m_tiles;
  c.tw = m % pl.tiles_w;
  int r2 = m / pl.tiles_w;
  c.th = r2 % pl.tiles_h;
  c.tn = r2 / pl.tiles_h;
  return c;
}

template <int BLOCK_N>
__global__ void __launch_bounds__(conv_threads<BLOCK_N>(), 1)
    conv_igemm_kernel(const __grid_constant__ ConvKernelParams p) {
  using Cfg = ConvCfg<BLOCK_N>;
  constexpr int kEpiGroups = epi_groups<BLOCK_N>();
  __shared__ float sbias[512];
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

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
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4 * kEpiGroups);
    }
    fence_barrier_init();
  }
  if (warp == 0 && lane == 0) {
    for (int i = 0; i < pl.n_amaps; ++i) tma_prefetch_desc(&p.amap[i]);
    tma_prefetch_desc(&p.bmap);
  }
  if (p.bias) {
    const int nb = p.out_f32 ? p.cs : pl.Cout;
    for (int i = threadIdx.x; i < nb && i < 512; i += blockDim.x) sbias[i] = p.bias[i];
  }
  if (warp == 1) tmem_alloc(tmem_slot, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const TileCoord tc = decode_tile(pl, p.n_tiles, tile);
        const int w0 = tc.tw * pl.bw, h0 = tc.th * pl.bh, n0 = tc.tn * pl.bn;
        const int brow = tc.phase * p.brows + tc.n_tile * BLOCK_N;
        for (int ks = 0; ks < k_steps; ++ks) {
          const int tap = ks / p.cin_chunks;
          const int chunk = ks - tap * p.cin_chunks;
          const int e = tc.phase * pl.taps + tap;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* a_dst = smem + stage * Cfg::kStageBytes;
          uint8_t* b_dst = a_dst + Cfg::kABytes;
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          tma_load_4d(a_dst, &p.amap[pl.tap_map[e]], &full_bar[stage], chunk * 64, w0 + pl.tap_dx[e],
                      h0 + pl.tap_dy[e], n0);
          tma_load_2d(b_dst, &p.bmap, &full_bar[stage], (tap * p.cin_chunks + chunk) * 64, brow);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(/*bf16*/ 1, 0, 0, 128, BLOCK_N);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty_bar[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BLOCK_N;
        for (int ks = 0; ks < k_steps; ++ks) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            umma_f16(d_tmem, umma_smem_desc(a_addr + k * 32, 16, 1024), umma_smem_desc(b_addr + k * 32, 16, 1024),
                     idesc, (ks | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[as]);
      }
    }
  } else {
    // ------------------------------------------------ epilogue (warp w owns TMEM lanes 32*(w%4)..+31)
    const int q = warp & 3;
    const int grp = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const int wl = row % pl.bw;
    const int hl = (row / pl.bw) % pl.bh;
    const int nl = row / (pl.bw * pl.bh);
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const TileCoord tc = decode_tile(pl, p.n_tiles, tile);
      const int gw = tc.tw * pl.bw + wl, gh = tc.th * pl.bh + hl, gn = tc.tn * pl.bn + nl;
      const bool valid = (gh < pl.Hg) && (gn < pl.N);
      const int oy = gh * pl.os + pl.ph_y[tc.phase], ox = gw * pl.os + pl.ph_x[tc.phase];
      const size_t pix = (static_cast<size_t>(gn) * pl.Hout + oy) * pl.Wout + ox;
      const int col0 = tc.n_tile * BLOCK_N;
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
#pragma unroll 1
      for (int c = grp; c < BLOCK_N / 32; c += kEpiGroups) {
        uint32_t r[32];
        __syncwarp();
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BLOCK_N + c * 32, r);
        tmem_ld_wait();
        const int cb = col0 + c * 32;
        if (p.out_f32) {  // narrow fp32 output (3 real channels): tanh / tanh' epilogues of the generator's last layer
          if (valid && cb < p.cs) {
            float* of = reinterpret_cast<float*>(p.out) + pix * p.cs + cb;
            const float* yf = reinterpret_cast<const float*>(p.yprev) + pix * p.cs + cb;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (cb + j < p.cs) {
                float o = __uint_as_float(r[j]);
                if (p.bias) o += sbias[cb + j];
                if (p.act == B2G_ACT_TANH) o = tanhf(o);
                if (p.dact == B2G_ACT_TANH) {
                  const float y = __ldg(yf + j);
                  o *= 1.f - y * y;
                }
                of[j] = o;
              }
            }
          }
          continue;
        }
        const size_t off = pix * pl.Cout + cb;  // element offset of this thread's 32 outputs
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        if (p.bias) {
#pragma unroll
          for (int j4 = 0; j4 < 8; ++j4) {
            const float4 b4 = *reinterpret_cast<const float4*>(&sbias[cb + j4 * 4]);
            v[j4 * 4 + 0] += b4.x; v[j4 * 4 + 1] += b4.y; v[j4 * 4 + 2] += b4.z; v[j4 * 4 + 3] += b4.w;
          }
        }
        if (p.act == B2G_ACT_LEAKY) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : v[j] * p.alpha;
        } else if (p.act == B2G_ACT_RELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        }
        if (p.drop_thresh8 > 0 && valid) {
          const uint32_t g0 = static_cast<uint32_t>(off >> 2);
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const uint32_t bits = b2g_dropout_bits(g0 + g, p.seed_lo, p.seed_hi);
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              const int byte = (bits >> (8 * b)) & 0xFF;
              v[g * 4 + b] = byte < p.drop_thresh8 ? 0.f : v[g * 4 + b] * p.drop_scale;
            }
          }
        }
        if (p.dact != B2G_ACT_NONE && valid) {
          const uint4* yp = reinterpret_cast<const uint4*>(p.yprev + off);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint4 yv = __ldg(yp + g);
            const uint32_t w4[4] = {yv.x, yv.y, yv.z, yv.w};
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              const float y0 = bf16_lo(w4[b]), y1 = bf16_hi(w4[b]);
              float f0, f1;
              if (p.dact == B2G_ACT_LEAKY) {
                f0 = y0 > 0.f ? p.dact_scale : (y0 < 0.f ? p.dact_alpha * p.dact_scale : p.dact_zero);
                f1 = y1 > 0.f ? p.dact_scale : (y1 < 0.f ? p.dact_alpha * p.dact_scale : p.dact_zero);
              } else if (p.dact == B2G_ACT_RELU) {
                f0 = y0 > 0.f ? 1.f : 0.f;
                f1 = y1 > 0.f ? 1.f : 0.f;
              } else {  // tanh
                f0 = 1.f - y0 * y0;
                f1 = 1.f - y1 * y1;
              }
              v[g * 8 + b * 2] *= f0;
              v[g * 8 + b * 2 + 1] *= f1;
            }
          }
        }
        if (valid) {
          uint4* op = reinterpret_cast<uint4*>(p.out + off);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint4 o;
            o.x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]);
            o.y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]);
            o.z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]);
            o.w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]);
            op[g] = o;
          }
        }
      }
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

// ------------------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------------------
template <int BLOCK_N>
static int launch_conv(const ConvKernelParams& kp, int total_tiles, cudaStream_t stream) {
  using Cfg = ConvCfg<BLOCK_N>;
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(conv_igemm_kernel<BLOCK_N>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::kSmemBytes));
    attr_set = true;
  }
  int grid = total_tiles < num_sms() ? total_tiles : num_sms();
  conv_igemm_kernel<BLOCK_N><<<grid, conv_threads<BLOCK_N>(), Cfg::kSmemBytes, stream>>>(kp);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

}  // namespace b2g

using namespace b2g;

extern "C" int b2g_conv_igemm(const b2g_conv_args* a, b2g_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!a || !a->in || !a->wpacked || !a->out) return fail(B2G_E_INVALID_ARG, "b2g_conv_igemm: null pointer");
  if (a->Cin % 64 || a->Cin <= 0) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: Cin=%d not a multiple of 64", a->Cin);
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

  ConvKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  if (!build_conv_plan(a->kind, a->N, a->Hin, a->Win, a->Cin, a->Cout, 128, &kp.plan))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_igemm: kind %d cannot tile N=%d H=%d W=%d", a->kind, a->N, a->Hin, a->Win);
  const ConvPlan& pl = kp.plan;
  const int block_n = a->out_f32 ? 32 : (a->Cout % 256 == 0 ? 256 : (a->Cout % 128 == 0 ? 128 : 64));
  kp.out_f32 = a->out_f32;
  kp.cs = a->out_f32 ? a->cout_store : a->Cout;
  kp.brows = a->out_f32 ? a->cout_store : a->Cout;
  kp.cin_chunks = a->Cin / 64;
  kp.n_tiles = a->Cout / block_n;

  // A tensor maps (innermost first: C, W, H, N)
  const size_t es = 2;
  const char* base = static_cast<const char*>(a->in);
  const uint32_t box[4] = {64u, (uint32_t)pl.bw, (uint32_t)pl.bh, (uint32_t)pl.bn};
  if (pl.n_amaps == 1) {
    const uint64_t dims[4] = {(uint64_t)a->Cin, (uint64_t)a->Win, (uint64_t)a->Hin, (uint64_t)a->N};
    const uint64_t str[3] = {(uint64_t)a->Cin * es, (uint64_t)a->Win * a->Cin * es, (uint64_t)a->Hin * a->Win * a->Cin * es};
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
  kp.out = static_cast<__nv_bfloat16*>(a->out);
  kp.bias = a->bias;
  kp.yprev = static_cast<const __nv_bfloat16*>(a->yprev);
  kp.act = a->act;
  kp.alpha = a->alpha;
  kp.drop_thresh8 = a->dropout_thresh8;
  kp.drop_scale = 256.f / (256.f - (float)a->dropout_thresh8);
  kp.seed_lo = (uint32_t)(a->dropout_seed & 0xFFFFFFFFull);
  kp.seed_hi = (uint32_t)(a->dropout_seed >> 32);
  kp.dact = a->dact;
  kp.dact_alpha = a->dact_alpha;
  kp.dact_scale = a->dact_scale;
  kp.dact_zero = a->dact_zero;

  const int total_tiles = pl.phases * pl.m_tiles * kp.n_tiles;
  switch (block_n) {
    case 256: return launch_conv<256>(kp, total_tiles, stream);
    case 128: return launch_conv<128>(kp, total_tiles, stream);
    case 32: return launch_conv<32>(kp, total_tiles, stream);
    default: return launch_conv<64>(kp, total_tiles, stream);
  }
}
