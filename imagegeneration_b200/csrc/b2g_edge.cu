// b2g_edge.cu — the two 3x3 stride-1 edge layers with a 3-channel side: D.conv1 (3 -> 64,
// sndcgan/SNDCGAN.py:73-78) and G.convT4 (64 -> 3 + tanh, :57-61), all three roles each.  These are
// HBM/CUDA-core bound (arithmetic intensity ~26 flop/B), so they are direct convolutions, not tensor-core GEMMs.
//
// All six roles share one weight index formula over the Keras kernel w[ky][kx][c3][c64]
// (Conv2D HWIO with Cin=3, Conv2DTranspose HWOI with Cout=3):
//   c3 -> c64, taps not flipped : D.conv1 forward ; G.convT4 input-gradient
//   c64 -> c3, taps flipped     : G.convT4 forward (+tanh) ; D.conv1 input-gradient (+tanh' of the generator)
//   wgrad [3][3][3][64]         : sum_pix t3[pix + (ky-1,kx-1)][c3] * t64[pix][c64]   (both layers)
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

// ---------------------------------------------------------------------------------------------------------
// 3 -> 64 : thread per output pixel, 64 fp32 accumulators, weights broadcast from shared memory
// ---------------------------------------------------------------------------------------------------------
template <typename T64>
__global__ void __launch_bounds__(128) conv3x3_c3_to_c64_kernel(
    const float* __restrict__ in,  // [N,H,W,3] fp32
    const float* __restrict__ w,   // [3][3][3][64] fp32
    const float* __restrict__ wscale, const float* __restrict__ bias, T64* __restrict__ out, int N, int H,
    int W, int act, float alpha, int drop_thresh8, float drop_scale, uint32_t seed_lo, uint32_t seed_hi,
    const T64* __restrict__ yprev, int dact) {
  __shared__ float4 ws[27 * 16];  // [tap*3+c3][16 float4 = 64 c64]
  __shared__ float bs[64];
  const float sc = wscale ? *wscale : 1.f;
  for (int i = threadIdx.x; i < 27 * 64; i += blockDim.x) reinterpret_cast<float*>(ws)[i] = w[i] * sc;
  if (threadIdx.x < 64) bs[threadIdx.x] = bias ? bias[threadIdx.x] : 0.f;
  __syncthreads();
  const long total = (long)N * H * W;
  for (long pix = blockIdx.x * (long)blockDim.x + threadIdx.x; pix < total; pix += (long)gridDim.x * blockDim.x) {
    const int x = (int)(pix % W);
    const int y = (int)((pix / W) % H);
    const long nbase = (pix / ((long)W * H)) * H * W;
    float acc[64];
#pragma unroll
    for (int c = 0; c < 64; ++c) acc[c] = bs[c];
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int yy = y + ky - 1;
      if (yy < 0 || yy >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int xx = x + kx - 1;
        if (xx < 0 || xx >= W) continue;
        const float* ip = in + (nbase + (long)yy * W + xx) * 3;
        const float v0 = __ldg(ip), v1 = __ldg(ip + 1), v2 = __ldg(ip + 2);
        const float4* wt = ws + (ky * 3 + kx) * 3 * 16;
#pragma unroll
        for (int g = 0; g < 16; ++g) {
          const float4 a = wt[g], b = wt[16 + g], c = wt[32 + g];
          acc[g * 4 + 0] += v0 * a.x + v1 * b.x + v2 * c.x;
          acc[g * 4 + 1] += v0 * a.y + v1 * b.y + v2 * c.y;
          acc[g * 4 + 2] += v0 * a.z + v1 * b.z + v2 * c.z;
          acc[g * 4 + 3] += v0 * a.w + v1 * b.w + v2 * c.w;
        }
      }
    }
    if (act == B2G_ACT_LEAKY) {
#pragma unroll
      for (int c = 0; c < 64; ++c) acc[c] = acc[c] > 0.f ? acc[c] : acc[c] * alpha;
    }
    const size_t off = (size_t)pix * 64;
    if (drop_thresh8 == 128) {  // rate 0.5: one bit per element, one hash word per 32 elements (b2g_keep128)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t w = b2g_dropout_bits((uint32_t)(off >> 5) + h, seed_lo, seed_hi);
#pragma unroll
        for (int e = 0; e < 32; ++e) acc[h * 32 + e] = b2g_keep128(w, e) ? acc[h * 32 + e] * drop_scale : 0.f;
      }
    } else if (drop_thresh8 > 0) {
      const uint32_t g0 = (uint32_t)(off >> 2);
#pragma unroll
      for (int g = 0; g < 16; ++g) {
        const uint32_t bits = b2g_dropout_bits(g0 + g, seed_lo, seed_hi);
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int byte = (bits >> (8 * b)) & 0xFF;
          acc[g * 4 + b] = byte < drop_thresh8 ? 0.f : acc[g * 4 + b] * drop_scale;
        }
      }
    }
    if (dact == B2G_ACT_RELU) {
#pragma unroll
      for (int g = 0; g < 8; ++g) {
        float y[8];
        edge_load8(yprev + off + g * 8, y);
#pragma unroll
        for (int b = 0; b < 8; ++b)
          if (!(y[b] > 0.f)) acc[g * 8 + b] = 0.f;
      }
    }
#pragma unroll
    for (int g = 0; g < 8; ++g) edge_store8(out + off + g * 8, &acc[g * 8]);
  }
}

// ---------------------------------------------------------------------------------------------------------
// 64 -> 3 (flipped taps): thread per output pixel, bf16 input read as 16-byte vectors
// ---------------------------------------------------------------------------------------------------------
template <typename T64>
__global__ void __launch_bounds__(128) conv3x3_c64_to_c3_kernel(
    const T64* __restrict__ in,            // [N,H,W,64] bf16 or fp32
    const float* __restrict__ w,           // [3][3][3][64] fp32
    const float* __restrict__ wscale, float* __restrict__ out, int N, int H, int W, int act,
    const float* __restrict__ yprev, int dact) {
  __shared__ float4 ws[9 * 64];  // [tap'][c64] -> (w[c3=0], w[c3=1], w[c3=2], 0) with the taps flipped
  const float sc = wscale ? *wscale : 1.f;
  for (int i = threadIdx.x; i < 9 * 64; i += blockDim.x) {
    const int tp = i / 64, c = i % 64;
    const int t = 8 - tp;  // flipped tap (ky,kx) = (2-ky', 2-kx')
    ws[i] = make_float4(w[(t * 3 + 0) * 64 + c] * sc, w[(t * 3 + 1) * 64 + c] * sc, w[(t * 3 + 2) * 64 + c] * sc, 0.f);
  }
  __syncthreads();
  const long total = (long)N * H * W;
  for (long pix = blockIdx.x * (long)blockDim.x + threadIdx.x; pix < total; pix += (long)gridDim.x * blockDim.x) {
    const int x = (int)(pix % W);
    const int y = (int)((pix / W) % H);
    const long nbase = (pix / ((long)W * H)) * H * W;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int yy = y + ky - 1;
      if (yy < 0 || yy >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int xx = x + kx - 1;
        if (xx < 0 || xx >= W) continue;
        const T64* ip = in + (nbase + (long)yy * W + xx) * 64;
        const float4* wt = ws + (ky * 3 + kx) * 64;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          float v[8];
          edge_load8(ip + g * 8, v);
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const float lo = v[b * 2], hi = v[b * 2 + 1];
            const float4 wl = wt[g * 8 + b * 2], wh = wt[g * 8 + b * 2 + 1];
            a0 += lo * wl.x + hi * wh.x;
            a1 += lo * wl.y + hi * wh.y;
            a2 += lo * wl.z + hi * wh.z;
          }
        }
      }
    }
    if (act == B2G_ACT_TANH) {
      a0 = tanhf(a0);
      a1 = tanhf(a1);
      a2 = tanhf(a2);
    }
    if (dact == B2G_ACT_TANH) {
      const float y0 = yprev[pix * 3], y1 = yprev[pix * 3 + 1], y2 = yprev[pix * 3 + 2];
      a0 *= 1.f - y0 * y0;
      a1 *= 1.f - y1 * y1;
      a2 *= 1.f - y2 * y2;
    }
    out[pix * 3] = a0;
    out[pix * 3 + 1] = a1;
    out[pix * 3 + 2] = a2;
  }
}

// ---------------------------------------------------------------------------------------------------------
// wgrad [27][64] (+ optional per-c64 column sum = bias gradient): register-tiled CUDA-core reduction.
// Block = 256 threads = 8 warps; each warp reduces its own pixels; lane (tg, cg) owns an 8-row x 8-col tile of
// the 32(27 padded) x 64 result; pixels are staged in shared memory as im2col rows [pix][32] + [pix][64].
// ---------------------------------------------------------------------------------------------------------
constexpr int kWgPix = 128;  // pixels staged per block iteration

template <typename T64>
__global__ void __launch_bounds__(256) wgrad3x3_c3_c64_kernel(const float* __restrict__ t3,  // [N,H,W,3] fp32
                                                              const T64* __restrict__ t64,   // [N,H,W,64] bf16 / fp32
                                                              float* __restrict__ dw,     // [27][64] accumulate
                                                              float* __restrict__ dbias,  // [64] accumulate or NULL
                                                              int N, int H, int W) {
  __shared__ __align__(16) float a_s[kWgPix][32];  // im2col of t3 (27 used, col 27 = 1.0 for the bias sum)
  __shared__ __align__(16) float b_s[kWgPix][64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tg = lane >> 3, cg = lane & 7;  // rows tg*8..+7 of 32, cols cg*8..+7 of 64
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  const long total = (long)N * H * W;
  const long nchunks = (total + kWgPix - 1) / kWgPix;
  for (long chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
    const long p0 = chunk * kWgPix;
    __syncthreads();
    for (int i = threadIdx.x; i < kWgPix * 32; i += blockDim.x) {
      const int pl = i >> 5, col = i & 31;
      const long pix = p0 + pl;
      float v = 0.f;
      if (pix < total) {
        if (col < 27) {
          const int tap = col / 3, c3 = col % 3;
          const int x = (int)(pix % W) + tap % 3 - 1;
          const int y = (int)((pix / W) % H) + tap / 3 - 1;
          if (x >= 0 && x < W && y >= 0 && y < H) {
            const long nb = (pix / ((long)W * H)) * H * W;
            v = __ldg(t3 + (nb + (long)y * W + x) * 3 + c3);
          }
        } else if (col == 27) {
          v = 1.f;
        }
      }
      a_s[pl][col] = v;
    }
    for (int i = threadIdx.x; i < kWgPix * 8; i += blockDim.x) {
      const int pl = i >> 3, g = i & 7;
      const long pix = p0 + pl;
      float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (pix < total) edge_load8(t64 + pix * 64 + g * 8, f);
      *reinterpret_cast<float4*>(&b_s[pl][g * 8]) = make_float4(f[0], f[1], f[2], f[3]);
      *reinterpret_cast<float4*>(&b_s[pl][g * 8 + 4]) = make_float4(f[4], f[5], f[6], f[7]);
    }
    __syncthreads();
    for (int pl = warp; pl < kWgPix; pl += 8) {
      const float4 a0 = *reinterpret_cast<const float4*>(&a_s[pl][tg * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&a_s[pl][tg * 8 + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&b_s[pl][cg * 8]);
      const float4 b1 = *reinterpret_cast<const float4*>(&b_s[pl][cg * 8 + 4]);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] += av[i] * bv[j];
    }
  }
  // cross-warp reduction through shared memory (reuse b_s: 8 warps x 32 x 64 floats = 64 KB is too much,
  // so reduce in two halves of 4 warps through a_s/b_s sized buffers)
  __syncthreads();
  float* red = &b_s[0][0];  // 128*64 floats = 4 warps x (32 x 64)
  for (int half = 0; half < 2; ++half) {
    if ((warp >> 2) == half) {
      float* dst = red + (warp & 3) * 2048;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) dst[(tg * 8 + i) * 64 + cg * 8 + j] = acc[i][j];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) {
      const float s = red[i] + red[2048 + i] + red[4096 + i] + red[6144 + i];
      const int row = i >> 6, col = i & 63;
      if (row < 27) atomicAdd(dw + row * 64 + col, s);
      else if (row == 27 && dbias) atomicAdd(dbias + col, s);
    }
    __syncthreads();
  }
}

}  // namespace b2g

using namespace b2g;

static int edge_grid(long pixels, int threads) {
  long blocks = (pixels + threads - 1) / threads;
  const long cap = (long)num_sms() * 16;
  return (int)(blocks < cap ? blocks : cap);
}

static int conv3x3_c3_to_c64_any(const float* in, const float* w, const float* wscale, const float* bias,
                                 void* out, int N, int H, int W, int act, float alpha, int dropout_thresh8,
                                 unsigned long long dropout_seed, const void* yprev, int dact, bool f32,
                                 b2g_stream_t stream_) {
  if (!in || !w || !out) return fail(B2G_E_INVALID_ARG, "b2g_conv3x3_c3_to_c64: null pointer");
  if (dact != B2G_ACT_NONE && (dact != B2G_ACT_RELU || !yprev))
    return fail(B2G_E_INVALID_ARG, "b2g_conv3x3_c3_to_c64: dact must be NONE or RELU with yprev");
  if (dropout_thresh8 < 0 || dropout_thresh8 > 255) return fail(B2G_E_INVALID_ARG, "bad dropout_thresh8");
  const long pixels = (long)N * H * W;
  const float ds = 256.f / (256.f - (float)dropout_thresh8);
  const uint32_t slo = (uint32_t)(dropout_seed & 0xFFFFFFFFull), shi = (uint32_t)(dropout_seed >> 32);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (f32)
    conv3x3_c3_to_c64_kernel<float><<<edge_grid(pixels, 128), 128, 0, stream>>>(
        in, w, wscale, bias, static_cast<float*>(out), N, H, W, act, alpha, dropout_thresh8, ds, slo, shi,
        static_cast<const float*>(yprev), dact);
  else
    conv3x3_c3_to_c64_kernel<__nv_bfloat16><<<edge_grid(pixels, 128), 128, 0, stream>>>(
        in, w, wscale, bias, static_cast<__nv_bfloat16*>(out), N, H, W, act, alpha, dropout_thresh8, ds, slo, shi,
        static_cast<const __nv_bfloat16*>(yprev), dact);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
extern "C" int b2g_conv3x3_c3_to_c64(const float* in, const float* w, const float* wscale, const float* bias,
                                     void* out, int N, int H, int W, int act, float alpha, int dropout_thresh8,
                                     unsigned long long dropout_seed, const void* yprev, int dact,
                                     b2g_stream_t stream) {
  return conv3x3_c3_to_c64_any(in, w, wscale, bias, out, N, H, W, act, alpha, dropout_thresh8, dropout_seed, yprev, dact,
                               false, stream);
}
extern "C" int b2g_conv3x3_c3_to_c64_f32(const float* in, const float* w, const float* wscale, const float* bias,
                                         float* out, int N, int H, int W, int act, float alpha, int dropout_thresh8,
                                         unsigned long long dropout_seed, const float* yprev, int dact,
                                         b2g_stream_t stream) {
  return conv3x3_c3_to_c64_any(in, w, wscale, bias, out, N, H, W, act, alpha, dropout_thresh8, dropout_seed, yprev, dact,
                               true, stream);
}

static int conv3x3_c64_to_c3_any(const void* in, const float* w, const float* wscale, float* out, int N, int H,
                                 int W, int act, const float* yprev, int dact, bool f32, b2g_stream_t stream_) {
  if (!in || !w || !out) return fail(B2G_E_INVALID_ARG, "b2g_conv3x3_c64_to_c3: null pointer");
  if (dact != B2G_ACT_NONE && (dact != B2G_ACT_TANH || !yprev))
    return fail(B2G_E_INVALID_ARG, "b2g_conv3x3_c64_to_c3: dact must be NONE or TANH with yprev");
  const long pixels = (long)N * H * W;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (f32)
    conv3x3_c64_to_c3_kernel<float><<<edge_grid(pixels, 128), 128, 0, stream>>>(static_cast<const float*>(in), w, wscale,
                                                                               out, N, H, W, act, yprev, dact);
  else
    conv3x3_c64_to_c3_kernel<__nv_bfloat16><<<edge_grid(pixels, 128), 128, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(in), w, wscale, out, N, H, W, act, yprev, dact);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
extern "C" int b2g_conv3x3_c64_to_c3(const void* in, const float* w, const float* wscale, float* out, int N, int H,
                                     int W, int act, const float* yprev, int dact, b2g_stream_t stream) {
  return conv3x3_c64_to_c3_any(in, w, wscale, out, N, H, W, act, yprev, dact, false, stream);
}
extern "C" int b2g_conv3x3_c64_to_c3_f32(const float* in, const float* w, const float* wscale, float* out, int N, int H,
                                         int W, int act, const float* yprev, int dact, b2g_stream_t stream) {
  return conv3x3_c64_to_c3_any(in, w, wscale, out, N, H, W, act, yprev, dact, true, stream);
}

static int wgrad3x3_c3_c64_any(const float* t3, const void* t64, float* dw, float* dbias, int N, int H, int W, bool f32,
                               b2g_stream_t stream_) {
  if (!t3 || !t64 || !dw) return fail(B2G_E_INVALID_ARG, "b2g_wgrad3x3_c3_c64: null pointer");
  const long chunks = ((long)N * H * W + kWgPix - 1) / kWgPix;
  const long cap = (long)num_sms() * 2;
  const int grid = (int)(chunks < cap ? chunks : cap);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (f32) wgrad3x3_c3_c64_kernel<float><<<grid, 256, 0, stream>>>(t3, static_cast<const float*>(t64), dw, dbias, N, H, W);
  else wgrad3x3_c3_c64_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(t3, static_cast<const __nv_bfloat16*>(t64), dw, dbias, N, H, W);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
extern "C" int b2g_wgrad3x3_c3_c64(const float* t3, const void* t64, float* dw, float* dbias, int N, int H, int W,
                                   b2g_stream_t stream) {
  return wgrad3x3_c3_c64_any(t3, t64, dw, dbias, N, H, W, false, stream);
}
extern "C" int b2g_wgrad3x3_c3_c64_f32(const float* t3, const float* t64, float* dw, float* dbias, int N, int H, int W,
                                       b2g_stream_t stream) {
  return wgrad3x3_c3_c64_any(t3, t64, dw, dbias, N, H, W, true, stream);
}

// ---------------------------------------------------------------------------------------------------------
// im2col-27: fp32 [N,H,W,3] -> bf16 [N,H,W,64] so that the 3-channel 3x3 layers run as K=64 GEMMs on the
// tcgen05 kernels (conv_igemm K1S1 forward / input-gradient, conv_wgrad K1S1 weight gradient).
// One thread per (pixel, 8-channel group); groups 4..7 are zero padding.
// ---------------------------------------------------------------------------------------------------------
namespace b2g {
// One block per image row (n, y): the three input rows y-1..y+1 are staged in shared memory with a zero halo, then
// every thread emits 16-byte groups of 8 output channels (coalesced stores; the 27 taps*channels of a pixel are three
// contiguous runs of 9 floats in the staged rows).  HBM-bound: 12 B read + 128 B written per pixel.
constexpr int kIm2colMaxW = 1024;
__global__ void __launch_bounds__(256) im2col27_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                       int N, int H, int W) {
  extern __shared__ float rows[];  // [3][(W + 2) * 3]
  const int row = blockIdx.x;      // n * H + y
  const int n = row / H, y = row - n * H;
  const int rs = (W + 2) * 3;
  for (int i = threadIdx.x; i < 3 * rs; i += 256) {
    const int ky = i / rs, j = i - ky * rs;  // j = (x + 1) * 3 + c
    const int yy = y + ky - 1;
    float v = 0.f;
    if (yy >= 0 && yy < H && j >= 3 && j < rs - 3) v = __ldg(in + ((size_t)(n * H + yy) * W) * 3 + (j - 3));
    rows[i] = v;
  }
  __syncthreads();
  uint4* orow = reinterpret_cast<uint4*>(out) + (size_t)row * W * 8;
  for (int idx = threadIdx.x; idx < W * 8; idx += 256) {
    const int x = idx >> 3, g = idx & 7;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (g < 4) {
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int col = g * 8 + k;         // < 32
        const int ky = (col * 57) >> 9;    // col / 9 for col < 32
        const int ky_c = ky > 2 ? 2 : ky;
        const float t = rows[ky_c * rs + x * 3 + (col - ky_c * 9)];
        v[k] = col < 27 ? t : (col == 27 ? 1.f : 0.f);
      }
      o.x = pack_bf16x2(v[0], v[1]);
      o.y = pack_bf16x2(v[2], v[3]);
      o.z = pack_bf16x2(v[4], v[5]);
      o.w = pack_bf16x2(v[6], v[7]);
    }
    orow[idx] = o;
  }
}
}  // namespace b2g

extern "C" int b2g_im2col27(const float* in, void* out, int N, int H, int W, b2g_stream_t stream_) {
  if (!in || !out) return fail(B2G_E_INVALID_ARG, "b2g_im2col27: null pointer");
  if (N <= 0 || H <= 0 || W <= 0 || W > b2g::kIm2colMaxW)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_im2col27: bad shape N=%d H=%d W=%d (W <= %d)", N, H, W, b2g::kIm2colMaxW);
  const size_t smem = (size_t)3 * (W + 2) * 3 * sizeof(float);
  b2g::im2col27_kernel<<<(unsigned)(N * H), 256, smem, static_cast<cudaStream_t>(stream_)>>>(
      in, static_cast<__nv_bfloat16*>(out), N, H, W);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
