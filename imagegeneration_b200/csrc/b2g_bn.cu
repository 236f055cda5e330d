// b2g_bn.cu — training-mode BatchNormalization and RMSprop for the WGAN path (SURVEY.md section 8(f) row N1;
// reference: wasserstein_gan/WGAN.py:60-96 critic, :112-132 generator, :99,:150 RMSprop, :33-44 ClipConstraint).
//
// All activation tensors are bf16 NHWC viewed as [P][C] (P = B*H*W pixels, C channels innermost, C % 8 == 0 and
// 256 % (C/8) == 0 so that a thread of a 256-thread block owns the same 8 channels on every grid-stride step and
// keeps their per-channel constants in registers).  Every kernel is a pure HBM stream (16-byte loads/stores):
//   bn_stats      : sums[q][c] = sum x, sums[q][C+c] = sum x^2 (8 partial copies q)          reads  2 B/elt
//   bn_finalize   : mean, 1/sqrt(var_biased+eps); moving stats (momentum, Bessel-corrected variance)
//   bn_apply_act  : y = leaky(gamma (x-mean) invstd + beta)                            reads 2, writes 2 B/elt
//   bn_bwd_reduce : red[c] = sum g, red[C+c] = sum g (x-mean)                          reads  4 B/elt
//   bn_bwd_apply  : dx = gamma invstd (g - mean(g) - xhat mean(g xhat)); dgamma, dbeta reads 4, writes 2 B/elt
// (g = dL/d(BN output) with the LeakyReLU derivative already applied by the producing conv epilogue.)
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  return make_uint4(pack2(f[0], f[1]), pack2(f[2], f[3]), pack2(f[4], f[5]), pack2(f[6], f[7]));
}

constexpr int kBnCopies = 8;  // the reduction outputs are [kBnCopies][2C]; their consumers add the copies

// two per-channel sums per element stream; MODE 0: (x, x^2) ; MODE 1: (g, g*(x-mean))
template <int MODE>
__global__ void __launch_bounds__(256) bn_reduce_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b,
                                                        const float* __restrict__ stat, float* __restrict__ out,
                                                        long n8, int C) {
  extern __shared__ float sacc[];  // [2C]
  for (int i = threadIdx.x; i < 2 * C; i += 256) sacc[i] = 0.f;
  __syncthreads();
  const int groups = C / 8;
  const int cg = threadIdx.x % groups;
  float mean[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) mean[k] = MODE == 1 ? stat[cg * 8 + k] : 0.f;
  float s0[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, s1[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const long stride = (long)gridDim.x * 256;
  constexpr int U = 4;  // 16-byte loads per stream per iteration; the NEXT iteration's loads are issued before this
                        // iteration's values are consumed (a plain unrolled loop is re-serialised into load -> use pairs)
  long i0 = (long)blockIdx.x * 256 + threadIdx.x;
  uint4 na[U], nb[U];
#pragma unroll
  for (int q = 0; q < U; ++q) {
    const long i = i0 + q * stride;
    const long ic = i < n8 ? i : n8 - 1;   // clamped: unconditional loads; slots past the end are zeroed below
    na[q] = __ldg(a + ic);
    if (MODE == 1) nb[q] = __ldg(b + ic);
  }
  while (i0 < n8) {
    uint4 va[U], vb[U];
#pragma unroll
    for (int q = 0; q < U; ++q) {
      va[q] = (i0 + q * stride < n8) ? na[q] : make_uint4(0u, 0u, 0u, 0u);
      if (MODE == 1) vb[q] = nb[q];
    }
    const long i1 = i0 + U * stride;
    if (i1 < n8) {
#pragma unroll
      for (int q = 0; q < U; ++q) {
        const long i = i1 + q * stride;
        const long ic = i < n8 ? i : n8 - 1;
        na[q] = __ldg(a + ic);
        if (MODE == 1) nb[q] = __ldg(b + ic);
      }
    }
#pragma unroll
    for (int q = 0; q < U; ++q) {
      float fa[8];
      unpack8(va[q], fa);   // out-of-range slots are zeros: they add nothing to either sum
      if (MODE == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          s0[k] += fa[k];
          s1[k] = fmaf(fa[k], fa[k], s1[k]);
        }
      } else {
        float fb[8];
        unpack8(vb[q], fb);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          s0[k] += fa[k];
          s1[k] = fmaf(fa[k], fb[k] - mean[k], s1[k]);
        }
      }
    }
    i0 = i1;
  }
  // lanes of a warp that own the same channel group (C < 256) are summed by shuffles first; the block's totals then go
  // to one of kBnCopies copies of the output (same-address reds serialise in L2: spread them)
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    if (off >= groups) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        s0[k] += __shfl_xor_sync(0xffffffffu, s0[k], off);
        s1[k] += __shfl_xor_sync(0xffffffffu, s1[k], off);
      }
    }
  }
  if ((threadIdx.x & 31) < groups) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      atomicAdd(&sacc[cg * 8 + k], s0[k]);
      atomicAdd(&sacc[C + cg * 8 + k], s1[k]);
    }
  }
  __syncthreads();
  float* dst = out + (size_t)(blockIdx.x % kBnCopies) * 2 * C;
  for (int i = threadIdx.x; i < 2 * C; i += 256) atomicAdd(dst + i, sacc[i]);
}

__global__ void bn_finalize_kernel(const float* __restrict__ sums, float inv_p, float bessel, int C, float eps,
                                   float momentum, float* __restrict__ stat, float* __restrict__ moving_mean,
                                   float* __restrict__ moving_var, const float* __restrict__ bias) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float sx = 0.f, sxx = 0.f;
#pragma unroll
  for (int q = 0; q < kBnCopies; ++q) {
    sx += sums[q * 2 * C + c];
    sxx += sums[q * 2 * C + C + c];
  }
  const float mean = sx * inv_p;
  const float var = fmaxf(sxx * inv_p - mean * mean, 0.f);
  stat[c] = mean;
  stat[C + c] = rsqrtf(var + eps);
  if (moving_mean) {
    // `bias`: the statistics were taken of the conv output WITHOUT its bias (BatchNorm cancels it); the layer's
    // true batch mean, which the moving mean tracks, includes it
    moving_mean[c] = moving_mean[c] * momentum + (mean + (bias ? bias[c] : 0.f)) * (1.f - momentum);
    moving_var[c] = moving_var[c] * momentum + var * bessel * (1.f - momentum);
  }
}

__global__ void bn_inference_stat_kernel(const float* __restrict__ moving_mean, const float* __restrict__ moving_var,
                                         int C, float eps, float* __restrict__ stat, const float* __restrict__ bias) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  stat[c] = moving_mean[c] - (bias ? bias[c] : 0.f);
  stat[C + c] = rsqrtf(moving_var[c] + eps);
}

__global__ void __launch_bounds__(256) bn_apply_act_kernel(const uint4* __restrict__ x, const float* __restrict__ stat,
                                                           const float* __restrict__ gamma,
                                                           const float* __restrict__ beta, float alpha,
                                                           uint4* __restrict__ y, long n8, int C) {
  const int cg = threadIdx.x % (C / 8);
  float sc[8], sh[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cg * 8 + k;
    sc[k] = gamma[c] * stat[C + c];
    sh[k] = beta[c] - stat[c] * sc[k];
  }
  const long stride = (long)gridDim.x * 256;
  constexpr int U = 4;
  for (long i0 = (long)blockIdx.x * 256 + threadIdx.x; i0 < n8; i0 += U * stride) {
    uint4 v[U];
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const long i = i0 + q * stride;
      if (i < n8) v[q] = __ldg(x + i);   // (measured: faster than clamped unconditional loads here, 96 vs 116 us)
    }
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const long i = i0 + q * stride;
      if (i >= n8) break;
      float f[8];
      unpack8(v[q], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float t = fmaf(f[k], sc[k], sh[k]);
        f[k] = fmaxf(t, t * alpha);
      }
      y[i] = pack8(f);
    }
  }
}

// g and dx MAY ALIAS (wgan.py applies the backward in place over the gradient buffer): neither is __restrict__ and g is
// read with a plain load, not through the read-only (.nc) path; a thread reads element i before it writes element i.
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const uint4* g, const uint4* __restrict__ x,
                                                           const float* __restrict__ stat,
                                                           const float* __restrict__ gamma,
                                                           const float* __restrict__ red, float inv_p,
                                                           uint4* dx, float* __restrict__ dgamma,
                                                           float* __restrict__ dbeta, long n8, int C) {
  // per-channel constants once per block (threads 0..C-1 add the partial copies), then 24 shared loads per thread
  extern __shared__ float sk[];  // [3][C]: a, c1, c0
  for (int c = threadIdx.x; c < C; c += 256) {
    const float mean = stat[c], invstd = stat[C + c];
    float sg = 0.f, sgx = 0.f;  // sum g, sum g*xhat
#pragma unroll
    for (int q = 0; q < kBnCopies; ++q) {
      sg += red[q * 2 * C + c];
      sgx += red[q * 2 * C + C + c];
    }
    sgx *= invstd;
    const float ak = gamma[c] * invstd;
    const float c1k = -ak * sgx * inv_p * invstd;
    sk[c] = ak;
    sk[C + c] = c1k;
    sk[2 * C + c] = -ak * sg * inv_p - c1k * mean;
    if (blockIdx.x == 0) {
      if (dgamma) dgamma[c] = sgx;
      if (dbeta) dbeta[c] = sg;
    }
  }
  __syncthreads();
  const int cg = threadIdx.x % (C / 8);
  float a[8], c1[8], c0[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    a[k] = sk[cg * 8 + k];
    c1[k] = sk[C + cg * 8 + k];
    c0[k] = sk[2 * C + cg * 8 + k];
  }
  // plain grid-stride loop: measured 0.96 of the copy peak as is (a 2x unrolled variant with clamped loads was slower)
  for (long i = (long)blockIdx.x * 256 + threadIdx.x; i < n8; i += (long)gridDim.x * 256) {
    float fg[8], fx[8];
    unpack8(g[i], fg);
    unpack8(__ldg(x + i), fx);
#pragma unroll
    for (int k = 0; k < 8; ++k) fg[k] = fmaf(a[k], fg[k], fmaf(c1[k], fx[k], c0[k]));
    dx[i] = pack8(fg);
  }
}

// Keras-2 RMSprop (rho, no momentum, not centred) on a flat arena; the first clip_n elements are clamped to
// [-clip, clip] after the update (ClipConstraint on the critic's conv kernels, which the caller stores first).
__global__ void __launch_bounds__(256) rmsprop_kernel(float4* __restrict__ w, const float4* __restrict__ g,
                                                      float4* __restrict__ rms, long n4, float lr, float rho,
                                                      float eps, float gscale, long clip_n4, float clip) {
  const float c = 1.f - rho;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n4; i += gridDim.x * 256L) {
    float4 gw = g[i], rw = rms[i], ww = w[i];
    float* gp = &gw.x; float* rp = &rw.x; float* wp = &ww.x;
    const bool do_clip = i < clip_n4;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gg = gp[k] * gscale;
      rp[k] = rho * rp[k] + c * gg * gg;
      wp[k] -= lr * gg / (sqrtf(rp[k]) + eps);
      if (do_clip) wp[k] = fminf(fmaxf(wp[k], -clip), clip);
    }
    rms[i] = rw;
    w[i] = ww;
  }
}

static int check_pc(const char* fn, long P, int C) {
  if (P <= 0 || C < 8 || C % 8 || C > 2048 || 256 % (C / 8))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "%s: P=%ld C=%d unsupported (C in 8,16,...,2048, powers of two)", fn, P, C);
  return B2G_OK;
}
static int stream_grid(long n8, int per_sm = 8) {
  long blocks = (n8 + 255) / 256;
  const long cap = (long)num_sms() * per_sm;
  if (blocks > cap) blocks = cap;
  return (int)(blocks < 1 ? 1 : blocks);
}

}  // namespace b2g

using namespace b2g;

extern "C" int b2g_bn_stats(const void* x, float* sums, long P, int C, b2g_stream_t stream_) {
  if (!x || !sums) return fail(B2G_E_INVALID_ARG, "b2g_bn_stats: null pointer");
  if (int rc = check_pc("b2g_bn_stats", P, C)) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  B2G_CUDA_CHECK(cudaMemsetAsync(sums, 0, kBnCopies * 2 * C * sizeof(float), stream));
  const long n8 = P * (C / 8);
  bn_reduce_kernel<0><<<stream_grid(n8, 4), 256, 2 * C * sizeof(float), stream>>>(
      static_cast<const uint4*>(x), nullptr, nullptr, sums, n8, C);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_bn_finalize(const float* sums, long P, int C, float eps, float momentum, float* stat,
                               float* moving_mean, float* moving_var, const float* bias, b2g_stream_t stream_) {
  if (!sums || !stat || (!moving_mean) != (!moving_var)) return fail(B2G_E_INVALID_ARG, "b2g_bn_finalize: null pointer");
  if (P <= 0 || C <= 0) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_bn_finalize: P=%ld C=%d", P, C);
  const float bessel = P > 1 ? (float)((double)P / (double)(P - 1)) : 1.f;
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream_)>>>(
      sums, (float)(1.0 / (double)P), bessel, C, eps, momentum, stat, moving_mean, moving_var, bias);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_bn_inference_stat(const float* moving_mean, const float* moving_var, int C, float eps, float* stat,
                                     const float* bias, b2g_stream_t stream_) {
  if (!moving_mean || !moving_var || !stat || C <= 0) return fail(B2G_E_INVALID_ARG, "b2g_bn_inference_stat: bad argument");
  bn_inference_stat_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream_)>>>(moving_mean, moving_var, C,
                                                                                            eps, stat, bias);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_bn_apply_act(const void* x, const float* stat, const float* gamma, const float* beta, float alpha,
                                void* y, long P, int C, b2g_stream_t stream_) {
  if (!x || !stat || !gamma || !beta || !y) return fail(B2G_E_INVALID_ARG, "b2g_bn_apply_act: null pointer");
  if (alpha < 0.f || alpha > 1.f) return fail(B2G_E_INVALID_ARG, "b2g_bn_apply_act: alpha must be in [0, 1]");
  if (int rc = check_pc("b2g_bn_apply_act", P, C)) return rc;
  const long n8 = P * (C / 8);
  bn_apply_act_kernel<<<stream_grid(n8), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const uint4*>(x), stat, gamma, beta, alpha, static_cast<uint4*>(y), n8, C);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_bn_bwd(const void* g, const void* x, const float* stat, const float* gamma, float* red, void* dx,
                          float* dgamma, float* dbeta, long P, int C, b2g_stream_t stream_) {
  if (!g || !x || !stat || !gamma || !red || !dx) return fail(B2G_E_INVALID_ARG, "b2g_bn_bwd: null pointer");
  if (int rc = check_pc("b2g_bn_bwd", P, C)) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  B2G_CUDA_CHECK(cudaMemsetAsync(red, 0, kBnCopies * 2 * C * sizeof(float), stream));
  const long n8 = P * (C / 8);
  bn_reduce_kernel<1><<<stream_grid(n8, 2), 256, 2 * C * sizeof(float), stream>>>(
      static_cast<const uint4*>(g), static_cast<const uint4*>(x), stat, red, n8, C);
  B2G_LAUNCH_CHECK();
  bn_bwd_apply_kernel<<<stream_grid(n8), 256, 3 * C * sizeof(float), stream>>>(static_cast<const uint4*>(g), static_cast<const uint4*>(x),
                                                           stat, gamma, red, (float)(1.0 / (double)P),
                                                           static_cast<uint4*>(dx), dgamma, dbeta, n8, C);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_rmsprop(float* w, const float* g, float* rms, long n, float lr, float rho, float eps, float gscale,
                           long clip_n, float clip, b2g_stream_t stream_) {
  if (!w || !g || !rms) return fail(B2G_E_INVALID_ARG, "b2g_rmsprop: null pointer");
  if (n % 4 || clip_n % 4 || clip_n < 0 || clip_n > n)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_rmsprop: n and clip_n must be multiples of 4 (pad the arena), clip_n <= n");
  if (n == 0) return B2G_OK;
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  const long cap = (long)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  rmsprop_kernel<<<(int)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<float4*>(w), reinterpret_cast<const float4*>(g), reinterpret_cast<float4*>(rms), n4, lr, rho, eps,
      gscale, clip_n / 4, clip);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

// ---------------------------------------------------------------------------------------------------------
// AveragePooling2D(pool_size=(8, 8)) (stride 8, 'valid') over a bf16 NHWC activation -> fp32 [N][H/8][W/8][C]:
// the feature extractor of the reference's FID-like score (sndcgan/generator_evaluation.py:134-138), SURVEY 8(f) N3.
// One thread per (output pixel, 8-channel group): 64 16-byte loads, fp32 accumulation.
// ---------------------------------------------------------------------------------------------------------
namespace b2g {
__global__ void __launch_bounds__(256) avgpool8_kernel(const uint4* __restrict__ x, float4* __restrict__ out, int N,
                                                       int H, int W, int C) {
  const int groups = C / 8, ph = H / 8, pw = W / 8;
  const long total = (long)N * ph * pw * groups;
  for (long t = blockIdx.x * 256L + threadIdx.x; t < total; t += gridDim.x * 256L) {
    const int cg = (int)(t % groups);
    long r = t / groups;
    const int ox = (int)(r % pw);
    r /= pw;
    const int oy = (int)(r % ph);
    const int n = (int)(r / ph);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int dy = 0; dy < 8; ++dy) {
      const uint4* row = x + (((long)n * H + oy * 8 + dy) * W + ox * 8) * groups + cg;
#pragma unroll
      for (int dx = 0; dx < 8; ++dx) {
        float f[8];
        unpack8(__ldg(row + (long)dx * groups), f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += f[k];
      }
    }
    const float s = 1.f / 64.f;
    out[t * 2] = make_float4(acc[0] * s, acc[1] * s, acc[2] * s, acc[3] * s);
    out[t * 2 + 1] = make_float4(acc[4] * s, acc[5] * s, acc[6] * s, acc[7] * s);
  }
}
}  // namespace b2g

extern "C" int b2g_avgpool8_bf16(const void* x, float* out, int N, int H, int W, int C, b2g_stream_t stream_) {
  if (!x || !out) return fail(B2G_E_INVALID_ARG, "b2g_avgpool8_bf16: null pointer");
  if (N <= 0 || H < 8 || W < 8 || C < 8 || C % 8)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_avgpool8_bf16: N=%d H=%d W=%d C=%d (H, W >= 8; C %% 8 == 0)", N, H, W, C);
  const long total = (long)N * (H / 8) * (W / 8) * (C / 8);
  long blocks = (total + 255) / 256;
  if (blocks > num_sms() * 8L) blocks = num_sms() * 8L;
  avgpool8_kernel<<<(int)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const uint4*>(x), reinterpret_cast<float4*>(out), N, H, W, C);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
