// b2g_cyclegan.cu — the non-GEMM kernels of the CycleGAN path (cyclegan/CycleGAN.py:60-208).
//
// Every convolution of that model (7x7 s1 'same', 3x3 s2 'valid' behind a reflection pad, 3x3 s2 'same' transposed,
// 3x3 s1 residual convs, 4x4 s2 / s1 'valid' PatchGAN convs, on image sizes like 63, 30, 14, 6 that no 128-pixel TMA box
// tiles) runs as   gather (im2col, bf16, this file)  ->  tcgen05 GEMM (b2g_conv_igemm K1S1 / b2g_conv_wgrad K1S1).
// One gather kernel covers forward convs, transposed convs and both kinds of input gradients:
//   mode 0 (correlation):  row (n, oy, ox), column (ky, kx, c)  <-  in[n, oy*s + ky - pad_t, ox*s + kx - pad_l, c]
//   mode 1 (transposed):   row (n, oy, ox), column (ky, kx, c)  <-  in[n, (oy + pad_t - ky)/s, (ox + pad_l - kx)/s, c]
//                          when both divisions are exact; zero otherwise (and outside the image).
// This is the FIRST correct CUDA path of that row, not its speed-of-light form: the gathered matrix makes one HBM round
// trip (k*k times the activation); the SNDCGAN kernels show what the implicit-GEMM form buys.
// Also here: the matching weight packing, ReflectionPadding2D and its gradient, the row-wise "instance" normalisation of
// the reference (InstanceNormalization(axis=1) on NHWC = statistics per image ROW) forward / backward with the activations
// and the ResBlock's add + ReLU fused in, and the L1 (cycle / identity) loss.
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

struct Im2colParams {
  const void* in;
  __nv_bfloat16* out;
  int in_f32;
  int N, Hin, Win, C, ldc;  // ldc: channel stride of the input tensor (>= C: GEMM outputs padded to 64 columns)
  int Hout, Wout, kh, kw;
  int mode, stride, pad_t, pad_l;
  int reflect;  // mode 0 only: out-of-image taps mirror back inside (ReflectionPadding2D fused into the gather)
  int K, Kpad;
  long M, Mpad;
};

__device__ __forceinline__ bool im2col_src(const Im2colParams& p, int oy, int ox, int ky, int kx, int* iy, int* ix) {
  if (p.mode == 0) {
    *iy = oy * p.stride + ky - p.pad_t;
    *ix = ox * p.stride + kx - p.pad_l;
    if (p.reflect) {
      *iy = *iy < 0 ? -*iy : (*iy >= p.Hin ? 2 * p.Hin - 2 - *iy : *iy);
      *ix = *ix < 0 ? -*ix : (*ix >= p.Win ? 2 * p.Win - 2 - *ix : *ix);
    }
  } else {
    const int ty = oy + p.pad_t - ky, tx = ox + p.pad_l - kx;
    if (ty < 0 || tx < 0 || ty % p.stride || tx % p.stride) return false;
    *iy = ty / p.stride;
    *ix = tx / p.stride;
  }
  return *iy >= 0 && *iy < p.Hin && *ix >= 0 && *ix < p.Win;
}

// one thread per 16-byte group (8 consecutive columns) of the output matrix [Mpad][Kpad] (padding written as zeros)
__global__ void __launch_bounds__(256) im2col_kernel(const Im2colParams p) {
  const int gk = p.Kpad / 8;
  const long total = p.Mpad * gk;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
    const long m = i / gk;
    const int k0 = (int)(i % gk) * 8;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (m < p.M && k0 < p.K) {
      const int ox = (int)(m % p.Wout);
      const long r = m / p.Wout;
      const int oy = (int)(r % p.Hout), n = (int)(r / p.Hout);
      if (!p.in_f32 && (p.C & 7) == 0 && (p.ldc & 7) == 0) {  // 8 columns = 8 channels of one tap: one 16-byte load
        const int tap = k0 / p.C, c0 = k0 % p.C;
        int iy, ix;
        if (im2col_src(p, oy, ox, tap / p.kw, tap % p.kw, &iy, &ix))
          o = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.in) +
                                                   (((size_t)n * p.Hin + iy) * p.Win + ix) * p.ldc + c0));
      } else {
        float f[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int k = k0 + j;
          f[j] = 0.f;
          if (k < p.K) {
            const int tap = k / p.C, c = k % p.C;
            int iy, ix;
            if (im2col_src(p, oy, ox, tap / p.kw, tap % p.kw, &iy, &ix)) {
              const size_t idx = (((size_t)n * p.Hin + iy) * p.Win + ix) * p.ldc + c;
              f[j] = p.in_f32 ? __ldg(static_cast<const float*>(p.in) + idx)
                              : __bfloat162float(static_cast<const __nv_bfloat16*>(p.in)[idx]);
            }
          }
        }
        o = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
      }
    }
    reinterpret_cast<uint4*>(p.out)[i] = o;
  }
}

// GEMM B operand of a gathered convolution: out[n][tap * Ck + c] (bf16, row stride Kpad, Npad rows, zero padding) from
// the fp32 Keras kernel w[tap][D2][D3]; (n, c) = (d3, d2) when n_is_d3 (Conv2D forward, Conv2DTranspose input
// gradient), else (d2, d3) (Conv2D input gradient, Conv2DTranspose forward).
__global__ void __launch_bounds__(256) pack_gemm_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out,
                                                               int taps, int D2, int D3, int n_is_d3, int Npad, int Kpad) {
  const int Cn = n_is_d3 ? D3 : D2, Ck = n_is_d3 ? D2 : D3;
  const long total = (long)Npad * Kpad;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
    const int n = (int)(i / Kpad), k = (int)(i % Kpad);
    float v = 0.f;
    if (n < Cn && k < taps * Ck) {
      const int tap = k / Ck, c = k % Ck;
      const int d2 = n_is_d3 ? c : n, d3 = n_is_d3 ? n : c;
      v = w[((size_t)tap * D2 + d2) * D3 + d3];
    }
    out[i] = __float2bfloat16_rn(v);
  }
}

// gradient of the reflection pad: dx[y, x] = sum of dpad over the (up to 4) padded positions that mirror onto (y, x)
__global__ void __launch_bounds__(256) reflect_fold_kernel(const __nv_bfloat16* __restrict__ dpad,
                                                           __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C,
                                                           int ldc, int pad) {
  const int Ho = H + 2 * pad, Wo = W + 2 * pad;
  const long total = (long)N * H * W * C;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
    const int c = (int)(i % C);
    long r = i / C;
    const int x = (int)(r % W);
    r /= W;
    const int y = (int)(r % H), n = (int)(r / H);
    int ys[3], xs[3], ny = 0, nx = 0;
    ys[ny++] = y + pad;
    if (y >= 1 && y <= pad) ys[ny++] = pad - y;
    if (y <= H - 2 && y >= H - 1 - pad) ys[ny++] = 2 * H - 2 - y + pad;
    xs[nx++] = x + pad;
    if (x >= 1 && x <= pad) xs[nx++] = pad - x;
    if (x <= W - 2 && x >= W - 1 - pad) xs[nx++] = 2 * W - 2 - x + pad;
    float s = 0.f;
    for (int a = 0; a < ny; ++a)
      for (int b = 0; b < nx; ++b)
        s += __bfloat162float(dpad[(((size_t)n * Ho + ys[a]) * Wo + xs[b]) * ldc + c]);
    dx[i] = __float2bfloat16_rn(s);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Row-wise "instance" normalisation (InstanceNormalization(axis=1) on NHWC, cyclegan/CycleGAN.py:65,69,99,132,143,153):
// statistics over the W x C values of one image ROW (n, h), gamma / beta indexed by h, epsilon 1e-3.
//   forward : u = a (+ res, then ReLU: the ResBlock's add + ReLU in front of its second norm, :89-91);
//             y = act(gamma[h] * (u - mean) * rstd + beta[h]);   stat[2 * (n*H + h)] = {mean, rstd}
//   backward: dz = dy * act'(y); dgamma[h] += sum dz * xhat; dbeta[h] += sum dz;
//             du = rstd * gamma * (dz - mean(dz) - xhat * mean(dz * xhat))   (* [u > 0] with the fused ReLU)
// One 256-thread block per row; the row lives in shared memory between the passes (one HBM read per operand).
// `a` / `da` have a channel stride ldc >= C (GEMM outputs padded to 64 columns: the 3-channel image layers); y / dy are
// bf16 [.., C] or, for those image layers, fp32.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float block_sum_256(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) t += red[i];
  return t;
}

__device__ __forceinline__ float act_apply(int act, float z, float alpha) {
  if (act == B2G_ACT_RELU) return fmaxf(z, 0.f);
  if (act == B2G_ACT_LEAKY) return z > 0.f ? z : z * alpha;
  if (act == B2G_ACT_TANH) return tanhf(z);
  return z;
}
__device__ __forceinline__ float act_deriv_from_y(int act, float y, float alpha) {
  if (act == B2G_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (act == B2G_ACT_LEAKY) return y > 0.f ? 1.f : alpha;
  if (act == B2G_ACT_TANH) return 1.f - y * y;
  return 1.f;
}

__global__ void __launch_bounds__(256) rownorm_fwd_kernel(const __nv_bfloat16* __restrict__ a,
                                                          const __nv_bfloat16* __restrict__ res,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          void* __restrict__ y, float* __restrict__ stat, int H, int W,
                                                          int C, int ldc, int act, float alpha, float eps, int y_f32) {
  extern __shared__ float su[];  // [W * C]
  __shared__ float red[8];
  const int row = blockIdx.x, h = row % H, L = W * C;
  const size_t base_a = (size_t)row * W * ldc, base_y = (size_t)row * W * C;
  float s = 0.f;
  for (int i = threadIdx.x; i < L; i += 256) {
    const int w = i / C, c = i % C;
    float u = __bfloat162float(a[base_a + (size_t)w * ldc + c]);
    if (res) u = fmaxf(u + __bfloat162float(res[base_y + i]), 0.f);
    su[i] = u;
    s += u;
  }
  const float mean = block_sum_256(s, red) / L;
  float q = 0.f;
  for (int i = threadIdx.x; i < L; i += 256) {
    const float d = su[i] - mean;
    q += d * d;
  }
  const float rstd = rsqrtf(block_sum_256(q, red) / L + eps);
  if (threadIdx.x == 0) {
    stat[2 * row] = mean;
    stat[2 * row + 1] = rstd;
  }
  const float g = gamma[h] * rstd, b = beta[h];
  for (int i = threadIdx.x; i < L; i += 256) {
    const float o = act_apply(act, (su[i] - mean) * g + b, alpha);
    if (y_f32) static_cast<float*>(y)[base_y + i] = o;
    else static_cast<__nv_bfloat16*>(y)[base_y + i] = __float2bfloat16_rn(o);
  }
}

__global__ void __launch_bounds__(256) rownorm_bwd_kernel(const void* __restrict__ dy, const void* __restrict__ y,
                                                          const __nv_bfloat16* __restrict__ a,
                                                          const __nv_bfloat16* __restrict__ res,
                                                          const float* __restrict__ gamma, const float* __restrict__ stat,
                                                          __nv_bfloat16* __restrict__ da, float* __restrict__ dgamma,
                                                          float* __restrict__ dbeta, int H, int W, int C, int ldc, int act,
                                                          float alpha, int y_f32) {
  extern __shared__ float sm[];  // xhat [L], dz [L]
  __shared__ float red[8];
  const int row = blockIdx.x, h = row % H, L = W * C;
  float* sx = sm;
  float* sd = sm + L;
  const size_t base_a = (size_t)row * W * ldc, base_y = (size_t)row * W * C;
  const float mean = stat[2 * row], rstd = stat[2 * row + 1];
  float s1 = 0.f, s2 = 0.f;
  for (int i = threadIdx.x; i < L; i += 256) {
    const int w = i / C, c = i % C;
    float u = __bfloat162float(a[base_a + (size_t)w * ldc + c]);
    bool pos = true;
    if (res) {
      u += __bfloat162float(res[base_y + i]);
      pos = u > 0.f;
      u = fmaxf(u, 0.f);
    }
    const float xh = (u - mean) * rstd;
    const float yv = y_f32 ? static_cast<const float*>(y)[base_y + i]
                           : __bfloat162float(static_cast<const __nv_bfloat16*>(y)[base_y + i]);
    const float g = y_f32 ? static_cast<const float*>(dy)[base_y + i]
                          : __bfloat162float(static_cast<const __nv_bfloat16*>(dy)[base_y + i]);
    const float dz = g * act_deriv_from_y(act, yv, alpha);
    sx[i] = xh;
    sd[i] = pos ? dz : -dz - 1e30f;  // the ReLU mask of the fused add rides on dz: values <= -1e29 mark u <= 0
    s1 += dz;
    s2 += dz * xh;
  }
  const float t1 = block_sum_256(s1, red), t2 = block_sum_256(s2, red);
  if (threadIdx.x == 0) {
    atomicAdd(dgamma + h, t2);
    atomicAdd(dbeta + h, t1);
  }
  const float g = gamma[h] * rstd, m1 = t1 / L, m2 = t2 / L;
  for (int i = threadIdx.x; i < L; i += 256) {
    const int w = i / C, c = i % C;
    float dz = sd[i];
    const bool pos = dz > -1e29f;
    if (!pos) dz = -(dz + 1e30f);
    const float du = pos ? g * (dz - m1 - sx[i] * m2) : 0.f;
    da[base_a + (size_t)w * ldc + c] = __float2bfloat16_rn(du);
  }
  if (ldc > C)  // padding columns of the GEMM operand must be zero
    for (int i = threadIdx.x; i < W * (ldc - C); i += 256) {
      const int w = i / (ldc - C), c = C + i % (ldc - C);
      da[base_a + (size_t)w * ldc + c] = __float2bfloat16_rn(0.f);
    }
}

// loss_out[0] += scale * mean|a - b| ;  da = scale * sign(a - b) / n   (calc_cycle_loss / identity_loss, :199-207)
__global__ void __launch_bounds__(256) l1_loss_kernel(const float* __restrict__ a, const float* __restrict__ b, long n,
                                                      float scale, float* __restrict__ loss_out, float* __restrict__ da) {
  __shared__ float red[8];
  const float gs = scale / (float)n;
  float s = 0.f;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n; i += gridDim.x * 256L) {
    const float d = a[i] - b[i];
    s += fabsf(d);
    if (da) da[i] = d > 0.f ? gs : (d < 0.f ? -gs : 0.f);
  }
  const float t = block_sum_256(s, red);
  if (threadIdx.x == 0) atomicAdd(loss_out, t * gs);
}

}  // namespace b2g

using namespace b2g;

static int cg_grid(long items) {
  long blocks = (items + 255) / 256;
  const long cap = (long)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

extern "C" int b2g_im2col(const void* in, int in_f32, int N, int Hin, int Win, int C, int ldc, int kh, int kw, int mode,
                          int stride, int pad_t, int pad_l, int reflect, int Hout, int Wout, void* out, long Mpad,
                          int Kpad, b2g_stream_t stream_) {
  if (!in || !out) return fail(B2G_E_INVALID_ARG, "b2g_im2col: null pointer");
  if (N < 1 || Hin < 1 || Win < 1 || C < 1 || ldc < C || kh < 1 || kw < 1 || stride < 1 || Hout < 1 || Wout < 1 ||
      (mode != 0 && mode != 1) || (reflect && mode != 0))
    return fail(B2G_E_INVALID_ARG, "b2g_im2col: bad geometry");
  if (reflect && (pad_t >= Hin || pad_l >= Win)) return fail(B2G_E_INVALID_ARG, "b2g_im2col: reflection wider than the image");
  Im2colParams p;
  p.in = in; p.out = static_cast<__nv_bfloat16*>(out); p.in_f32 = in_f32;
  p.N = N; p.Hin = Hin; p.Win = Win; p.C = C; p.ldc = ldc;
  p.Hout = Hout; p.Wout = Wout; p.kh = kh; p.kw = kw;
  p.mode = mode; p.stride = stride; p.pad_t = pad_t; p.pad_l = pad_l; p.reflect = reflect;
  p.K = kh * kw * C; p.Kpad = Kpad;
  p.M = (long)N * Hout * Wout; p.Mpad = Mpad;
  if (Kpad % 8 || Kpad < p.K || Mpad < p.M) return fail(B2G_E_INVALID_ARG, "b2g_im2col: Kpad / Mpad too small or Kpad % 8 != 0");
  im2col_kernel<<<cg_grid(Mpad * (Kpad / 8)), 256, 0, static_cast<cudaStream_t>(stream_)>>>(p);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_pack_gemm_weight(const float* w, int taps, int D2, int D3, int n_is_d3, void* out, int Npad, int Kpad,
                                    b2g_stream_t stream_) {
  if (!w || !out) return fail(B2G_E_INVALID_ARG, "b2g_pack_gemm_weight: null pointer");
  const int Cn = n_is_d3 ? D3 : D2, Ck = n_is_d3 ? D2 : D3;
  if (Npad < Cn || Kpad < taps * Ck) return fail(B2G_E_INVALID_ARG, "b2g_pack_gemm_weight: padded sizes too small");
  pack_gemm_weight_kernel<<<cg_grid((long)Npad * Kpad), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      w, static_cast<__nv_bfloat16*>(out), taps, D2, D3, n_is_d3, Npad, Kpad);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_reflect_fold(const void* dpad, void* dx, int N, int H, int W, int C, int ldc, int pad,
                                b2g_stream_t stream_) {
  if (!dpad || !dx) return fail(B2G_E_INVALID_ARG, "b2g_reflect_fold: null pointer");
  if (pad < 1 || pad >= H || pad >= W || ldc < C) return fail(B2G_E_INVALID_ARG, "b2g_reflect_fold: bad geometry");
  reflect_fold_kernel<<<cg_grid((long)N * H * W * C), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(dpad), static_cast<__nv_bfloat16*>(dx), N, H, W, C, ldc, pad);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

constexpr int kRownormMaxSmem = 200 * 1024;

extern "C" int b2g_rownorm_fwd(const void* a, const void* res, const float* gamma, const float* beta, void* y, float* stat,
                               int N, int H, int W, int C, int ldc, int act, float alpha, float eps, int y_f32,
                               b2g_stream_t stream_) {
  if (!a || !gamma || !beta || !y || !stat) return fail(B2G_E_INVALID_ARG, "b2g_rownorm_fwd: null pointer");
  if (ldc < C || (res && ldc != C)) return fail(B2G_E_INVALID_ARG, "b2g_rownorm_fwd: bad channel stride");
  const size_t smem = (size_t)W * C * sizeof(float);
  if (smem > (size_t)kRownormMaxSmem) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_rownorm_fwd: a row of %d x %d values does not fit shared memory", W, C);
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(rownorm_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRownormMaxSmem));
    attr_set = true;
  }
  rownorm_fwd_kernel<<<N * H, 256, smem, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(res), gamma, beta, y, stat, H, W, C, ldc, act,
      alpha, eps, y_f32);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_rownorm_bwd(const void* dy, const void* y, const void* a, const void* res, const float* gamma,
                               const float* stat, void* da, float* dgamma, float* dbeta, int N, int H, int W, int C, int ldc,
                               int act, float alpha, int y_f32, b2g_stream_t stream_) {
  if (!dy || !y || !a || !gamma || !stat || !da || !dgamma || !dbeta) return fail(B2G_E_INVALID_ARG, "b2g_rownorm_bwd: null pointer");
  if (ldc < C || (res && ldc != C)) return fail(B2G_E_INVALID_ARG, "b2g_rownorm_bwd: bad channel stride");
  const size_t smem = 2 * (size_t)W * C * sizeof(float);
  if (smem > (size_t)kRownormMaxSmem) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_rownorm_bwd: a row of %d x %d values does not fit shared memory", W, C);
  static bool attr_set = false;
  if (!attr_set) {
    B2G_CUDA_CHECK(cudaFuncSetAttribute(rownorm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRownormMaxSmem));
    attr_set = true;
  }
  rownorm_bwd_kernel<<<N * H, 256, smem, static_cast<cudaStream_t>(stream_)>>>(
      dy, y, static_cast<const __nv_bfloat16*>(a), static_cast<const __nv_bfloat16*>(res), gamma, stat,
      static_cast<__nv_bfloat16*>(da), dgamma, dbeta, H, W, C, ldc, act, alpha, y_f32);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_l1_loss(const float* a, const float* b, long n, float scale, float* loss_out, float* da,
                           b2g_stream_t stream_) {
  if (!a || !b || !loss_out || n < 1) return fail(B2G_E_INVALID_ARG, "b2g_l1_loss: bad arguments");
  long blocks = (n + 255) / 256;
  if (blocks > num_sms() * 4L) blocks = num_sms() * 4L;
  l1_loss_kernel<<<(int)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(a, b, n, scale, loss_out, da);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
