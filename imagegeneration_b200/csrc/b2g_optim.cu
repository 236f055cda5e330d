// b2g_optim.cu — HBM-bound parameter-side kernels:
//   * fused Keras-form Adam over a flat parameter arena (replaces the 13 + 2x16 per-variable
//     Adam._resource_apply_dense ops behind optimizer.apply_gradients, sndcgan/SNDCGAN.py:249,:257,:267)
//   * per-channel column sums of a bf16 activation gradient (BiasAddGrad / BatchNorm beta gradient)
//   * finalisation of raw conv weight gradients for the folded BatchNorm scale (gamma gradient) and for
//     spectral normalisation (north-star feature, not in the reference)
//   * multi-tensor spectral-norm power iteration that reads every weight matrix exactly once.
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

// ---------------------------------------------------------------------------------------------------------
// Adam (Keras 2): m += (g-m)(1-b1); v += (g^2-v)(1-b2); w -= lr_t * m / (sqrt(v) + eps),
// lr_t = lr*sqrt(1-b2^t)/(1-b1^t) computed on the host in double.  gscale multiplies g first (1/world for DP).
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) adam_kernel(float4* __restrict__ w, const float4* __restrict__ g,
                                                   float4* __restrict__ m, float4* __restrict__ v, long n4,
                                                   float lr_t_val, const float* __restrict__ lr_t_dev, float b1,
                                                   float b2, float eps, float gscale) {
  const float lr_t = lr_t_dev ? *lr_t_dev : lr_t_val;
  const float c1 = 1.f - b1, c2 = 1.f - b2;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    float4 gw = g[i], mw = m[i], vw = v[i], ww = w[i];
    float* gp = &gw.x; float* mp = &mw.x; float* vp = &vw.x; float* wp = &ww.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gg = gp[k] * gscale;
      mp[k] += (gg - mp[k]) * c1;
      vp[k] += (gg * gg - vp[k]) * c2;
      wp[k] -= lr_t * mp[k] / (sqrtf(vp[k]) + eps);
    }
    m[i] = mw; v[i] = vw; w[i] = ww;
  }
}

// ---------------------------------------------------------------------------------------------------------
// out[c] += sum_p x[p][c]   (x bf16 [P][C], C % 8 == 0, C <= 2048)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) colsum_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out,
                                                     long P, int C) {
  extern __shared__ float sacc[];  // [C]
  const int groups = C / 8;
  const int rows_par = 256 / groups > 0 ? 256 / groups : 1;
  for (int i = threadIdx.x; i < C; i += 256) sacc[i] = 0.f;
  __syncthreads();
  const int cg = threadIdx.x % groups;
  const int r = threadIdx.x / groups;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (r < rows_par) {
    for (long p = (long)blockIdx.x * rows_par + r; p < P; p += (long)gridDim.x * rows_par) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(x + p * C) + cg);
      acc[0] += bf16_lo(v.x); acc[1] += bf16_hi(v.x); acc[2] += bf16_lo(v.y); acc[3] += bf16_hi(v.y);
      acc[4] += bf16_lo(v.z); acc[5] += bf16_hi(v.z); acc[6] += bf16_lo(v.w); acc[7] += bf16_hi(v.w);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) atomicAdd(&sacc[cg * 8 + k], acc[k]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += 256) atomicAdd(out + i, sacc[i]);
}

// ---------------------------------------------------------------------------------------------------------
// Folded-BatchNorm finalisation for a Conv2DTranspose kernel W[taps][Co][Ci] whose forward used
// W' = W * gscale[co] (gscale = gamma*inv_std):  given the raw gradient G = dL/dW' in place,
//   dgamma[co] = inv_std * sum_{tap,ci} W*G ;   G <- G * gscale[co]  (= dL/dW).   One block per co.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bn_fold_finalize_kernel(const float* __restrict__ W, float* __restrict__ G,
                                                               const float* __restrict__ gamma, float inv_std,
                                                               float* __restrict__ dgamma, int taps, int Co, int Ci) {
  const int co = blockIdx.x;
  const float gs = gamma[co] * inv_std;
  float acc = 0.f;
  for (int i = threadIdx.x; i < taps * Ci; i += 256) {
    const int t = i / Ci, ci = i % Ci;
    const size_t idx = ((size_t)t * Co + co) * Ci + ci;
    const float g = G[idx];
    acc += W[idx] * g;
    G[idx] = g * gs;
  }
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    dgamma[co] = t * inv_std;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Spectral normalisation, multi-tensor.  Layer l: W_l [K_l][C_l] row-major (Keras reshape(W, (-1, Cout))).
// Pass 1 (one launch, all layers): v_raw = W u ; vsq += |v_raw|^2 ; t_raw += W^T v_raw  — W read ONCE.
// Pass 2 (one launch, one block per layer): nv = |v_raw| ; t = t_raw/nv ; sigma = |t| ; u <- t/sigma.
// (v = v_raw/nv and W^T v = t, so sigma = v^T W u_new = |t|.)
// ---------------------------------------------------------------------------------------------------------
struct SnLayer {
  const float* W;
  float* u;       // [C]   persistent, updated by pass 2
  float* v_raw;   // [K]
  float* t_raw;   // [C]   scratch, zero on entry, re-zeroed by pass 2
  float* scal;    // [4]   {vsq scratch (zero on entry), sigma, 1/sigma, nv}
  int K, C;
};
constexpr int kSnMaxLayers = 8;
struct SnParams {
  SnLayer layer[kSnMaxLayers];
  int n_layers;
  int blocks_per_layer;
};

__global__ void __launch_bounds__(256) sn_power_kernel(const __grid_constant__ SnParams p) {
  const SnLayer& L = p.layer[blockIdx.y];
  const int K = L.K, C = L.C;
  __shared__ float tsh[512];
  __shared__ float vsq_sh;
  for (int i = threadIdx.x; i < 512; i += 256) tsh[i] = 0.f;
  if (threadIdx.x == 0) vsq_sh = 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float vsq = 0.f;
  if (C >= 32) {
    const int per = C / 32;  // columns per lane (C in {64,128,256,512} -> 2..16)
    float ureg[16], tacc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      ureg[i] = (i < per) ? L.u[lane + 32 * i] : 0.f;
      tacc[i] = 0.f;
    }
    for (int k = blockIdx.x * 8 + warp; k < K; k += gridDim.x * 8) {
      float wreg[16];
      float dot = 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        wreg[i] = (i < per) ? __ldg(L.W + (size_t)k * C + lane + 32 * i) : 0.f;
        dot += wreg[i] * ureg[i];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
      if (lane == 0) L.v_raw[k] = dot;
      vsq += (lane == 0) ? dot * dot : 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) tacc[i] += wreg[i] * dot;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (i < per) atomicAdd(&tsh[lane + 32 * i], tacc[i]);
  } else {  // C == 1 (Dense head): thread per row
    const float u0 = L.u[0];
    float t = 0.f;
    for (int k = blockIdx.x * 256 + threadIdx.x; k < K; k += gridDim.x * 256) {
      const float wv = __ldg(L.W + k);
      const float d = wv * u0;
      L.v_raw[k] = d;
      vsq += d * d;
      t += wv * d;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) atomicAdd(&tsh[0], t);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) vsq += __shfl_xor_sync(0xffffffffu, vsq, o);
  if (lane == 0) atomicAdd(&vsq_sh, vsq);
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += 256) atomicAdd(L.t_raw + i, tsh[i]);
  if (threadIdx.x == 0) atomicAdd(L.scal, vsq_sh);
}

__global__ void __launch_bounds__(256) sn_finalize_kernel(const __grid_constant__ SnParams p) {
  const SnLayer& L = p.layer[blockIdx.x];
  __shared__ float red[8];
  __shared__ float sig_sh;
  const float nv = sqrtf(L.scal[0]) + 1e-12f;
  float acc = 0.f;
  for (int i = threadIdx.x; i < L.C; i += 256) {
    const float t = L.t_raw[i] / nv;
    acc += t * t;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    sig_sh = sqrtf(t);
  }
  __syncthreads();
  const float sigma = sig_sh;
  for (int i = threadIdx.x; i < L.C; i += 256) {
    L.u[i] = L.t_raw[i] / nv / (sigma + 1e-12f);
    L.t_raw[i] = 0.f;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    L.scal[0] = 0.f;
    L.scal[1] = sigma;
    L.scal[2] = 1.f / sigma;
    L.scal[3] = nv;
  }
}

// Gradient through W_sn = W / sigma with sigma = v^T W u (u, v constants):
//   dW = G/sigma - (<G,W>/sigma^2) v u^T,  G = dL/dW_sn (raw).  Pass A: dot = <G,W> ; pass B: in-place update.
__global__ void __launch_bounds__(256) sn_grad_dot_kernel(const float* __restrict__ W, const float* __restrict__ G,
                                                          long n, float* __restrict__ dot) {
  float acc = 0.f;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n; i += gridDim.x * 256L) acc += W[i] * G[i];
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    atomicAdd(dot, t);
  }
}
__global__ void __launch_bounds__(256) sn_grad_apply_kernel(float* __restrict__ G, const float* __restrict__ u,
                                                            const float* __restrict__ v_raw,
                                                            const float* __restrict__ scal, float* __restrict__ dot,
                                                            int K, int C) {
  const float sigma = scal[1], inv = scal[2], nv = scal[3];
  const float coef = *dot * inv * inv / nv;
  (void)sigma;
  const long n = (long)K * C;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n; i += gridDim.x * 256L) {
    const int k = (int)(i / C), c = (int)(i % C);
    G[i] = G[i] * inv - coef * v_raw[k] * u[c];
  }
}
__global__ void zero_scalar_kernel(float* p) { *p = 0.f; }

}  // namespace b2g

using namespace b2g;

__global__ void step_prologue_kernel(long long* step_ctr, unsigned long long seed_base, unsigned long long* seeds,
                                     float* lr3, float lr_g, float lr_d, float b1, float b2) {
  const long long s = *step_ctr;
  const int i = threadIdx.x;
  if (i < 21) {
    const int pass = i / 7, layer = i % 7;
    unsigned long long x = seed_base ^ (unsigned long long)((3 * s + pass) * 8 + layer);
    x += 0x9E3779B97F4A7C15ull;  // splitmix64
    unsigned long long z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    seeds[i] = z ^ (z >> 31);
  } else if (i < 24) {
    const int k = i - 21;  // 0: generator, 1 / 2: first / second discriminator update of the step
    const double t = k == 0 ? (double)(s + 1) : (double)(2 * s + k);
    const double lr = k == 0 ? lr_g : lr_d;
    lr3[k] = (float)(lr * sqrt(1.0 - pow((double)b2, t)) / (1.0 - pow((double)b1, t)));
  }
  __syncthreads();
  if (i == 0) *step_ctr = s + 1;
}

static int adam_any(float* w, const float* g, float* m, float* v, long n, float lr_t, const float* lr_t_dev,
                    float beta1, float beta2, float eps, float gscale, b2g_stream_t stream_) {
  if (!w || !g || !m || !v) return fail(B2G_E_INVALID_ARG, "b2g_adam: null pointer");
  if (n % 4) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_adam: arena length must be a multiple of 4 (pad it)");
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  const long cap = (long)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  adam_kernel<<<(int)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<float4*>(w), reinterpret_cast<const float4*>(g), reinterpret_cast<float4*>(m),
      reinterpret_cast<float4*>(v), n4, lr_t, lr_t_dev, beta1, beta2, eps, gscale);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_adam(float* w, const float* g, float* m, float* v, long n, float lr_t, float beta1, float beta2,
                        float eps, float gscale, b2g_stream_t stream) {
  return adam_any(w, g, m, v, n, lr_t, nullptr, beta1, beta2, eps, gscale, stream);
}
extern "C" int b2g_adam_dev(float* w, const float* g, float* m, float* v, long n, const float* lr_t_dev, float beta1,
                            float beta2, float eps, float gscale, b2g_stream_t stream) {
  if (!lr_t_dev) return fail(B2G_E_INVALID_ARG, "b2g_adam_dev: null lr_t pointer");
  return adam_any(w, g, m, v, n, 0.f, lr_t_dev, beta1, beta2, eps, gscale, stream);
}
extern "C" int b2g_step_prologue(long long* step_ctr, unsigned long long seed_base, unsigned long long* seeds21,
                                 float* lr3, float lr_g, float lr_d, float beta1, float beta2, b2g_stream_t stream_) {
  if (!step_ctr || !seeds21 || !lr3) return fail(B2G_E_INVALID_ARG, "b2g_step_prologue: null pointer");
  step_prologue_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream_)>>>(step_ctr, seed_base, seeds21, lr3, lr_g, lr_d,
                                                                         beta1, beta2);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_colsum_bf16(const void* x, float* out, long P, int C, b2g_stream_t stream_) {
  if (!x || !out) return fail(B2G_E_INVALID_ARG, "b2g_colsum_bf16: null pointer");
  if (C % 8 || C > 2048 || C < 8) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_colsum_bf16: C=%d unsupported", C);
  const int groups = C / 8;
  const int rows_par = 256 / groups > 0 ? 256 / groups : 1;
  long blocks = (P + rows_par * 16 - 1) / (rows_par * 16);
  const long cap = (long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  colsum_kernel<<<(int)blocks, 256, C * sizeof(float), static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const __nv_bfloat16*>(x), out, P, C);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_bn_fold_finalize(const float* W, float* G, const float* gamma, float inv_std, float* dgamma,
                                    int taps, int Co, int Ci, b2g_stream_t stream_) {
  if (!W || !G || !gamma || !dgamma) return fail(B2G_E_INVALID_ARG, "b2g_bn_fold_finalize: null pointer");
  bn_fold_finalize_kernel<<<Co, 256, 0, static_cast<cudaStream_t>(stream_)>>>(W, G, gamma, inv_std, dgamma, taps, Co, Ci);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_spectral_norm_multi(int n_layers, const float* const* W, float* const* u, float* const* v_raw,
                                       float* const* t_raw, float* const* scal, const int* K, const int* C,
                                       b2g_stream_t stream_) {
  if (n_layers < 1 || n_layers > kSnMaxLayers) return fail(B2G_E_INVALID_ARG, "b2g_spectral_norm_multi: 1..8 layers");
  SnParams p;
  memset(&p, 0, sizeof(p));
  for (int i = 0; i < n_layers; ++i) {
    if (!W[i] || !u[i] || !v_raw[i] || !t_raw[i] || !scal[i]) return fail(B2G_E_INVALID_ARG, "b2g_spectral_norm_multi: null pointer");
    if (!(C[i] == 1 || (C[i] % 32 == 0 && C[i] <= 512)))
      return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_spectral_norm_multi: C=%d unsupported", C[i]);
    p.layer[i] = SnLayer{W[i], u[i], v_raw[i], t_raw[i], scal[i], K[i], C[i]};
  }
  p.n_layers = n_layers;
  p.blocks_per_layer = (num_sms() * 4 + n_layers - 1) / n_layers;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  sn_power_kernel<<<dim3(p.blocks_per_layer, n_layers), 256, 0, stream>>>(p);
  B2G_LAUNCH_CHECK();
  sn_finalize_kernel<<<n_layers, 256, 0, stream>>>(p);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_sn_grad_finalize(const float* W, float* G, const float* u, const float* v_raw, const float* scal,
                                    float* dot_scratch, int K, int C, b2g_stream_t stream_) {
  if (!W || !G || !u || !v_raw || !scal || !dot_scratch) return fail(B2G_E_INVALID_ARG, "b2g_sn_grad_finalize: null pointer");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const long n = (long)K * C;
  long blocks = (n + 255) / 256;
  if (blocks > num_sms() * 4L) blocks = num_sms() * 4L;
  zero_scalar_kernel<<<1, 1, 0, stream>>>(dot_scratch);
  sn_grad_dot_kernel<<<(int)blocks, 256, 0, stream>>>(W, G, n, dot_scratch);
  B2G_LAUNCH_CHECK();
  sn_grad_apply_kernel<<<(int)blocks, 256, 0, stream>>>(G, u, v_raw, scal, dot_scratch, K, C);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
