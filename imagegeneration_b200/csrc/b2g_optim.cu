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
// ONE launch for all layers; every weight matrix is read exactly once:
//   v_raw = W u ; vsq += |v_raw|^2 ; t_raw += W^T v_raw   (row blocks, byte-balanced across the layers)
//   the last block of a layer to finish (ticket in scal[4]) finalises it:
//   nv = |v_raw| ; t = t_raw/nv ; sigma = |t| ; u <- t/sigma.   (v = v_raw/nv and W^T v = t, so sigma = v^T W u_new = |t|.)
// Rows are read as float4 per lane (C in {128,256,512}: 4 KiB per warp iteration, the next 4 KiB prefetched), partial
// W^T v sums go registers -> per-warp shared slabs -> one red per column per block.
// ---------------------------------------------------------------------------------------------------------
struct SnLayer {
  const float* W;
  float* u;       // [C]   persistent, updated by the finalising block
  float* v_raw;   // [K]
  float* t_raw;   // [C]   scratch, zero on entry, re-zeroed on exit
  float* scal;    // [8]   {vsq scratch (zero on entry), sigma, 1/sigma, nv, ticket (uint, zero on entry), -, -, -}
  int K, C;
};
constexpr int kSnMaxLayers = 8;
constexpr int kSnThreads = 256;
struct SnParams {
  SnLayer layer[kSnMaxLayers];
  int blk_start[kSnMaxLayers + 1];  // first block of each layer; blk_start[n_layers] = grid size
  int n_layers;
};

template <int NV, int R>
__device__ __forceinline__ void sn_load_rows(const float4* __restrict__ W4, int k0, int K, int lane, float4 (&w)[R][NV]) {
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int j = 0; j < NV; ++j)
      w[r][j] = (k0 + r < K) ? __ldg(W4 + (size_t)(k0 + r) * (NV * 32) + j * 32 + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
}

// R rows of C = NV*128 columns per warp iteration, the next iteration's rows in flight while this one is reduced.
// Partial W^T v sums: registers -> this warp's slab of shared memory (no shared atomics: fp32 atomicAdd on shared
// memory is a compare-and-swap loop).
template <int NV, int R>
__device__ __forceinline__ void sn_rows_vec(const SnLayer& L, int lb, int nb, float* slab, float& vsq) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kWarps = kSnThreads / 32;
  const float4* __restrict__ W4 = reinterpret_cast<const float4*>(L.W);
  float4 u4[NV], tacc[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    u4[j] = __ldg(reinterpret_cast<const float4*>(L.u) + j * 32 + lane);
    tacc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int K = L.K;
  const int stride = nb * kWarps * R;
  int k0 = (lb * kWarps + warp) * R;
  float4 wn[R][NV];
  if (k0 < K) sn_load_rows<NV, R>(W4, k0, K, lane, wn);
  while (k0 < K) {
    float4 w[R][NV];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int j = 0; j < NV; ++j) w[r][j] = wn[r][j];
    const int k1 = k0 + stride;
    if (k1 < K) sn_load_rows<NV, R>(W4, k1, K, lane, wn);
    float dot[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float d = 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j)
        d += w[r][j].x * u4[j].x + w[r][j].y * u4[j].y + w[r][j].z * u4[j].z + w[r][j].w * u4[j].w;
      dot[r] = d;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int r = 0; r < R; ++r) dot[r] += __shfl_xor_sync(0xffffffffu, dot[r], o);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (lane == 0 && k0 + r < K) {
        L.v_raw[k0 + r] = dot[r];
        vsq += dot[r] * dot[r];
      }
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        tacc[j].x += w[r][j].x * dot[r];
        tacc[j].y += w[r][j].y * dot[r];
        tacc[j].z += w[r][j].z * dot[r];
        tacc[j].w += w[r][j].w * dot[r];
      }
    }
    k0 = k1;
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) reinterpret_cast<float4*>(slab + warp * 512)[j * 32 + lane] = tacc[j];
}

// Generic columns (C a multiple of 32, <= 512): one row per warp iteration, scalar loads (the 27 x 64 layer).
__device__ __forceinline__ void sn_rows_scalar(const SnLayer& L, int lb, int nb, float* slab, float& vsq) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kWarps = kSnThreads / 32;
  const int K = L.K, C = L.C;
  const int per = C / 32;
  float ureg[16], tacc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    ureg[i] = (i < per) ? L.u[lane + 32 * i] : 0.f;
    tacc[i] = 0.f;
  }
  for (int k = lb * kWarps + warp; k < K; k += nb * kWarps) {
    float wreg[16];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      wreg[i] = (i < per) ? __ldg(L.W + (size_t)k * C + lane + 32 * i) : 0.f;
      dot += wreg[i] * ureg[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) {
      L.v_raw[k] = dot;
      vsq += dot * dot;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) tacc[i] += wreg[i] * dot;
  }
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if (i < per) slab[warp * 512 + lane + 32 * i] = tacc[i];
}

// C == 1 (Dense head, K x 1): v_raw = W u0, thread per 4 rows when K % 4 == 0.
__device__ __forceinline__ void sn_rows_c1(const SnLayer& L, int lb, int nb, float* slab, float& vsq) {
  const float u0 = L.u[0];
  const int K = L.K;
  float t = 0.f;
  if ((K & 3) == 0) {
    const float4* __restrict__ W4 = reinterpret_cast<const float4*>(L.W);
    float4* V4 = reinterpret_cast<float4*>(L.v_raw);
    const int n4 = K / 4, stride = nb * kSnThreads;
    for (int k = lb * kSnThreads + threadIdx.x; k < n4; k += 4 * stride) {
      float4 wv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        wv[q] = (k + q * stride < n4) ? __ldg(W4 + k + q * stride) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 d = make_float4(wv[q].x * u0, wv[q].y * u0, wv[q].z * u0, wv[q].w * u0);
        if (k + q * stride < n4) V4[k + q * stride] = d;
        vsq += d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w;
        t += wv[q].x * d.x + wv[q].y * d.y + wv[q].z * d.z + wv[q].w * d.w;
      }
    }
  } else {
    for (int k = lb * kSnThreads + threadIdx.x; k < K; k += nb * kSnThreads) {
      const float wv = __ldg(L.W + k);
      const float d = wv * u0;
      L.v_raw[k] = d;
      vsq += d * d;
      t += wv * d;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0) slab[(threadIdx.x >> 5) * 512] = t;
}

__device__ __forceinline__ void sn_finalize_layer(const SnLayer& L, float* red, float* sig_sh) {
  const float nv = sqrtf(__ldcg(L.scal)) + 1e-12f;
  float acc = 0.f;
  for (int i = threadIdx.x; i < L.C; i += kSnThreads) {
    const float t = __ldcg(L.t_raw + i) / nv;
    acc += t * t;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < kSnThreads / 32; ++i) t += red[i];
    *sig_sh = sqrtf(t);
  }
  __syncthreads();
  const float sigma = *sig_sh;
  for (int i = threadIdx.x; i < L.C; i += kSnThreads) {
    L.u[i] = __ldcg(L.t_raw + i) / nv / (sigma + 1e-12f);
    L.t_raw[i] = 0.f;
  }
  if (threadIdx.x == 0) {
    L.scal[0] = 0.f;
    L.scal[1] = sigma;
    L.scal[2] = 1.f / sigma;
    L.scal[3] = nv;
    *reinterpret_cast<unsigned*>(L.scal + 4) = 0u;
  }
}

__global__ void __launch_bounds__(kSnThreads, 2) sn_power_kernel(const __grid_constant__ SnParams p) {
  int li = 0;
  while (li + 1 < p.n_layers && (int)blockIdx.x >= p.blk_start[li + 1]) ++li;
  const SnLayer& L = p.layer[li];
  const int lb = blockIdx.x - p.blk_start[li], nb = p.blk_start[li + 1] - p.blk_start[li];
  const int C = L.C;
  constexpr int kWarps = kSnThreads / 32;
  __shared__ __align__(16) float slab[kWarps * 512];  // per-warp partial W^T v
  __shared__ float red[kWarps];
  __shared__ float sig_sh;
  __shared__ int last_sh;
  for (int i = threadIdx.x; i < kWarps * 512; i += kSnThreads) slab[i] = 0.f;
  __syncthreads();
  float vsq = 0.f;
  if (C == 512) sn_rows_vec<4, 2>(L, lb, nb, slab, vsq);
  else if (C == 256) sn_rows_vec<2, 4>(L, lb, nb, slab, vsq);
  else if (C == 128) sn_rows_vec<1, 8>(L, lb, nb, slab, vsq);
  else if (C >= 32) sn_rows_scalar(L, lb, nb, slab, vsq);
  else sn_rows_c1(L, lb, nb, slab, vsq);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) vsq += __shfl_xor_sync(0xffffffffu, vsq, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = vsq;
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += kSnThreads) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += slab[w * 512 + i];
    atomicAdd(L.t_raw + i, t);
  }
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kWarps; ++w) t += red[w];
    atomicAdd(L.scal, t);
  }
  // last block of this layer finalises it
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(L.scal + 4), 1u);
    last_sh = (ticket == (unsigned)(nb - 1));
  }
  __syncthreads();
  if (!last_sh) return;
  __threadfence();
  sn_finalize_layer(L, red, &sig_sh);
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

// ---------------------------------------------------------------------------------------------------------
// Adam FUSED with spectral normalisation: ONE pass over the discriminator arena per apply_gradients replaces the
// stand-alone power iteration (24.6 MB read again, 3 x per step), the 8 x 3 gradient-finalisation launches and the
// plain Adam launch.  For every spectrally-normalised matrix W_l inside the arena a warp owns whole rows and
//   (1) turns the raw weight gradient G = dL/d(W/sigma) into dL/dW = G/sigma - (<G,W>/sigma^2) v u^T   (u, v_raw, sigma
//       of the power iteration the forward pass used; <G,W> comes from sn_dot_multi_kernel),
//   (2) applies Keras Adam to the row (registers),
//   (3) runs the NEXT power iteration on the updated row while it is still in registers:
//       v_raw[k] = w'_k . u,  t_raw += w'_k * v_raw[k]   (Miyato et al. Alg. 1: one iteration per weight update),
// and the last block of a layer (ticket) finalises sigma and u exactly like sn_power_kernel.  Biases and alignment
// padding between the matrices are updated element-wise by one trailing block.  Extra HBM traffic over plain Adam:
// the K floats of v_raw per layer.
// ---------------------------------------------------------------------------------------------------------
struct AdamSnLayer {
  long off;  // element offset of W_l in the arena
  int K, C;
  float *u, *v_raw, *t_raw, *scal, *dot;
};
struct AdamSnRest {
  long off;
  int n;
};
struct AdamSnParams {
  AdamSnLayer layer[kSnMaxLayers];
  AdamSnRest rest[kSnMaxLayers + 1];
  int blk_start[kSnMaxLayers + 1];
  int n_layers, n_rest;
  float *w, *m, *v;
  const float* g;
  float lr_t_val;
  const float* lr_t_dev;
  float b1, b2, eps, gscale;
};

__device__ __forceinline__ SnLayer L_as_sn(const AdamSnLayer& L) {
  return SnLayer{nullptr, L.u, L.v_raw, L.t_raw, L.scal, L.K, L.C};
}

struct AdamCoef {
  float lr_t, c1, c2, eps;
};
__device__ __forceinline__ void adam_elt(const AdamCoef& a, float gg, float& w, float& m, float& v) {
  m += (gg - m) * a.c1;
  v += (gg * gg - v) * a.c2;
  w -= a.lr_t * m / (sqrtf(v) + a.eps);
}

template <int NV, int R>
__device__ __forceinline__ void adam_sn_rows_vec(const AdamSnParams& p, const AdamSnLayer& L, const AdamCoef& a, int lb,
                                                 int nb, float* slab, float& vsq) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kWarps = kSnThreads / 32;
  float4* W4 = reinterpret_cast<float4*>(p.w + L.off);
  float4* M4 = reinterpret_cast<float4*>(p.m + L.off);
  float4* V4 = reinterpret_cast<float4*>(p.v + L.off);
  const float4* G4 = reinterpret_cast<const float4*>(p.g + L.off);
  float4 u4[NV], tacc[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    u4[j] = __ldcg(reinterpret_cast<const float4*>(L.u) + j * 32 + lane);
    tacc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float inv = __ldcg(L.scal + 2), nv = __ldcg(L.scal + 3);
  const float coef = nv > 0.f ? __ldcg(L.dot) * inv * inv / nv : 0.f;
  const float gs = p.gscale;
  const int K = L.K;
  for (int k0 = (lb * kWarps + warp) * R; k0 < K; k0 += nb * kWarps * R) {
    float4 w[R][NV], g[R][NV], m[R][NV], v[R][NV];
    float vold[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const bool ok = k0 + r < K;
      vold[r] = ok ? __ldcg(L.v_raw + k0 + r) : 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const size_t i = (size_t)(k0 + r) * (NV * 32) + j * 32 + lane;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        w[r][j] = ok ? W4[i] : z;
        g[r][j] = ok ? __ldcs(G4 + i) : z;
        m[r][j] = ok ? M4[i] : z;
        v[r][j] = ok ? V4[i] : z;
      }
    }
    float dot[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float cv = coef * vold[r];
      float d = 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        float* wp = &w[r][j].x; float* mp = &m[r][j].x; float* vp = &v[r][j].x;
        const float* gp = &g[r][j].x; const float* up = &u4[j].x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          adam_elt(a, gs * (gp[e] * inv - cv * up[e]), wp[e], mp[e], vp[e]);
          d += wp[e] * up[e];
        }
      }
      dot[r] = d;
      if (k0 + r < K) {
#pragma unroll
        for (int j = 0; j < NV; ++j) {
          const size_t i = (size_t)(k0 + r) * (NV * 32) + j * 32 + lane;
          W4[i] = w[r][j]; M4[i] = m[r][j]; V4[i] = v[r][j];
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int r = 0; r < R; ++r) dot[r] += __shfl_xor_sync(0xffffffffu, dot[r], o);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (lane == 0 && k0 + r < K) {
        L.v_raw[k0 + r] = dot[r];
        vsq += dot[r] * dot[r];
      }
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        tacc[j].x += w[r][j].x * dot[r];
        tacc[j].y += w[r][j].y * dot[r];
        tacc[j].z += w[r][j].z * dot[r];
        tacc[j].w += w[r][j].w * dot[r];
      }
    }
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) reinterpret_cast<float4*>(slab + warp * 512)[j * 32 + lane] = tacc[j];
}

// any C that is a multiple of 32 (<= 512): one row per warp iteration, scalar accesses (the 27 x 64 layer)
__device__ __forceinline__ void adam_sn_rows_scalar(const AdamSnParams& p, const AdamSnLayer& L, const AdamCoef& a,
                                                    int lb, int nb, float* slab, float& vsq) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kWarps = kSnThreads / 32;
  const int K = L.K, C = L.C, per = C / 32;
  const float inv = __ldcg(L.scal + 2), nv = __ldcg(L.scal + 3);
  const float coef = nv > 0.f ? __ldcg(L.dot) * inv * inv / nv : 0.f;
  float ureg[16], tacc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    ureg[i] = (i < per) ? __ldcg(L.u + lane + 32 * i) : 0.f;
    tacc[i] = 0.f;
  }
  for (int k = lb * kWarps + warp; k < K; k += nb * kWarps) {
    const float cv = coef * __ldcg(L.v_raw + k);
    float wreg[16];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      wreg[i] = 0.f;
      if (i < per) {
        const size_t idx = L.off + (size_t)k * C + lane + 32 * i;
        float w = p.w[idx], m = p.m[idx], v = p.v[idx];
        adam_elt(a, p.gscale * (p.g[idx] * inv - cv * ureg[i]), w, m, v);
        p.w[idx] = w; p.m[idx] = m; p.v[idx] = v;
        wreg[i] = w;
        dot += w * ureg[i];
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) {
      L.v_raw[k] = dot;
      vsq += dot * dot;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) tacc[i] += wreg[i] * dot;
  }
#pragma unroll
  for (int i = 0; i < 16; ++i)
    if (i < per) slab[warp * 512 + lane + 32 * i] = tacc[i];
}

// C == 1 (Dense head, K x 1): a thread owns rows (float4 = 4 rows when K % 4 == 0); v_raw = w' u0, t = sum w' v_raw
__device__ __forceinline__ void adam_sn_rows_c1(const AdamSnParams& p, const AdamSnLayer& L, const AdamCoef& a, int lb,
                                                int nb, float* slab, float& vsq) {
  const float u0 = __ldcg(L.u);
  const float inv = __ldcg(L.scal + 2), nv = __ldcg(L.scal + 3);
  const float coef = nv > 0.f ? __ldcg(L.dot) * inv * inv / nv : 0.f;
  const float cu = coef * u0, gs = p.gscale;
  const int K = L.K;
  float t = 0.f;
  if ((K & 3) == 0) {
    float4* W4 = reinterpret_cast<float4*>(p.w + L.off);
    float4* M4 = reinterpret_cast<float4*>(p.m + L.off);
    float4* V4 = reinterpret_cast<float4*>(p.v + L.off);
    const float4* G4 = reinterpret_cast<const float4*>(p.g + L.off);
    float4* R4 = reinterpret_cast<float4*>(L.v_raw);
    const int n4 = K / 4, stride = nb * kSnThreads;
    for (int k = lb * kSnThreads + threadIdx.x; k < n4; k += 2 * stride) {
      float4 w[2], g[2], m[2], v[2], vo[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int i = k + q * stride;
        const bool ok = i < n4;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        w[q] = ok ? W4[i] : z; g[q] = ok ? __ldcs(G4 + i) : z; m[q] = ok ? M4[i] : z; v[q] = ok ? V4[i] : z;
        vo[q] = ok ? __ldcg(R4 + i) : z;
      }
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int i = k + q * stride;
        float* wp = &w[q].x; float* mp = &m[q].x; float* vp = &v[q].x;
        const float* gp = &g[q].x; const float* op = &vo[q].x;
        float4 d4;
        float* dp = &d4.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          adam_elt(a, gs * (gp[e] * inv - cu * op[e]), wp[e], mp[e], vp[e]);
          dp[e] = wp[e] * u0;
        }
        if (i < n4) {
          W4[i] = w[q]; M4[i] = m[q]; V4[i] = v[q]; R4[i] = d4;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            vsq += dp[e] * dp[e];
            t += wp[e] * dp[e];
          }
        }
      }
    }
  } else {
    for (int k = lb * kSnThreads + threadIdx.x; k < K; k += nb * kSnThreads) {
      const size_t idx = L.off + k;
      float w = p.w[idx], m = p.m[idx], v = p.v[idx];
      adam_elt(a, gs * (p.g[idx] * inv - cu * L.v_raw[k]), w, m, v);
      p.w[idx] = w; p.m[idx] = m; p.v[idx] = v;
      const float d = w * u0;
      L.v_raw[k] = d;
      vsq += d * d;
      t += w * d;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0) slab[(threadIdx.x >> 5) * 512] = t;
}

__global__ void __launch_bounds__(kSnThreads, 2) adam_sn_kernel(const __grid_constant__ AdamSnParams p) {
  AdamCoef a;
  a.lr_t = p.lr_t_dev ? *p.lr_t_dev : p.lr_t_val;
  a.c1 = 1.f - p.b1; a.c2 = 1.f - p.b2; a.eps = p.eps;
  if ((int)blockIdx.x >= p.blk_start[p.n_layers]) {  // trailing block: biases + padding, plain Adam
    for (int s = 0; s < p.n_rest; ++s) {
      const long off = p.rest[s].off;
      for (int i = threadIdx.x; i < p.rest[s].n; i += kSnThreads) {
        float w = p.w[off + i], m = p.m[off + i], v = p.v[off + i];
        adam_elt(a, p.g[off + i] * p.gscale, w, m, v);
        p.w[off + i] = w; p.m[off + i] = m; p.v[off + i] = v;
      }
    }
    return;
  }
  int li = 0;
  while (li + 1 < p.n_layers && (int)blockIdx.x >= p.blk_start[li + 1]) ++li;
  const AdamSnLayer& L = p.layer[li];
  const int lb = blockIdx.x - p.blk_start[li], nb = p.blk_start[li + 1] - p.blk_start[li];
  const int C = L.C;
  constexpr int kWarps = kSnThreads / 32;
  __shared__ __align__(16) float slab[kWarps * 512];
  __shared__ float red[kWarps];
  __shared__ float sig_sh;
  __shared__ int last_sh;
  for (int i = threadIdx.x; i < kWarps * 512; i += kSnThreads) slab[i] = 0.f;
  __syncthreads();
  float vsq = 0.f;
  if (C == 512) adam_sn_rows_vec<4, 1>(p, L, a, lb, nb, slab, vsq);
  else if (C == 256) adam_sn_rows_vec<2, 2>(p, L, a, lb, nb, slab, vsq);
  else if (C == 128) adam_sn_rows_vec<1, 4>(p, L, a, lb, nb, slab, vsq);
  else if (C >= 32) adam_sn_rows_scalar(p, L, a, lb, nb, slab, vsq);
  else adam_sn_rows_c1(p, L, a, lb, nb, slab, vsq);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) vsq += __shfl_xor_sync(0xffffffffu, vsq, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = vsq;
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += kSnThreads) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += slab[w * 512 + i];
    atomicAdd(L.t_raw + i, t);
  }
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < kWarps; ++w) t += red[w];
    atomicAdd(L.scal, t);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(L.scal + 4), 1u);
    last_sh = (ticket == (unsigned)(nb - 1));
  }
  __syncthreads();
  if (!last_sh) return;
  __threadfence();
  sn_finalize_layer(L_as_sn(L), red, &sig_sh);
  if (threadIdx.x == 0) *L.dot = 0.f;  // consumed: the next sn_dot_multi accumulates from zero
}

// dot[l] += <G_l, W_l> for every spectrally-normalised matrix, one launch (blocks balanced by bytes)
struct SnDotParams {
  const float* W[kSnMaxLayers];
  const float* G[kSnMaxLayers];
  float* dot[kSnMaxLayers];
  long n[kSnMaxLayers];
  int blk_start[kSnMaxLayers + 1];
  int n_layers;
};
__global__ void __launch_bounds__(256) sn_dot_multi_kernel(const __grid_constant__ SnDotParams p) {
  int li = 0;
  while (li + 1 < p.n_layers && (int)blockIdx.x >= p.blk_start[li + 1]) ++li;
  const int lb = blockIdx.x - p.blk_start[li], nb = p.blk_start[li + 1] - p.blk_start[li];
  const float* __restrict__ W = p.W[li];
  const float* __restrict__ G = p.G[li];
  const long n = p.n[li];
  float acc = 0.f;
  if ((n & 3) == 0) {
    const float4* W4 = reinterpret_cast<const float4*>(W);
    const float4* G4 = reinterpret_cast<const float4*>(G);
    const long n4 = n / 4, stride = (long)nb * 256;
    for (long i = lb * 256L + threadIdx.x; i < n4; i += 2 * stride) {
      const float4 a0 = __ldg(W4 + i), b0 = __ldg(G4 + i);
      const bool ok = i + stride < n4;
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 a1 = ok ? __ldg(W4 + i + stride) : z, b1 = ok ? __ldg(G4 + i + stride) : z;
      acc += a0.x * b0.x + a0.y * b0.y + a0.z * b0.z + a0.w * b0.w + a1.x * b1.x + a1.y * b1.y + a1.z * b1.z + a1.w * b1.w;
    }
  } else {
    for (long i = lb * 256L + threadIdx.x; i < n; i += (long)nb * 256) acc += W[i] * G[i];
  }
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    atomicAdd(p.dot[li], t);
  }
}

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

// ---------------------------------------------------------------------------------------------------------
// Data-parallel Adam over NVSwitch multicast memory (NVLS): gradient all-reduce + update + weight broadcast in ONE pass.
// Rank r owns elements [begin, end) of the arena.  For those it reads the SUM of every replica's gradient with
// multimem.ld_reduce (the switch adds the N copies on the way), applies Adam with its LOCAL m / v (optimizer state is
// sharded: each rank keeps only its slice current) and writes the new weights to EVERY replica with multimem.st.
// Per GPU that is |arena| bytes out + |arena| bytes in over NVLink instead of a ring / tree all-reduce followed by N
// identical full-arena Adam passes; replicas stay bit-identical because each element is computed once.
// The caller brackets the launch with cross-rank barriers (all gradients written before / all weights written after).
// No counterpart in the reference (single-GPU TensorFlow); replaces NCCL all-reduce + b2g_adam on the data-parallel path.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 mc_ld_reduce_add(const float4* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void mc_st(float4* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__global__ void __launch_bounds__(256) adam_nvls_kernel(float4* __restrict__ w_mc, const float4* __restrict__ g_mc,
                                                        const float4* __restrict__ w_loc, float4* __restrict__ m,
                                                        float4* __restrict__ v, long i0, long i1, float lr_t_val,
                                                        const float* __restrict__ lr_t_dev, float b1, float b2, float eps,
                                                        float gscale) {
  const float lr_t = lr_t_dev ? *lr_t_dev : lr_t_val;
  const float c1 = 1.f - b1, c2 = 1.f - b2;
  for (long i = i0 + blockIdx.x * (long)blockDim.x + threadIdx.x; i < i1; i += (long)gridDim.x * blockDim.x) {
    float4 gw = mc_ld_reduce_add(g_mc + i), mw = m[i], vw = v[i], ww = w_loc[i];
    float* gp = &gw.x; float* mp = &mw.x; float* vp = &vw.x; float* wp = &ww.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gg = gp[k] * gscale;
      mp[k] += (gg - mp[k]) * c1;
      vp[k] += (gg * gg - vp[k]) * c2;
      wp[k] -= lr_t * mp[k] / (sqrtf(vp[k]) + eps);
    }
    m[i] = mw; v[i] = vw;
    mc_st(w_mc + i, ww);
  }
}

extern "C" int b2g_adam_nvls(void* w_mc, const void* g_mc, const float* w_local, float* m, float* v, long begin, long end,
                             float lr_t, const float* lr_t_dev, float beta1, float beta2, float eps, float gscale,
                             b2g_stream_t stream_) {
  if (!w_mc || !g_mc || !w_local || !m || !v) return fail(B2G_E_INVALID_ARG, "b2g_adam_nvls: null pointer");
  if (begin < 0 || end < begin || begin % 4 || end % 4)
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_adam_nvls: slice [%ld, %ld) must be non-negative multiples of 4", begin, end);
  if (end == begin) return B2G_OK;
  const long n4 = (end - begin) / 4;
  long blocks = (n4 + 255) / 256;
  const long cap = (long)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  adam_nvls_kernel<<<(int)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<float4*>(w_mc), reinterpret_cast<const float4*>(g_mc), reinterpret_cast<const float4*>(w_local),
      reinterpret_cast<float4*>(m), reinterpret_cast<float4*>(v), begin / 4, end / 4, lr_t, lr_t_dev, beta1, beta2, eps,
      gscale);
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
  // blocks per layer in proportion to the bytes of W (at least one), two 256-thread blocks per SM in total
  double total = 0.0;
  for (int i = 0; i < n_layers; ++i) total += (double)K[i] * C[i];
  const int budget = num_sms() * 2;
  int start = 0;
  for (int i = 0; i < n_layers; ++i) {
    p.blk_start[i] = start;
    const double share = (double)K[i] * C[i] / total * budget;
    int nb = (int)(share + 0.5);
    // no more blocks than 4 KiB row groups per warp can use
    const long units = ((long)K[i] * C[i] + 1023) / 1024;
    const long max_nb = (units + (kSnThreads / 32) - 1) / (kSnThreads / 32);
    if (nb > max_nb) nb = (int)max_nb;
    if (nb < 1) nb = 1;
    start += nb;
  }
  p.blk_start[n_layers] = start;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  sn_power_kernel<<<start, kSnThreads, 0, stream>>>(p);
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

// ---- Adam fused with spectral normalisation (see adam_sn_kernel) ----------------------------------------------------
extern "C" int b2g_sn_dot_multi(const float* w, const float* g, int n_layers, const long* off, const int* K,
                                const int* C, float* dots, b2g_stream_t stream_) {
  if (!w || !g || !off || !K || !C || !dots) return fail(B2G_E_INVALID_ARG, "b2g_sn_dot_multi: null pointer");
  if (n_layers < 1 || n_layers > kSnMaxLayers) return fail(B2G_E_INVALID_ARG, "b2g_sn_dot_multi: 1..8 layers");
  SnDotParams p;
  memset(&p, 0, sizeof(p));
  double total = 0.0;
  for (int i = 0; i < n_layers; ++i) total += (double)K[i] * C[i];
  const int budget = num_sms() * 4;
  int start = 0;
  for (int i = 0; i < n_layers; ++i) {
    if (off[i] % 4) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_sn_dot_multi: matrix offsets must be multiples of 4");
    p.W[i] = w + off[i];
    p.G[i] = g + off[i];
    p.dot[i] = dots + i;
    p.n[i] = (long)K[i] * C[i];
    p.blk_start[i] = start;
    int nb = (int)((double)p.n[i] / total * budget + 0.5);
    const long max_nb = (p.n[i] + 2047) / 2048;  // at least two float4 per thread
    if (nb > max_nb) nb = (int)max_nb;
    if (nb < 1) nb = 1;
    start += nb;
  }
  p.blk_start[n_layers] = start;
  p.n_layers = n_layers;
  sn_dot_multi_kernel<<<start, 256, 0, static_cast<cudaStream_t>(stream_)>>>(p);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_adam_sn(float* w, const float* g, float* m, float* v, long n, int n_layers, const long* off,
                           const int* K, const int* C, float* const* u, float* const* v_raw, float* const* t_raw,
                           float* const* scal, float* dots, float lr_t, const float* lr_t_dev, float beta1, float beta2,
                           float eps, float gscale, b2g_stream_t stream_) {
  if (!w || !g || !m || !v || !off || !K || !C || !u || !v_raw || !t_raw || !scal || !dots)
    return fail(B2G_E_INVALID_ARG, "b2g_adam_sn: null pointer");
  if (n_layers < 1 || n_layers > kSnMaxLayers) return fail(B2G_E_INVALID_ARG, "b2g_adam_sn: 1..8 layers");
  AdamSnParams p;
  memset(&p, 0, sizeof(p));
  p.w = w; p.g = g; p.m = m; p.v = v;
  p.lr_t_val = lr_t; p.lr_t_dev = lr_t_dev; p.b1 = beta1; p.b2 = beta2; p.eps = eps; p.gscale = gscale;
  double total = 0.0;
  long pos = 0;
  int n_rest = 0;
  for (int i = 0; i < n_layers; ++i) {
    const long len = (long)K[i] * C[i];
    if (off[i] % 4 || off[i] < pos || off[i] + len > n)
      return fail(B2G_E_INVALID_ARG, "b2g_adam_sn: matrices must be 4-aligned, ascending and inside the arena");
    if (!(C[i] == 1 || (C[i] % 32 == 0 && C[i] <= 512)))
      return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_adam_sn: C=%d unsupported", C[i]);
    if (!u[i] || !v_raw[i] || !t_raw[i] || !scal[i]) return fail(B2G_E_INVALID_ARG, "b2g_adam_sn: null layer pointer");
    if (off[i] > pos) p.rest[n_rest++] = AdamSnRest{pos, (int)(off[i] - pos)};
    pos = off[i] + len;
    p.layer[i] = AdamSnLayer{off[i], K[i], C[i], u[i], v_raw[i], t_raw[i], scal[i], dots + i};
    total += (double)len * (C[i] == 1 ? 4.0 : 1.0);
  }
  if (pos < n) p.rest[n_rest++] = AdamSnRest{pos, (int)(n - pos)};
  p.n_rest = n_rest;
  p.n_layers = n_layers;
  const int budget = num_sms() * 2;
  int start = 0;
  for (int i = 0; i < n_layers; ++i) {
    p.blk_start[i] = start;
    const double wgt = (double)K[i] * C[i] * (C[i] == 1 ? 4.0 : 1.0);
    int nb = (int)(wgt / total * budget + 0.5);
    const long units = C[i] == 1 ? ((long)K[i] + 1023) / 1024 : ((long)K[i] * C[i] + 4095) / 4096;  // work per block
    if (nb > units) nb = (int)units;
    if (nb < 1) nb = 1;
    start += nb;
  }
  p.blk_start[n_layers] = start;
  adam_sn_kernel<<<start + 1, kSnThreads, 0, static_cast<cudaStream_t>(stream_)>>>(p);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
