// b2g_dense.cu — the two dense layers of the SNDCGAN step and the loss.
//   G stem  : Dense(128 -> 512*h*w, no bias) + inference BatchNorm + ReLU   (sndcgan/SNDCGAN.py:31-34):
//             tensor-core kernels in b2g_dense_tc.cu
//   D head  : Flatten + Dense(1) (:122-123), BinaryCrossentropy(from_logits) (:170,:246,:254,:262) or hinge,
//             and the head's backward which also applies conv7's LeakyReLU/Dropout derivative.
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

// ---------------------------------------------------------------------------------------------------------
// D head: partial dot products  part[b][s] = sum_{j in slice s} x[b][j] * w[j]
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) dhead_partial_kernel(const T* __restrict__ x,
                                                            const float* __restrict__ w, float* __restrict__ part,
                                                            int F, int S) {
  const int b = blockIdx.y, s = blockIdx.x;
  const int per = ((F / 8 + S - 1) / S);  // 8-element groups per slice
  const int g0 = s * per, g1 = min(g0 + per, F / 8);
  const T* xp = x + (size_t)b * F;
  const float4* wp = reinterpret_cast<const float4*>(w);
  float acc = 0.f;
  for (int g = g0 + threadIdx.x; g < g1; g += 256) {
    float v[8];
    edge_load8(xp + (size_t)g * 8, v);
    const float4 wa = __ldg(wp + 2 * g), wb = __ldg(wp + 2 * g + 1);
    acc += v[0] * wa.x + v[1] * wa.y + v[2] * wa.z + v[3] * wa.w + v[4] * wb.x + v[5] * wb.y + v[6] * wb.z + v[7] * wb.w;
  }
  __shared__ float red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    part[b * S + s] = t;
  }
}

// logits[b] = wscale * sum_s part[b][s] + bias;  loss and dL/dlogit.  One block.
//   loss_kind 0: BCE-with-logits against `label`;  1: D hinge (label 1: relu(1-x), 0: relu(1+x));  2: G hinge (-x);
//   3: Wasserstein label*x
//   loss_out[0] = inv_batch * sum_b loss_b ;  dlogit[b] = inv_batch * dloss_b/dx ; loss_out[1] = inv_batch*sum sigmoid
__global__ void dhead_loss_kernel(const float* __restrict__ part, int S, const float* __restrict__ wscale,
                                  const float* __restrict__ bias, int B, int loss_kind, float label, float inv_batch,
                                  float* __restrict__ logits, float* __restrict__ dlogit,
                                  float* __restrict__ loss_out) {
  __shared__ float red[2][32];
  const float sc = wscale ? *wscale : 1.f;
  float lsum = 0.f, psum = 0.f;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    float t = 0.f;
    for (int s = 0; s < S; ++s) t += part[b * S + s];
    const float x = t * sc + bias[0];
    logits[b] = x;
    const float sig = 1.f / (1.f + __expf(-x));
    float l, d;
    if (loss_kind == 0) {
      l = fmaxf(x, 0.f) - x * label + log1pf(__expf(-fabsf(x)));
      d = sig - label;
    } else if (loss_kind == 1) {
      if (label > 0.5f) { l = fmaxf(1.f - x, 0.f); d = x < 1.f ? -1.f : 0.f; }
      else              { l = fmaxf(1.f + x, 0.f); d = x > -1.f ? 1.f : 0.f; }
    } else if (loss_kind == 2) {
      l = -x;
      d = -1.f;
    } else {  // 3: Wasserstein, mean(y_true * y_pred) with y_true = label (wasserstein_gan/WGAN.py:48-49)
      l = label * x;
      d = label;
    }
    dlogit[b] = d * inv_batch;
    lsum += l;
    psum += sig;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
    psum += __shfl_xor_sync(0xffffffffu, psum, o);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = lsum; red[1][threadIdx.x >> 5] = psum; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, c = 0.f;
    for (int i = 0; i < (int)(blockDim.x + 31) / 32; ++i) { a += red[0][i]; c += red[1][i]; }
    loss_out[0] = a * inv_batch;
    loss_out[1] = c * inv_batch;
  }
}

// Head backward: dx[b][j] = dlogit[b] * wscale * w[j] * act'(x[b][j]);  dw[j] (+)= sum_b dlogit[b] * x[b][j];
// db (+)= sum_b dlogit[b].   act' = LeakyReLU/Dropout derivative recovered from the saved conv7 output x.
// Grid (F / 2048, S): blockIdx.y takes every S-th sample, four loads in flight per thread — the first version walked all
// B samples serially in 144 blocks (one 16-byte load in flight per thread: 1.9 TB/s on a 75 MB stream, 40 us; round 2).
template <typename T>
__global__ void __launch_bounds__(256) dhead_bwd_kernel(const T* __restrict__ x,
                                                        const float* __restrict__ w, const float* __restrict__ wscale,
                                                        const float* __restrict__ dlogit, int B, int F, float alpha,
                                                        float dscale, float dzero, T* __restrict__ dx,
                                                        float* __restrict__ dw, float* __restrict__ db) {
  const int g = blockIdx.x * 256 + threadIdx.x;
  const int S = gridDim.y, s0 = blockIdx.y;
  if (g == 0 && s0 == 0 && db) {
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dlogit[b];
    db[0] += s;
  }
  if (g >= F / 8) return;
  const float sc = wscale ? *wscale : 1.f;
  const float4 wa = __ldg(reinterpret_cast<const float4*>(w) + 2 * g), wb = __ldg(reinterpret_cast<const float4*>(w) + 2 * g + 1);
  const float wv[8] = {wa.x * sc, wa.y * sc, wa.z * sc, wa.w * sc, wb.x * sc, wb.y * sc, wb.z * sc, wb.w * sc};
  const float fneg = alpha * dscale;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  constexpr int U = 4;
  for (int b0 = s0; b0 < B; b0 += U * S) {
    float yv[U][8], dl[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int b = b0 + u * S;
      dl[u] = b < B ? __ldg(dlogit + b) : 0.f;
      if (b < B) edge_load8(x + (size_t)b * F + (size_t)g * 8, yv[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int b = b0 + u * S;
      if (b >= B) break;
      float o[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float y = yv[u][i];
        acc[i] += dl[u] * y;
        const float f = y > 0.f ? dscale : (y < 0.f ? fneg : dzero);
        o[i] = dl[u] * wv[i] * f;
      }
      edge_store8(dx + (size_t)b * F + (size_t)g * 8, o);
    }
  }
  if (dw) {
    float* dp = dw + (size_t)g * 8;
    if (S == 1) {
      float4* d4 = reinterpret_cast<float4*>(dp);
      float4 a = d4[0], c = d4[1];
      a.x += acc[0]; a.y += acc[1]; a.z += acc[2]; a.w += acc[3];
      c.x += acc[4]; c.y += acc[5]; c.z += acc[6]; c.w += acc[7];
      d4[0] = a;
      d4[1] = c;
    } else {
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dp), "f"(acc[0]), "f"(acc[1]), "f"(acc[2]), "f"(acc[3])
                   : "memory");
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dp + 4), "f"(acc[4]), "f"(acc[5]), "f"(acc[6]),
                   "f"(acc[7])
                   : "memory");
    }
  }
}

}  // namespace b2g

using namespace b2g;

static int dhead_fwd_loss_any(const void* x, const float* w, const float* wscale, const float* bias, int B,
                              int F, int loss_kind, float label, float inv_batch, float* workspace,
                              int workspace_floats, float* logits, float* dlogit, float* loss_out, bool f32,
                              b2g_stream_t stream_) {
  if (!x || !w || !bias || !workspace || !logits || !dlogit || !loss_out)
    return fail(B2G_E_INVALID_ARG, "b2g_dhead_fwd_loss: null pointer");
  if (F % 8) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_dhead_fwd_loss: F must be a multiple of 8");
  int S = (4 * num_sms() + B - 1) / B;
  if (S < 1) S = 1;
  if (S > 64) S = 64;
  if (workspace_floats < B * S) return fail(B2G_E_INVALID_ARG, "b2g_dhead_fwd_loss: workspace needs %d floats", B * S);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (f32) dhead_partial_kernel<float><<<dim3(S, B), 256, 0, stream>>>(static_cast<const float*>(x), w, workspace, F, S);
  else dhead_partial_kernel<__nv_bfloat16><<<dim3(S, B), 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), w, workspace, F, S);
  B2G_LAUNCH_CHECK();
  dhead_loss_kernel<<<1, 256, 0, stream>>>(workspace, S, wscale, bias, B, loss_kind, label, inv_batch, logits, dlogit,
                                           loss_out);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_dhead_fwd_loss(const void* x, const float* w, const float* wscale, const float* bias, int B,
                                  int F, int loss_kind, float label, float inv_batch, float* workspace,
                                  int workspace_floats, float* logits, float* dlogit, float* loss_out,
                                  b2g_stream_t stream) {
  return dhead_fwd_loss_any(x, w, wscale, bias, B, F, loss_kind, label, inv_batch, workspace, workspace_floats, logits,
                            dlogit, loss_out, false, stream);
}
extern "C" int b2g_dhead_fwd_loss_f32(const float* x, const float* w, const float* wscale, const float* bias, int B,
                                      int F, int loss_kind, float label, float inv_batch, float* workspace,
                                      int workspace_floats, float* logits, float* dlogit, float* loss_out,
                                      b2g_stream_t stream) {
  return dhead_fwd_loss_any(x, w, wscale, bias, B, F, loss_kind, label, inv_batch, workspace, workspace_floats, logits,
                            dlogit, loss_out, true, stream);
}

static int dhead_bwd_any(const void* x, const float* w, const float* wscale, const float* dlogit, int B, int F,
                         float alpha, float dscale, float dzero, void* dx, float* dw, float* db, bool f32,
                         b2g_stream_t stream_) {
  if (!x || !w || !dlogit || !dx) return fail(B2G_E_INVALID_ARG, "b2g_dhead_bwd: null pointer");
  if (F % 8) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_dhead_bwd: F must be a multiple of 8");
  const int gx = (F / 8 + 255) / 256;
  int S = (4 * num_sms() + gx - 1) / gx;      // ~4 blocks per SM; every slice keeps at least 4 samples
  if (S > B / 4) S = B / 4;
  if (S < 1) S = 1;
  const dim3 grid(gx, S);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (f32)
    dhead_bwd_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(x), w, wscale, dlogit, B, F, alpha, dscale,
                                                      dzero, static_cast<float*>(dx), dw, db);
  else
    dhead_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), w, wscale, dlogit, B, F,
                                                              alpha, dscale, dzero, static_cast<__nv_bfloat16*>(dx), dw, db);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
extern "C" int b2g_dhead_bwd(const void* x, const float* w, const float* wscale, const float* dlogit, int B, int F,
                             float alpha, float dscale, float dzero, void* dx, float* dw, float* db,
                             b2g_stream_t stream) {
  return dhead_bwd_any(x, w, wscale, dlogit, B, F, alpha, dscale, dzero, dx, dw, db, false, stream);
}
extern "C" int b2g_dhead_bwd_f32(const float* x, const float* w, const float* wscale, const float* dlogit, int B, int F,
                                 float alpha, float dscale, float dzero, float* dx, float* dw, float* db,
                                 b2g_stream_t stream) {
  return dhead_bwd_any(x, w, wscale, dlogit, B, F, alpha, dscale, dzero, dx, dw, db, true, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of b2g_dhead_fwd_loss (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  The partial-sum scratch (part[b][s], at most 64 slices per sample) lives in a file-scope device array, so
// the caller passes no workspace.  Every GPU that loads the library holds the kB2gHeadScratchFloats * 4 = 256 KiB until
// the process exits.  Nothing in the array has to start at zero: dhead_partial_kernel writes every part[b][s] that
// dhead_loss_kernel reads.  ALL calls share the one array: two calls that overlap on the GPU (different streams) race.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kB2gHeadScratchFloats = 64 * 1024;  // 64 slices x up to 1024 samples
static __device__ __align__(256) float g_b2g_head_scratch[kB2gHeadScratchFloats];

// In host code `g_b2g_head_scratch` is only a stand-in symbol, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, once per device, with that device current.
static int head_scratch_ptr(float** out) {
  constexpr int kMaxDev = 64;
  static float* ptrs[kMaxDev] = {nullptr};
  int dev = 0;
  B2G_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) return fail(B2G_E_CUDA, "b2g_dhead_fwd_loss_static: device index %d out of range", dev);
  float* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    void* sym = nullptr;
    B2G_CUDA_CHECK(cudaGetSymbolAddress(&sym, g_b2g_head_scratch));
    if (!sym) return fail(B2G_E_CUDA, "b2g_dhead_fwd_loss_static: cudaGetSymbolAddress returned a null address");
    p = static_cast<float*>(sym);
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);  // racing threads store the same value
  }
  *out = p;
  return B2G_OK;
}

extern "C" int b2g_dhead_fwd_loss_static(const void* x, const float* w, const float* wscale, const float* bias, int B,
                                         int F, int loss_kind, float label, float inv_batch, float* logits,
                                         float* dlogit, float* loss_out, b2g_stream_t stream) {
  if (B < 1) return fail(B2G_E_INVALID_ARG, "b2g_dhead_fwd_loss_static: B must be positive");
  int S = (4 * num_sms() + B - 1) / B;  // same slicing as b2g_dhead_fwd_loss
  if (S < 1) S = 1;
  if (S > 64) S = 64;
  if ((long)B * S > kB2gHeadScratchFloats)
    return fail(B2G_E_UNSUPPORTED_SHAPE,
                "b2g_dhead_fwd_loss_static: batch %d needs %ld scratch floats, the static scratch holds %d (use "
                "b2g_dhead_fwd_loss with a caller workspace)", B, (long)B * S, kB2gHeadScratchFloats);
  float* scratch = nullptr;
  const int rc = head_scratch_ptr(&scratch);
  if (rc) return rc;
  return dhead_fwd_loss_any(x, w, wscale, bias, B, F, loss_kind, label, inv_batch, scratch, kB2gHeadScratchFloats, logits,
                            dlogit, loss_out, false, stream);
}
