// b2g_misc.cu — small HBM-bound kernels either side of the training step:
//   * dact:  dpre = g * act'(y) from the saved forward output — the activation/dropout derivative as a stand-alone pass.
//            train_step never launches it (its derivative epilogues are fused into the input-gradient kernels); the
//            torch.library autograd formulas (imagegeneration_b200/torch_ops.py) do, because autograd hands a layer the
//            gradient of its OUTPUT.
//   * bce_logits: tf.keras.losses.BinaryCrossentropy(from_logits=True) (sndcgan/SNDCGAN.py:170) on a logits vector with
//            a label vector: loss scalar and dL/dlogit, one block.
//   * gather_u8: batch assembly from a device-resident uint8 image cache: out[b] = images[idx[b]] / 127.5 - 1
//            (`Rescaling(1/127.5, offset=-1)` + `cache().shuffle()` of get_dataset, sndcgan/SNDCGAN.py:141-143).
//   * colsum_f32: out[c] += sum_p x[p][c] for fp32 activations (bias gradient on the fp32/tf32 path).
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

__device__ __forceinline__ float dact_factor(int kind, float y, float alpha, float scale, float zero) {
  if (kind == B2G_ACT_LEAKY) return y > 0.f ? scale : (y < 0.f ? alpha * scale : zero);
  if (kind == B2G_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (kind == B2G_ACT_TANH) return 1.f - y * y;
  return 1.f;
}

__global__ void __launch_bounds__(256) dact_bf16_kernel(const uint4* __restrict__ g, const uint4* __restrict__ y,
                                                        uint4* __restrict__ out, long n8, int kind, float alpha,
                                                        float scale, float zero) {
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n8; i += gridDim.x * 256L) {
    const uint4 gv = __ldg(g + i), yv = __ldg(y + i);
    const uint32_t gw[4] = {gv.x, gv.y, gv.z, gv.w}, yw[4] = {yv.x, yv.y, yv.z, yv.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      o[k] = pack_bf16x2(bf16_lo(gw[k]) * dact_factor(kind, bf16_lo(yw[k]), alpha, scale, zero),
                         bf16_hi(gw[k]) * dact_factor(kind, bf16_hi(yw[k]), alpha, scale, zero));
    out[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

__global__ void __launch_bounds__(256) dact_f32_kernel(const float* __restrict__ g, const float* __restrict__ y,
                                                       float* __restrict__ out, long n, int kind, float alpha,
                                                       float scale, float zero) {
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n; i += gridDim.x * 256L)
    out[i] = g[i] * dact_factor(kind, y[i], alpha, scale, zero);
}

__global__ void bce_logits_kernel(const float* __restrict__ logits, const float* __restrict__ labels, int B,
                                  float* __restrict__ loss_out, float* __restrict__ dlogit) {
  __shared__ float red[32];
  float lsum = 0.f;
  const float inv = 1.f / (float)B;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const float x = logits[b], t = labels[b];
    lsum += fmaxf(x, 0.f) - x * t + log1pf(__expf(-fabsf(x)));
    if (dlogit) dlogit[b] = (1.f / (1.f + __expf(-x)) - t) * inv;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = lsum;
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f;
    for (int i = 0; i < (int)(blockDim.x + 31) / 32; ++i) a += red[i];
    loss_out[0] = a * inv;
  }
}

// one image = per_image bytes (multiple of 16); a block streams 16-byte groups: 16 uint8 -> 4 float4
__global__ void __launch_bounds__(256) gather_u8_kernel(const uint8_t* __restrict__ images, const long* __restrict__ idx,
                                                        float* __restrict__ out, int B, long per_image) {
  const long groups = per_image / 16;
  const long total = groups * B;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
    const long b = i / groups, gidx = i % groups;
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(images + idx[b] * per_image) + gidx);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    float4* o = reinterpret_cast<float4*>(out + b * per_image) + gidx * 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float s = 1.f / 127.5f;
      __stcs(o + k, make_float4((float)(w[k] & 0xFF) * s - 1.f, (float)((w[k] >> 8) & 0xFF) * s - 1.f,
                                (float)((w[k] >> 16) & 0xFF) * s - 1.f, (float)(w[k] >> 24) * s - 1.f));
    }
  }
}

__global__ void __launch_bounds__(256) colsum_f32_kernel(const float* __restrict__ x, float* __restrict__ out, long P,
                                                         int C) {
  // thread t owns channels (t % (C/4)) * 4 .. +3 of rows t / (C/4), t / (C/4) + rows_par, ...
  extern __shared__ float sacc[];
  const int groups = C / 4;
  const int rows_par = 256 / groups > 0 ? 256 / groups : 1;
  for (int i = threadIdx.x; i < C; i += 256) sacc[i] = 0.f;
  __syncthreads();
  const int cg = threadIdx.x % groups, r = threadIdx.x / groups;
  if (r < rows_par) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long p = (long)blockIdx.x * rows_par + r; p < P; p += (long)gridDim.x * rows_par) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(x + p * C) + cg);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    atomicAdd(&sacc[cg * 4 + 0], acc.x); atomicAdd(&sacc[cg * 4 + 1], acc.y);
    atomicAdd(&sacc[cg * 4 + 2], acc.z); atomicAdd(&sacc[cg * 4 + 3], acc.w);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += 256) atomicAdd(out + i, sacc[i]);
}

}  // namespace b2g

using namespace b2g;

static int grid_for(long work_items) {
  long blocks = (work_items + 255) / 256;
  const long cap = (long)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

extern "C" int b2g_dact(const void* g, const void* y, void* out, long n, int f32, int kind, float alpha, float scale,
                        float zero, b2g_stream_t stream_) {
  if (!g || !y || !out) return fail(B2G_E_INVALID_ARG, "b2g_dact: null pointer");
  if (kind < B2G_ACT_NONE || kind > B2G_ACT_TANH) return fail(B2G_E_INVALID_ARG, "b2g_dact: unknown activation %d", kind);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (f32) {
    dact_f32_kernel<<<grid_for(n), 256, 0, stream>>>(static_cast<const float*>(g), static_cast<const float*>(y),
                                                     static_cast<float*>(out), n, kind, alpha, scale, zero);
  } else {
    if (n % 8) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_dact: bf16 tensors need a multiple of 8 elements");
    dact_bf16_kernel<<<grid_for(n / 8), 256, 0, stream>>>(static_cast<const uint4*>(g), static_cast<const uint4*>(y),
                                                          static_cast<uint4*>(out), n / 8, kind, alpha, scale, zero);
  }
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_bce_logits(const float* logits, const float* labels, int B, float* loss_out, float* dlogit,
                              b2g_stream_t stream_) {
  if (!logits || !labels || !loss_out || B < 1) return fail(B2G_E_INVALID_ARG, "b2g_bce_logits: bad arguments");
  bce_logits_kernel<<<1, 256, 0, static_cast<cudaStream_t>(stream_)>>>(logits, labels, B, loss_out, dlogit);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_gather_u8_to_f32(const void* images_u8, const long* idx, float* out, int B, long per_image,
                                    b2g_stream_t stream_) {
  if (!images_u8 || !idx || !out || B < 1) return fail(B2G_E_INVALID_ARG, "b2g_gather_u8_to_f32: bad arguments");
  if (per_image % 16) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_gather_u8_to_f32: image bytes must be a multiple of 16");
  gather_u8_kernel<<<grid_for(per_image / 16 * B), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const uint8_t*>(images_u8), idx, out, B, per_image);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_colsum_f32(const float* x, float* out, long P, int C, b2g_stream_t stream_) {
  if (!x || !out) return fail(B2G_E_INVALID_ARG, "b2g_colsum_f32: null pointer");
  if (C % 4 || C > 1024 || C < 4) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_colsum_f32: C=%d unsupported", C);
  const int groups = C / 4;
  const int rows_par = 256 / groups > 0 ? 256 / groups : 1;
  long blocks = (P + rows_par * 16 - 1) / (rows_par * 16);
  const long cap = (long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  colsum_f32_kernel<<<(int)blocks, 256, C * sizeof(float), static_cast<cudaStream_t>(stream_)>>>(x, out, P, C);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}
