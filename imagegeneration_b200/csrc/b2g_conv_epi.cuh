// b2g_conv_epi.cuh — fused epilogue shared by the implicit-GEMM conv kernels (b2g_conv.cu, b2g_conv_rows.cu).
//
// One call handles the 32 fp32 accumulator columns a thread just read from TMEM for ONE output pixel:
// bias -> LeakyReLU/ReLU/tanh -> dropout -> activation derivative recovered from the saved forward output ->
// bf16 NHWC store (or the narrow fp32 store of the 3-channel edge layers).
// Replaces the BiasAdd / LeakyRelu / Relu / Tanh / Dropout ops and their gradients that TensorFlow runs as
// separate kernels after each convolution (sndcgan/SNDCGAN.py:40-41,:61,:77-78).
#pragma once
#include "b2g_ptx.cuh"
#include "../../include/b200gan.h"

namespace b2g {

struct EpiParams {
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
  int out_f32;  // narrow fp32 output: `cs` real channels, out/yprev are float [pixels][cs]
  int cs;       // channel stride of the output tensor (== Cout unless out_f32)
  int Cout;
  int f32_io;   // tf32 path: out / yprev are full-width fp32 [pixels][Cout]
};

// r: 32 accumulator columns [cb, cb+32) of output pixel `pix` (flat N*H*W index); `valid` = pixel inside the tensor.
__device__ __forceinline__ void conv_epilogue_chunk(const EpiParams& p, const float* sbias, const uint32_t (&r)[32],
                                                    size_t pix, int cb, bool valid) {
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
    return;
  }
  const size_t off = pix * p.Cout + cb;  // element offset of this thread's 32 outputs
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
  if (p.dact != B2G_ACT_NONE && valid && p.f32_io) {
    const float4* yp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.yprev) + off);
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      const float4 yv = __ldg(yp + g);
      const float y4[4] = {yv.x, yv.y, yv.z, yv.w};
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const float y0 = y4[b];
        float f0;
        if (p.dact == B2G_ACT_LEAKY) f0 = y0 > 0.f ? p.dact_scale : (y0 < 0.f ? p.dact_alpha * p.dact_scale : p.dact_zero);
        else if (p.dact == B2G_ACT_RELU) f0 = y0 > 0.f ? 1.f : 0.f;
        else f0 = 1.f - y0 * y0;
        v[g * 4 + b] *= f0;
      }
    }
  } else if (p.dact != B2G_ACT_NONE && valid) {
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
  if (valid && p.f32_io) {
    float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + off);
#pragma unroll
    for (int g = 0; g < 8; ++g) op[g] = make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
  } else if (valid) {
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

// Fills the epilogue parameters from the C-ABI arguments.
inline void fill_epi_params(const b2g_conv_args* a, EpiParams* e) {
  e->out = static_cast<__nv_bfloat16*>(a->out);
  e->bias = a->bias;
  e->yprev = static_cast<const __nv_bfloat16*>(a->yprev);
  e->act = a->act;
  e->alpha = a->alpha;
  e->drop_thresh8 = a->dropout_thresh8;
  e->drop_scale = 256.f / (256.f - (float)a->dropout_thresh8);
  e->seed_lo = (uint32_t)(a->dropout_seed & 0xFFFFFFFFull);
  e->seed_hi = (uint32_t)(a->dropout_seed >> 32);
  e->dact = a->dact;
  e->dact_alpha = a->dact_alpha;
  e->dact_scale = a->dact_scale;
  e->dact_zero = a->dact_zero;
  e->out_f32 = a->out_f32;
  e->cs = a->out_f32 ? a->cout_store : a->Cout;
  e->Cout = a->Cout;
  e->f32_io = a->tf32;
}

}  // namespace b2g
