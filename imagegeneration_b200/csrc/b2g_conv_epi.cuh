// b2g_conv_epi.cuh — fused epilogue shared by the implicit-GEMM conv kernels (b2g_conv.cu, b2g_conv_rows.cu).
//
// One call handles the 32 fp32 accumulator columns a thread just read from TMEM for ONE output pixel:
// bias -> LeakyReLU/ReLU/tanh -> dropout -> activation derivative recovered from the saved forward output ->
// bf16 NHWC store (or the narrow fp32 store of the 3-channel edge layers).
// Replaces the BiasAdd / LeakyRelu / Relu / Tanh / Dropout ops and their gradients that TensorFlow runs as
// separate kernels after each convolution (sndcgan/SNDCGAN.py:40-41,:61,:77-78).
#pragma once
#include <stdlib.h>

#include "b2g_conv.cuh"
#include "b2g_host.h"
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
  float pre_scale;  // dropout scale 1/(1-rate) applied BEFORE the (positively homogeneous) activation, folded into
                    // the bias FFMA; the kernels' shared bias tables hold bias * pre_scale (epi_load_bias)
  int leaky_max;    // 0 <= alpha <= 1: LeakyReLU(v) = max(v, alpha * v)
  uint32_t seed_lo, seed_hi;
  const unsigned long long* seed_dev;  // non-null: the seed lives in device memory (graph replay), see epi_resolve
  int dact;
  float dact_alpha, dact_scale, dact_zero;
  int out_f32;  // narrow fp32 output: `cs` real channels, out/yprev are float [pixels][cs]
  int cs;       // channel stride of the output tensor (== Cout unless out_f32)
  int Cout;
  int f32_io;   // tf32 path: out / yprev are full-width fp32 [pixels][Cout]
  float* colsum;  // optional fp32 [Cout]: += per-channel sums of the stored (bf16-rounded) outputs over all pixels
                  // = bias / BatchNorm-beta gradient of the layer whose output gradient this kernel produces
  float* colsumsq;  // optional fp32 [Cout] (with colsum, bias == NULL): += per-channel sums of SQUARES of the stored
                    // outputs — with colsum the batch statistics of a training-mode BatchNorm behind this conv
                    // (WGAN path).  The per-CTA partials live in the kernels' bias table, unused without a bias.
  int cs_copies;  // the per-CTA column-sum tables hold this many copies of [Cout] (512 / Cout in the 512-float tables):
                  // warp w adds into copy w % cs_copies — fp32 shared-memory atomicAdd is a compare-and-swap loop,
                  // and with Cout = 64 eight epilogue warps would otherwise fight over the same 32 words per chunk
  int dbg;      // measurement switch (env B2G_DBG_EPI, never set in production): bit 0 = epilogue warps drain the
                // accumulators but skip math, saved-output loads and stores -> the launch time is the MMA/load pipeline alone
  int tma;      // 1: bf16 output (and yprev) move through per-warp shared-memory boxes and TMA (EpiTmaMaps)
  int qw, qh, qn;  // the 32 pixels one epilogue warp owns, as a box of the tile (qw * qh * qn == 32)
  int lean;     // TMA epilogue only (fill_epi_tma): 0 = general math (conv_epilogue_math); 1 = lean forward (bias, none /
                // ReLU / max-form LeakyReLU, dropout off or rate 0.5); 2 = lean ReLU'; 3 = lean LeakyReLU' with a zero
                // saved output -> 0 (dropout layers); 4 = lean LeakyReLU' with zero counted as negative.  Same results
                // bit for bit, a third of the instructions (epi_lean_fwd / epi_lean_bwd)
};

// The dropout seed of a call, resolved once per epilogue warp (it may live in device memory for graph replay).  Kept
// OUT of EpiParams on purpose: copying the whole parameter block into a per-thread struct to patch two words moved
// every epilogue constant from the constant bank into registers and cost the backward (dact) epilogues 15-25 %.
struct EpiSeed {
  uint32_t lo, hi;
};

// Tensor maps of the TMA epilogue: bf16 NHWC output / saved forward output, one map per output phase (transposed
// conv: phase (py, px) is the stride-2 sub-grid starting at (py, px)); box {32 channels, qw, qh, qn}, 64-byte swizzle.
struct alignas(64) EpiTmaMaps {
  CUtensorMap omap[4];
  CUtensorMap ymap[4];
};
constexpr int kEpiStageBytes = 2048;  // one warp's [32 pixels][32 channels] bf16 box

// bias -> activation -> dropout -> activation derivative -> bf16 (or full-width fp32) store for 32 channels of one
// pixel.  yload(g) returns 16-byte group g (8 bf16) of the saved forward output, ostore(g, o) stores output group g;
// both are only called for bf16 tensors and valid pixels.
template <class YLoad, class OStore>
__device__ __forceinline__ void conv_epilogue_math(const EpiParams& p, const float* sbias, const uint32_t (&r)[32],
                                                   size_t pix, int cb, bool valid, YLoad yload, OStore ostore,
                                                   float* s_colsum = nullptr, bool colsum_here = false,
                                                   float* s_colsq = nullptr, EpiSeed sd = EpiSeed{0u, 0u}) {
  const size_t off = pix * p.Cout + cb;  // element offset of this thread's 32 outputs
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  if (p.bias) {  // (acc + bias) * pre_scale as one FFMA per element
    const float ps = p.pre_scale;
#pragma unroll
    for (int j4 = 0; j4 < 8; ++j4) {
      const float4 b4 = *reinterpret_cast<const float4*>(&sbias[cb + j4 * 4]);
      v[j4 * 4 + 0] = fmaf(v[j4 * 4 + 0], ps, b4.x); v[j4 * 4 + 1] = fmaf(v[j4 * 4 + 1], ps, b4.y);
      v[j4 * 4 + 2] = fmaf(v[j4 * 4 + 2], ps, b4.z); v[j4 * 4 + 3] = fmaf(v[j4 * 4 + 3], ps, b4.w);
    }
  } else if (p.pre_scale != 1.f) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= p.pre_scale;
  }
  if (p.act == B2G_ACT_LEAKY && p.leaky_max) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], v[j] * p.alpha);
  } else if (p.act == B2G_ACT_LEAKY) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.f ? v[j] : v[j] * p.alpha;
  } else if (p.act == B2G_ACT_RELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  if (p.drop_thresh8 == 128 && valid && p.dact == B2G_ACT_NONE && !p.f32_io && !colsum_here) {
    // rate 0.5 (the reference's default, Trainer.py:22) on a bf16 forward epilogue: ONE hash word for the thread's 32
    // elements (b2g_keep128); word << s carries the keep bits of elements 4s .. 4s+3 at the byte tops, PRMT in
    // sign-replication mode makes the bf16x2 AND masks: 1.25 instructions per element plus one hash per 32.
    const uint32_t w = b2g_dropout_bits(static_cast<uint32_t>(off >> 5), sd.lo, sd.hi);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const uint32_t y0 = w << (2 * g), y1 = w << (2 * g + 1);
      uint4 o;
      o.x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]) & b2g_keep_mask_pair(y0, 0);
      o.y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]) & b2g_keep_mask_pair(y0, 1);
      o.z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]) & b2g_keep_mask_pair(y1, 0);
      o.w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]) & b2g_keep_mask_pair(y1, 1);
      ostore(g, o);
    }
    return;
  }
  if (p.drop_thresh8 == 128 && valid) {  // same one-bit masks, fp32 selects (fp32/tf32 path, fused column sums)
    const uint32_t w = b2g_dropout_bits(static_cast<uint32_t>(off >> 5), sd.lo, sd.hi);
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = b2g_keep128(w, j) ? v[j] : 0.f;
  } else if (p.drop_thresh8 > 0 && valid) {
    // general rates: element i keeps its value iff byte (i & 3) of the hash of group i >> 2 is >= thresh8: the byte is
    // moved to the top of the word and compared with thresh8 << 24 (the junk below it cannot change the outcome); the
    // 1/(1-rate) scale was applied before the activation (pre_scale)
    const uint32_t g0 = static_cast<uint32_t>(off >> 2);
    const uint32_t t24 = static_cast<uint32_t>(p.drop_thresh8) << 24;
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      const uint32_t bits = b2g_dropout_bits(g0 + g, sd.lo, sd.hi);
      v[g * 4 + 0] = (bits << 24) >= t24 ? v[g * 4 + 0] : 0.f;
      v[g * 4 + 1] = (bits << 16) >= t24 ? v[g * 4 + 1] : 0.f;
      v[g * 4 + 2] = (bits << 8) >= t24 ? v[g * 4 + 2] : 0.f;
      v[g * 4 + 3] = bits >= t24 ? v[g * 4 + 3] : 0.f;
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
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const uint4 yv = yload(g);
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
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      uint4 o;
      o.x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]);
      o.y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]);
      o.z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]);
      o.w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]);
      ostore(g, o);
      if (colsum_here) {  // keep the ROUNDED values for the column sums below
        v[g * 8 + 0] = bf16_lo(o.x); v[g * 8 + 1] = bf16_hi(o.x); v[g * 8 + 2] = bf16_lo(o.y); v[g * 8 + 3] = bf16_hi(o.y);
        v[g * 8 + 4] = bf16_lo(o.z); v[g * 8 + 5] = bf16_hi(o.z); v[g * 8 + 6] = bf16_lo(o.w); v[g * 8 + 7] = bf16_hi(o.w);
      }
    }
  }
  if (colsum_here) {
    // column sums over the warp's 32 pixels by a transposing butterfly: after the step with offset `off` every
    // lane holds half as many partial sums; lane l ends with the sum of channel cb + l (31 shuffles in total)
    if (!valid) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = 0.f;
    }
    const int lane = threadIdx.x & 31;
    const int cs_off = (static_cast<int>(threadIdx.x >> 5) % p.cs_copies) * p.Cout;
    if (s_colsq) {  // sums of squares first (the butterfly below consumes v)
      float q[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) q[j] = v[j] * v[j];
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
          const float mine = up ? q[i + off] : q[i];
          const float theirs = up ? q[i] : q[i + off];
          q[i] = mine + __shfl_xor_sync(0xffffffffu, theirs, off);
        }
      }
      atomicAdd(&s_colsq[cs_off + cb + lane], q[0]);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      const bool up = (lane & off) != 0;
#pragma unroll
      for (int i = 0; i < off; ++i) {
        const float mine = up ? v[i + off] : v[i];
        const float theirs = up ? v[i] : v[i + off];
        v[i] = mine + __shfl_xor_sync(0xffffffffu, theirs, off);
      }
    }
    atomicAdd(&s_colsum[cs_off + cb + lane], v[0]);
  }
}

// r: 32 accumulator columns [cb, cb+32) of output pixel `pix` (flat N*H*W index); `valid` = pixel inside the tensor.
__device__ __forceinline__ void conv_epilogue_chunk(const EpiParams& p, const float* sbias, const uint32_t (&r)[32],
                                                    size_t pix, int cb, bool valid, float* s_colsum = nullptr,
                                                    EpiSeed sd = EpiSeed{0u, 0u}) {
  if (p.dbg & 1) return;
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
  conv_epilogue_math(
      p, sbias, r, pix, cb, valid,
      [&](int g) { return __ldg(reinterpret_cast<const uint4*>(p.yprev + pix * p.Cout + cb) + g); },
      [&](int g, const uint4& o) { reinterpret_cast<uint4*>(p.out + pix * p.Cout + cb)[g] = o; }, s_colsum,
      p.colsum != nullptr, p.colsumsq ? const_cast<float*>(sbias) : nullptr, sd);
}

// Block-wide helpers of the fused column sums: per-CTA partial sums live in a shared table (zeroed before the
// block sync of the prologue, flushed with one atomic per channel after the final block sync).
__device__ __forceinline__ void epi_colsum_zero(const EpiParams& p, float* s_colsum, int n, float* sbias = nullptr) {
  const int m = n * p.cs_copies;
  if (p.colsum)
    for (int i = threadIdx.x; i < m; i += blockDim.x) s_colsum[i] = 0.f;
  if (p.colsumsq && sbias)  // sums of squares use the bias table (no bias in that mode)
    for (int i = threadIdx.x; i < m; i += blockDim.x) sbias[i] = 0.f;
}
__device__ __forceinline__ void epi_colsum_flush(const EpiParams& p, const float* s_colsum, int n,
                                                 const float* sbias = nullptr) {
  if (p.colsum)
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      float t = 0.f;
      for (int c = 0; c < p.cs_copies; ++c) t += s_colsum[c * n + i];
      atomicAdd(p.colsum + i, t);
    }
  if (p.colsumsq && sbias)
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      float t = 0.f;
      for (int c = 0; c < p.cs_copies; ++c) t += sbias[c * n + i];
      atomicAdd(p.colsumsq + i, t);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// TMA-staged variant.  A thread-per-pixel epilogue writes 16 bytes per lane into 32 different 128-byte lines per
// store instruction: 32 LSU wavefronts each, 128 per 32x32 chunk (and as many again for the saved-output loads of a
// backward epilogue) — measured as THE bottleneck of the layers with <= 128 output channels (profiles/r1_v6_*).
// Here each epilogue warp owns one 2 KiB shared-memory box [32 pixels][32 channels] (64-byte swizzle, conflict
// free 16-byte accesses): the saved output arrives in it by TMA, the results leave from it by TMA.
// ---------------------------------------------------------------------------------------------------------------
// ---------------------------------------------------------------------------------------------------------------
// Lean epilogue math of the TMA epilogue (EpiParams::lean).  ncu of round 2 showed the epilogue warps of every layer with
// <= 128 output channels issue-bound: 600-1000 instructions per 32x32 item, most of them one-fp32-lane-at-a-time math,
// a dropout hash per 4 elements and compare/select chains.  Here: packed fp32 pairs (FFMA2 / FMUL2), one dropout hash
// per 32 elements (b2g_keep128), derivative selects as bf16x2 compare masks (HSET2) + one LOP3 on packed words, ReLU
// folded into the bf16 conversion.  Bit-identical to conv_epilogue_math for the cases fill_epi_tma routes here.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_bf16x2_relu(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// s += bf16 half of w, exact widening inside the add (FHADD.BF16 takes the register half directly: no unpack)
__device__ __forceinline__ void bf2_acc(uint32_t w, float& s_lo, float& s_hi) {
  asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %2;\n\tadd.rn.f32.bf16 %0, l, %0;\n\tadd.rn.f32.bf16 %1, h, %1;\n\t}"
      : "+f"(s_lo), "+f"(s_hi)
      : "r"(w));
}
__device__ __forceinline__ void bf2_acc_sq(uint32_t w, float& q_lo, float& q_hi) {
  asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %2;\n\tfma.rn.f32.bf16 %0, l, l, %0;\n\tfma.rn.f32.bf16 %1, h, h, %1;\n\t}"
      : "+f"(q_lo), "+f"(q_hi)
      : "r"(w));
}

// forward: r = the thread's 32 accumulator columns [cb, cb+32) of its pixel; `grp32` = (flat element index of the first
// one) >> 5; results go to the thread's 64-byte row of the warp's box (64-byte swizzle `sw`).
__device__ __forceinline__ void epi_lean_fwd(const EpiParams& p, const float* sbias, const uint32_t (&r)[32], uint32_t grp32,
                                             int cb, uint8_t* row, uint32_t sw, EpiSeed sd) {
  uint64_t v[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) v[k] = f2_pack(__uint_as_float(r[2 * k]), __uint_as_float(r[2 * k + 1]));
  const uint64_t ps2 = f2_pack(p.pre_scale, p.pre_scale);
  if (p.bias) {  // (acc * pre_scale + bias * pre_scale): the shared table holds the scaled bias
#pragma unroll
    for (int k4 = 0; k4 < 8; ++k4) {
      const float4 b4 = *reinterpret_cast<const float4*>(&sbias[cb + k4 * 4]);
      v[2 * k4] = f2_fma(v[2 * k4], ps2, f2_pack(b4.x, b4.y));
      v[2 * k4 + 1] = f2_fma(v[2 * k4 + 1], ps2, f2_pack(b4.z, b4.w));
    }
  } else if (p.pre_scale != 1.f) {
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = f2_mul(v[k], ps2);
  }
  uint32_t o[16];
  if (p.act == B2G_ACT_LEAKY) {  // max(v, alpha * v), 0 <= alpha <= 1
    const uint64_t al2 = f2_pack(p.alpha, p.alpha);
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float v0, v1, a0, a1;
      f2_unpack(v[k], v0, v1);
      f2_unpack(f2_mul(v[k], al2), a0, a1);
      o[k] = pack_bf16x2(fmaxf(v0, a0), fmaxf(v1, a1));
    }
  } else if (p.act == B2G_ACT_RELU) {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float v0, v1;
      f2_unpack(v[k], v0, v1);
      o[k] = pack_bf16x2_relu(v0, v1);
    }
  } else {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float v0, v1;
      f2_unpack(v[k], v0, v1);
      o[k] = pack_bf16x2(v0, v1);
    }
  }
  if (p.drop_thresh8 == 128) {
    const uint32_t w = b2g_dropout_bits(grp32, sd.lo, sd.hi);
#pragma unroll
    for (int s = 0; s < 8; ++s) {
      const uint32_t y = w << s;
      o[2 * s] &= b2g_keep_mask_pair(y, 0);
      o[2 * s + 1] &= b2g_keep_mask_pair(y, 1);
    }
  }
#pragma unroll
  for (int g = 0; g < 4; ++g)
    *reinterpret_cast<uint4*>(row + ((g ^ sw) << 4)) = make_uint4(o[4 * g], o[4 * g + 1], o[4 * g + 2], o[4 * g + 3]);
}

// backward: the thread's row of the box holds the saved forward output y; it is replaced by acc * act'(y).
// MODE 2: ReLU' (y > 0);  3: LeakyReLU' scale / alpha*scale / 0 for y > 0 / < 0 / == 0;  4: y <= 0 -> alpha*scale.
template <int MODE>
__device__ __forceinline__ void epi_lean_bwd(const EpiParams& p, const uint32_t (&r)[32], uint8_t* row, uint32_t sw) {
  const float as = p.dact_alpha * p.dact_scale;
  const uint64_t s2 = f2_pack(p.dact_scale, p.dact_scale), as2 = f2_pack(as, as);
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint4* slot = reinterpret_cast<uint4*>(row + ((g ^ sw) << 4));
    const uint4 yv = *slot;
    const uint32_t y4[4] = {yv.x, yv.y, yv.z, yv.w};
    uint32_t o[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const float v0 = __uint_as_float(r[g * 8 + b * 2]), v1 = __uint_as_float(r[g * 8 + b * 2 + 1]);
      const uint32_t pos = bf2_gt0_mask(y4[b]);
      if (MODE == 2) {
        o[b] = pack_bf16x2(v0, v1) & pos;
      } else {
        const uint64_t v = f2_pack(v0, v1);
        float a0, a1, b0, b1;
        f2_unpack(f2_mul(v, s2), a0, a1);
        f2_unpack(f2_mul(v, as2), b0, b1);
        const uint32_t pa = pack_bf16x2(a0, a1), pb = pack_bf16x2(b0, b1);
        if (MODE == 4) o[b] = (pa & pos) | (pb & ~pos);
        else o[b] = (pa & pos) | (pb & bf2_lt0_mask(y4[b]));
      }
    }
    *slot = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// Every epilogue warp owns NB boxes.  Item k (the k-th 32-pixel x 32-channel chunk the warp processes, counted over the
// whole kernel) lives in box k % NB: a backward (dact) epilogue first loads the SAVED forward output into it by TMA
// (epi_tma_begin), the math runs in place, the result leaves by TMA (epi_tma_chunk).  With NB = 2 (where shared memory
// allows) begin(k + 1) is issued while item k is still being computed.
//
// What round 2 measured about this epilogue (tools/conv_probe.py with B2G_DBG_EPI, ncu source pages, profiles/r2_*):
//  * a backward item is 600-1000 instructions of ONE warp's dependent work; 8 warps with 3-box rings and run-ahead loads
//    were 45 % SLOWER than 16 warps with one box — the epilogue needs warps and fewer instructions first;
//  * fetching the saved output with ordinary global loads (registers, one item ahead) or cp.async instead of TMA was
//    slower still: fence.proxy.async — needed before every TMA store — waits for the thread's outstanding loads, so the
//    "prefetch" of item k + 1 stalled item k's store (ncu: FENCE.VIEW.ASYNC / ELECT with long-scoreboard stalls);
//    TMA loads are not on the thread's scoreboard;
//  * tile decoding by runtime division, per item, was 12 % of the instructions (now multiply-high, FastDiv), the column
//    sums read back from the box 17 % (now 16 LDS.32 per lane instead of 32 LDS.U16 + unconditional squares).
template <int NB>
struct EpiWarpStageN {
  uint8_t* buf;      // this warp's NB boxes (1 KiB aligned), box b at buf + b * kEpiStageBytes
  uint64_t* ybar;    // NB mbarriers of the saved-output loads
  uint32_t k_begin;  // items begun
  uint32_t k_chunk;  // items processed
  // running column sums of this lane's two channels (plain sums only; with colsumsq the per-item path is used): kept in
  // registers ACROSS items while the warp stays on the same channels (it does: a warp's chunk is fixed by its group),
  // combined across the lane halves and added to the CTA's table once, in epi_tma_colsum_flush
  float cs0, cs1;
  int cs_cb;         // first channel of the 32-channel chunk the running sums belong to; -1: none
};
using EpiWarpStage = EpiWarpStageN<1>;

// Shared memory the TMA epilogue needs behind a kernel's own buffers (1 KiB aligned): boxes, then mbarriers.
__host__ __device__ constexpr int epi_tma_smem_bytes(int warps, int nb = 1) {
  return warps * nb * kEpiStageBytes + ((warps * nb * 8 + 255) / 256) * 256;
}

template <int NB = 1>
__device__ __forceinline__ EpiWarpStageN<NB> epi_tma_stage(uint8_t* base, int warps, int w) {
  EpiWarpStageN<NB> st;
  st.buf = base + w * NB * kEpiStageBytes;
  st.ybar = reinterpret_cast<uint64_t*>(base + warps * NB * kEpiStageBytes) + w * NB;
  st.k_begin = 0;
  st.k_chunk = 0;
  st.cs0 = 0.f;
  st.cs1 = 0.f;
  st.cs_cb = -1;
  return st;
}
// by one thread, before fence_barrier_init() and the block-wide sync
__device__ __forceinline__ void epi_tma_init(uint8_t* base, int warps, int nb = 1) {
  for (int w = 0; w < warps * nb; ++w) mbar_init(reinterpret_cast<uint64_t*>(base + warps * nb * kEpiStageBytes) + w, 1);
}

template <int N>
__device__ __forceinline__ void tma_store_wait_read_n() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// Whole warp: begins the warp's next item — frees its box (the store that last used it has read it) and starts loading
// the saved forward output of channels [cb, cb+32) of the warp's pixels (box origin bx, by, bn in the phase grid).
// Call it as early as possible (before waiting for the accumulator; with NB = 2 one item ahead).  begin(k) must come
// after chunk(k - NB) and before chunk(k).
template <int NB>
__device__ __forceinline__ void epi_tma_begin(const EpiParams& p, const CUtensorMap* ymap, EpiWarpStageN<NB>& st,
                                              int cb, int bx, int by, int bn) {
  if (p.dbg & 1) return;
  const uint32_t b = st.k_begin % NB;
  ++st.k_begin;
  if (p.dact == B2G_ACT_NONE) return;  // forward epilogues have nothing to load; chunk() frees the box itself
  if (elect_one_sync()) {  // elect.sync picks the same lane for the same mask: the bulk groups stay with one thread
    // NB = 1: the previous item's store; NB = 2 (one item ahead): the store of item k - 2 is the newest one issued
    tma_store_wait_read_n<0>();
    mbar_expect_tx(&st.ybar[b], kEpiStageBytes);
    tma_load_4d(st.buf + b * kEpiStageBytes, ymap, &st.ybar[b], cb, bx, by, bn);
  }
  __syncwarp();
}

// Whole warp: adds the running column sums of the warp (EpiWarpStageN::cs0 / cs1) into the CTA's shared table.  Call
// when the warp moves to other channels and before the kernel's final block-wide sync.
template <int NB>
__device__ __forceinline__ void epi_tma_colsum_flush(const EpiParams& p, EpiWarpStageN<NB>& st, float* s_colsum, int lane) {
  if (st.cs_cb < 0) return;
  const uint32_t cp = lane & 15, h = lane >> 4;
  const int cs_off = (static_cast<int>(threadIdx.x >> 5) % p.cs_copies) * p.Cout;
  float s0 = st.cs0, s1 = st.cs1;
  s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
  atomicAdd(&s_colsum[cs_off + st.cs_cb + 2 * cp + h], h ? s1 : s0);
  st.cs0 = 0.f;
  st.cs1 = 0.f;
  st.cs_cb = -1;
}

// Whole warp: epilogue math of one chunk (lane = pixel) through the box, then the TMA store.
template <int NB>
__device__ __forceinline__ void epi_tma_chunk(const EpiParams& p, const CUtensorMap* omap, EpiWarpStageN<NB>& st,
                                              const float* sbias, const uint32_t (&r)[32], size_t pix, int cb, int lane,
                                              int bx, int by, int bn, float* s_colsum = nullptr,
                                              EpiSeed sd = EpiSeed{0u, 0u}) {
  if (p.dbg & 1) return;
  const uint32_t k = st.k_chunk++;
  const uint32_t b = k % NB;
  uint8_t* box = st.buf + b * kEpiStageBytes;
  uint8_t* row = box + lane * 64;
  const uint32_t sw = (lane >> 1) & 3;  // 64-byte swizzle: 16-byte group g of row i lives at g ^ ((i >> 1) & 3)
  if (p.dact != B2G_ACT_NONE) {
    mbar_wait(&st.ybar[b], (k / NB) & 1);
  } else {
    // forward: the box is free once the store of item k - NB has read it (NB - 1 newer stores may still be pending)
    if (elect_one_sync()) tma_store_wait_read_n<NB - 1>();
    __syncwarp();
  }
  if (p.lean == 1)
    epi_lean_fwd(p, sbias, r, static_cast<uint32_t>((pix * p.Cout + cb) >> 5), cb, row, sw, sd);
  else if (p.lean == 2)
    epi_lean_bwd<2>(p, r, row, sw);
  else if (p.lean == 3)
    epi_lean_bwd<3>(p, r, row, sw);
  else if (p.lean == 4)
    epi_lean_bwd<4>(p, r, row, sw);
  else
    conv_epilogue_math(
        p, sbias, r, pix, cb, true,
        [&](int g) { return *reinterpret_cast<const uint4*>(row + ((g ^ sw) << 4)); },
        [&](int g, const uint4& o) { *reinterpret_cast<uint4*>(row + ((g ^ sw) << 4)) = o; }, nullptr, false, nullptr,
        sd);
  fence_proxy_async_smem();
  __syncwarp();
  if (elect_one_sync()) tma_store_4d(omap, box, cb, bx, by, bn);
  __syncwarp();
  if (p.colsum) {
    // Column sums straight from the box (the TMA store only READS it: no hazard).  Lane = (channel pair cp = lane & 15,
    // pixel parity h = lane >> 4): 16 LDS.32 per lane — the 16 lanes of a half warp cover one 64-byte pixel row, the two
    // halves rows 2j and 2j + 1 = all 32 banks, conflict free — then one shuffle joins the two parities.  Round 1 read
    // one bf16 per lane and pixel (32 LDS.U16 + sum-of-squares FMAs even without BatchNorm statistics): 185 instructions
    // per item, 17 % of a backward epilogue (ncu r2).  Rows of pixels outside the tensor hold exact zeros: masked by the
    // zero-filled saved output when act'(0) = 0, else the callers clear those pixels' accumulators (zero_oob).
    const uint32_t cp = lane & 15, h = lane >> 4;
    const uint8_t* base = box + h * 64 + (cp & 3) * 4;
    const uint32_t cg = cp >> 2;
    float s0 = 0.f, s1 = 0.f;
    const int cs_off = (static_cast<int>(threadIdx.x >> 5) % p.cs_copies) * p.Cout;
    if (p.colsumsq) {
      float q0 = 0.f, q1 = 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(base + j * 128 + ((cg ^ (j & 3)) << 4));
        bf2_acc(w, s0, s1);
        bf2_acc_sq(w, q0, q1);
      }
      q0 += __shfl_xor_sync(0xffffffffu, q0, 16);
      q1 += __shfl_xor_sync(0xffffffffu, q1, 16);
      atomicAdd(&const_cast<float*>(sbias)[cs_off + cb + 2 * cp + h], h ? q1 : q0);
    } else {
      // plain sums: accumulate into the warp's running registers (no shuffle, no shared atomic per item)
      if (st.cs_cb != cb) {
        epi_tma_colsum_flush(p, st, s_colsum, lane);
        st.cs_cb = cb;
      }
      s0 = st.cs0;
      s1 = st.cs1;
#pragma unroll
      for (int j = 0; j < 16; ++j)  // one mixed-precision add per element (FHADD.BF16), no unpacking
        bf2_acc(*reinterpret_cast<const uint32_t*>(base + j * 128 + ((cg ^ (j & 3)) << 4)), s0, s1);
      st.cs0 = s0;
      st.cs1 = s1;
      return;
    }
    s0 += __shfl_xor_sync(0xffffffffu, s0, 16);
    s1 += __shfl_xor_sync(0xffffffffu, s1, 16);
    atomicAdd(&s_colsum[cs_off + cb + 2 * cp + h], h ? s1 : s0);
  }
}

// Whole warp, before the kernel's final block-wide sync: the last stores must have read their boxes.
__device__ __forceinline__ void epi_tma_end() {
  if (elect_one_sync()) tma_store_wait_read();
  __syncwarp();
}
template <int NB>
__device__ __forceinline__ void epi_tma_end(const EpiParams& p, EpiWarpStageN<NB>& st, float* s_colsum, int lane) {
  if (p.colsum) epi_tma_colsum_flush(p, st, s_colsum, lane);
  epi_tma_end();
}

// The dropout seed of this call: from the arguments, or read from device memory when the call asked for it
// (b2g_conv_args.dropout_seed_dev: graph replay).
__device__ __forceinline__ EpiSeed epi_resolve_seed(const EpiParams& p) {
  EpiSeed sd{p.seed_lo, p.seed_hi};
  if (p.seed_dev) {
    const unsigned long long s = *p.seed_dev;
    sd.lo = static_cast<uint32_t>(s & 0xFFFFFFFFull);
    sd.hi = static_cast<uint32_t>(s >> 32);
  }
  return sd;
}

// Block-wide: shared bias table = bias[0..n) * pre_scale (see EpiParams::pre_scale); call before the block sync.
__device__ __forceinline__ void epi_load_bias(const EpiParams& p, float* sbias, int n) {
  if (p.bias)
    for (int i = threadIdx.x; i < n; i += blockDim.x) sbias[i] = p.bias[i] * p.pre_scale;
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
  e->pre_scale = a->dropout_thresh8 > 0 ? e->drop_scale : 1.f;  // act is NONE / LEAKY / RELU here (validated)
  e->leaky_max = (a->alpha >= 0.f && a->alpha <= 1.f) ? 1 : 0;
  e->seed_lo = (uint32_t)(a->dropout_seed & 0xFFFFFFFFull);
  e->seed_hi = (uint32_t)(a->dropout_seed >> 32);
  e->seed_dev = a->dropout_seed_dev;
  e->dact = a->dact;
  e->dact_alpha = a->dact_alpha;
  e->dact_scale = a->dact_scale;
  e->dact_zero = a->dact_zero;
  e->out_f32 = a->out_f32;
  e->cs = a->out_f32 ? a->cout_store : a->Cout;
  e->Cout = a->Cout;
  e->f32_io = a->tf32;
  e->colsum = a->colsum;
  e->colsumsq = a->colsumsq;
  e->cs_copies = (a->Cout > 0 && a->Cout <= 512) ? (512 / a->Cout > 0 ? 512 / a->Cout : 1) : 1;  // 512-float tables
  e->tma = 0;
  e->lean = 0;
  {
    const char* d = getenv("B2G_DBG_EPI");  // read per call: measurement runs flip it between launches
    e->dbg = d ? atoi(d) : 0;
  }
}

// Host: decides whether the TMA epilogue applies and builds its tensor maps (b2g_conv_epi_host.cu-free: inline).
// `enabled` lets a kernel without shared-memory room opt out.
// `row32`: the kernel's warps own 32 consecutive pixels of one row (row-slab kernel) whatever the plan's box says.
inline int fill_epi_tma(const b2g_conv_args* a, const ConvPlan& pl, bool enabled, EpiParams* e, EpiTmaMaps* m,
                        bool row32 = false) {
  static int env = []() { const char* s = getenv("B2G_EPI_TMA"); return s ? atoi(s) : 1; }();
  e->tma = 0;
  if (!enabled || !env || a->out_f32 || a->tf32 || a->Cout % 32) return B2G_OK;
  const int qw = row32 ? 32 : (pl.bw < 32 ? pl.bw : 32);
  const int qh = row32 ? 1 : (pl.bh < 32 / qw ? pl.bh : 32 / qw);
  const int qn = 32 / (qw * qh);
  if (qw * qh * qn != 32 || (!row32 && qn > pl.bn)) return B2G_OK;
  const int os = pl.os;
  const size_t es = 2;
  for (int ph = 0; ph < pl.phases; ++ph) {
    const size_t boff = ((size_t)pl.ph_y[ph] * pl.Wout + pl.ph_x[ph]) * a->Cout * es;
    const uint64_t dims[4] = {(uint64_t)a->Cout, (uint64_t)pl.Wout / os, (uint64_t)pl.Hout / os, (uint64_t)pl.N};
    const uint64_t str[3] = {(uint64_t)os * a->Cout * es, (uint64_t)os * pl.Wout * a->Cout * es,
                             (uint64_t)pl.Hout * pl.Wout * a->Cout * es};
    const uint32_t box[4] = {32u, (uint32_t)qw, (uint32_t)qh, (uint32_t)qn};
    int rc = make_tmap(&m->omap[ph], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, static_cast<char*>(a->out) + boff, dims, str,
                       box, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
    if (a->dact != B2G_ACT_NONE) {
      rc = make_tmap(&m->ymap[ph], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, static_cast<const char*>(a->yprev) + boff, dims,
                     str, box, CU_TENSOR_MAP_SWIZZLE_64B);
      if (rc) return rc;
    }
  }
  e->tma = 1;
  e->qw = qw; e->qh = qh; e->qn = qn;
  {
    static int lean_env = []() { const char* s = getenv("B2G_EPI_LEAN"); return s ? atoi(s) : 1; }();
    const bool act_ok = a->act == B2G_ACT_NONE || a->act == B2G_ACT_RELU || (a->act == B2G_ACT_LEAKY && e->leaky_max);
    const bool drop_ok = a->dropout_thresh8 == 0 || a->dropout_thresh8 == 128;
    e->lean = 0;
    if (!lean_env) {
    } else if (a->dact == B2G_ACT_NONE && !a->colsum && !a->colsumsq && act_ok && drop_ok) {
      e->lean = 1;
    } else if (a->dact != B2G_ACT_NONE && !a->bias && a->act == B2G_ACT_NONE && !a->dropout_thresh8 && !a->colsumsq) {
      if (a->dact == B2G_ACT_RELU) e->lean = 2;
      else if (a->dact == B2G_ACT_LEAKY && a->dact_zero == 0.f) e->lean = 3;
      else if (a->dact == B2G_ACT_LEAKY && a->dact_zero == a->dact_alpha * a->dact_scale) e->lean = 4;
    }
  }
  return B2G_OK;
}

}  // namespace b2g
