// b2g_api.cu — library-wide state of the C ABI: version, thread-local error text, driver entry points,
// and the weight re-packing kernel that feeds the implicit-GEMM convolutions.
#include <stdlib.h>
#include <string.h>

#include "b2g_conv.cuh"
#include "b2g_host.h"
#include "b2g_ptx.cuh"

namespace b2g {

char* error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess) return nullptr;
    if (q != cudaDriverEntryPointSuccess) return nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

static int g_sm_limit = 0;  // b2g_set_sm_limit: 0 = all SMs

int num_sms() {
  const int hw = hw_num_sms();
  const int lim = g_sm_limit;
  return (lim > 0 && lim < hw) ? lim : hw;
}

int hw_num_sms() {
  static int n = []() {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) return 148;
    return v;
  }();
  return n;
}

// One thread per packed element: wpacked[ph][n][tap*Ck + c] (bf16) <- w[ky][kx][d2][d3] (fp32) * scales.
template <typename OutT>
__global__ void pack_conv_weight_kernel(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad,
                                        const float* __restrict__ w, const float* __restrict__ scale_scalar,
                                        const float* __restrict__ scale_d2, const float* __restrict__ scale_d3,
                                        OutT* __restrict__ out, int phases, int taps, long total) {
  const float s0 = scale_scalar ? *scale_scalar : 1.f;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    int d2, d3;
    const long src = pack_source_index(kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad, taps, i, &d2, &d3);
    float v = 0.f;
    if (src >= 0) {
      v = w[src] * s0;
      if (scale_d2) v *= scale_d2[d2];
      if (scale_d3) v *= scale_d3[d3];
    }
    if constexpr (sizeof(OutT) == 4) {
      // round to nearest tf32 here: tcgen05 kind::tf32 truncates the low 13 mantissa bits of what it reads
      uint32_t t;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t) : "f"(v));
      out[i] = __uint_as_float(t);
    } else {
      out[i] = __float2bfloat16_rn(v);
    }
  }
}

// All packings of one model in ONE launch (the train step re-packs every conv weight after each optimizer update).
// The jobs travel as kernel parameters; block b works on tile (b - first_block[j]) of job j.  A tile is 32 output
// rows n x 64 packed columns c of one (phase, tap): the fp32 source is read along whichever of n / c is contiguous in
// the Keras layout (coalesced), transposed through shared memory and written as 128-byte bf16 rows.
constexpr int kMaxPackJobs = 24;
constexpr int kPackTileN = 32, kPackTileC = 64;
struct PackJobDev {
  int kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad, taps, phases, first_block;
  int Cn, Ckr, Ck, n_blocks, c_blocks;
  int f32;  // fp32 packing (rounded to tf32) instead of bf16
  const float *w, *scale_scalar, *scale_d2, *scale_d3;
  void* out;
};
struct alignas(16) PackMultiParams {
  PackJobDev job[kMaxPackJobs];
  int n_jobs;
};

struct PackTile {
  const PackJobDev* q;
  const float* src;  // [D2][D3] plane of this tile's tap
  int ph, tap, n0, c0;
};
__device__ __forceinline__ PackTile pack_tile_decode(const PackMultiParams& p, int t) {
  int j = 0;
  while (j + 1 < p.n_jobs && t >= p.job[j + 1].first_block) ++j;
  const PackJobDev& q = p.job[j];
  t -= q.first_block;
  PackTile o;
  o.q = &q;
  const int cb = t % q.c_blocks; t /= q.c_blocks;
  const int nb = t % q.n_blocks; t /= q.n_blocks;
  o.tap = t % q.taps;
  o.ph = t / q.taps;
  o.n0 = nb * kPackTileN;
  o.c0 = cb * kPackTileC;
  int ky, kx;
  if (q.kind == B2G_CONVT_K4S2) {
    int d;
    convt_tap(o.ph >> 1, o.tap >> 1, &ky, &d);
    convt_tap(o.ph & 1, o.tap & 1, &kx, &d);
  } else {
    ky = o.tap / q.kw;
    kx = o.tap % q.kw;
    if (q.flip) {
      ky = q.kh - 1 - ky;
      kx = q.kw - 1 - kx;
    }
  }
  o.src = q.w + ((size_t)ky * q.kw + kx) * q.D2 * q.D3;
  return o;
}
// the 8 scaled fp32 values this thread contributes to a tile (coalesced along whichever of n / c is contiguous)
__device__ __forceinline__ void pack_tile_load(const PackTile& t, int tid, float (&v)[8]) {
  const PackJobDev& q = *t.q;
  const float s0 = q.scale_scalar ? *q.scale_scalar : 1.f;
  if (q.n_is_d3) {  // n = d3 is the contiguous source index: a warp reads 32 consecutive n of one c
    const int nx = tid & 31, cy = tid >> 5;
    const int n = t.n0 + nx;
    const float sn = (q.scale_d3 && n < q.Cn) ? q.scale_d3[n] : 1.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int c = t.c0 + cy + 8 * r;
      float x = 0.f;
      if (n < q.Cn && c < q.Ckr) {
        x = t.src[(size_t)c * q.D3 + n] * s0;  // same multiplication order as the single-tensor kernel
        if (q.scale_d2) x *= q.scale_d2[c];
        x *= sn;
      }
      v[r] = x;
    }
  } else {  // c = d3 is contiguous: 64 consecutive c of one n
    const int cx = tid & 63, ny = tid >> 6;
    const int c = t.c0 + cx;
    const float sc = (q.scale_d3 && c < q.Ckr) ? q.scale_d3[c] : 1.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int n = t.n0 + ny + 4 * r;
      float x = 0.f;
      if (n < q.Cn && c < q.Ckr) {
        x = t.src[(size_t)n * q.D3 + c] * s0;
        if (q.scale_d2) x *= q.scale_d2[n];
        x *= sc;
      }
      v[r] = x;
    }
  }
}

// Persistent blocks: tile t, t + gridDim.x, ...; the NEXT tile's loads are in flight while the current one goes through
// shared memory (round 1 ran one tile per block: ~6 000 short-lived blocks, 34 us for 75 MB of traffic).
__global__ void __launch_bounds__(256) pack_conv_weight_multi_kernel(const __grid_constant__ PackMultiParams p,
                                                                     int total_tiles) {
  __shared__ float tile[kPackTileN][kPackTileC + 1];
  const int tid = threadIdx.x;
  int t = blockIdx.x;
  if (t >= total_tiles) return;
  PackTile cur = pack_tile_decode(p, t), nxt = cur;
  float v[8], vn[8];
  pack_tile_load(cur, tid, v);
  for (; t < total_tiles; t += gridDim.x) {
    const int tn = t + gridDim.x;
    if (tn < total_tiles) {
      nxt = pack_tile_decode(p, tn);
      pack_tile_load(nxt, tid, vn);
    }
    const PackJobDev& q = *cur.q;
    __syncthreads();  // the previous tile's readers are done with the shared tile
    if (q.n_is_d3) {
      const int nx = tid & 31, cy = tid >> 5;
#pragma unroll
      for (int r = 0; r < 8; ++r) tile[nx][cy + 8 * r] = v[r];
    } else {
      const int cx = tid & 63, ny = tid >> 6;
#pragma unroll
      for (int r = 0; r < 8; ++r) tile[ny + 4 * r][cx] = v[r];
    }
    __syncthreads();
    // out[ph][n][tap * Ck + c]: 8 threads write the 64 columns of one row (16 bytes each)
    const int nl = tid >> 3, c8 = (tid & 7) * 8;
    const int n = cur.n0 + nl;
    if (n < q.Cn && cur.c0 + c8 < q.Ck) {
      const float* tr = &tile[nl][c8];
      const size_t off = ((size_t)cur.ph * q.Cn + n) * ((size_t)q.taps * q.Ck) + (size_t)cur.tap * q.Ck + cur.c0 + c8;
      if (q.f32) {  // rounded to nearest tf32 here: tcgen05 kind::tf32 truncates the low 13 mantissa bits of what it reads
        uint32_t t[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(t[k]) : "f"(tr[k]));
        uint4* dst = reinterpret_cast<uint4*>(static_cast<float*>(q.out) + off);
        dst[0] = make_uint4(t[0], t[1], t[2], t[3]);
        dst[1] = make_uint4(t[4], t[5], t[6], t[7]);
      } else {
        uint4 o;
        o.x = pack_bf16x2(tr[0], tr[1]);
        o.y = pack_bf16x2(tr[2], tr[3]);
        o.z = pack_bf16x2(tr[4], tr[5]);
        o.w = pack_bf16x2(tr[6], tr[7]);
        *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(q.out) + off) = o;
      }
    }
    cur = nxt;
#pragma unroll
    for (int r = 0; r < 8; ++r) v[r] = vn[r];
  }
}

}  // namespace b2g

using namespace b2g;

extern "C" int b2g_version(void) { return 100; }

extern "C" int b2g_set_sm_limit(int n) {
  if (n < 0) return fail(B2G_E_INVALID_ARG, "b2g_set_sm_limit: n must be >= 0 (0 = all SMs)");
  const int prev = b2g::g_sm_limit;
  b2g::g_sm_limit = (n > 0 && n < 8) ? 8 : n;
  (void)prev;
  return B2G_OK;
}
extern "C" int b2g_num_sms(void) { return b2g::num_sms(); }
extern "C" const char* b2g_last_error(void) { return error_buffer(); }

// phases / taps of a packing kind; rejects kernel sizes and flips the kind does not define
static int pack_geometry(int kind, int kh, int kw, int flip, int* phases_out, int* taps_out) {
  int phases, taps;
  if (kind == B2G_CONV_K3S1) {
    if (kh != 3 || kw != 3) return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight: K3S1 needs a 3x3 kernel");
    phases = 1; taps = 9;
  } else if (kind == B2G_CONV_K1S1) {
    if (kh != 1 || kw != 1 || flip) return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight: K1S1 needs 1x1, no flip");
    phases = 1; taps = 1;
  } else if (kind == B2G_CONV_K4S2) {
    if (kh != 4 || kw != 4 || flip) return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight: K4S2 needs 4x4, no flip");
    phases = 1; taps = 16;
  } else if (kind == B2G_CONVT_K4S2) {
    if (kh != 4 || kw != 4 || flip) return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight: CONVT needs 4x4, no flip");
    phases = 4; taps = 4;
  } else {
    return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight: unknown kind %d", kind);
  }
  *phases_out = phases;
  *taps_out = taps;
  return B2G_OK;
}

static int pack_conv_weight_any(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad,
                                const float* w, const float* scale_scalar, const float* scale_d2,
                                const float* scale_d3, void* wpacked, b2g_stream_t stream_, bool f32) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!w || !wpacked) return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight: null pointer");
  int phases, taps;
  int rc = pack_geometry(kind, kh, kw, flip, &phases, &taps);
  if (rc) return rc;
  const int ckr = n_is_d3 ? D2 : D3, cn = n_is_d3 ? D3 : D2;
  const int ck = ck_pad > ckr ? ck_pad : ckr;
  const long total = (long)phases * taps * cn * ck;
  const int threads = 256;
  long blocks = (total + threads - 1) / threads;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (f32)
    pack_conv_weight_kernel<float><<<(int)blocks, threads, 0, stream>>>(
        kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad, w, scale_scalar, scale_d2, scale_d3, static_cast<float*>(wpacked),
        phases, taps, total);
  else
    pack_conv_weight_kernel<__nv_bfloat16><<<(int)blocks, threads, 0, stream>>>(
        kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad, w, scale_scalar, scale_d2, scale_d3,
        static_cast<__nv_bfloat16*>(wpacked), phases, taps, total);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

extern "C" int b2g_pack_conv_weight(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad,
                                    const float* w, const float* scale_scalar, const float* scale_d2,
                                    const float* scale_d3, void* wpacked, b2g_stream_t stream) {
  return pack_conv_weight_any(kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad, w, scale_scalar, scale_d2, scale_d3, wpacked,
                              stream, false);
}
extern "C" int b2g_pack_conv_weight_f32(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad,
                                        const float* w, const float* scale_scalar, const float* scale_d2,
                                        const float* scale_d3, void* wpacked, b2g_stream_t stream) {
  return pack_conv_weight_any(kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad, w, scale_scalar, scale_d2, scale_d3, wpacked,
                              stream, true);
}

extern "C" int b2g_pack_conv_weight_multi(const b2g_pack_job* jobs, int n_jobs, b2g_stream_t stream_) {
  if (!jobs || n_jobs <= 0 || n_jobs > kMaxPackJobs)
    return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight_multi: 1..%d jobs", kMaxPackJobs);
  PackMultiParams p;
  memset(&p, 0, sizeof(p));
  int blocks = 0;
  for (int j = 0; j < n_jobs; ++j) {
    const b2g_pack_job& s = jobs[j];
    if (!s.w || !s.wpacked) return fail(B2G_E_INVALID_ARG, "b2g_pack_conv_weight_multi: null pointer in job %d", j);
    int phases, taps;
    int rc = pack_geometry(s.kind, s.kh, s.kw, s.flip, &phases, &taps);
    if (rc) return rc;
    const int ckr = s.n_is_d3 ? s.D2 : s.D3, cn = s.n_is_d3 ? s.D3 : s.D2;
    const int ck = s.ck_pad > ckr ? s.ck_pad : ckr;
    if (ck % 8) return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_pack_conv_weight_multi: job %d packs %d columns per tap (not a multiple of 8)", j, ck);
    PackJobDev& d = p.job[j];
    d.kind = s.kind; d.kh = s.kh; d.kw = s.kw; d.D2 = s.D2; d.D3 = s.D3;
    d.n_is_d3 = s.n_is_d3; d.flip = s.flip; d.ck_pad = s.ck_pad; d.taps = taps; d.phases = phases;
    d.Cn = cn; d.Ckr = ckr; d.Ck = ck;
    d.n_blocks = (cn + kPackTileN - 1) / kPackTileN;
    d.c_blocks = (ck + kPackTileC - 1) / kPackTileC;
    d.first_block = blocks;
    d.w = s.w; d.scale_scalar = s.scale_scalar; d.scale_d2 = s.scale_d2; d.scale_d3 = s.scale_d3;
    d.out = s.wpacked;
    d.f32 = s.f32;
    blocks += phases * taps * d.n_blocks * d.c_blocks;
  }
  p.n_jobs = n_jobs;
  // one tile per block: measured (tools/sn_probe.py) 34-40 us for the discriminator's 14 packings, against 42-57 us for
  // persistent blocks with a register-prefetched next tile (grid = 6 blocks per SM) — the tiles are short and the
  // block scheduler overlaps them better than one block's own prefetch does.  B2G_PACK_PERSISTENT=1 keeps the other form.
  static int persistent = []() { const char* e = getenv("B2G_PACK_PERSISTENT"); return e ? atoi(e) : 0; }();
  const int cap = persistent ? num_sms() * 6 : blocks;
  pack_conv_weight_multi_kernel<<<blocks < cap ? blocks : cap, 256, 0, static_cast<cudaStream_t>(stream_)>>>(p, blocks);
  B2G_LAUNCH_CHECK();
  return B2G_OK;
}

// ---- host-only introspection (no CUDA calls): lets the CPU test-suite check the tiling / tap / packing logic ----
extern "C" int b2g_conv_plan_query(int kind, int N, int Hin, int Win, int tile_pixels, int* out, int out_len) {
  ConvPlan p;
  memset(&p, 0, sizeof(p));
  if (out_len < 80) return fail(B2G_E_INVALID_ARG, "b2g_conv_plan_query: out needs 80 ints");
  if (!build_conv_plan(kind, N, Hin, Win, 64, 64, tile_pixels, &p))
    return fail(B2G_E_UNSUPPORTED_SHAPE, "b2g_conv_plan_query: cannot tile kind %d N=%d H=%d W=%d", kind, N, Hin, Win);
  const int hdr[16] = {p.phases, p.taps, p.Hg, p.Wg, p.bw, p.bh, p.bn, p.tiles_w, p.tiles_h, p.tiles_n, p.m_tiles,
                       p.Hout, p.Wout, p.os, p.n_amaps, 0};
  for (int i = 0; i < 16; ++i) out[i] = hdr[i];
  for (int i = 0; i < 16; ++i) {
    out[16 + i] = p.tap_map[i];
    out[32 + i] = p.tap_dy[i];
    out[48 + i] = p.tap_dx[i];
  }
  for (int i = 0; i < 4; ++i) {
    out[64 + i] = p.ph_y[i];
    out[68 + i] = p.ph_x[i];
  }
  return B2G_OK;
}

extern "C" long b2g_pack_source_index(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad,
                                      long i) {
  const int taps = kind == B2G_CONV_K3S1 ? 9 : (kind == B2G_CONV_K4S2 ? 16 : (kind == B2G_CONVT_K4S2 ? 4 : 1));
  int d2, d3;
  return pack_source_index(kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad, taps, i, &d2, &d3);
}
