// b2g_conv.cuh — tile plan shared by the implicit-GEMM conv kernels (fprop / dgrad / convT) and wgrad.
//
// Every conv role is one GEMM  D[M, Cout] = A[M, taps*Cin] * Wp[Cout, taps*Cin]^T  per output phase, where the
// M index runs over a "phase grid" (N, Hg, Wg) of pixels and the A tile of tap t is the input tensor read at
// grid position + (dy_t, dx_t) through one of up to 4 TMA tensor maps (zero fill outside = SAME padding).
//   K3S1   : 1 phase, 9 taps, grid = input = output.                       in(y+ky-1, x+kx-1)
//   K4S2   : 1 phase, 16 taps, grid = output; the input is seen as 4 parity sub-grids (one tensor map each):
//            in(2y+ky-1, 2x+kx-1) = sub[(ky+1)&1][(kx+1)&1](y + floor((ky-1)/2), x + floor((kx-1)/2))
//   K1S1   : 1 phase, 1 tap, no shift (plain GEMM over pixels; edge layers after im2col27)
//   CONVT  : 4 phases (py,px) x 4 taps (jy,jx), grid = input, output pixel (2y+py, 2x+px):
//            py=0: ky = 1 (dy 0), 3 (dy -1);  py=1: ky = 0 (dy +1), 2 (dy 0)   [Y = 2*iy + ky - 1]
#pragma once
#include <cuda.h>
#include <stdint.h>

#include "../../include/b200gan.h"

namespace b2g {

constexpr int kMaxTapEntries = 16;  // phases * taps <= 16 for all kinds

struct ConvPlan {
  int kind;
  int phases, taps;        // taps per phase
  int N, Hg, Wg;           // phase grid (GEMM M space)
  int bw, bh, bn;          // pixel tile box, bw*bh*bn == tile_pixels
  int tiles_w, tiles_h, tiles_n, m_tiles;
  int Cin, Cout;           // GEMM K per tap, GEMM N
  int Hout, Wout, os;      // output tensor dims and output stride (1 or 2)
  int8_t tap_map[kMaxTapEntries];  // which A tensor map
  int8_t tap_dy[kMaxTapEntries];
  int8_t tap_dx[kMaxTapEntries];
  int8_t ph_y[4], ph_x[4];         // output phase offsets
  int n_amaps;                     // 1 or 4
};

// ky (0..3) of a 4x4 stride-2 conv -> (parity sub-grid row, row offset inside the sub-grid)
__host__ __device__ constexpr void k4s2_tap(int k, int* parity, int* off) {
  // input coordinate 2*o + k - 1:  k=0 -> 2(o-1)+1 ; k=1 -> 2o ; k=2 -> 2o+1 ; k=3 -> 2(o+1)
  *parity = (k + 1) & 1;
  *off = (k == 0) ? -1 : (k == 3 ? 1 : 0);
}
// convT phase p (0/1), tap j (0/1) -> (kernel index k, input offset d) with Y = 2*iy + k - 1, Y = 2*y + p
__host__ __device__ constexpr void convt_tap(int p, int j, int* k, int* d) {
  if (p == 0) { *k = j == 0 ? 1 : 3; *d = j == 0 ? 0 : -1; }
  else        { *k = j == 0 ? 0 : 2; *d = j == 0 ? 1 : 0; }
}

// Source element of packed weight element i (layout [phase][n][tap*Ck + c], Ck zero-padded to ck_pad) inside the
// fp32 Keras kernel w[kh][kw][D2][D3]; returns -1 for a padding element.  *d2 / *d3 receive the channel indices
// (for the per-channel scales).  Shared by the pack kernel and the host-side plan tests.
__host__ __device__ inline long pack_source_index(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip,
                                                  int ck_pad, int taps, long i, int* d2_out, int* d3_out) {
  const int Cn = n_is_d3 ? D3 : D2;
  const int Ckr = n_is_d3 ? D2 : D3;
  const int Ck = ck_pad > Ckr ? ck_pad : Ckr;
  const int c = (int)(i % Ck);
  long r = i / Ck;
  if (c >= Ckr) return -1;
  const int tap = (int)(r % taps);
  r /= taps;
  const int n = (int)(r % Cn);
  const int ph = (int)(r / Cn);
  int ky, kx;
  if (kind == B2G_CONVT_K4S2) {
    int d;
    convt_tap(ph >> 1, tap >> 1, &ky, &d);
    convt_tap(ph & 1, tap & 1, &kx, &d);
  } else {
    ky = tap / kw;
    kx = tap % kw;
    if (flip) {
      ky = kh - 1 - ky;
      kx = kw - 1 - kx;
    }
  }
  const int d2 = n_is_d3 ? c : n;
  const int d3 = n_is_d3 ? n : c;
  *d2_out = d2;
  *d3_out = d3;
  return (((long)ky * kw + kx) * D2 + d2) * D3 + d3;
}

constexpr bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }

// Chooses the pixel tile (bw, bh, bn) with bw*bh*bn == tile_pixels that wastes the fewest rows.
// Returns false when the grid width cannot be tiled (Wg must divide tile_pixels or be a multiple of it).
constexpr bool choose_tile(int tile_pixels, int N, int Hg, int Wg, int* bw, int* bh, int* bn) {
  int w = 0;
  if (Wg >= tile_pixels) {
    if (Wg % tile_pixels) return false;
    w = tile_pixels;
  } else {
    if (!is_pow2(Wg)) return false;
    w = Wg;
  }
  int rest = tile_pixels / w;
  long best_cost = -1;
  int best_h = 1;
  for (int h = 1; h <= rest; h *= 2) {
    int n = rest / h;
    long cost = (long)((Hg + h - 1) / h) * h * (long)((N + n - 1) / n) * n;
    if (best_cost < 0 || cost < best_cost || (cost == best_cost && h > best_h)) {
      best_cost = cost;
      best_h = h;
    }
  }
  *bw = w;
  *bh = best_h;
  *bn = rest / best_h;
  return true;
}

// Fills the geometry + tap tables for a conv role. Returns false on unsupported shapes.
constexpr bool build_conv_plan(int kind, int N, int Hin, int Win, int Cin, int Cout, int tile_pixels, ConvPlan* p) {
  p->kind = kind;
  p->N = N;
  p->Cin = Cin;
  p->Cout = Cout;
  for (int i = 0; i < 4; ++i) p->ph_y[i] = p->ph_x[i] = 0;
  if (kind == B2G_CONV_K3S1) {
    p->phases = 1; p->taps = 9; p->n_amaps = 1;
    p->Hg = Hin; p->Wg = Win; p->Hout = Hin; p->Wout = Win; p->os = 1;
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx) {
        int t = ky * 3 + kx;
        p->tap_map[t] = 0; p->tap_dy[t] = (int8_t)(ky - 1); p->tap_dx[t] = (int8_t)(kx - 1);
      }
  } else if (kind == B2G_CONV_K1S1 || kind == B2G_CONV_K3S1_C3) {
    p->phases = 1; p->taps = 1; p->n_amaps = 1;
    p->Hg = Hin; p->Wg = Win; p->Hout = Hin; p->Wout = Win; p->os = 1;
    p->tap_map[0] = 0; p->tap_dy[0] = 0; p->tap_dx[0] = 0;
  } else if (kind == B2G_CONV_K4S2) {
    if ((Hin & 1) || (Win & 1)) return false;
    p->phases = 1; p->taps = 16; p->n_amaps = 4;
    p->Hg = Hin / 2; p->Wg = Win / 2; p->Hout = Hin / 2; p->Wout = Win / 2; p->os = 1;
    for (int ky = 0; ky < 4; ++ky)
      for (int kx = 0; kx < 4; ++kx) {
        int py = 0, dy = 0, px = 0, dx = 0;
        k4s2_tap(ky, &py, &dy);
        k4s2_tap(kx, &px, &dx);
        int t = ky * 4 + kx;
        p->tap_map[t] = (int8_t)(py * 2 + px); p->tap_dy[t] = (int8_t)dy; p->tap_dx[t] = (int8_t)dx;
      }
  } else if (kind == B2G_CONVT_K4S2) {
    p->phases = 4; p->taps = 4; p->n_amaps = 1;
    p->Hg = Hin; p->Wg = Win; p->Hout = Hin * 2; p->Wout = Win * 2; p->os = 2;
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        int ph = py * 2 + px;
        p->ph_y[ph] = (int8_t)py; p->ph_x[ph] = (int8_t)px;
        for (int jy = 0; jy < 2; ++jy)
          for (int jx = 0; jx < 2; ++jx) {
            int ky = 0, dy = 0, kx = 0, dx = 0;
            convt_tap(py, jy, &ky, &dy);
            convt_tap(px, jx, &kx, &dx);
            int e = ph * 4 + jy * 2 + jx;
            p->tap_map[e] = 0; p->tap_dy[e] = (int8_t)dy; p->tap_dx[e] = (int8_t)dx;
          }
      }
  } else {
    return false;
  }
  if (!choose_tile(tile_pixels, N, p->Hg, p->Wg, &p->bw, &p->bh, &p->bn)) return false;
  p->tiles_w = p->Wg / p->bw;
  p->tiles_h = (p->Hg + p->bh - 1) / p->bh;
  p->tiles_n = (N + p->bn - 1) / p->bn;
  p->m_tiles = p->tiles_w * p->tiles_h * p->tiles_n;
  return true;
}

// Division by a launch constant as one multiply-high: q = umulhi(x, ceil(2^32 / d)) is exact while x * d < 2^32 (the
// host checks the largest x it will be used on; otherwise `m` = 0 and the device divides).  Tile decoding runs per item in
// every epilogue warp: five runtime divisions cost ~125 instructions, 12 % of an epilogue-bound kernel (ncu r2).
struct FastDiv {
  uint32_t d, m;  // m == 0: plain division (d == 1, or the exactness bound does not hold)
};
inline FastDiv make_fastdiv(int d, long x_max) {
  FastDiv f;
  f.d = (uint32_t)(d > 0 ? d : 1);
  f.m = (d > 1 && (long)x_max * d < (1L << 32)) ? (uint32_t)(((1ULL << 32) + d - 1) / d) : 0u;
  return f;
}
#ifdef __CUDACC__
__device__ __forceinline__ void fast_divmod(int x, FastDiv f, int& q, int& r) {
  q = f.m ? (int)__umulhi((uint32_t)x, f.m) : (f.d == 1u ? x : x / (int)f.d);
  r = x - q * (int)f.d;
}
#endif

// Tile index -> (phase, N tile, pixel box) for the per-tap kernels; device code only.
struct TileCoord {
  int phase, n_tile, tn, th, tw;
};
__device__ __forceinline__ TileCoord decode_tile(const ConvPlan& pl, int n_tiles, int tile) {
  TileCoord c;
  c.n_tile = tile % n_tiles;
  int rest = tile / n_tiles;
  int m = rest % pl.m_tiles;
  c.phase = rest / pl.m_tiles;
  c.tw = m % pl.tiles_w;
  int r2 = m / pl.tiles_w;
  c.th = r2 % pl.tiles_h;
  c.tn = r2 / pl.tiles_h;
  return c;
}

// Division-free walk over the pixel tiles of a persistent CTA (tile = blockIdx.x, += gridDim.x): the mixed-radix
// coordinate (tw, th, tn) is decoded once and then advanced by the digits of gridDim.x with carries.  decode_tile's three
// runtime divisions per tile and per warp were 10 % of the executed instructions of the issue-bound 3-channel conv
// (ncu r2 source page).
struct TileWalk {
  int tw, th, tn;     // current tile
  int dw, dh, dn;     // digits of the stride
};
__device__ __forceinline__ TileWalk tile_walk_begin(const ConvPlan& pl, int first, int stride) {
  TileWalk w;
  w.tw = first % pl.tiles_w;
  int r = first / pl.tiles_w;
  w.th = r % pl.tiles_h;
  w.tn = r / pl.tiles_h;
  w.dw = stride % pl.tiles_w;
  r = stride / pl.tiles_w;
  w.dh = r % pl.tiles_h;
  w.dn = r / pl.tiles_h;
  return w;
}
__device__ __forceinline__ void tile_walk_next(const ConvPlan& pl, TileWalk& w) {
  w.tw += w.dw;
  int c = 0;
  if (w.tw >= pl.tiles_w) { w.tw -= pl.tiles_w; c = 1; }
  w.th += w.dh + c;
  c = 0;
  if (w.th >= pl.tiles_h) { w.th -= pl.tiles_h; c = 1; }
  w.tn += w.dn + c;
}

// CTA-pair (cta_group::2) kernel (b2g_conv_pair.cu); *handled tells the caller whether the role was taken
int launch_conv_pair(const b2g_conv_args* a, const ConvPlan& pl, int block_n, void* stream, bool* handled);
// weight gradient with the shifted operand read once per pixel row (b2g_wgrad_slab.cu)
int launch_wgrad_slab(const b2g_wgrad_args* a, void* stream, bool* handled);
// weight + bias gradient of the 3-channel edge layers, im2col operand built in shared memory (b2g_wgrad_c3.cu)
int launch_wgrad_c3(const b2g_wgrad_args* a, void* stream);
// 3-channel 3x3 conv with the im2col patch built in shared memory (b2g_conv_c3.cu)
int launch_conv_c3(const b2g_conv_args* a, void* stream);
// 64 -> 3 channel 3x3 conv as one 64 -> 27 GEMM per pixel + a gather over a ring of rows (b2g_conv_o3.cu)
int launch_conv_o3(const b2g_conv_args* a, void* stream, bool* handled);

}  // namespace b2g
