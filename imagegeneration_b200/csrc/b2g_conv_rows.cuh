// b2g_conv_rows.cuh — schedule of the row-slab implicit-GEMM conv kernel (b2g_conv_rows.cu).
//
// For layers whose phase grid is >= 128 pixels wide, an M tile is 128 consecutive pixels of ONE image row.  The
// activation operand of tap (dy, dx) is then the input row y+dy shifted by dx pixels, so instead of one TMA box
// per tap the kernel loads each needed input row ONCE as a "slab" of 130 pixels x 64 channels (1 halo pixel each
// side, zero-filled outside the image) and addresses the dx taps by moving the UMMA smem descriptor by
// dx * 128 bytes.  A super-tile is R consecutive rows x P output phases (R*P accumulators in TMEM), so a slab is
// shared by up to 3 (dx) x R (rows) x P (phases) MMAs and a weight tile by R MMAs: L2->SM traffic per MMA drops
// 2.3-4x against the per-tap kernel, which is what bounds the Cout <= 128 layers.
//
// The schedule is a static program (identical for every super-tile of a phase group) built on the host:
//   slabs[i] : input map, row offset relative to the first row of the super-tile, channel chunk — loaded in order
//   steps[j] : one weight tile (k-block `bcol`, local phase) + R MMAs: row r uses slab slab0 + r at pixel shift
//              dxi and accumulator acc0 + r*P; after the step the oldest `rel` slabs are released (FIFO).
#pragma once
#include "b2g_conv.cuh"

namespace b2g {

constexpr int kRowsMaxSlabs = 24;
constexpr int kRowsMaxSteps = 64;
constexpr int kRowsSlabPixels = 130;

struct RowsStep {
  int16_t bcol;    // k-block (64 columns) of the packed weight matrix
  int8_t phase_l;  // local phase inside the phase group
  int8_t slab0;    // slab used by row r = 0
  int8_t dxi;      // pixel shift inside the slab: dx + 1
  int8_t acc0;     // accumulator of row r = 0
  int8_t rel;      // slabs released after this step
  int8_t wide;     // 1: merged step — the SAME shifted slab feeds local phases phase_l and phase_l + 1 with ONE MMA of
                   // N = 2 * BLOCK_N (weight tiles bcol | bcol2 stacked in one stage, accumulators acc0 and acc0 + 1)
  int16_t bcol2;   // wide steps: k-block of the second phase's weight tile
};

struct RowsSched {
  int n_slabs, n_steps;
  int8_t slab_map[kRowsMaxSlabs], slab_dy[kRowsMaxSlabs], slab_chunk[kRowsMaxSlabs];
  RowsStep steps[kRowsMaxSteps];
};

struct RowsGeom {
  int R, P, PG, n_acc;  // rows per super-tile, phases per phase group, phase groups, accumulators
  int row_blocks;       // ceil(Hg / R)
};

// One attempt with a fixed number of rows per super-tile.
// merge (transposed conv, P = 2): the two sub-pixel phases of a phase group (px = 0, 1) both use the dx = 0 shift of every
// input row — with N = 64 tiles each MMA group reads 16 KB of A and 8 KB of B from shared memory for 128 tensor cycles
// of math (192 B/clk against the 128 B/clk the operand path delivers, measured in round 1 and 2).  One N = 128 MMA for
// both phases reads the A operand once: 80 instead of 96 KB per (row, dy, channel chunk).  The wide step is issued FIRST
// of its dy so that at the first step of a super-tile both accumulators are still unwritten (one accumulate flag).
constexpr bool try_rows_sched(const ConvPlan& pl, int ring, int R, int P, RowsGeom* g, RowsSched* sched /*[2]*/,
                              bool merge = false) {
  const int chunks = pl.Cin / 64;
  g->R = R;
  g->P = P;
  g->PG = pl.phases / P;
  g->n_acc = R * P;
  g->row_blocks = (pl.Hg + R - 1) / R;
  if (g->PG > 2) return false;
  for (int pg = 0; pg < g->PG; ++pg) {
    RowsSched& s = sched[pg];
    s.n_slabs = 0;
    s.n_steps = 0;
    int last_use[kRowsMaxSlabs] = {};
    for (int c = 0; c < chunks; ++c) {
      for (int m = 0; m < pl.n_amaps; ++m) {
        int dmin = 127, dmax = -128;
        for (int pl_ = 0; pl_ < P; ++pl_)
          for (int t = 0; t < pl.taps; ++t) {
            const int e = (pg * P + pl_) * pl.taps + t;
            if (pl.tap_map[e] != m) continue;
            if (pl.tap_dx[e] < -1 || pl.tap_dx[e] > 1) return false;
            if (pl.tap_dy[e] < dmin) dmin = pl.tap_dy[e];
            if (pl.tap_dy[e] > dmax) dmax = pl.tap_dy[e];
          }
        if (dmin > dmax) continue;
        const int base = s.n_slabs;
        const int cnt = dmax - dmin + R;
        if (base + cnt > kRowsMaxSlabs) return false;
        for (int i = 0; i < cnt; ++i) {
          s.slab_map[base + i] = (int8_t)m;
          s.slab_dy[base + i] = (int8_t)(dmin + i);
          s.slab_chunk[base + i] = (int8_t)c;
          last_use[base + i] = -1;
        }
        s.n_slabs += cnt;
        for (int dy = dmin; dy <= dmax; ++dy) {
          // merged step: taps t0 of local phase 0 and t1 of local phase 1 with this dy and the same dx
          int mt0 = -1, mt1 = -1;
          if (merge && P == 2)
            for (int t0 = 0; t0 < pl.taps && mt0 < 0; ++t0)
              for (int t1 = 0; t1 < pl.taps; ++t1) {
                const int e0 = (pg * P) * pl.taps + t0, e1 = (pg * P + 1) * pl.taps + t1;
                if (pl.tap_map[e0] == m && pl.tap_map[e1] == m && pl.tap_dy[e0] == dy && pl.tap_dy[e1] == dy &&
                    pl.tap_dx[e0] == pl.tap_dx[e1]) {
                  mt0 = t0;
                  mt1 = t1;
                  break;
                }
              }
          if (mt0 >= 0) {
            if (s.n_steps >= kRowsMaxSteps) return false;
            RowsStep& st = s.steps[s.n_steps];
            st.bcol = (int16_t)(mt0 * chunks + c);
            st.bcol2 = (int16_t)(mt1 * chunks + c);
            st.phase_l = 0;
            st.slab0 = (int8_t)(base + dy - dmin);
            st.dxi = (int8_t)(pl.tap_dx[(pg * P) * pl.taps + mt0] + 1);
            st.acc0 = 0;
            st.rel = 0;
            st.wide = 1;
            for (int r = 0; r < R; ++r) last_use[st.slab0 + r] = s.n_steps;
            ++s.n_steps;
          }
          for (int pl_ = 0; pl_ < P; ++pl_)
            for (int t = 0; t < pl.taps; ++t) {
              const int e = (pg * P + pl_) * pl.taps + t;
              if (pl.tap_map[e] != m || pl.tap_dy[e] != dy) continue;
              if (mt0 >= 0 && ((pl_ == 0 && t == mt0) || (pl_ == 1 && t == mt1))) continue;  // covered by the wide step
              if (s.n_steps >= kRowsMaxSteps) return false;
              RowsStep& st = s.steps[s.n_steps];
              st.bcol = (int16_t)(t * chunks + c);
              st.bcol2 = 0;
              st.phase_l = (int8_t)pl_;
              st.slab0 = (int8_t)(base + dy - dmin);
              st.dxi = (int8_t)(pl.tap_dx[e] + 1);
              st.acc0 = (int8_t)pl_;
              st.rel = 0;
              st.wide = 0;
              for (int r = 0; r < R; ++r) last_use[st.slab0 + r] = s.n_steps;
              ++s.n_steps;
            }
        }
      }
    }
    // FIFO release counts; the live span (needed slab - released slabs) must fit the ring
    int released = 0;
    for (int i = 0; i + 1 < s.n_slabs; ++i)
      if (last_use[i] > last_use[i + 1]) return false;
    for (int j = 0; j < s.n_steps; ++j) {
      const int need = s.steps[j].slab0 + R - 1;
      if (need + 1 - released > ring) return false;
      int rel = 0;
      while (released + rel < s.n_slabs && last_use[released + rel] == j) ++rel;
      s.steps[j].rel = (int8_t)rel;
      released += rel;
    }
    if (released != s.n_slabs) return false;
  }
  return true;
}

// Builds the geometry and one schedule per phase group: the largest R in {4, 2, 1} whose accumulators fit 256 TMEM
// columns (x2 buffers) and whose schedule fits the tables and the slab ring.  `ring` = slab stages the kernel owns.
// Returns false when the role does not fit (the caller falls back to the per-tap kernel).
constexpr bool build_rows_sched(const ConvPlan& pl, int block_n, int ring, RowsGeom* g, RowsSched* sched /*[2]*/,
                                bool merge = false) {
  if (pl.kind == B2G_CONV_K1S1) return false;
  if (pl.bw != 128 || pl.bh != 1 || pl.bn != 1) return false;
  if (pl.Cin % 64 || block_n > 256) return false;
  int P = pl.kind == B2G_CONVT_K4S2 ? 2 : 1;
  if (P * block_n > 256) P = 1;
  int R = 256 / (P * block_n);
  if (R > 4) R = 4;
  while (R > 1 && R > pl.Hg) R /= 2;
  for (; R >= 1; R /= 2)
    if (try_rows_sched(pl, ring, R, P, g, sched, merge && 2 * block_n <= 256)) return true;
  return false;
}

// ---------------------------------------------------------------------------------------------------------------
// Compile-time schedules for the model's own roles.  The generic kernel walks the schedule as a table, which costs
// the single issuing thread ~130 instructions per step (ring arithmetic, table loads, R2UR moves) — more than the
// 2 x 256 tensor-core cycles a step is worth.  For a fixed (kind, Cin/64, R, P) and ring sizes that DIVIDE the number
// of slabs / weight tiles per super-tile, every smem address, barrier and accumulator of every step is a constant:
// the MMA warp's loop is fully unrolled into waits, UTCHMMA groups and commits with immediate operands.
// ---------------------------------------------------------------------------------------------------------------
struct RowsSchedPair {
  RowsGeom g;
  RowsSched s[2];
  bool ok;
};

struct RowsDyn {  // generic kernel: table-driven issue loop, default ring sizes
  static constexpr bool kStatic = false;
  static constexpr bool kMerge = false;
  static constexpr int kS = 0, kSB = 0;
};

template <int KIND, int CH, int R_, int P_, int BN, int S_, int SB_, bool MERGE_ = false>
struct RowsStatic {
  static constexpr bool kStatic = true;
  static constexpr bool kMerge = MERGE_;
  static constexpr int kKind = KIND, kCh = CH, kR = R_, kP = P_, kS = S_, kSB = SB_;
  static constexpr RowsSchedPair make() {
    ConvPlan pl{};
    // any geometry whose phase grid is 128 pixels wide yields the same schedule
    const bool ok_plan = build_conv_plan(KIND, 1, KIND == B2G_CONV_K4S2 ? 16 : 8, KIND == B2G_CONV_K4S2 ? 256 : 128,
                                         CH * 64, BN, 128, &pl);
    RowsSchedPair sp{};
    sp.ok = ok_plan && try_rows_sched(pl, S_, R_, P_, &sp.g, sp.s, MERGE_);
    return sp;
  }
  static constexpr bool ok() { return make().ok; }
  static constexpr int n_steps() { return make().s[0].n_steps; }
  static constexpr int n_slabs() { return make().s[0].n_slabs; }
  static constexpr RowsStep step(int j) { return make().s[0].steps[j]; }
  static constexpr int need_before(int J) {  // highest slab index any step j < J has waited for
    const RowsSchedPair sp = make();
    int m = -1;
    for (int j = 0; j < J; ++j)
      if (sp.s[0].steps[j].slab0 + R_ - 1 > m) m = sp.s[0].steps[j].slab0 + R_ - 1;
    return m;
  }
  static constexpr int released_before(int J) {
    const RowsSchedPair sp = make();
    int n = 0;
    for (int j = 0; j < J; ++j) n += sp.s[0].steps[j].rel;
    return n;
  }
  static constexpr bool written_before(int J, int acc) {
    const RowsSchedPair sp = make();
    for (int j = 0; j < J; ++j)
      for (int r = 0; r < R_; ++r) {
        if (sp.s[0].steps[j].acc0 + r * P_ == acc) return true;
        if (sp.s[0].steps[j].wide && sp.s[0].steps[j].acc0 + 1 + r * P_ == acc) return true;
      }
    return false;
  }
  static_assert(S_ >= 1 && SB_ >= 1, "ring sizes");
};

// true when the host-built schedule (both phase groups) equals the compile-time one on the MMA side
template <class SS>
inline bool rows_static_matches(const RowsGeom& g, const RowsSched* sched) {
  constexpr RowsSchedPair sp = SS::make();
  if (!sp.ok || g.R != SS::kR || g.P != SS::kP) return false;
  for (int pg = 0; pg < g.PG; ++pg) {
    if (sched[pg].n_slabs != sp.s[0].n_slabs || sched[pg].n_steps != sp.s[0].n_steps) return false;
    for (int j = 0; j < sp.s[0].n_steps; ++j) {
      const RowsStep &a = sched[pg].steps[j], &b = sp.s[0].steps[j];
      if (a.slab0 != b.slab0 || a.dxi != b.dxi || a.acc0 != b.acc0 || a.rel != b.rel || a.wide != b.wide) return false;
    }
  }
  return true;
}

// Launches the row-slab kernel when the role fits it; *handled tells the caller whether it did.
// static_only: only the compile-time-scheduled variants (tried before the CTA-pair kernel; the generic table-driven
// variant is tried after it).
int launch_conv_rows(const b2g_conv_args* a, const ConvPlan& pl, int block_n, void* stream, bool* handled,
                     bool static_only);

}  // namespace b2g
