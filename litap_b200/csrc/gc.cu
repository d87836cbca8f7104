// gc.cu — slope_gc on the device (R/flow_05_slope_gc.R:73-164): gradients
// towards east / north (sgre, sgcn), their magnitudes (sgr, sgc, 0 -> 0.00001), row / column
// curvatures (scr, scc) and the hillslope run labels of every row and every column
// (hill_r_dir/n/cell, hill_c_dir/n/cell).
//
// What the reference does, and how it maps:
//   * :103-135  zero gradients take 0.00001 x the sign of the previous cell (west for sgre, SOUTH for
//     sgcn), found by synchronous rounds over the whole grid with a "no progress -> +0.00001" exit.
//     Its fixed point is a sweep along the line: a run of zeros takes the sign of the non-zero value
//     before it; a run whose predecessor is NA is never resolved (-> +0.00001) unless it is a single
//     cell, which then looks at the cell after it (fix_step below; checked against the literal
//     Jacobi rounds of the oracle).
//   * :145-160  cumsum / factor / 1:n() per row and per column are a second sweep with a handful of
//     counters (label_step): NA gradients poison the rest of the line exactly like cumsum(NA).
// All lines run in parallel:
//   k_gc_rows   a warp owns a row and walks along it 32 columns at a time; the sweeps become warp scans
//               (ballot / popc / clz / shuffle) with the carried state in warp-uniform registers
//   k_gc_cols_stencil / _fix / _sum / _label  a plain stencil, then one thread per (column, row segment):
//               south -> north (zero signs), north -> south (hill_c_dir / n / cell), see "columns" below
// The NA ring of add_buffer() is required (like everywhere on this path): it is what makes the
// reference's linear seqno +- 1 addressing row-local.
#include "cellmath.h"
#include "litap_internal.h"

namespace {

constexpr double GC_TINY = 0.00001;
__device__ __forceinline__ double gc_na() { return __longlong_as_double(0x7ff8000000000000LL); }

struct GcPtrs {
  double *sgre, *sgr, *sgcn, *sgc, *scr, *scc;
  uint8_t *rdir, *cdir;
  int32_t *rn, *rcell, *cn, *ccell;
};

// ---- the two sweeps ------------------------------------------------------------------------------
struct FixState {
  double prev;    // fixed value of the previous cell of the line (NA before the first)
  double runval;  // what the cells of the current zero run take
  bool in_run;
};
__device__ __forceinline__ void fix_init(FixState& s) { s.prev = gc_na(); s.runval = 0; s.in_run = false; }
// g: this cell's gradient as computed; nxt: the next cell's (NA past the end).  Returns the fixed value.
// Written with selects only: the lanes of a warp are in different states all the time, branches would
// serialise them and sit on the critical path of the sweep.
__device__ __forceinline__ double fix_step(FixState& s, double g, double nxt) {
  const bool z = g == 0.0;
  const double rv_prev = s.prev > 0 ? GC_TINY : -GC_TINY;                        // 0.00001 * sign(previous)
  const double rv_next = nxt != nxt ? gc_na() : (nxt > 0 ? GC_TINY : -GC_TINY);  // single cell: 0.00001 * sign(next)
  const double rv_start = s.prev == s.prev ? rv_prev : (nxt == 0.0 ? GC_TINY : rv_next);  // never resolved: positive
  s.runval = (z && !s.in_run) ? rv_start : s.runval;
  const double out = z ? s.runval : g;
  s.in_run = z;
  s.prev = out;
  return out;
}

struct LabelState {
  int cum, dense, last_val, cnt, na_cnt;
  bool lag_na, lag_pos, poisoned, have_last;
};
__device__ __forceinline__ void label_init(LabelState& s) {
  s.cum = s.dense = s.last_val = s.cnt = s.na_cnt = 0;
  s.lag_na = true; s.lag_pos = false; s.poisoned = false; s.have_last = false;
}
// v: fixed gradient of the cell (never 0; NaN = NA), elev_na: the cell has no elevation (selects only, see fix_step)
__device__ __forceinline__ void label_step(LabelState& s, double v, bool elev_na, int& n_out, int& cell_out) {
  const bool s_na = v != v, pos = v > 0;
  s.poisoned = s.poisoned || (!s.lag_na && s_na);                // (NA | FALSE) = NA: cumsum is NA from here on
  const int inc = (s.lag_na || pos != s.lag_pos) ? 1 : 0;
  s.cum += s.poisoned ? 0 : inc;
  const bool valid = !elev_na && !s.poisoned;
  const bool newlab = valid && (!s.have_last || s.cum != s.last_val);
  s.dense += newlab ? 1 : 0;
  s.last_val = newlab ? s.cum : s.last_val;
  s.have_last = s.have_last || newlab;
  s.cnt = (newlab ? 0 : s.cnt) + (valid ? 1 : 0);
  s.na_cnt += valid ? 0 : 1;                                     // NA is one more group of the line
  n_out = valid ? s.dense : 0;
  cell_out = valid ? s.cnt : (elev_na ? 0 : s.na_cnt);
  s.lag_na = s_na;
  s.lag_pos = pos;
}

// ---- rows ------------------------------------------------------------------------------------------
// One warp per row, 32 columns per step, lane = column.  Both sweeps are scans whose operators are
// cheap on a warp: every state variable of fix_step / label_step is a prefix count, a prefix OR or
// "the value at the last marked lane", i.e. ballot + popc / clz + one shuffle, with the state at the
// end of the chunk carried in warp-uniform registers.  (fix_step / label_step stay the definition
// and run as they are in the column kernels; test_gpu_slope_gc checks both against the oracle.)
constexpr int GR_WARPS = 4;     // rows per block

// elevation of linear neighbour seqno - 1 / seqno + 1 exactly as nb_values addresses it (no wrap check)
__device__ __forceinline__ double gc_elev_lin(const double* __restrict__ elev, const GridDims& d, int r, int c) {
  if (c < 0) { c = d.NC - 1; --r; }
  else if (c >= d.NC) { c = 0; ++r; }
  if (r < 0 || r >= d.NR) return gc_na();
  return __ldg(elev + (size_t)r * d.pitch + c);
}
// (sgre, scr) of a cell from its elevation and its west / east neighbours (R/flow_05_slope_gc.R:80-90)
__device__ __forceinline__ void gc_row_cell(double e5, double n4, double n6, double two_g, double g2, double& ge, double& cr) {
  if (n4 != n4) n4 = e5;
  if (n6 != n6) n6 = e5;
  ge = (n4 - n6) / two_g;
  cr = (2 * e5 - n4 - n6) / g2;
}
__device__ __forceinline__ int gc_top(unsigned m) { return 31 - __clz(m); }   // highest set bit, m != 0

// class of a gradient value as three warp masks (bit = lane): NA, zero, positive
struct GcMasks { unsigned na, z, pos; };

__global__ void __launch_bounds__(GR_WARPS * 32) k_gc_rows(const double* __restrict__ elev, GridDims d, double grid_m, GcPtrs P) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * GR_WARPS + (threadIdx.x >> 5);
  if (r >= d.NR) return;
  const double two_g = 2 * grid_m, g2 = grid_m * grid_m;
  const double* __restrict__ row = elev + (size_t)r * d.pitch;
  const unsigned le = (2u << lane) - 1u;   // lanes <= mine
  const unsigned lt = le >> 1;             // lanes <  mine
  auto ld = [&](int cb) { const int c = cb + lane; return c < d.NC ? __ldg(row + c) : gc_na(); };
  const double e_west = gc_elev_lin(elev, d, r, -1), e_east = gc_elev_lin(elev, d, r, d.NC);
  // gradient / curvature of the chunk at cb from its elevations, the last one before it and the chunk after it;
  // lanes past the last column come out NA (that is what the cell after the last one reads as)
  auto grad = [&](int cb, double Ec, double e_before, double En, double& ge, double& cr, GcMasks& m) {
    const int c = cb + lane;
    double n4 = __shfl_up_sync(FULL, Ec, 1), n6 = __shfl_down_sync(FULL, Ec, 1);
    const double en0 = __shfl_sync(FULL, En, 0);
    if (lane == 0) n4 = e_before;
    if (lane == 31) n6 = en0;
    if (c == d.NC - 1) n6 = e_east;
    gc_row_cell(Ec, n4, n6, two_g, g2, ge, cr);
    if (c >= d.NC) ge = gc_na();
    m.na = __ballot_sync(FULL, ge != ge);
    m.z = __ballot_sync(FULL, ge == 0.0);
    m.pos = __ballot_sync(FULL, ge > 0);
  };
  // carried state, identical in all lanes (bits 0/1)
  unsigned c_prev_na = 1, c_prev_pos = 0, c_inrun = 0;      // class of the raw gradient before the chunk
  unsigned c_run_na = 0, c_run_pos = 0;                     // value of the zero run that is open (while c_inrun)
  unsigned c_lag_na = 1, c_lag_pos = 0;
  bool c_poisoned = false, c_have_last = false;
  int c_cum = 0, c_dense = 0, c_last_val = 0, c_cnt = 0, c_na = 0;
  double E0 = ld(0), E1 = ld(32), E2 = ld(64), E3 = ld(96);
  double ge, cr;
  GcMasks m;
  grad(0, E0, e_west, E1, ge, cr, m);
  for (int cb = 0; cb < d.NC; cb += 32) {
    const double E4 = ld(cb + 128);   // four chunks ahead of the arithmetic
    const int c = cb + lane;
    const bool active = c < d.NC;
    const unsigned colmask = d.NC - cb >= 32 ? FULL : (1u << (d.NC - cb)) - 1u;
    // the next chunk's gradients: the sign rule looks one cell ahead
    double ge_n, cr_n;
    GcMasks mn;
    grad(cb + 32, E1, __shfl_sync(FULL, E0, 31), E2, ge_n, cr_n, mn);
    const size_t o = (size_t)r * d.pitch + c;
    if (active) {
      double ar = fabs(ge);
      if (ar == 0) ar = GC_TINY;
      P.sgr[o] = ar;
      P.scr[o] = cr;
    }
    // ---- zero signs (fix_step as mask arithmetic) ----
    const unsigned p_na = (m.na << 1) | c_prev_na, p_pos = (m.pos << 1) | c_prev_pos, p_z = (m.z << 1) | c_inrun;
    const unsigned n_na = (m.na >> 1) | (mn.na << 31), n_pos = (m.pos >> 1) | (mn.pos << 31), n_z = (m.z >> 1) | (mn.z << 31);
    const unsigned rs = m.z & ~p_z;                                             // run starts
    const unsigned rv_na = p_na & ~n_z & n_na;                                  // value a run starting here takes: NA ...
    const unsigned rv_pos = (~p_na & p_pos) | (p_na & (n_z | (~n_na & n_pos)));  // ... or +0.00001 (else -0.00001)
    const bool z = (m.z >> lane) & 1u;
    const unsigned sb = rs & le;
    const int s_at = sb ? gc_top(sb) : 0;
    const bool run_na = sb ? (rv_na >> s_at) & 1u : c_run_na != 0;
    const bool run_pos = sb ? (rv_pos >> s_at) & 1u : c_run_pos != 0;
    const double v = z ? (run_na ? gc_na() : (run_pos ? GC_TINY : -GC_TINY)) : ge;
    // ---- run labels (label_step as mask arithmetic) ----
    const bool ena = E0 != E0;
    const bool s_na = z ? run_na : ((m.na >> lane) & 1u) != 0;
    const bool pos = z ? (!run_na && run_pos) : ((m.pos >> lane) & 1u) != 0;
    const unsigned v_na = __ballot_sync(FULL, s_na), v_pos = __ballot_sync(FULL, pos), e_na = __ballot_sync(FULL, ena);
    const unsigned lag_na = (v_na << 1) | c_lag_na, lag_pos = (v_pos << 1) | c_lag_pos;
    const unsigned pmask = colmask & ~lag_na & v_na;                            // (NA | FALSE) = NA: cumsum is NA from here on
    const unsigned pz = c_poisoned ? FULL : (pmask ? ~((pmask & (0u - pmask)) - 1u) : 0u);   // poisoned lanes
    const unsigned imask = colmask & ~pz & (lag_na | (v_pos ^ lag_pos));        // cumsum increments
    const unsigned vmask = colmask & ~e_na & ~pz;                               // cells with a label
    const bool valid = (vmask >> lane) & 1u;
    const int inc_le = __popc(imask & le);
    const int cum = c_cum + inc_le;
    const unsigned vlow = vmask & lt;
    // a label starts where the cumsum moved since the previous labelled cell
    const bool newlab = valid && (vlow ? (imask & le & ~((2u << gc_top(vlow)) - 1u)) != 0 : (!c_have_last || cum != c_last_val));
    const unsigned nlmask = __ballot_sync(FULL, newlab);
    const unsigned nlb = nlmask & le;
    const int dense = c_dense + __popc(nlb);
    const int cnt = nlb ? __popc(vmask & le & ~((1u << gc_top(nlb)) - 1u)) : c_cnt + __popc(vmask & le);
    const int na_cnt = c_na + __popc(~vmask & colmask & le);
    if (active) {
      P.sgre[o] = v;
      P.rdir[o] = s_na ? 0 : (pos ? 2 : 4);
      P.rn[o] = valid ? dense : 0;
      P.rcell[o] = valid ? cnt : (ena ? 0 : na_cnt);
    }
    // ---- carry to the next chunk ----
    c_prev_na = m.na >> 31; c_prev_pos = m.pos >> 31; c_inrun = m.z >> 31;
    if (rs) { const int t = gc_top(rs); c_run_na = (rv_na >> t) & 1u; c_run_pos = (rv_pos >> t) & 1u; }
    c_lag_na = v_na >> 31; c_lag_pos = v_pos >> 31;
    if (vmask) { c_last_val = c_cum + __popc(imask & ((2u << gc_top(vmask)) - 1u)); c_have_last = true; }
    c_cum += __popc(imask);
    c_poisoned = c_poisoned || pmask != 0;
    c_dense += __popc(nlmask);
    c_cnt = nlmask ? __popc(vmask & ~((1u << gc_top(nlmask)) - 1u)) : c_cnt + __popc(vmask);
    c_na += __popc(~vmask & colmask);
    E0 = E1; E1 = E2; E2 = E3; E3 = E4;
    ge = ge_n; cr = cr_n; m = mn;
  }
}

// ---- columns ---------------------------------------------------------------------------------------
// A column is cut into segments of GC_SEG rows and a thread owns one (column, segment), so that there
// are NC x NR / GC_SEG threads instead of NC; lanes are consecutive columns (coalesced).  What a segment
// needs from the rest of its column:
//   k_gc_cols_stencil  data-parallel: sgcn as computed, sgc, scc; hill_c_dir temporarily holds two flags,
//                      0x80 "elev is NA" and 0x40 "gradient is 0" (zero cells are rewritten in place by
//                      the next kernel, the flag is what other threads look at)
//   k_gc_cols_fix      south -> north.  The state entering a segment is settled by the first non-zero
//                      cell south of it: the thread starts there and runs fix_step over the zeros in
//                      between without storing (usually no cell at all; a lake costs its extent).
//   k_gc_cols_sum      north -> south.  label_step's counters are prefix sums, so a segment is first run
//                      from a blank state (with its true lag, the cell above it) and summarised ...
//   k_gc_cols_label    ... then every thread folds the summaries of the segments above its own (gc_fold)
//                      and runs label_step for real: hill_c_dir / n / cell.
constexpr int GC_THREADS = 128;
constexpr int GC_PF = 8;  // rows per group; the next group is in flight while one is worked on
constexpr int GC_NSUM = 8;

__device__ __forceinline__ void gc_col_cell(double e5, double n2, double n8, double two_g, double g2, double& gn, double& cc) {
  if (n2 != n2) n2 = e5;
  if (n8 != n8) n8 = e5;
  gn = (n2 - n8) / two_g;
  cc = (2 * e5 - n2 - n8) / g2;
}

__global__ void __launch_bounds__(256) k_gc_cols_stencil(const double* __restrict__ elev, GridDims d, double grid_m, GcPtrs P) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d.NC) return;
  const size_t o = (size_t)r * d.pitch + c;
  const double e5 = __ldg(elev + o);
  const double n2 = r + 1 < d.NR ? __ldg(elev + o + d.pitch) : gc_na();   // seqno + max_cols past the end: NA
  const double n8 = r > 0 ? __ldg(elev + o - d.pitch) : gc_na();
  double gn, cc;
  gc_col_cell(e5, n2, n8, 2 * grid_m, grid_m * grid_m, gn, cc);
  double ac = fabs(gn);
  if (ac == 0) ac = GC_TINY;
  P.sgcn[o] = gn;
  P.sgc[o] = ac;
  P.scc[o] = cc;
  P.cdir[o] = (e5 != e5 ? 0x80 : 0) | (gn == 0.0 ? 0x40 : 0);
}

__global__ void __launch_bounds__(GC_THREADS) k_gc_cols_fix(GridDims d, GcPtrs P, int seg_rows) {
  const int c = blockIdx.x * GC_THREADS + threadIdx.x;
  if (c >= d.NC) return;
  const int r_lo = blockIdx.y * seg_rows, r_hi = min(d.NR, r_lo + seg_rows) - 1;
  double* g = P.sgcn + c;
  const uint8_t* f = P.cdir + c;
  // gradient as computed: zero cells are read through their flag (their value may already be rewritten)
  auto G = [&](int r) {
    if (r < 0 || r >= d.NR) return gc_na();
    const size_t o = (size_t)r * d.pitch;
    const double v = g[o];
    return (f[o] & 0x40) ? 0.0 : v;
  };
  FixState fs;
  fix_init(fs);
  {
    int q = r_hi + 1;
    while (q < d.NR && (f[(size_t)q * d.pitch] & 0x40)) ++q;
    if (q < d.NR) fs.prev = g[(size_t)q * d.pitch];
    for (int r = q - 1; r > r_hi; --r) fix_step(fs, 0.0, G(r - 1));
  }
  // rows are visited from r_hi down to r_lo; a[k] = row rt-k, n[k] = row rt-GC_PF-k
  double a[GC_PF];
#pragma unroll
  for (int k = 0; k < GC_PF; ++k) a[k] = G(r_hi - k);
  for (int rt = r_hi; rt >= r_lo; rt -= GC_PF) {
    double n[GC_PF];
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) n[k] = G(rt - GC_PF - k);
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) {
      const int r = rt - k;
      if (r >= r_lo) {
        const double cur = a[k], nxt = k + 1 < GC_PF ? a[k + 1] : n[0];  // (rows below 0 read as NA)
        const double v = fix_step(fs, cur, nxt);
        if (cur == 0.0) g[(size_t)r * d.pitch] = v;
      }
    }
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) a[k] = n[k];
  }
}

// label_step over rows [r_lo, r_hi] of a column; STORE = false: only the state is wanted.  first_cum = cum at
// the first valid cell (-1: none)
template <bool STORE>
__device__ __forceinline__ void gc_col_label_run(const GridDims& d, const GcPtrs& P, int c, int r_lo, int r_hi, LabelState& ls,
                                                 int& first_cum) {
  const double* g = P.sgcn + c;
  uint8_t* dir = P.cdir + c;
  auto G = [&](int r) { return r <= r_hi ? g[(size_t)r * d.pitch] : gc_na(); };
  auto D = [&](int r) { return r <= r_hi ? dir[(size_t)r * d.pitch] : (uint8_t)0; };
  if (r_lo > 0) {  // lag: the fixed gradient of the cell above the segment
    const double vl = g[(size_t)(r_lo - 1) * d.pitch];
    ls.lag_na = vl != vl;
    ls.lag_pos = vl > 0;
  }
  first_cum = -1;
  double ga[GC_PF];
  uint8_t da[GC_PF];
#pragma unroll
  for (int k = 0; k < GC_PF; ++k) { ga[k] = G(r_lo + k); da[k] = D(r_lo + k); }
  for (int rt = r_lo; rt <= r_hi; rt += GC_PF) {
    double gn[GC_PF];
    uint8_t dn[GC_PF];
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) { gn[k] = G(rt + GC_PF + k); dn[k] = D(rt + GC_PF + k); }
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) {
      const int r = rt + k;
      if (r <= r_hi) {
        int hn, hc;
        const double v = ga[k];
        label_step(ls, v, (da[k] & 0x80) != 0, hn, hc);
        if (STORE) {
          const size_t o = (size_t)r * d.pitch + c;
          dir[(size_t)r * d.pitch] = v != v ? 0 : (v > 0 ? 1 : 3);
          P.cn[o] = hn;
          P.ccell[o] = hc;
        } else if (hn > 0 && first_cum < 0) first_cum = ls.cum;
      }
    }
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) { ga[k] = gn[k]; da[k] = dn[k]; }
  }
}

// summary of a segment run from a blank state: [0] cum at the first valid cell (-1: no valid cell), [1] labels
// started (the first valid cell always starts one here), [2] cnt at the end, [3] cum at the last valid cell,
// [4] cum at the end, [5] poisoned, [6] cells counted as NA
__global__ void __launch_bounds__(GC_THREADS) k_gc_cols_sum(GridDims d, GcPtrs P, int seg_rows, int32_t* __restrict__ sum) {
  const int c = blockIdx.x * GC_THREADS + threadIdx.x;
  if (c >= d.NC) return;
  const int r_lo = blockIdx.y * seg_rows, r_hi = min(d.NR, r_lo + seg_rows) - 1;
  LabelState ls;
  label_init(ls);
  int first_cum;
  gc_col_label_run<false>(d, P, c, r_lo, r_hi, ls, first_cum);
  int32_t* o = sum + ((size_t)blockIdx.y * GC_NSUM) * d.NC + c;
  o[0 * (size_t)d.NC] = first_cum;
  o[1 * (size_t)d.NC] = ls.dense;
  o[2 * (size_t)d.NC] = ls.cnt;
  o[3 * (size_t)d.NC] = ls.last_val;
  o[4 * (size_t)d.NC] = ls.cum;
  o[5 * (size_t)d.NC] = ls.poisoned ? 1 : 0;
  o[6 * (size_t)d.NC] = ls.na_cnt;
}

// state after a segment, from the state before it and the segment's summary (rows = its length)
__device__ __forceinline__ void gc_fold(LabelState& s, const int32_t* __restrict__ o, size_t stride, int rows) {
  if (s.poisoned) { s.na_cnt += rows; return; }   // cumsum is NA: every cell below is in the NA group
  const int first_cum = o[0], dense = o[stride], cnt = o[2 * stride], last_val = o[3 * stride];
  if (first_cum >= 0) {
    // only the first valid cell's "new label?" depends on what came before; all later ones compare inside
    const bool first_new = !s.have_last || s.cum + first_cum != s.last_val;
    s.cnt = (dense == 1 && !first_new) ? s.cnt + cnt : cnt;
    s.dense += dense - (first_new ? 0 : 1);
    s.last_val = s.cum + last_val;
    s.have_last = true;
  }
  s.cum += o[4 * stride];
  s.poisoned = o[5 * stride] != 0;
  s.na_cnt += o[6 * stride];
}

__global__ void __launch_bounds__(GC_THREADS) k_gc_cols_label(GridDims d, GcPtrs P, int seg_rows, const int32_t* __restrict__ sum) {
  const int c = blockIdx.x * GC_THREADS + threadIdx.x;
  if (c >= d.NC) return;
  const int r_lo = blockIdx.y * seg_rows, r_hi = min(d.NR, r_lo + seg_rows) - 1;
  LabelState ls;
  label_init(ls);
  for (int j = 0; j < (int)blockIdx.y; ++j)
    gc_fold(ls, sum + ((size_t)j * GC_NSUM) * d.NC + c, (size_t)d.NC, seg_rows);
  int first_cum;
  gc_col_label_run<true>(d, P, c, r_lo, r_hi, ls, first_cum);
}

}  // namespace

int run_slope_gc(litap_grid* g, double grid_m) {
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  for (int k = 0; k < 6; ++k) if (!g->gc_f64[k]) g->gc_f64[k] = grid_alloc_array<double>(d, -1);
  for (int k = 0; k < 2; ++k) if (!g->gc_u8[k]) g->gc_u8[k] = grid_alloc_array<uint8_t>(d, -1);
  for (int k = 0; k < 4; ++k) if (!g->gc_i32[k]) g->gc_i32[k] = grid_alloc_array<int32_t>(d, -1);
  for (int k = 0; k < 6; ++k) if (!g->gc_f64[k]) return set_error(LITAP_ENOMEM, "device allocation failed (slope_gc outputs)");
  for (int k = 0; k < 2; ++k) if (!g->gc_u8[k]) return set_error(LITAP_ENOMEM, "device allocation failed (slope_gc outputs)");
  for (int k = 0; k < 4; ++k) if (!g->gc_i32[k]) return set_error(LITAP_ENOMEM, "device allocation failed (slope_gc outputs)");
  GcPtrs P;
  P.sgre = g->gc_f64[0]; P.sgr = g->gc_f64[1]; P.sgcn = g->gc_f64[2]; P.sgc = g->gc_f64[3];
  P.scr = g->gc_f64[4]; P.scc = g->gc_f64[5];
  P.rdir = g->gc_u8[0]; P.cdir = g->gc_u8[1];
  P.rn = g->gc_i32[0]; P.rcell = g->gc_i32[1]; P.cn = g->gc_i32[2]; P.ccell = g->gc_i32[3];
  KERNEL("k_gc_rows", k_gc_rows<<<(d.NR + GR_WARPS - 1) / GR_WARPS, GR_WARPS * 32, 0, st>>>(g->elev, d, grid_m, P));
  const int seg_rows = d.NR < 2048 ? 64 : 512;
  const int nseg = (d.NR + seg_rows - 1) / seg_rows;
  if (!g->gc_sum) g->gc_sum = (int32_t*)pool_alloc((size_t)nseg * GC_NSUM * d.NC * sizeof(int32_t));
  if (!g->gc_sum) return set_error(LITAP_ENOMEM, "device allocation failed (slope_gc summaries)");
  const dim3 cgrid((d.NC + GC_THREADS - 1) / GC_THREADS, nseg);
  KERNEL("k_gc_cols_stencil", k_gc_cols_stencil<<<dim3((d.NC + 255) / 256, d.NR), 256, 0, st>>>(g->elev, d, grid_m, P));
  KERNEL("k_gc_cols_fix", k_gc_cols_fix<<<cgrid, GC_THREADS, 0, st>>>(d, P, seg_rows));
  KERNEL("k_gc_cols_sum", k_gc_cols_sum<<<cgrid, GC_THREADS, 0, st>>>(d, P, seg_rows, g->gc_sum));
  KERNEL("k_gc_cols_label", k_gc_cols_label<<<cgrid, GC_THREADS, 0, st>>>(d, P, seg_rows, g->gc_sum));
  return LITAP_OK;
}
