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
// A line is inherently sequential in this formulation; all lines run in parallel:
//   k_gc_rows   a block owns 16 rows and walks along them in 32-column chunks staged through shared
//               memory (coalesced loads / stores by the whole block, one lane per row in between)
//   k_gc_cols_stencil / _fix / _label  a plain stencil, then one thread per column south -> north (zero signs)
//               and north -> south (hill_c_dir / n / cell)
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
constexpr int GR_ROWS = 8;      // rows per block = lanes of the sweep; many small blocks per SM overlap their phases
constexpr int GR_CH = 32;       // columns per chunk
constexpr int GR_THREADS = 64;

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

constexpr int GR_LD = (GR_ROWS * (GR_CH + 3) + GR_THREADS - 1) / GR_THREADS;  // elevations a thread stages per chunk

__global__ void __launch_bounds__(GR_THREADS) k_gc_rows(const double* __restrict__ elev, GridDims d, double grid_m, GcPtrs P) {
  __shared__ double s_e[GR_ROWS][GR_CH + 5];     // columns cb-1 .. cb+GR_CH+1 of the block's rows
  __shared__ double s_ge[GR_ROWS][GR_CH + 3];    // gradient of columns cb .. cb+GR_CH (one ahead of the sign rule)
  __shared__ double s_v[GR_ROWS][GR_CH + 1];     // gradient with its zero signs settled
  __shared__ int s_n[GR_ROWS][GR_CH + 1], s_cell[GR_ROWS][GR_CH + 1];
  const int r0 = blockIdx.x * GR_ROWS;
  const double two_g = 2 * grid_m, g2 = grid_m * grid_m;
  FixState fs;
  LabelState ls;
  fix_init(fs);
  label_init(ls);
  // the chunk after the current one travels in registers while the current one is worked on
  double pre[GR_LD];
  auto fetch = [&](int cb) {
#pragma unroll
    for (int q = 0; q < GR_LD; ++q) {
      const int i = threadIdx.x + q * GR_THREADS;
      const int rr = i / (GR_CH + 3), k = i - rr * (GR_CH + 3);
      const int r = r0 + rr;
      pre[q] = (rr < GR_ROWS && r < d.NR && cb < d.NC) ? gc_elev_lin(elev, d, r, cb - 1 + k) : gc_na();
    }
  };
  fetch(0);
  for (int cb = 0; cb < d.NC; cb += GR_CH) {
    __syncthreads();  // the previous chunk's tiles have been consumed
#pragma unroll
    for (int q = 0; q < GR_LD; ++q) {
      const int i = threadIdx.x + q * GR_THREADS;
      const int rr = i / (GR_CH + 3), k = i - rr * (GR_CH + 3);
      if (rr < GR_ROWS) s_e[rr][k] = pre[q];
    }
    __syncthreads();
    fetch(cb + GR_CH);
    // data-parallel part, all threads: gradient (one column ahead for the sign rule), magnitude, curvature
    for (int i = threadIdx.x; i < GR_ROWS * (GR_CH + 1); i += GR_THREADS) {
      const int rr = i / (GR_CH + 1), j = i - rr * (GR_CH + 1);
      const int r = r0 + rr, c = cb + j;
      double ge = gc_na(), cr = gc_na();
      if (r < d.NR && c < d.NC) gc_row_cell(s_e[rr][j + 1], s_e[rr][j], s_e[rr][j + 2], two_g, g2, ge, cr);
      s_ge[rr][j] = ge;
      if (j < GR_CH && r < d.NR && c < d.NC) {
        const size_t o = (size_t)r * d.pitch + c;
        double ar = fabs(ge);
        if (ar == 0) ar = GC_TINY;
        P.sgr[o] = ar;
        P.scr[o] = cr;
      }
    }
    __syncthreads();
    // sequential part, one lane per row: the two sweeps over the chunk
    if (threadIdx.x < GR_ROWS && r0 + (int)threadIdx.x < d.NR) {
      const int rr = threadIdx.x;
      const int nc = min(GR_CH, d.NC - cb);
      double g_cur = s_ge[rr][0];
#pragma unroll 4
      for (int j = 0; j < nc; ++j) {
        const double g_nxt = s_ge[rr][j + 1];   // NA past the last column
        const double e5 = s_e[rr][j + 1];
        const double v = fix_step(fs, g_cur, g_nxt);
        int hn, hc;
        label_step(ls, v, e5 != e5, hn, hc);
        s_v[rr][j] = v;
        s_n[rr][j] = hn;
        s_cell[rr][j] = hc;
        g_cur = g_nxt;
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < GR_ROWS * GR_CH; i += GR_THREADS) {
      const int rr = i / GR_CH, j = i - rr * GR_CH;
      const int r = r0 + rr, c = cb + j;
      if (r >= d.NR || c >= d.NC) continue;
      const size_t o = (size_t)r * d.pitch + c;
      const double v = s_v[rr][j];
      P.sgre[o] = v;
      P.rdir[o] = v != v ? 0 : (v > 0 ? 2 : 4);
      P.rn[o] = s_n[rr][j];
      P.rcell[o] = s_cell[rr][j];
    }
  }
}

// ---- columns ---------------------------------------------------------------------------------------
// k_gc_cols_stencil  data-parallel: sgcn as computed, sgc, scc; hill_c_dir temporarily holds "elev is NA"
// k_gc_cols_fix      one thread per column, south -> north: zero gradients take their sign (only those cells
//                    are rewritten)
// k_gc_cols_label    one thread per column, north -> south: hill_c_dir / n / cell
// The sweeps read 8-9 bytes per cell and keep GC_PF x 2 rows in flight per thread: with one thread per
// column there are only NC/32 warps, so latency has to be covered inside the thread.
constexpr int GC_THREADS = 32;
constexpr int GC_PF = 16;  // rows per group; two groups are in flight

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
  P.cdir[o] = e5 != e5 ? 0x80 : 0;
}

__global__ void __launch_bounds__(GC_THREADS) k_gc_cols_fix(GridDims d, GcPtrs P) {
  const int c = blockIdx.x * GC_THREADS + threadIdx.x;
  if (c >= d.NC) return;
  double* g = P.sgcn + c;
  auto G = [&](int r) { return (r < 0 || r >= d.NR) ? gc_na() : g[(size_t)r * d.pitch]; };
  FixState fs;
  fix_init(fs);
  // rows are visited from NR-1 down to 0; a[k] = row rt-k, b[k] = row rt-GC_PF-k, then one more group is fetched
  double a[GC_PF], b[GC_PF];
#pragma unroll
  for (int k = 0; k < GC_PF; ++k) { a[k] = G(d.NR - 1 - k); b[k] = G(d.NR - 1 - GC_PF - k); }
  for (int rt = d.NR - 1; rt >= 0; rt -= GC_PF) {
    double n[GC_PF];
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) n[k] = G(rt - 2 * GC_PF - k);
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) {
      const int r = rt - k;
      if (r >= 0) {
        const double cur = a[k], nxt = k + 1 < GC_PF ? a[k + 1] : b[0];  // (rows below 0 read as NA)
        const double v = fix_step(fs, cur, nxt);
        if (cur == 0.0) g[(size_t)r * d.pitch] = v;
      }
    }
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) { a[k] = b[k]; b[k] = n[k]; }
  }
}

__global__ void __launch_bounds__(GC_THREADS) k_gc_cols_label(GridDims d, GcPtrs P) {
  const int c = blockIdx.x * GC_THREADS + threadIdx.x;
  if (c >= d.NC) return;
  const double* g = P.sgcn + c;
  uint8_t* dir = P.cdir + c;
  auto G = [&](int r) { return r < d.NR ? g[(size_t)r * d.pitch] : gc_na(); };
  auto D = [&](int r) { return r < d.NR ? dir[(size_t)r * d.pitch] : (uint8_t)0; };
  LabelState ls;
  label_init(ls);
  double ga[GC_PF], gb[GC_PF];
  uint8_t da[GC_PF], db[GC_PF];
#pragma unroll
  for (int k = 0; k < GC_PF; ++k) { ga[k] = G(k); da[k] = D(k); gb[k] = G(GC_PF + k); db[k] = D(GC_PF + k); }
  for (int rt = 0; rt < d.NR; rt += GC_PF) {
    double gn[GC_PF];
    uint8_t dn[GC_PF];
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) { gn[k] = G(rt + 2 * GC_PF + k); dn[k] = D(rt + 2 * GC_PF + k); }
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) {
      const int r = rt + k;
      if (r < d.NR) {
        int hn, hc;
        const double v = ga[k];
        label_step(ls, v, (da[k] & 0x80) != 0, hn, hc);
        const size_t o = (size_t)r * d.pitch + c;
        dir[(size_t)r * d.pitch] = v != v ? 0 : (v > 0 ? 1 : 3);
        P.cn[o] = hn;
        P.ccell[o] = hc;
      }
    }
#pragma unroll
    for (int k = 0; k < GC_PF; ++k) { ga[k] = gb[k]; gb[k] = gn[k]; da[k] = db[k]; db[k] = dn[k]; }
  }
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
  KERNEL("k_gc_rows", k_gc_rows<<<(d.NR + GR_ROWS - 1) / GR_ROWS, GR_THREADS, 0, st>>>(g->elev, d, grid_m, P));
  const int cblk = (d.NC + GC_THREADS - 1) / GC_THREADS;
  KERNEL("k_gc_cols_stencil", k_gc_cols_stencil<<<dim3((d.NC + 255) / 256, d.NR), 256, 0, st>>>(g->elev, d, grid_m, P));
  KERNEL("k_gc_cols_fix", k_gc_cols_fix<<<cblk, GC_THREADS, 0, st>>>(d, P));
  KERNEL("k_gc_cols_label", k_gc_cols_label<<<cblk, GC_THREADS, 0, st>>>(d, P));
  return LITAP_OK;
}
