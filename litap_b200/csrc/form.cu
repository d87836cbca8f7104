// form.cu — calc_form on the device (R/form_01_calc_form.R:4-67): slope %, slope deg, aspect,
// profile and plan curvature as one fused 3x3 stencil, 48 B/cell (8 read + 5 x 8 written).
//   k_form       : every cell from its own window, with the final NA -> 0 step (:60-65)
//   k_form_edges : the four replicated lines (:54-57) recomputed from their source cells
#include <cmath>
#include "cellmath.h"
#include "litap_internal.h"
#include <cstdlib>
#include "rowring.cuh"
#include "stencil.cuh"

namespace {

constexpr int FORM_CPT = 2;       // columns per thread: 16-byte loads and stores
constexpr int FORM_ROWS = 32;
constexpr int FORM_THREADS = 128;  // 256 columns per block

struct FormPtrs { double* o[5]; };
// replicated lines: out(r, c) takes the window of (mapr(r), mapc(c)); the maps differ from the
// identity on at most two rows and two columns, simulated on the host in the reference's order.
struct EdgeMap { int dst_c[2], src_c[2], dst_r[2], src_r[2]; };

// NA -> 0 where the cell itself has an elevation; aspect keeps its value (:60-65)
__device__ __forceinline__ void finalise(FormOut& o, bool elev_ok) {
  if (elev_ok) {
    if (is_nan(o.slope_pct)) o.slope_pct = 0.0;
    if (is_nan(o.slope_deg)) o.slope_deg = 0.0;
    if (is_nan(o.prof)) o.prof = 0.0;
    if (is_nan(o.plan)) o.plan = 0.0;
  }
}
__device__ __forceinline__ void store_final(const FormPtrs& P, size_t i, FormOut o, bool elev_ok) {
  finalise(o, elev_ok);
  P.o[0][i] = o.slope_pct;
  P.o[1][i] = o.slope_deg;
  P.o[2][i] = o.aspect;
  P.o[3][i] = o.prof;
  P.o[4][i] = o.plan;
}

// One block = FORM_ROWS rows x 256 columns.  Every cell first takes form_cell_quick (cellmath.h);
// the few it leaves in doubt (~0.2 % on terrain; every cell next to an NA) are queued in shared memory
// and resolved densely at the end of the block by form_cell_fast / form_cell, so the exact paths
// never run inside a divergent warp.
constexpr int FORM_QCAP = 2048;
#ifndef FORM_MINB
#define FORM_MINB 5
#endif

__device__ __noinline__ void form_exact_cell(const double* __restrict__ elev, const GridDims& d, const FormK& K,
                                                const double* __restrict__ trig, const FormPtrs& P, int r, int c) {
  double e[10];
#pragma unroll
  for (int n = 1; n <= 9; ++n) {
    const int rr = r + nb_dr(n), cc = c + nb_dc(n);  // cc >= 0: column 0 is never queued; pad columns are NaN
    e[n] = (rr < 0 || rr >= d.NR) ? sw_nan() : __ldg(elev + (size_t)rr * d.pitch + cc);
  }
  FormOut o;
  if (!form_cell_fast(e, K, trig, trig + 181, &o)) o = form_cell(e, K.grid, trig, trig + 181);
  store_final(P, (size_t)r * d.pitch + c, o, !is_nan(e[5]));
}

constexpr int FORM_STAGES = 12;  // ring slots (rows in flight: 8..12)
constexpr int FORM_GROUP = 4;    // slots recycled per block barrier
using FormRing = RowRing<double, FORM_THREADS * FORM_CPT, FORM_STAGES>;

// z = elev * 100 of the thread's window columns in ring slot `row`: own columns 2+2t, 3+2t, the
// neighbours either side; an invalid row (outside the grid) is NA
__device__ __forceinline__ void form_take_row(RowWin<FORM_CPT>& w, FormRing& ring, int k, bool valid) {
  if (valid) {
    const int s = k % FORM_STAGES;
    rr_wait(&ring.bar[s], (unsigned)(k / FORM_STAGES) & 1u);
    const double* row = ring.row[s] + 2 * threadIdx.x;
    const double2 own = *reinterpret_cast<const double2*>(row + 2);
    w.v[0] = row[1] * 100;
    w.v[1] = own.x * 100;
    w.v[2] = own.y * 100;
    w.v[3] = row[4] * 100;
  } else {
#pragma unroll
    for (int j = 0; j < FORM_CPT + 2; ++j) w.v[j] = sw_nan();
  }
}

__global__ void __launch_bounds__(FORM_THREADS, FORM_MINB) k_form(const double* __restrict__ elev, GridDims d, FormK K, FormQ Q,
                                                       const double* __restrict__ trig,
                                                       const double* __restrict__ qtab, FormPtrs P, int row_lo, int row_hi) {
  static_assert(FORM_CPT == 2, "window layout of form_take_row");
  __shared__ FormRing ring;
  __shared__ double s_qw[181 * 3], s_qat[QK_ATAN_N];
  __shared__ unsigned short s_queue[FORM_QCAP];
  __shared__ unsigned s_qn;
  const int cb = blockIdx.x * FORM_THREADS * FORM_CPT + 1;  // first column of the block (column 1 is the aligned one)
  const int c0 = cb + threadIdx.x * FORM_CPT;
  const int r0 = row_lo + blockIdx.y * FORM_ROWS;  // rows [row_lo, row_hi) of the grid; row row_hi must be resident
  const int rend = min(r0 + FORM_ROWS, row_hi);
  const bool live = c0 < d.pitch - LITAP_XOFF;
  // ring row k = grid row r0 - 1 + k, k = 0 .. kmax
  const int kmax = rend - r0 + 1;
  // the segment starts at column cb - 2 and must stay inside pitch + 32 elements of its row
  const unsigned seg_bytes = (unsigned)min(FormRing::SEG, ((d.pitch + 32 - (cb - 2)) / 2) * 2) * sizeof(double);
  int next_k = 0;  // thread 0: first ring row not yet requested
  if (threadIdx.x == 0) {
    for (int s = 0; s < FORM_STAGES; ++s) rr_bar_init(&ring.bar[s]);
    rr_fence_init();
    s_qn = 0;
    for (; next_k < FORM_STAGES && next_k <= kmax; ++next_k) {
      const int rr = r0 - 1 + next_k;
      if (rr >= 0 && rr < d.NR)
        rr_issue(ring.row[next_k % FORM_STAGES], elev + (size_t)rr * d.pitch + (cb - 2), seg_bytes, &ring.bar[next_k % FORM_STAGES]);
      else rr_skip(&ring.bar[next_k % FORM_STAGES]);
    }
  }
  for (int i = threadIdx.x; i < 181 * 3 + QK_ATAN_N; i += FORM_THREADS) {
    const double v = __ldg(qtab + i);
    if (i < 181 * 3) s_qw[i] = v; else s_qat[i - 181 * 3] = v;
  }
  __syncthreads();
  RowWin<FORM_CPT> p, q, n;  // rows r-1, r, r+1 as z = elev * 100
  form_take_row(p, ring, 0, r0 - 1 >= 0);
  form_take_row(q, ring, 1, true);
  for (int r = r0; r < rend; ++r) {
    const int j = r - r0;
    if (j > 0 && (j % FORM_GROUP) == 0) {
      __syncthreads();  // every thread has taken ring rows <= j + 1: their slots are free
      if (threadIdx.x == 0) {
        for (; next_k <= kmax && next_k - FORM_STAGES <= j + 1; ++next_k) {
          const int rr = r0 - 1 + next_k;
          if (rr < d.NR)
            rr_issue(ring.row[next_k % FORM_STAGES], elev + (size_t)rr * d.pitch + (cb - 2), seg_bytes,
                     &ring.bar[next_k % FORM_STAGES]);
          else rr_skip(&ring.bar[next_k % FORM_STAGES]);
        }
      }
    }
    form_take_row(n, ring, j + 2, r + 1 < d.NR);
    FormOut o[FORM_CPT];
    bool okj[FORM_CPT];
#pragma unroll
    for (int jj = 0; jj < FORM_CPT; ++jj) {
      double z[10];
      z[7] = p.v[jj]; z[8] = p.v[jj + 1]; z[9] = p.v[jj + 2];
      z[4] = q.v[jj]; z[5] = q.v[jj + 1]; z[6] = q.v[jj + 2];
      z[1] = n.v[jj]; z[2] = n.v[jj + 1]; z[3] = n.v[jj + 2];
      okj[jj] = form_cell_quick(z, Q, s_qw, s_qat, &o[jj]);
    }
    if (live) {
      const size_t i = (size_t)r * d.pitch + c0;
      *reinterpret_cast<double2*>(P.o[0] + i) = make_double2(o[0].slope_pct, o[1].slope_pct);
      *reinterpret_cast<double2*>(P.o[1] + i) = make_double2(o[0].slope_deg, o[1].slope_deg);
      *reinterpret_cast<double2*>(P.o[2] + i) = make_double2(o[0].aspect, o[1].aspect);
      *reinterpret_cast<double2*>(P.o[3] + i) = make_double2(o[0].prof, o[1].prof);
      *reinterpret_cast<double2*>(P.o[4] + i) = make_double2(o[0].plan, o[1].plan);
      if (c0 == 1) {  // ring column 0: NA everywhere
#pragma unroll
        for (int k = 0; k < 5; ++k) P.o[k][i - 1] = sw_nan();
      }
#pragma unroll
      for (int jj = 0; jj < FORM_CPT; ++jj) {
        if (c0 + jj >= d.NC) {  // pad columns: NA
#pragma unroll
          for (int k = 0; k < 5; ++k) P.o[k][i + jj] = sw_nan();
        } else if (!okj[jj]) {
          const unsigned slot = atomicAdd(&s_qn, 1u);
          if (slot < FORM_QCAP) s_queue[slot] = (unsigned short)((j << 8) | (threadIdx.x * FORM_CPT + jj));
          else form_exact_cell(elev, d, K, trig, P, r, c0 + jj);  // queue full (NA-heavy block): resolve in place
        }
      }
    }
    p = q;
    q = n;
  }
  __syncthreads();  // queued cells overwrite what the quick path stored for them
  const unsigned nq = min(s_qn, (unsigned)FORM_QCAP);
  for (unsigned i = threadIdx.x; i < nq; i += FORM_THREADS) {
    const unsigned w = s_queue[i];
    form_exact_cell(elev, d, K, trig, P, r0 + (int)(w >> 8), cb + (int)(w & 255u));
  }
}

__device__ __forceinline__ int map1(int x, const int* dst, const int* src) {
  // the second assignment reads the state left by the first (R executes them in sequence)
  if (x == dst[1]) { int s = src[1]; return s == dst[0] ? src[0] : s; }
  if (x == dst[0]) return src[0];
  return x;
}

__global__ void k_form_edges(const double* __restrict__ elev, GridDims d, double grid,
                             const double* __restrict__ trig, FormPtrs P, EdgeMap em, int row_lo, int row_hi) {
  // line 0,1: destination columns (all rows); line 2,3: destination rows (all columns)
  const int line = blockIdx.y;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  int r, c;
  if (line < 2) { c = em.dst_c[line]; r = row_lo + t; if (c < 0 || r >= row_hi) return; }
  else { r = em.dst_r[line - 2]; c = t; if (r < row_lo || r >= row_hi || c >= d.NC) return; }
  const int sr = map1(r, em.dst_r, em.src_r), sc = map1(c, em.dst_c, em.src_c);
  const double NA = __longlong_as_double(0x7ff8000000000000LL);
  double e[10];
#pragma unroll
  for (int n = 1; n <= 9; ++n) {
    int rr = sr + nb_dr(n), cc = sc + nb_dc(n);
    e[n] = (rr < 0 || rr >= d.NR || cc < 0 || cc >= d.NC) ? NA : elev[(size_t)rr * d.pitch + cc];
  }
  FormOut o = form_cell(e, grid, trig, trig + 181);
  const size_t i = (size_t)r * d.pitch + c;
  store_final(P, i, o, !is_nan(elev[i]));
}

double* g_trig = nullptr;  // sin((A*pi)/180), cos((A*pi)/180), A = 0..180, from the host libm
double* g_qtab = nullptr;  // form_fill_quick_tables(): 543 curvature weights + 513 arctangents
double g_qtab_grid = 0.0;  // the grid size the weights were built for
double g_trig_host[362];
}  // namespace

int run_form(litap_grid* g, double grid_m) { return run_form_rows(g, grid_m, 0, g->d.NR); }

// rows [row_lo, row_hi) of every output (row_hi < NR: row row_hi of elev must already be on the device)
int run_form_rows(litap_grid* g, double grid_m, int row_lo, int row_hi) {
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  if (row_lo >= row_hi) return LITAP_OK;
  if (!g_trig) {
    double h[362];
    for (int A = 0; A <= 180; ++A) {
      double t = ((double)A * M_PI) / 180;  // deg_rad(), R/form_01_calc_form.R:115
      h[A] = std::sin(t);
      h[181 + A] = std::cos(t);
    }
    g_trig = (double*)pool_alloc(sizeof(h));
    if (!g_trig) return set_error(LITAP_ENOMEM, "device allocation failed (trig table)");
    CUDA_TRY(cudaMemcpyAsync(g_trig, h, sizeof(h), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    for (int i = 0; i < 362; ++i) g_trig_host[i] = h[i];
  }
  if (!g_qtab || g_qtab_grid != grid_m) {
    double h[181 * 3 + QK_ATAN_N];
    form_fill_quick_tables(grid_m, g_trig_host, g_trig_host + 181, h, h + 181 * 3);
    if (!g_qtab) g_qtab = (double*)pool_alloc(sizeof(h));
    if (!g_qtab) return set_error(LITAP_ENOMEM, "device allocation failed (form tables)");
    CUDA_TRY(cudaMemcpyAsync(g_qtab, h, sizeof(h), cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));  // h is a stack array; also orders the update after earlier launches
    g_qtab_grid = grid_m;
  }
  FormPtrs P;
  for (int k = 0; k < 5; ++k) {
    if (!g->form[k]) g->form[k] = grid_alloc_array<double>(d, -1);
    if (!g->form[k]) return set_error(LITAP_ENOMEM, "device allocation failed (form outputs)");
    P.o[k] = g->form[k];
  }
  {
    const int cols = (d.pitch - LITAP_XOFF + FORM_CPT - 1) / FORM_CPT;  // threads across a row
    dim3 grid((cols + FORM_THREADS - 1) / FORM_THREADS, (row_hi - row_lo + FORM_ROWS - 1) / FORM_ROWS);
    FormK K;
    K.grid = grid_m;
    K.inv2g = 1.0 / (2 * grid_m);
    K.invg2 = 1.0 / (grid_m * grid_m);
    FormQ Q;
    Q.inv2g = K.inv2g;
    Q.gz = 8.0 * 4.440892098500626e-16 * 57295.78 / (grid_m * grid_m);
    // LITAP_FORM_BPS (tuning): cap k_form's resident blocks per SM with unused dynamic shared memory, so that the flow
    // stages of litap_*_flow_form (other pipes of the SM: ALU / shared memory, k_form is fp64-bound) stay co-resident
    static int dyn_smem = -1;
    if (dyn_smem < 0) {
      dyn_smem = 0;
      if (const char* e = getenv("LITAP_FORM_BPS")) {
        const int bps = atoi(e);
        cudaFuncAttributes fa;
        if (bps > 0 && cudaFuncGetAttributes(&fa, k_form) == cudaSuccess) {
          const int per_block = (227 * 1024) / bps - 1024;   // shared memory of one block such that bps + 1 do not fit
          const int share = (227 * 1024) / (bps + 1);
          int want = share + 1024 - (int)fa.sharedSizeBytes;
          if (want > per_block - (int)fa.sharedSizeBytes) want = per_block - (int)fa.sharedSizeBytes;
          if (want > 0 && cudaFuncSetAttribute(k_form, cudaFuncAttributeMaxDynamicSharedMemorySize, want) == cudaSuccess) dyn_smem = want;
        }
      }
    }
    KERNEL("k_form", k_form<<<grid, FORM_THREADS, dyn_smem, st>>>(g->elev, d, K, Q, g_trig, g_qtab, P, row_lo, row_hi));
  }
  // R/form_01_calc_form.R:54-57 in buffered 0-based coordinates, in this order:
  //   col 1 <- col 2 ; col NC-2 <- col NC-3 ; row 1 <- row 2 ; row NR-2 <- row NR-3
  EdgeMap em;
  em.dst_c[0] = 1; em.src_c[0] = 2 < d.NC ? 2 : -1;
  em.dst_c[1] = d.NC - 2; em.src_c[1] = d.NC - 3;
  // (a row band replicates rows only where it holds the first / last row of the DEM)
  em.dst_r[0] = g->edge_top_row; em.src_r[0] = g->edge_top_row + 1 < d.NR ? g->edge_top_row + 1 : -1;
  em.dst_r[1] = g->edge_bot_row; em.src_r[1] = g->edge_bot_row - 1;
  {
    int m = (row_hi - row_lo) > d.NC ? (row_hi - row_lo) : d.NC;
    dim3 grid((m + 127) / 128, 4);
    KERNEL("k_form_edges", k_form_edges<<<grid, 128, 0, st>>>(g->elev, d, grid_m, g_trig, P, em, row_lo, row_hi));
  }
  return LITAP_OK;
}

// called by litap_shutdown through pool_release_all(): the table is pool-owned
void form_forget_trig() { g_trig = nullptr; g_qtab = nullptr; g_qtab_grid = 0.0; }

// ---- packed downloads (litap_flow_form): every output of calc_form is a rounded number -- k / 1000, whole degrees for
// aspect --, so rows [row_lo, row_hi) can cross PCIe as the int32 k (half the bytes) and be widened on the host with the
// same correctly rounded quotient (hostexpand.cpp).  NA -> INT32_MIN, -0 -> INT32_MIN + 1.  A value that its integer does
// not reproduce bit for bit (|k| >= 2^31: slopes above 2 million per cent) raises the output's flag; the caller then
// downloads that output as fp64.
struct PackPtrs { const double* in[5]; int32_t* out[5]; };
__global__ void __launch_bounds__(256) k_form_pack(GridDims d, PackPtrs P, int row_lo, int row_hi, int* __restrict__ flags) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = row_lo + blockIdx.y;
  if (c >= d.NC || r >= row_hi) return;
  const size_t id = (size_t)r * d.pitch + c;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    if (!P.out[i]) continue;
    const double v = P.in[i][id];
    int32_t k;
    bool ok = true;
    if (is_nan(v)) k = INT32_MIN;
    else if (v == 0.0) k = __double_as_longlong(v) < 0 ? INT32_MIN + 1 : 0;
    else {
      const double kd = i == 2 ? v : rint(v * 1000.0);
      ok = fabs(kd) < 2147483000.0;
      k = ok ? (int32_t)kd : 0;
      const double back = i == 2 ? (double)k : (double)k / 1000.0;
      ok = ok && __double_as_longlong(back) == __double_as_longlong(v);
    }
    if (!ok) flags[i] = 1;
    P.out[i][id] = k;
  }
}
int run_form_pack(litap_grid* g, int32_t* const* out, int row_lo, int row_hi, int* flags) {
  const GridDims d = g->d;
  if (row_lo >= row_hi) return LITAP_OK;
  PackPtrs P;
  for (int i = 0; i < 5; ++i) { P.in[i] = g->form[i]; P.out[i] = out[i]; }
  dim3 grid((d.NC + 255) / 256, row_hi - row_lo);
  KERNEL("k_form_pack", k_form_pack<<<grid, 256, 0, ctx().stream>>>(d, P, row_lo, row_hi, flags));
  return LITAP_OK;
}
