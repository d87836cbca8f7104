// ddir.cu — calc_ddir2 on the device (R/flow_01_calc_ddir.R:2-118).
//   k_elev_max / k_invert : invert()                      R/flow_06_invert.R:1-5
//   k_d8                  : finddir2 + worklist of flats  R/LITAP_functions.R:257-273  (9 B/cell)
//   k_plateau_all         : all synchronous plateau rounds in one cooperative launch
//                                                         R/flow_01_calc_ddir.R:12-41
//   k_plateau_probe/apply : one round (row-band mode: the host exchanges halo rows between rounds)
//   k_patch_* + cub sort  : flat-pit patches: components (union-find), median centre, get_dir
//                                                         R/flow_01_calc_ddir.R:48-107
// No raster and no result ever goes through the host: it reads three counters per call.
#include <cooperative_groups.h>
#include <cstdlib>
#include <cub/cub.cuh>
#include "cellmath.h"
#include "litap_internal.h"
#include "rowring.cuh"
#include "stencil.cuh"

namespace cg = cooperative_groups;

namespace {

__device__ __forceinline__ unsigned long long f64_orderable(double x) {
  unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double f64_from_orderable(unsigned long long k) {
  unsigned long long b = (k & 0x8000000000000000ull) ? (k & 0x7FFFFFFFFFFFFFFFull) : ~k;
  return __longlong_as_double((long long)b);
}

// device counters of one calc_ddir2 call
struct DdirCounters {
  unsigned long long nflat;    // flat cells after the initial stencil (may exceed the list capacity)
  unsigned long long elev_max; // orderable max(elev) for invert()
  unsigned npatch;             // cells of flat-pit patches
  unsigned rounds;             // plateau rounds that changed something
  unsigned changed[2];         // per-round "any change" flags (double buffered)
  unsigned pad[2];
};

// ---- invert ------------------------------------------------------------------------------------
__global__ void k_elev_max(const double* __restrict__ elev, GridDims d, int r_lo, int r_hi, unsigned long long* out) {
  unsigned long long best = 0;  // below every orderable finite value
  for (int r = r_lo + blockIdx.y; r < r_hi; r += gridDim.y)
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < d.NC; c += gridDim.x * blockDim.x) {
      double v = elev[(size_t)r * d.pitch + c];
      if (!is_nan(v)) { unsigned long long k = f64_orderable(v); best = k > best ? k : best; }
    }
  for (int o = 16; o; o >>= 1) { unsigned long long t = __shfl_xor_sync(0xffffffffu, best, o); best = t > best ? t : best; }
  if ((threadIdx.x & 31) == 0 && best) atomicMax(out, best);
}
__global__ void k_invert(double* __restrict__ elev, GridDims d, const unsigned long long* mx) {
  const double m = f64_from_orderable(*mx);
  int c = blockIdx.x * blockDim.x + threadIdx.x + 1 - 32;  // warp-aligned on interior column 1
  int r = blockIdx.y;
  if (c >= 0 && c < d.NC) {
    size_t i = (size_t)r * d.pitch + c;
    elev[i] = m - elev[i];  // NaN stays NaN
  }
}

// ---- D8 stencil + flat worklist ------------------------------------------------------------------
constexpr int D8_CPT = 4;        // columns per thread: 32-byte loads, uchar4 stores
constexpr int D8_THREADS = 128;  // 512 columns per block
#ifndef D8_ROWS_PER_BLOCK
#define D8_ROWS_PER_BLOCK 32
#endif
constexpr int D8_ROWS = D8_ROWS_PER_BLOCK;      // rows per block
constexpr int D8_STAGE = 1024;   // flats a block stages in shared memory before appending

// the reference's loop for the (very rare) windows d8_fast does not decide: out of line, and fed from global memory
// (four arguments), so that the stencil body stays small - the unrolled body with the window passed in registers was
// 45 KB of code and the kernel waited for instructions (stall "no_instruction" 2.2 per issue)
__device__ __noinline__ int d8_slow_at(const double* __restrict__ elev, int NR, int pitch, int r, int c) {
  double nb[10];
#pragma unroll
  for (int n = 1; n <= 9; ++n) {
    const int rr = r + nb_dr(n);
    nb[n] = (rr < 0 || rr >= NR) ? sw_nan() : elev[(size_t)rr * pitch + c + nb_dc(n)];   // column pads are readable
  }
  return d8_slow(nb[5], nb);
}
// Worklist entries: cell id, with bit 31 (LIST_DEAD) set when the flat cell has no neighbour at its own elevation - no
// plateau round can ever give such a cell a direction (96 % of the flat cells of the bench DEM are single-cell pits),
// so the rounds skip it without touching the rasters; the patch stage looks at every entry.
constexpr uint32_t LIST_DEAD = 0x80000000u, LIST_ID = 0x7FFFFFFFu;

// flats of one thread-row that do not fit the block's staging buffer (a very flat block)
__device__ __noinline__ void d8_flats_direct(unsigned fm, uint4 ids, unsigned k, uint32_t* s_ids,
                                             uint32_t* __restrict__ list, unsigned long long cap, DdirCounters* ctr) {
  const uint32_t v[4] = {ids.x, ids.y, ids.z, ids.w};
  for (int j = 0; j < 4; ++j) {
    if (!((fm >> (8 * j)) & 1u)) continue;
    if (k < D8_STAGE) s_ids[k] = v[j];
    else {
      unsigned long long g = atomicAdd(&ctr->nflat, 1ull);
      if (g < cap) list[g] = v[j];
    }
    ++k;
  }
}

// one row of the block: directions of the thread's four cells from the three row windows, stores, flats.
// colmask: 0x01 in byte j where column c0 + j < NC
__device__ __forceinline__ void d8_row(const RowWin<D8_CPT>& p, const RowWin<D8_CPT>& q, const RowWin<D8_CPT>& n,
                                       const double* __restrict__ elev, const GridDims& d, int r, int c0, bool live,
                                       bool own_row, unsigned colmask, size_t i, uint8_t* __restrict__ ddir,
                                       uint8_t* __restrict__ flat, uint32_t* __restrict__ list, unsigned long long cap,
                                       DdirCounters* ctr, uint32_t* s_ids, unsigned* s_n) {
  unsigned od = 0, rare = 0;
#pragma unroll
  for (int j = 0; j < D8_CPT; ++j) {
    double nb[10];
    nb[7] = p.v[j]; nb[8] = p.v[j + 1]; nb[9] = p.v[j + 2];
    nb[4] = q.v[j]; nb[5] = q.v[j + 1]; nb[6] = q.v[j + 2];
    nb[1] = n.v[j]; nb[2] = n.v[j + 1]; nb[3] = n.v[j + 2];
    const int dd = d8_fast(nb[5], nb);
    od |= ((unsigned)dd & 0xFFu) << (8 * j);   // a rare window leaves 0xFF
    rare |= (unsigned)(dd < 0) << j;
  }
  if (rare && live) {   // (a thread beyond the row allocation works on whatever its ring columns hold and stores nothing)
    for (unsigned m = rare; m; m &= m - 1) {
      const int j = __ffs(m) - 1;
      const unsigned dd = (unsigned)d8_slow_at(elev, d.NR, d.pitch, r, c0 + j);
      od = (od & ~(0xFFu << (8 * j))) | (dd << (8 * j));
    }
  }
  // flatcell byte: 2 for an NA cell, 1 for a flat one, else 0; flats inside the grid columns
  const unsigned eq5 = __vcmpeq4(od, 0x05050505u) & 0x01010101u;
  const unsigned of = eq5 | (__vcmpeq4(od, 0u) & 0x02020202u);
  const unsigned fm = eq5 & colmask;
  if (live) {
    *reinterpret_cast<unsigned*>(ddir + i) = od;
    *reinterpret_cast<unsigned*>(flat + i) = of;
    if (c0 == 1) { ddir[i - 1] = 0; flat[i - 1] = 2; }  // ring column 0 (always NA)
  }
  if (fm && own_row) {  // halo rows of a band are not ours
    const unsigned cnt = (unsigned)__popc(fm);
    unsigned k = atomicAdd(s_n, cnt);
    uint32_t ids[D8_CPT];
#pragma unroll
    for (int j = 0; j < D8_CPT; ++j) {
      ids[j] = (uint32_t)i + j;
      if ((fm >> (8 * j)) & 1u) {   // a neighbour at the cell's own elevation? (none of a flat cell's neighbours is lower)
        const double e = q.v[j + 1];
        const bool level = (p.v[j] == e) | (p.v[j + 1] == e) | (p.v[j + 2] == e) | (q.v[j] == e) | (q.v[j + 2] == e) |
                           (n.v[j] == e) | (n.v[j + 1] == e) | (n.v[j + 2] == e);
        if (!level) ids[j] |= LIST_DEAD;
      }
    }
    if (k + cnt <= D8_STAGE) {
#pragma unroll
      for (int j = 0; j < D8_CPT; ++j)
        if ((fm >> (8 * j)) & 1u) s_ids[k++] = ids[j];
    } else d8_flats_direct(fm, make_uint4(ids[0], ids[1], ids[2], ids[3]), k, s_ids, list, cap, ctr);
  }
}

#ifndef D8_MINB
#define D8_MINB 6
#endif
#ifndef D8_STAGES
#define D8_STAGES 4
#endif
// Rows reach the block through the TMA row ring (rowring.cuh, as in k_form): segments of D8_COLS + 4 columns, fetched
// by 1-D bulk copies D8_STAGES - D8_GROUP .. D8_STAGES rows ahead of the arithmetic.  With the rows prefetched
// through registers (one row ahead per warp) the kernel had ~18 KB in flight per SM and ran at the latency, not at
// the issue rate, once the per-cell arithmetic was halved.
constexpr int D8_COLS = D8_THREADS * D8_CPT;
#ifndef D8_GROUP_ROWS
#define D8_GROUP_ROWS 2
#endif
constexpr int D8_GROUP = D8_GROUP_ROWS;      // ring slots recycled per block barrier
using D8Ring = RowRing<double, D8_COLS, D8_STAGES>;

// the thread's window columns of ring row k (grid row r0 - 1 + k): own columns at slot elements 2 + 4t .. 5 + 4t; a row
// outside the grid is read from a slot full of NaN
__device__ __forceinline__ void d8_take_row(RowWin<D8_CPT>& w, D8Ring& ring, const double* nanrow, int k, bool valid) {
  const int s = k % D8_STAGES;
  if (valid) rr_wait(&ring.bar[s], (unsigned)(k / D8_STAGES) & 1u);
  const double* row = (valid ? ring.row[s] : nanrow) + D8_CPT * threadIdx.x;
  const double2 a = *reinterpret_cast<const double2*>(row + 2);
  const double2 b = *reinterpret_cast<const double2*>(row + 4);
  w.v[0] = row[1];
  w.v[1] = a.x; w.v[2] = a.y; w.v[3] = b.x; w.v[4] = b.y;
  w.v[5] = row[6];
}

__global__ void __launch_bounds__(D8_THREADS, D8_MINB) k_d8(const double* __restrict__ elev, GridDims d,
                                                   uint8_t* __restrict__ ddir, uint8_t* __restrict__ flat,
                                                   uint32_t* __restrict__ list, unsigned long long cap,
                                                   int own_lo, int own_hi, DdirCounters* ctr) {
  __shared__ D8Ring ring;
  __shared__ __align__(16) double s_nan[D8Ring::SEG];
  __shared__ uint32_t s_ids[D8_STAGE];
  __shared__ unsigned s_n;
  __shared__ unsigned long long s_base;
  const int cb = blockIdx.x * D8_COLS + 1;                // first column of the block (column 1 is the aligned one)
  const int c0 = cb + threadIdx.x * D8_CPT;
  const int r0 = blockIdx.y * D8_ROWS;
  const int rend = min(r0 + D8_ROWS, d.NR);
  const bool live = c0 < d.pitch - LITAP_XOFF;  // the thread's columns lie inside the row allocation
  // ring row k = grid row r0 - 1 + k, k = 0 .. kmax; the segment starts at column cb - 2 and must stay inside
  // pitch + 32 elements of its row
  const int kmax = rend - r0 + 1;
  const unsigned seg_bytes = (unsigned)min(D8Ring::SEG, ((d.pitch + 32 - (cb - 2)) / 2) * 2) * sizeof(double);
  int next_k = 0;  // thread 0: first ring row not yet requested
  if (threadIdx.x == 0) {
    for (int s = 0; s < D8_STAGES; ++s) rr_bar_init(&ring.bar[s]);
    rr_fence_init();
    s_n = 0;
    for (; next_k < D8_STAGES && next_k <= kmax; ++next_k) {
      const int rr = r0 - 1 + next_k;
      if (rr >= 0 && rr < d.NR)
        rr_issue(ring.row[next_k % D8_STAGES], elev + (size_t)rr * d.pitch + (cb - 2), seg_bytes, &ring.bar[next_k % D8_STAGES]);
      else rr_skip(&ring.bar[next_k % D8_STAGES]);
    }
  }
  for (int k = threadIdx.x; k < D8Ring::SEG; k += D8_THREADS) s_nan[k] = sw_nan();
  __syncthreads();
  unsigned colmask = 0;
#pragma unroll
  for (int j = 0; j < D8_CPT; ++j) colmask |= (unsigned)(c0 + j < d.NC) << (8 * j);
  RowWin<D8_CPT> p, q, n;  // rows r-1, r, r+1
  d8_take_row(p, ring, s_nan, 0, r0 - 1 >= 0);
  d8_take_row(q, ring, s_nan, 1, true);
  size_t i = (size_t)r0 * d.pitch + c0;
  for (int r = r0; r < rend; ++r) {
    const int j = r - r0;
    if (j > 0 && (j % D8_GROUP) == 0) {
      __syncthreads();  // every thread has taken ring rows <= j + 1: their slots are free
      if (threadIdx.x == 0) {
        for (; next_k <= kmax && next_k - D8_STAGES <= j + 1; ++next_k) {
          const int rk = r0 - 1 + next_k;
          if (rk < d.NR)
            rr_issue(ring.row[next_k % D8_STAGES], elev + (size_t)rk * d.pitch + (cb - 2), seg_bytes,
                     &ring.bar[next_k % D8_STAGES]);
          else rr_skip(&ring.bar[next_k % D8_STAGES]);
        }
      }
    }
    d8_take_row(n, ring, s_nan, j + 2, r + 1 < d.NR);
    d8_row(p, q, n, elev, d, r, c0, live, r >= own_lo && r < own_hi, colmask, i, ddir, flat, list, cap, ctr, s_ids, &s_n);
    i += d.pitch;
    p = q;
    q = n;
  }
  __syncthreads();
  const unsigned ns = min(s_n, (unsigned)D8_STAGE);
  if (threadIdx.x == 0 && ns) s_base = atomicAdd(&ctr->nflat, (unsigned long long)ns);
  __syncthreads();
  for (unsigned k = threadIdx.x; k < ns; k += D8_THREADS)
    if (s_base + k < cap) list[s_base + k] = s_ids[k];
}

// ---- plateau rounds (Jacobi: probe reads the previous round's ddir, apply writes) ---------------
// R/flow_01_calc_ddir.R:17-41: a flat cell takes the lowest direction n whose neighbour has the
// same elevation (elev - elev_n >= 0 can only be equality: no neighbour is lower) and already drains.
// ddir is re-read across rounds of one launch while other SMs write it: L2 loads (ld.cg), never L1 / nc.
constexpr uint8_t PLATEAU_DEAD = 0x80;   // probe result / candidate byte: the worklist entry needs no further probe
__device__ __forceinline__ uint8_t plateau_probe_cell(uint32_t id, const uint8_t* ddir,
                                                      const double* __restrict__ elev, const GridDims& d) {
  if (__ldcg(ddir + id) != 5) return 0;
  const int r = id / d.pitch, c = id - r * d.pitch;
  // the cell and its eight neighbours in ONE round trip: sixteen independent loads (a neighbour outside the grid reads
  // the cell itself, which is flat and so never qualifies), then the lowest code that qualifies.  With the
  // neighbours visited one after the other and an early exit, a probe was a chain of up to sixteen dependent loads,
  // and the whole stage (three passes over 1 % of the cells) cost 0.23 ms at 8192^2.
  const double e = elev[id];
  double en[10];
  uint8_t dn[10];
#pragma unroll
  for (int k = 1; k <= 9; ++k) {
    if (k == 5) continue;
    const int rr = r + nb_dr(k), cc = c + nb_dc(k);
    const bool in = rr >= 0 && rr < d.NR && cc >= 0 && cc < d.NC;
    const size_t j = in ? (size_t)rr * d.pitch + cc : (size_t)id;
    en[k] = elev[j];
    dn[k] = __ldcg(ddir + j);
  }
  uint8_t res = 0;
  bool level = false;   // a neighbour at the cell's own elevation exists (no neighbour of a flat cell is lower)
#pragma unroll
  for (int k = 9; k >= 1; --k) {
    if (k == 5) continue;
    const bool eq = !is_nan(en[k]) && (e - en[k] >= 0);
    level |= eq;
    if (eq && dn[k] != 5 && dn[k] != 0) res = (uint8_t)k;
  }
  // PLATEAU_DEAD: no neighbour at the cell's elevation, so no round can ever give it a direction (a single-cell pit:
  // 96 % of the flat cells of the bench DEM) - the caller may drop it from the later rounds
  return level ? res : PLATEAU_DEAD;
}

// all rounds in one cooperative launch (single-device path): no host round trip per round
__global__ void __launch_bounds__(256) k_plateau_all(const uint32_t* __restrict__ list, unsigned long long cap,
                                                     uint8_t* ddir, const double* __restrict__ elev, GridDims d,
                                                     uint8_t* __restrict__ cand, DdirCounters* ctr) {
  cg::grid_group grid = cg::this_grid();
  const unsigned long long nf = ctr->nflat;
  const unsigned n = (unsigned)(nf < cap ? nf : cap);
  const unsigned stride = gridDim.x * blockDim.x, t0 = blockIdx.x * blockDim.x + threadIdx.x;
  for (unsigned round = 0;; ++round) {
    unsigned* flag = &ctr->changed[round & 1];
    bool any = false;
    // every pass re-reads only the candidate byte (coalesced) of an entry that is settled: single-cell pits after the
    // first probe, resolved cells once their direction is stored.  Each pass used to pull ~10 sectors per flat cell
    // from DRAM, three times over, for the 4 % of the entries that can change at all.
    for (unsigned i = t0; i < n; i += stride) {
      if (round && cand[i] == PLATEAU_DEAD) continue;
      const uint32_t v = list[i];
      uint8_t res = (v & LIST_DEAD) ? PLATEAU_DEAD : plateau_probe_cell(v & LIST_ID, ddir, elev, d);
      cand[i] = res;
      any |= (res & 15) != 0;
    }
    if (__syncthreads_or(any) && threadIdx.x == 0) atomicOr(flag, 1u);
    if (t0 == 0) ctr->changed[(round + 1) & 1] = 0;  // next round's flag (nobody reads it before the sync)
    grid.sync();
    if (!*reinterpret_cast<volatile unsigned*>(flag)) break;
    for (unsigned i = t0; i < n; i += stride) {
      const uint8_t cd = cand[i];
      if (cd & 15) { ddir[list[i] & LIST_ID] = cd; cand[i] = PLATEAU_DEAD; }
    }
    if (t0 == 0) ctr->rounds = round + 1;
    grid.sync();
  }
}

// one round as two launches (row-band mode)
__global__ void k_plateau_probe(const uint32_t* __restrict__ list, unsigned n, const uint8_t* ddir,
                                const double* __restrict__ elev, GridDims d, uint8_t* __restrict__ cand,
                                unsigned* changed) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  bool any = false;
  if (i < n) {
    const uint32_t v = list[i];   // (a band's halo rows change between rounds, its elevations do not: dead stays dead)
    uint8_t res = (v & LIST_DEAD) ? 0 : plateau_probe_cell(v & LIST_ID, ddir, elev, d) & 15;
    cand[i] = res;
    any = res != 0;
  }
  if (__syncthreads_or(any) && threadIdx.x == 0) atomicOr(changed, 1u);
}
__global__ void k_plateau_apply(const uint32_t* __restrict__ list, unsigned n, const uint8_t* __restrict__ cand,
                                uint8_t* __restrict__ ddir) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && cand[i]) ddir[list[i] & LIST_ID] = cand[i];
}

// ---- flat-pit patches ------------------------------------------------------------------------------
// A patch cell is a cell that is still ddir == 5 after the plateau rounds and has an 8-neighbour that
// is also still 5 (R/flow_01_calc_ddir.R:48-57).  Patch cells travel as 64-bit keys (row << 32 | col),
// rows counted in the whole DEM (row_off = first row of a band), so the resolver below is the same
// for one device and for the gathered patch cells of all row bands.
__device__ __forceinline__ unsigned long long pkey(long long r, long long c) {
  return ((unsigned long long)r << 32) | (unsigned long long)(uint32_t)c;
}

__global__ void k_patch_mark(const uint32_t* __restrict__ list, unsigned long long cap, const uint8_t* __restrict__ ddir,
                             GridDims d, long long row_off, unsigned long long* __restrict__ keys, DdirCounters* ctr) {
  const unsigned long long nf = ctr->nflat;
  const unsigned n = (unsigned)(nf < cap ? nf : cap);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    uint32_t id = list[i] & LIST_ID;
    if (ddir[id] != 5) continue;
    const int r = id / d.pitch, c = id - r * d.pitch;
    bool hit = false;
#pragma unroll
    for (int k = 1; k <= 9; ++k) {
      if (k == 5) continue;
      int rr = r + nb_dr(k), cc = c + nb_dc(k);
      if (rr < 0 || rr >= d.NR || cc < 0 || cc >= d.NC) continue;
      hit |= ddir[(size_t)rr * d.pitch + cc] == 5;
    }
    if (hit) keys[atomicAdd(&ctr->npatch, 1u)] = pkey(row_off + r, c);
  }
}

// index of `key` in the sorted array K[0..n), or -1
__device__ __forceinline__ int pfind(const unsigned long long* __restrict__ K, unsigned n, unsigned long long key) {
  unsigned lo = 0, hi = n;
  while (lo < hi) {
    unsigned mid = (lo + hi) >> 1;
    if (K[mid] < key) lo = mid + 1; else hi = mid;
  }
  return (lo < n && K[lo] == key) ? (int)lo : -1;
}

// lock-free union-find over list positions; the larger root is hooked under the smaller one, so the
// root of a component is its first cell in (row, col) order = lowest seqno
__device__ __forceinline__ uint32_t uf_find(uint32_t* parent, uint32_t x) {
  uint32_t p = *reinterpret_cast<volatile uint32_t*>(parent + x);
  while (p != x) {
    uint32_t gp = *reinterpret_cast<volatile uint32_t*>(parent + p);
    if (gp != p) parent[x] = gp;  // path halving (x is not a root and never becomes one again)
    x = p;
    p = gp;
  }
  return x;
}
__device__ __forceinline__ void uf_unite(uint32_t* parent, uint32_t a, uint32_t b) {
  while (true) {
    a = uf_find(parent, a);
    b = uf_find(parent, b);
    if (a == b) return;
    if (a < b) { uint32_t t = a; a = b; b = t; }
    if (atomicCAS(parent + a, a, b) == a) return;
  }
}

__global__ void k_patch_init(unsigned n, uint32_t* parent, unsigned long long* bestd, uint32_t* bestid) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { parent[i] = i; bestd[i] = ~0ull; bestid[i] = 0xFFFFFFFFu; }
}
__global__ void k_patch_union(const unsigned long long* __restrict__ K, unsigned n, uint32_t* parent,
                              uint16_t* __restrict__ opts) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long r = (long long)(K[i] >> 32), c = (long long)(uint32_t)K[i];
  unsigned o = 0;
#pragma unroll
  for (int k = 1; k <= 9; ++k) {
    if (k == 5) continue;
    long long rr = r + nb_dr(k), cc = c + nb_dc(k);
    if (rr < 0 || cc < 0) continue;
    int j = pfind(K, n, pkey(rr, cc));
    if (j < 0) continue;
    o |= 1u << k;
    if (k == 6 || k <= 3) uf_unite(parent, i, (uint32_t)j);  // E, SW, S, SE: every adjacency once
  }
  opts[i] = (uint16_t)o;
}
__global__ void k_patch_flatten(const unsigned long long* __restrict__ K, unsigned n, uint32_t* parent,
                                unsigned long long* __restrict__ kc, unsigned long long* __restrict__ kr) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t root = uf_find(parent, i);
  kc[i] = ((unsigned long long)root << 32) | (uint32_t)K[i];          // (component, col)
  kr[i] = ((unsigned long long)root << 32) | (uint32_t)(K[i] >> 32);  // (component, row); rows < 2^32
}
// ceiling(median(.)) of the cols and of the rows of every component (R/flow_01_calc_ddir.R:78)
__global__ void k_patch_median(const unsigned long long* __restrict__ kc, const unsigned long long* __restrict__ kr,
                               unsigned n, uint32_t* __restrict__ medc, uint32_t* __restrict__ medr) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t root = (uint32_t)(kc[i] >> 32);
  if (i > 0 && (uint32_t)(kc[i - 1] >> 32) == root) return;  // not the head of its segment
  unsigned lo = i, hi = n;  // first index whose component is > root
  while (lo < hi) {
    unsigned mid = (lo + hi) >> 1;
    if ((uint32_t)(kc[mid] >> 32) <= root) lo = mid + 1; else hi = mid;
  }
  const unsigned m = lo - i, a = i + (m - 1) / 2, b = i + m / 2;
  // median of an even count = mean of the middle two; coordinates are non-negative integers
  medc[root] = (uint32_t)(((unsigned long long)(uint32_t)kc[a] + (uint32_t)kc[b] + 1) / 2);
  medr[root] = (uint32_t)(((unsigned long long)(uint32_t)kr[a] + (uint32_t)kr[b] + 1) / 2);
}
__device__ __forceinline__ unsigned long long pdist2(unsigned long long key, uint32_t cc, uint32_t cr) {
  long long dc = (long long)(uint32_t)key - (long long)cc, dr = (long long)(key >> 32) - (long long)cr;
  return (unsigned long long)(dc * dc + dr * dr);
}
// centre cell = smallest Euclidean distance to (medc, medr), first in seqno order (R/flow_01_calc_ddir.R:80-86)
__global__ void k_patch_best(const unsigned long long* __restrict__ K, unsigned n, uint32_t* parent,
                             const uint32_t* __restrict__ medc, const uint32_t* __restrict__ medr,
                             unsigned long long* bestd, uint32_t* bestid, int pass) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t root = uf_find(parent, i);
  const unsigned long long d2 = pdist2(K[i], medc[root], medr[root]);
  if (pass == 0) atomicMin(bestd + root, d2);
  else if (d2 == bestd[root]) atomicMin(bestid + root, i);
}
__global__ void k_patch_dirs(const unsigned long long* __restrict__ K, unsigned n, uint32_t* parent,
                             const uint32_t* __restrict__ bestid, const uint16_t* __restrict__ opts,
                             uint8_t* __restrict__ newdir) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long kf = K[bestid[uf_find(parent, i)]];
  newdir[i] = (uint8_t)get_dir_int((long long)(K[i] >> 32) + 1, (long long)(uint32_t)K[i] + 1,
                                   (long long)(kf >> 32) + 1, (long long)(uint32_t)kf + 1, opts[i]);
}
__global__ void k_patch_apply(const unsigned long long* __restrict__ K, const uint8_t* __restrict__ newdir, unsigned n,
                              uint8_t* __restrict__ ddir, GridDims d, long long row_off, int own_lo, int own_hi) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long r = (long long)(K[i] >> 32) - row_off;
  if (r < own_lo || r >= own_hi) return;  // another band's cell
  ddir[(size_t)r * d.pitch + (uint32_t)K[i]] = newdir[i];
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// host side
namespace {
inline unsigned nblk(size_t n, unsigned t = 256) { return (unsigned)((n + t - 1) / t); }

// per-grid worklists of calc_ddir2 (kept on the grid between the steps of the row-band protocol)
int ddir_ensure(litap_grid* g, size_t cap) {
  const GridDims d = g->d;
  if (!g->ddir) g->ddir = grid_alloc_array<uint8_t>(d, 0);
  if (!g->flat) g->flat = grid_alloc_array<uint8_t>(d, 2);
  if (!g->ddir || !g->flat) return set_error(LITAP_ENOMEM, "device allocation failed (ddir)");
  if (!g->dd_ctr) g->dd_ctr = pool_alloc(sizeof(DdirCounters));
  if (!g->dd_ctr) return set_error(LITAP_ENOMEM, "device allocation failed (ddir counters)");
  if (cap > g->dd_cap) {
    if (g->dd_list) pool_free(g->dd_list);
    if (g->dd_cand) pool_free(g->dd_cand);
    if (g->dd_keys) pool_free(g->dd_keys);
    g->dd_list = (uint32_t*)pool_alloc(cap * sizeof(uint32_t));
    g->dd_cand = (uint8_t*)pool_alloc(cap);
    g->dd_keys = (unsigned long long*)pool_alloc(cap * sizeof(unsigned long long));
    g->dd_cap = (g->dd_list && g->dd_cand && g->dd_keys) ? cap : 0;
    if (!g->dd_cap) return set_error(LITAP_ENOMEM, "device allocation failed (flat worklist)");
  }
  return LITAP_OK;
}
inline size_t ddir_default_cap(const GridDims& d) { return (size_t)d.cells / 8 + 4096; }

int launch_d8(litap_grid* g) {
  const GridDims d = g->d;
  dim3 grid(nblk((size_t)(d.pitch - LITAP_XOFF + D8_CPT - 1) / D8_CPT, D8_THREADS), (d.NR + D8_ROWS - 1) / D8_ROWS);
  KERNEL("k_d8", k_d8<<<grid, D8_THREADS, 0, ctx().stream>>>(g->elev, d, g->ddir, g->flat, g->dd_list,
                                                              (unsigned long long)g->dd_cap, g->own_lo, g->own_hi,
                                                              (DdirCounters*)g->dd_ctr));
  return LITAP_OK;
}
}  // namespace

// Resolve flat-pit patches given the UNSORTED keys of all patch cells (device array, np entries);
// on return *K_out / *dirs_out (scratch-owned device arrays) hold the sorted keys and the new
// direction of every patch cell.
int ddir_patch_resolve(const unsigned long long* keys, unsigned np, PoolScratch& S, unsigned long long** K_out,
                       uint8_t** dirs_out) {
  cudaStream_t st = ctx().stream;
  unsigned long long* K = S.get<unsigned long long>(np);
  unsigned long long* kc = S.get<unsigned long long>(np);
  unsigned long long* kr = S.get<unsigned long long>(np);
  unsigned long long* kc2 = S.get<unsigned long long>(np);
  unsigned long long* kr2 = S.get<unsigned long long>(np);
  unsigned long long* bestd = S.get<unsigned long long>(np);
  uint32_t* parent = S.get<uint32_t>(np);
  uint32_t* bestid = S.get<uint32_t>(np);
  uint32_t* medc = S.get<uint32_t>(np);
  uint32_t* medr = S.get<uint32_t>(np);
  uint16_t* opts = S.get<uint16_t>(np);
  uint8_t* newdir = S.get<uint8_t>(np);
  size_t tmpn = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, tmpn, keys, K, (int)np, 0, 64, st);
  void* tmp = S.get<unsigned char>(tmpn + 256);
  if (!K || !kc || !kr || !kc2 || !kr2 || !bestd || !parent || !bestid || !medc || !medr || !opts || !newdir || !tmp)
    return set_error(LITAP_ENOMEM, "device allocation failed (patch scratch)");
  const unsigned nb = nblk(np);
  // key = (global row << 32 | col), rows < 2^24; (component < np) << 32 | row or col below: skip the empty top digits
  int cbits = 1;
  while ((1ull << cbits) < (unsigned long long)np) ++cbits;
  CUDA_TRY(cub::DeviceRadixSort::SortKeys(tmp, tmpn, keys, K, (int)np, 0, 56, st));
  KERNEL("k_patch_init", k_patch_init<<<nb, 256, 0, st>>>(np, parent, bestd, bestid));
  KERNEL("k_patch_union", k_patch_union<<<nb, 256, 0, st>>>(K, np, parent, opts));
  KERNEL("k_patch_flatten", k_patch_flatten<<<nb, 256, 0, st>>>(K, np, parent, kc, kr));
  CUDA_TRY(cub::DeviceRadixSort::SortKeys(tmp, tmpn, kc, kc2, (int)np, 0, 32 + cbits, st));
  CUDA_TRY(cub::DeviceRadixSort::SortKeys(tmp, tmpn, kr, kr2, (int)np, 0, 32 + cbits, st));
  KERNEL("k_patch_median", k_patch_median<<<nb, 256, 0, st>>>(kc2, kr2, np, medc, medr));
  KERNEL("k_patch_best", k_patch_best<<<nb, 256, 0, st>>>(K, np, parent, medc, medr, bestd, bestid, 0));
  KERNEL("k_patch_best", k_patch_best<<<nb, 256, 0, st>>>(K, np, parent, medc, medr, bestd, bestid, 1));
  KERNEL("k_patch_dirs", k_patch_dirs<<<nb, 256, 0, st>>>(K, np, parent, bestid, opts, newdir));
  ctx().launches += 3;  // cub sorts (at least)
  *K_out = K;
  *dirs_out = newdir;
  return LITAP_OK;
}

int run_ddir(litap_grid* g, int invert) {
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  int rc = ddir_ensure(g, g->dd_cap ? g->dd_cap : ddir_default_cap(d));
  if (rc) return rc;
  DdirCounters* ctr = (DdirCounters*)g->dd_ctr;
  if (invert) {
    CUDA_TRY(cudaMemsetAsync(ctr, 0, sizeof(DdirCounters), st));
    dim3 gb(nblk(d.NC) < 64 ? nblk(d.NC) : 64, d.NR < 1024 ? d.NR : 1024);
    KERNEL("k_elev_max", k_elev_max<<<gb, 256, 0, st>>>(g->elev, d, 0, d.NR, &ctr->elev_max));
    dim3 gi(nblk((size_t)d.NC + 31), d.NR);
    KERNEL("k_invert", k_invert<<<gi, 256, 0, st>>>(g->elev, d, &ctr->elev_max));
  }
  static int coop_blocks = 0;
  if (!coop_blocks) {
    int per_sm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_plateau_all, 256, 0));
    if (per_sm < 1) return set_error(LITAP_ECUDA, "k_plateau_all cannot be made resident");
    int want = 4;   // blocks per SM: enough threads for the worklist, few enough to start beside calc_form's blocks
    if (const char* e = getenv("LITAP_PLATEAU_BPS")) want = atoi(e) > 0 ? atoi(e) : want;
    coop_blocks = c.sm_count * (per_sm < want ? per_sm : want);
  }
  DdirCounters h;
  for (int attempt = 0;; ++attempt) {
    CUDA_TRY(cudaMemsetAsync(ctr, 0, sizeof(DdirCounters), st));
    rc = launch_d8(g);
    if (rc) return rc;
    {
      unsigned long long cap = g->dd_cap;
      GridDims dd = d;
      void* args[] = {(void*)&g->dd_list, (void*)&cap, (void*)&g->ddir, (void*)&g->elev, (void*)&dd,
                      (void*)&g->dd_cand, (void*)&ctr};
      ProfScope prof("k_plateau_all");
      CUDA_TRY(cudaLaunchCooperativeKernel((const void*)k_plateau_all, dim3(coop_blocks), dim3(256), args, 0, st));
      c.launches++;
    }
    KERNEL("k_patch_mark", k_patch_mark<<<c.sm_count * 4, 256, 0, st>>>(g->dd_list, (unsigned long long)g->dd_cap, g->ddir,
                                                                       d, g->row_off, g->dd_keys, ctr));
    CUDA_TRY(cudaMemcpyAsync(&h, ctr, sizeof(h), cudaMemcpyDeviceToHost, st));
    STREAM_SYNC(st);
    if (h.nflat <= g->dd_cap) break;
    if (attempt) return set_error(LITAP_ECUDA, "internal: flat worklist overflow after regrowth");
    rc = ddir_ensure(g, (size_t)h.nflat + 4096);  // unusually flat DEM: regrow the worklist and redo
    if (rc) return rc;
  }
  c.tm.flat_cells = (int64_t)h.nflat;
  g->pit_bound = (int64_t)h.nflat;  // pits are cells that stay flat: calc_shed4 sizes its pit arrays from this
  c.tm.plateau_rounds = (int)h.rounds;
  c.tm.patch_cells = h.npatch;
  if (h.npatch) {
    PoolScratch S;
    unsigned long long* K = nullptr;
    uint8_t* nd = nullptr;
    rc = ddir_patch_resolve(g->dd_keys, h.npatch, S, &K, &nd);
    if (rc) return rc;
    KERNEL("k_patch_apply", k_patch_apply<<<nblk(h.npatch), 256, 0, st>>>(K, nd, h.npatch, g->ddir, d, g->row_off,
                                                                         g->own_lo, g->own_hi));
    STREAM_SYNC(st);  // scratch goes back to the pool
  }
  return LITAP_OK;
}

// ------------------------------------------------------------------------------------------------
// steps of calc_ddir2 for the row-band driver (bands.cu): same kernels, with the exchanges between
// bands done by the caller between the steps
int ddir_band_grow(litap_grid* g, size_t cap) { return ddir_ensure(g, cap); }

int ddir_band_begin(litap_grid* g) {
  int rc = ddir_ensure(g, g->dd_cap ? g->dd_cap : ddir_default_cap(g->d));
  if (rc) return rc;
  CUDA_TRY(cudaMemsetAsync(g->dd_ctr, 0, sizeof(DdirCounters), ctx().stream));
  return launch_d8(g);
}
int ddir_band_counts(litap_grid* g, unsigned long long* nflat, unsigned* npatch) {
  DdirCounters h;
  CUDA_TRY(cudaMemcpyAsync(&h, g->dd_ctr, sizeof(h), cudaMemcpyDeviceToHost, ctx().stream));
  STREAM_SYNC(ctx().stream);
  if (nflat) *nflat = h.nflat;
  if (npatch) *npatch = h.npatch;
  return LITAP_OK;
}
int ddir_band_probe(litap_grid* g, unsigned nflat, unsigned* changed_dev) {
  if (!nflat) return LITAP_OK;
  KERNEL("k_plateau_probe", k_plateau_probe<<<nblk(nflat), 256, 0, ctx().stream>>>(g->dd_list, nflat, g->ddir, g->elev, g->d,
                                                                                 g->dd_cand, changed_dev));
  return LITAP_OK;
}
int ddir_band_apply(litap_grid* g, unsigned nflat) {
  if (!nflat) return LITAP_OK;
  KERNEL("k_plateau_apply", k_plateau_apply<<<nblk(nflat), 256, 0, ctx().stream>>>(g->dd_list, nflat, g->dd_cand, g->ddir));
  return LITAP_OK;
}
int ddir_band_mark(litap_grid* g) {
  KERNEL("k_patch_mark", k_patch_mark<<<ctx().sm_count * 4, 256, 0, ctx().stream>>>(
                             g->dd_list, (unsigned long long)g->dd_cap, g->ddir, g->d, g->row_off, g->dd_keys,
                             (DdirCounters*)g->dd_ctr));
  return LITAP_OK;
}
int ddir_band_patch_apply(litap_grid* g, const unsigned long long* K, const uint8_t* nd, unsigned np) {
  KERNEL("k_patch_apply", k_patch_apply<<<nblk(np), 256, 0, ctx().stream>>>(K, nd, np, g->ddir, g->d, g->row_off, g->own_lo,
                                                                          g->own_hi));
  return LITAP_OK;
}
int ddir_band_elev_max(litap_grid* g, unsigned long long* out_dev) {
  const GridDims d = g->d;
  dim3 gb(nblk(d.NC) < 64 ? nblk(d.NC) : 64, d.NR < 1024 ? d.NR : 1024);
  KERNEL("k_elev_max", k_elev_max<<<gb, 256, 0, ctx().stream>>>(g->elev, d, g->own_lo, g->own_hi, out_dev));
  return LITAP_OK;
}
int ddir_band_invert(litap_grid* g, const unsigned long long* max_dev) {
  const GridDims d = g->d;
  dim3 gi(nblk((size_t)d.NC + 31), d.NR);
  KERNEL("k_invert", k_invert<<<gi, 256, 0, ctx().stream>>>(g->elev, d, max_dev));
  return LITAP_OK;
}
