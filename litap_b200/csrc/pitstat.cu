// pitstat.cu — pit_stat1 on the device (second half): the pour point of every watershed — the lowest
// place on its rim where it tips into a neighbouring shed — and the statistics of the pit below it
// (R/flow_03_pit_stat_fast.R:1-162; called by flow_mapper at R/flow_mapper.R:181,222,378,416 and by every round of pit
// removal, R/flow_04_remove_pit_fast.R:19,44,90,136,240,270).
//
//   k_ps_cells     per cell: shed present / shed area; per RIDGE cell the one row of the reference's `pp` table that
//                  survives its per-cell selection (method "litap": the lowest neighbour in another shed, first in
//                  neighbour order; "landmapr": the last neighbour, in LandMapR's 7 8 9 4 6 1 2 3 order, that lies below
//                  the running minimum of the pour elevations before it), and a 64-bit atomicMin of the pour
//                  elevation's order-preserving image into the shed's slot                                    (:24-63)
//   k_ps_tied      candidates that attain their shed's minimum -> a compact list (a few per shed)
//   3 x cub sort   arrange(shedno, elev, upslope, seqno) as stable LSD passes over that list                   (:67)
//   k_ps_select    first row of every shed = its pour point; the other rows are LandMapR's "omits"            (:73-78)
//   cub sort       the cells at or below their shed's pour elevation, grouped by shed in seqno order
//   k_ps_pits      per shed: pit area, edge flag, pit cell, and pit_vol = sum(pour_elev - elev) with R's long
//                  double accumulator in the reference's row order (cellmath.h ld80_*)                         (:93-111)
//   k_ps_rows      the table, one row per shed that has a pour point, in shedno order                          (:112-123)
// A DEM with a single shed takes the reference's "only one watershed" branch (:127-151).
#include <cub/cub.cuh>
#include <algorithm>
#include <cstring>
#include "cellmath.h"
#include "litap_internal.h"

namespace pitstat {

constexpr int NCOL = 21;
constexpr int PT = 256;
enum { C_SHEDNO, C_SHED_AREA, C_EDGE_PIT, C_PIT_AREA, C_PIT_VOL, C_PIT_ELEV, C_PIT_SEQNO, C_PIT_ROW, C_PIT_COL, C_PRE_VOL,
       C_VARATIO, C_IN_SEQNO, C_IN_ROW, C_IN_COL, C_IN_ELEV, C_OUT_SEQNO, C_OUT_ROW, C_OUT_COL, C_OUT_ELEV, C_DRAINS_TO,
       C_POUR_ELEV };
const char* const NAMES[NCOL] = {"shedno", "shed_area", "edge_pit", "pit_area", "pit_vol", "pit_elev", "pit_seqno", "pit_row",
                                 "pit_col", "pre_vol", "varatio", "in_seqno", "in_row", "in_col", "in_elev", "out_seqno",
                                 "out_row", "out_col", "out_elev", "drains_to", "pour_elev"};

struct In {
  const double* elev;
  const uint8_t* ddir;
  const int32_t* shed;
  const uint32_t* upslope;
  const uint8_t* ridge;
  const uint8_t* edge;
  GridDims d;
};
struct Sheds {                   // indexed by shed number 0 .. maxid
  unsigned long long* minpour;   // image of the lowest pour elevation (all ones = none)
  uint32_t* area;                // cells of the shed
  uint32_t* sel_cell;            // focal cell of the pour point
  uint8_t* sel_n;                // neighbour code of its out cell
  uint32_t* omit_cell;
  double* omit_elev;
  uint32_t* pit_area;
  uint32_t* pit_cell;
  uint8_t* edge_pit;
  double* pit_vol;
  uint32_t* row;                 // row of the output table (exclusive scan of "has a pour point")
};

__device__ __forceinline__ unsigned long long f64_orderable(double x) {
  unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double f64_from_orderable(unsigned long long o) {
  unsigned long long b = (o & 0x8000000000000000ull) ? (o & 0x7FFFFFFFFFFFFFFFull) : ~o;
  return __longlong_as_double((long long)b);
}
__device__ __forceinline__ int nb_off(int n, int pitch) { return nb_dr(n) * pitch + nb_dc(n); }

struct NonZero {
  __host__ __device__ uint32_t operator()(uint32_t a) const { return a ? 1u : 0u; }
};

__global__ void __launch_bounds__(PT) k_ps_maxid(In I, unsigned* out /* [0] max shed number */) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  unsigned m = 0;
  if (c < I.d.NC) {
    const int32_t s = I.shed[(size_t)r * I.d.pitch + c];
    if (s > 0) m = (unsigned)s;
  }
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

// the per-cell selection; cand_n = neighbour code of the chosen row (0 = the cell contributes no row)
__global__ void __launch_bounds__(PT) k_ps_cells(In I, Sheds S, int method, uint8_t* __restrict__ cand_n) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= I.d.NC) return;
  const size_t id = (size_t)r * I.d.pitch + c;
  const int32_t s = I.shed[id];
  uint8_t pick = 0;
  if (s > 0) {
    atomicAdd(&S.area[s], 1u);
    if (I.ridge[id] == 1 && r > 0 && c > 0 && r < I.d.NR - 1 && c < I.d.NC - 1) {
      const double e = I.elev[id];
      double en[10];
      unsigned foreign = 0;
#pragma unroll
      for (int n = 1; n <= 9; ++n) {
        en[n] = 0;
        if (n == 5) continue;
        const size_t j = id + nb_off(n, I.d.pitch);
        const int32_t sn = I.shed[j];
        if (sn > 0 && sn != s) { foreign |= 1u << n; en[n] = I.elev[j]; }
      }
      if (foreign) {
        double pour = 0;
        if (method == 0) {          // "litap": elev_n == min(elev_n), first row = lowest neighbour code (:46-48)
          double mn = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
          for (int n = 1; n <= 9; ++n)
            if (((foreign >> n) & 1) && en[n] < mn) mn = en[n];
#pragma unroll
          for (int n = 9; n >= 1; --n)
            if (((foreign >> n) & 1) && en[n] == mn) pick = (uint8_t)n;
          pour = fmax(e, mn);
        } else {                    // "landmapr" (:53-63)
          const int order[8] = {7, 8, 9, 4, 6, 1, 2, 3};
          double cmin = __longlong_as_double(0x7ff0000000000000LL);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int n = order[k];
            if (!((foreign >> n) & 1)) continue;
            const double p = fmax(e, en[n]);
            if (en[n] < cmin) { pick = (uint8_t)n; pour = p; }   // below everything before it: kept, and the last kept wins
            cmin = fmin(cmin, p);
          }
        }
        if (pick) atomicMin(&S.minpour[s], f64_orderable(pour + 0.0));   // -0 and +0 are one elevation
      }
    }
  }
  cand_n[id] = pick;
}

__device__ __forceinline__ double cand_pour(const In& I, size_t id, int n) {
  return fmax(I.elev[id], I.elev[id + nb_off(n, I.d.pitch)]);
}

// candidates at their shed's lowest pour elevation: (shed, elev image, upslope << 32 | cell)
__global__ void __launch_bounds__(PT) k_ps_tied(In I, Sheds S, const uint8_t* __restrict__ cand_n, unsigned* cursor, unsigned cap,
                                                uint32_t* __restrict__ t_shed, unsigned long long* __restrict__ t_elev,
                                                unsigned long long* __restrict__ t_key) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= I.d.NC) return;
  const size_t id = (size_t)r * I.d.pitch + c;
  const int n = cand_n[id];
  if (!n) return;
  const int32_t s = I.shed[id];
  if (f64_orderable(cand_pour(I, id, n) + 0.0) != S.minpour[s]) return;
  const unsigned at = atomicAdd(cursor, 1u);
  if (at >= cap) return;
  t_shed[at] = (uint32_t)s;
  t_elev[at] = f64_orderable(I.elev[id] + 0.0);
  t_key[at] = ((unsigned long long)I.upslope[id] << 32) | (unsigned long long)id;
}
__global__ void k_ps_iota(uint32_t* a, unsigned n) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = i;
}
template <typename T>
__global__ void k_ps_gather(const T* __restrict__ src, const uint32_t* __restrict__ idx, T* __restrict__ dst, unsigned n) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[idx[i]];
}
// sorted tied list (perm = position -> original entry): the first row of a shed is its pour point, the rest are "omits"
__global__ void __launch_bounds__(PT) k_ps_select(In I, Sheds S, const uint8_t* __restrict__ cand_n, const uint32_t* __restrict__ s_shed,
                                                  const uint32_t* __restrict__ perm, const unsigned long long* __restrict__ t_key,
                                                  unsigned n) {
  const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t s = s_shed[i];
  if (i && s_shed[i - 1] == s) return;
  const uint32_t cell = (uint32_t)(t_key[perm[i]] & 0xFFFFFFFFull);
  S.sel_cell[s] = cell;
  S.sel_n[s] = cand_n[cell];
  unsigned k = 0;
  Lds sacc;
  sacc.m = 0; sacc.e = 0; sacc.neg = 0;
  for (unsigned j = i + 1; j < n && s_shed[j] == s; ++j) {   // sum(elev[-1]) in the arranged order, long double (:75-76)
    const uint32_t cj = (uint32_t)(t_key[perm[j]] & 0xFFFFFFFFull);
    sacc = lds_add(sacc, lds_from_double(I.elev[cj]));
    ++k;
  }
  S.omit_cell[s] = k;   // one row per focal cell: n_distinct(seqno) - 1
  S.omit_elev[s] = lds_to_double(sacc);
}

// sort key of every cell for the pit statistics: its shed if the cell lies at or below the shed's pour elevation
__global__ void __launch_bounds__(PT) k_ps_pitkeys(In I, Sheds S, uint32_t* __restrict__ keys, uint32_t* __restrict__ ids) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)I.d.cells) return;
  const int r = (int)(i / I.d.pitch), c = (int)(i % I.d.pitch);
  uint32_t k = 0xFFFFFFFFu;
  if (c < I.d.NC && r < I.d.NR) {
    const int32_t s = I.shed[i];
    if (s > 0 && S.minpour[s] != ~0ull && I.elev[i] <= f64_from_orderable(S.minpour[s])) k = (uint32_t)s;
  }
  keys[i] = k;
  ids[i] = (uint32_t)i;
}
__global__ void __launch_bounds__(PT) k_ps_pits(In I, Sheds S, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ ids, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t s = keys[i];
  if (s == 0xFFFFFFFFu || (i && keys[i - 1] == s)) return;
  const double pour = f64_from_orderable(S.minpour[s]);
  Ld80 acc;
  acc.m = 0; acc.e = 0;
  unsigned area = 0, pit = 0xFFFFFFFFu;
  bool edge = false;
  for (size_t j = i; j < n && keys[j] == s; ++j) {   // the shed's cells in seqno order (stable sort): sum(pour_elev - elev) (:103)
    const uint32_t cell = ids[j];
    acc = ld80_add(acc, ld80_from_double(pour - I.elev[cell]));
    ++area;
    edge = edge || I.edge[cell] == 1;
    if (pit == 0xFFFFFFFFu && I.ddir[cell] == 5) pit = cell;
  }
  S.pit_area[s] = area;
  S.pit_cell[s] = pit;
  S.edge_pit[s] = edge;
  S.pit_vol[s] = ld80_to_double(acc);
}
__global__ void k_ps_flags(Sheds S, unsigned maxid, uint32_t* __restrict__ flag) {
  const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s <= maxid) flag[s] = (s > 0 && S.minpour[s] != ~0ull) ? 1u : 0u;
}
struct Cols { double* c[NCOL]; };
__global__ void k_ps_rows(In I, Sheds S, unsigned maxid, const uint32_t* __restrict__ flag, int method, Cols O, unsigned cap) {
  const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s > maxid || !flag[s]) return;
  const unsigned row = S.row[s];
  if (row >= cap) return;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  const int NC = I.d.NC, pitch = I.d.pitch;
  const double pour = f64_from_orderable(S.minpour[s]);
  const uint32_t in = S.sel_cell[s];
  const int n = S.sel_n[s];
  const uint32_t outc = in + nb_off(n, pitch);
  const double shed_area = (double)S.area[s];
  double pit_area = (double)S.pit_area[s], pit_vol = S.pit_vol[s];
  if (method == 1) {   // LandMapR adds a cell to the pit only when it reaches a new LOW elevation (:114-123)
    const double oc = (double)S.omit_cell[s];
    pit_area = pit_area - oc;
    pit_vol = pit_vol - (pour * oc - S.omit_elev[s]);
  }
  const double varatio = shed_area > 0 ? pit_vol / shed_area * 1000 : 0;
  auto seq = [&](uint32_t cell) { return (double)((size_t)(cell / pitch) * NC + (cell % pitch) + 1); };
  auto rw = [&](uint32_t cell) { return (double)(cell / pitch + 1); };
  auto cl = [&](uint32_t cell) { return (double)(cell % pitch + 1); };
  const uint32_t pit = S.pit_cell[s];
  const bool hp = pit != 0xFFFFFFFFu;
  double v[NCOL];
  v[C_SHEDNO] = s; v[C_SHED_AREA] = shed_area; v[C_EDGE_PIT] = S.edge_pit[s]; v[C_PIT_AREA] = pit_area; v[C_PIT_VOL] = pit_vol;
  v[C_PIT_ELEV] = hp ? I.elev[pit] : nan; v[C_PIT_SEQNO] = hp ? seq(pit) : nan; v[C_PIT_ROW] = hp ? rw(pit) : nan;
  v[C_PIT_COL] = hp ? cl(pit) : nan; v[C_PRE_VOL] = 0; v[C_VARATIO] = varatio;
  v[C_IN_SEQNO] = seq(in); v[C_IN_ROW] = rw(in); v[C_IN_COL] = cl(in); v[C_IN_ELEV] = I.elev[in];
  v[C_OUT_SEQNO] = seq(outc); v[C_OUT_ROW] = rw(outc); v[C_OUT_COL] = cl(outc); v[C_OUT_ELEV] = I.elev[outc];
  v[C_DRAINS_TO] = I.shed[outc]; v[C_POUR_ELEV] = pour;
#pragma unroll
  for (int k = 0; k < NCOL; ++k)
    if (O.c[k]) O.c[k][row] = v[k];
}
// "only one watershed": the first pit cell of the DEM
__global__ void __launch_bounds__(PT) k_ps_first_pit(In I, unsigned* out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= I.d.NC) return;
  const size_t id = (size_t)r * I.d.pitch + c;
  if (I.ddir[id] == 5 && !is_nan(I.elev[id])) atomicMin(out, (unsigned)id);
}

// edge_map of load_file() (R/LITAP_load.R:255-267): cells with a buffer (ring) or missing neighbour
__global__ void __launch_bounds__(PT) k_ps_edge_map(const double* __restrict__ elev, GridDims d, uint8_t* __restrict__ edge) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d.NC) return;
  const size_t id = (size_t)r * d.pitch + c;
  uint8_t e = 0;
  if (r > 0 && c > 0 && r < d.NR - 1 && c < d.NC - 1 && !is_nan(elev[id])) {
#pragma unroll
    for (int n = 1; n <= 9; ++n)
      if (n != 5 && is_nan(elev[id + nb_off(n, d.pitch)])) e = 1;
  }
  edge[id] = e;
}

}  // namespace pitstat

namespace {
inline unsigned nblk(size_t n, unsigned t = 256) { return (unsigned)((n + t - 1) / t); }
#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)
}

extern "C" const char* const* litap_pit_stat_names(int* n) {
  if (n) *n = pitstat::NCOL;
  return pitstat::NAMES;
}

// the grid carries elev, ddir, shed (the shed column to use), upslope, ridge; edge = the edge_map column (pitched, device)
int run_pit_stat(litap_grid* g, const uint8_t* edge, int method, int64_t cap, double* const* out, int64_t* nrows) {
  using namespace pitstat;
  Ctx& c = ctx();
  cudaStream_t st = c.stream;
  const GridDims d = g->d;
  In I = {g->elev, g->ddir, g->shed, g->upslope, g->ridge, edge, d};
  PoolScratch P;
  unsigned* ctr = P.get<unsigned>(4);   // [0] max shed number, [1] tied cursor, [2] first pit
  if (!ctr) return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat)");
  CUDA_TRY(cudaMemsetAsync(ctr, 0, 16, st));
  CUDA_TRY(cudaMemsetAsync(ctr + 2, 0xFF, 4, st));
  dim3 cgrid(nblk(d.NC, PT), d.NR);
  KERNEL("k_ps_maxid", k_ps_maxid<<<cgrid, PT, 0, st>>>(I, ctr));
  unsigned h[4];
  CUDA_TRY(cudaMemcpyAsync(h, ctr, 16, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  const unsigned maxid = h[0];
  *nrows = 0;
  if (!maxid) return LITAP_OK;
  const size_t ns = (size_t)maxid + 1;
  Sheds S;
  S.minpour = P.get<unsigned long long>(ns);
  S.area = P.get<uint32_t>(ns);
  S.sel_cell = P.get<uint32_t>(ns);
  S.sel_n = P.get<uint8_t>(ns);
  S.omit_cell = P.get<uint32_t>(ns);
  S.omit_elev = P.get<double>(ns);
  S.pit_area = P.get<uint32_t>(ns);
  S.pit_cell = P.get<uint32_t>(ns);
  S.edge_pit = P.get<uint8_t>(ns);
  S.pit_vol = P.get<double>(ns);
  S.row = P.get<uint32_t>(ns);
  uint32_t* flag = P.get<uint32_t>(ns);
  uint8_t* cand_n = P.get<uint8_t>((size_t)d.cells + 64);
  if (!S.minpour || !S.area || !S.sel_cell || !S.sel_n || !S.omit_cell || !S.omit_elev || !S.pit_area || !S.pit_cell ||
      !S.edge_pit || !S.pit_vol || !S.row || !flag || !cand_n)
    return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat shed tables)");
  CUDA_TRY(cudaMemsetAsync(S.minpour, 0xFF, ns * 8, st));
  CUDA_TRY(cudaMemsetAsync(S.area, 0, ns * 4, st));
  CUDA_TRY(cudaMemsetAsync(S.omit_cell, 0, ns * 4, st));
  CUDA_TRY(cudaMemsetAsync(S.omit_elev, 0, ns * 8, st));
  CUDA_TRY(cudaMemsetAsync(S.pit_area, 0, ns * 4, st));
  CUDA_TRY(cudaMemsetAsync(S.pit_cell, 0xFF, ns * 4, st));
  CUDA_TRY(cudaMemsetAsync(S.edge_pit, 0, ns, st));
  CUDA_TRY(cudaMemsetAsync(S.pit_vol, 0, ns * 8, st));
  KERNEL("k_ps_cells", k_ps_cells<<<cgrid, PT, 0, st>>>(I, S, method, cand_n));
  // number of sheds present; the single-shed branch
  {
    cub::TransformInputIterator<uint32_t, NonZero, const uint32_t*> it(S.area, NonZero());
    size_t tb = 0;
    uint32_t* nsheds_d = P.get<uint32_t>(1);
    cub::DeviceReduce::Sum(nullptr, tb, it, nsheds_d, (int)ns, st);
    void* tmp = P.get<unsigned char>(tb + 256);
    if (!nsheds_d || !tmp) return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat)");
    CUDA_TRY(cub::DeviceReduce::Sum(tmp, tb, it, nsheds_d, (int)ns, st));
    uint32_t nsheds = 0;
    CUDA_TRY(cudaMemcpyAsync(&nsheds, nsheds_d, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (nsheds <= 1) {   // R/flow_03_pit_stat_fast.R:127-151
      KERNEL("k_ps_first_pit", k_ps_first_pit<<<cgrid, PT, 0, st>>>(I, ctr + 2));
      std::vector<uint32_t> area(ns);
      CUDA_TRY(cudaMemcpyAsync(area.data(), S.area, ns * 4, cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaMemcpyAsync(h, ctr, 16, cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaStreamSynchronize(st));
      unsigned sid = 0;
      for (size_t s = 1; s < ns; ++s) if (area[s]) sid = (unsigned)s;
      *nrows = 1;
      if (cap < 1) return LITAP_OK;
      const double nan = std::nan("");
      for (int k = 0; k < NCOL; ++k) if (out[k]) out[k][0] = nan;
      if (out[C_SHEDNO]) out[C_SHEDNO][0] = sid;
      if (out[C_SHED_AREA]) out[C_SHED_AREA][0] = 1;   // shed_area = seq_along(pit_vol)
      if (out[C_EDGE_PIT]) out[C_EDGE_PIT][0] = 1;
      if (h[2] != 0xFFFFFFFFu) {
        const unsigned pit = h[2];
        double e = 0;
        CUDA_TRY(cudaMemcpy(&e, g->elev + pit, 8, cudaMemcpyDeviceToHost));
        if (out[C_PIT_ELEV]) out[C_PIT_ELEV][0] = e;
        if (out[C_PIT_SEQNO]) out[C_PIT_SEQNO][0] = (double)((size_t)(pit / d.pitch) * d.NC + pit % d.pitch + 1);
        if (out[C_PIT_ROW]) out[C_PIT_ROW][0] = pit / d.pitch + 1;
        if (out[C_PIT_COL]) out[C_PIT_COL][0] = pit % d.pitch + 1;
      }
      return LITAP_OK;
    }
  }
  // ---- tied candidates, arranged
  // (a first pass only counts; the list is a few entries per shed)
  CUDA_TRY(cudaMemsetAsync(ctr + 1, 0, 4, st));
  KERNEL("k_ps_tied", k_ps_tied<<<cgrid, PT, 0, st>>>(I, S, cand_n, ctr + 1, 0u, nullptr, nullptr, nullptr));
  CUDA_TRY(cudaMemcpyAsync(h, ctr, 16, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  const unsigned nt = h[1];
  if (nt) {
    uint32_t* t_shed = P.get<uint32_t>(nt);
    unsigned long long* t_elev = P.get<unsigned long long>(nt);
    unsigned long long* t_key = P.get<unsigned long long>(nt);
    unsigned long long* k64a = P.get<unsigned long long>(nt);
    unsigned long long* k64b = P.get<unsigned long long>(nt);
    uint32_t* k32a = P.get<uint32_t>(nt);
    uint32_t* k32b = P.get<uint32_t>(nt);
    uint32_t* pa = P.get<uint32_t>(nt);
    uint32_t* pb = P.get<uint32_t>(nt);
    size_t t1 = 0, t2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, t1, k64a, k64b, pa, pb, (int)nt, 0, 64, st);
    cub::DeviceRadixSort::SortPairs(nullptr, t2, k32a, k32b, pa, pb, (int)nt, 0, 32, st);
    const size_t tn = std::max(t1, t2) + 256;
    void* tmp = P.get<unsigned char>(tn);
    if (!t_shed || !t_elev || !t_key || !k64a || !k64b || !k32a || !k32b || !pa || !pb || !tmp)
      return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat candidates)");
    CUDA_TRY(cudaMemsetAsync(ctr + 1, 0, 4, st));
    KERNEL("k_ps_tied", k_ps_tied<<<cgrid, PT, 0, st>>>(I, S, cand_n, ctr + 1, nt, t_shed, t_elev, t_key));
    KERNEL("k_ps_iota", k_ps_iota<<<nblk(nt), 256, 0, st>>>(pa, nt));
    size_t tc = tn;   // arrange(shedno, elev, upslope, seqno): stable LSD passes, last key first
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tc, t_key, k64b, pa, pb, (int)nt, 0, 64, st));
    KERNEL("k_ps_gather", k_ps_gather<<<nblk(nt), 256, 0, st>>>(t_elev, pb, k64a, nt));
    tc = tn;
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tc, k64a, k64b, pb, pa, (int)nt, 0, 64, st));
    KERNEL("k_ps_gather", k_ps_gather<<<nblk(nt), 256, 0, st>>>(t_shed, pa, k32a, nt));
    tc = tn;
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tc, k32a, k32b, pa, pb, (int)nt, 0, 32, st));
    c.launches += 3;
    KERNEL("k_ps_select", k_ps_select<<<nblk(nt), 256, 0, st>>>(I, S, cand_n, k32b, pb, t_key, nt));
  }
  // ---- the cells of every pit, grouped by shed in seqno order
  {
    const size_t ncell = (size_t)d.cells;
    uint32_t* ka = P.get<uint32_t>(ncell);
    uint32_t* kb = P.get<uint32_t>(ncell);
    uint32_t* ia = P.get<uint32_t>(ncell);
    uint32_t* ib = P.get<uint32_t>(ncell);
    size_t tb = 0;
    const int end_bit = 32;   // cells that do not count carry all ones: they must sort behind every shed
    cub::DeviceRadixSort::SortPairs(nullptr, tb, ka, kb, ia, ib, (int64_t)ncell, 0, end_bit, st);
    void* tmp = P.get<unsigned char>(tb + 256);
    if (!ka || !kb || !ia || !ib || !tmp) return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat cells)");
    KERNEL("k_ps_pitkeys", k_ps_pitkeys<<<nblk(ncell, PT), PT, 0, st>>>(I, S, ka, ia));
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tb, ka, kb, ia, ib, (int64_t)ncell, 0, end_bit, st));
    c.launches += 1;
    KERNEL("k_ps_pits", k_ps_pits<<<nblk(ncell, PT), PT, 0, st>>>(I, S, kb, ib, ncell));
  }
  // ---- the table
  KERNEL("k_ps_flags", k_ps_flags<<<nblk(ns), 256, 0, st>>>(S, maxid, flag));
  {
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, flag, S.row, (int)ns, st);
    void* tmp = P.get<unsigned char>(tb + 256);
    if (!tmp) return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat)");
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tb, flag, S.row, (int)ns, st));
  }
  uint32_t last_row = 0, last_flag = 0;
  CUDA_TRY(cudaMemcpyAsync(&last_row, S.row + maxid, 4, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(&last_flag, flag + maxid, 4, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  const unsigned rows = last_row + last_flag;
  *nrows = rows;
  const unsigned wr = (unsigned)std::min<int64_t>(rows, cap);
  if (!wr) return LITAP_OK;
  Cols O;
  double* dev = P.get<double>((size_t)wr * NCOL);
  if (!dev) return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat table)");
  for (int k = 0; k < NCOL; ++k) O.c[k] = dev + (size_t)k * wr;
  KERNEL("k_ps_rows", k_ps_rows<<<nblk(ns), 256, 0, st>>>(I, S, maxid, flag, method, O, wr));
  for (int k = 0; k < NCOL; ++k)
    if (out[k]) CUDA_TRY(cudaMemcpyAsync(out[k], O.c[k], (size_t)wr * 8, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  return LITAP_OK;
}

// pit_stat1 on a resident grid: the shed / upslope / ridge columns calc_shed4 left on it (shed = "initial_shed"), the
// edge_map derived from the elevations on the device
extern "C" int litap_grid_pit_stat(litap_grid* g, int method, int64_t cap, double* const* out, int64_t* nrows) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || !out || !nrows) return set_error(LITAP_EARG, "litap_grid_pit_stat: NULL");
  if (method != 0 && method != 1) return set_error(LITAP_EARG, "litap_grid_pit_stat: method must be 0 (litap) or 1 (landmapr)");
  if (g->nbands > 0) return set_error(LITAP_EARG, "litap_grid_pit_stat: not available on a row band (run it on a whole-DEM grid)");
  if (!g->ddir || !g->shed || !g->upslope || !g->ridge) return set_error(LITAP_EARG, "litap_grid_pit_stat: needs ddir, shed, upslope and ridge on the grid");
  Ctx& c = ctx();
  const int64_t l0 = c.launches;
  uint8_t* edge = grid_alloc_array<uint8_t>(g->d, 0);
  if (!edge) return set_error(LITAP_ENOMEM, "device allocation failed (edge_map)");
  struct EdgeFree { uint8_t* p; ~EdgeFree() { cudaStreamSynchronize(ctx().stream); grid_free_array(p); } } ef{edge};
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  dim3 cgrid(nblk(g->d.NC, pitstat::PT), g->d.NR);
  KERNEL("k_ps_edge_map", pitstat::k_ps_edge_map<<<cgrid, pitstat::PT, 0, c.stream>>>(g->elev, g->d, edge));
  int rc = run_pit_stat(g, edge, method, cap, out, nrows);
  if (rc) return rc;
  CUDA_TRY(cudaEventRecord(c.ev[7], c.stream));
  CUDA_TRY(cudaEventSynchronize(c.ev[7]));
  cudaEventElapsedTime(&c.tm.pitstat_ms, c.ev[6], c.ev[7]);
  c.tm.kernel_launches = c.launches - l0;
  return LITAP_OK;
}
