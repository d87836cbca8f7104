// relz.cu — calc_relz + calc_length on the device:
//   calc_stream3  R/form_03_calc_relz.R:194-276   first channel cell downstream (streams; ridges on the inverted DEM)
//   calc_pit3     R/form_03_calc_relz.R:424-479   pit of the cell's shed, steps to the end of the flow path
//   calc_relief   R/form_03_calc_relz.R:487-514   relief percentages
//   calc_length   R/form_04_calc_length.R:3-35    slope lengths
//
// The reference walks the flow path of every cell in a sorted order, cuts the walk at the first cell
// it has already visited or the first channel cell and copies from there.  What a cell ends up with
// does not depend on that order: it is always "the first cell on my flow path (me included) whose
// upslope >= str_val, else the cell the path ends in" and the number of steps to it (the oracle
// follows the R order literally; the two agree on the reference's golden and on synthetic DEMs).
// That is pointer jumping with a payload:
//   k_relz_init   pair[c] = (c, 0) if c stops the walk, else (drec[c], 1)           (target, steps) in one 64-bit word
//   k_relz_jump   pair[c] = (target of my target, my steps + its steps), in place: a pair is always a true
//                 statement about c's path, so reading a neighbour's pair mid-update is harmless; ~log2(longest
//                 path) rounds; a cell whose target is a stopping cell is marked final and costs one load after that.  The four traces (stream, path end, ridge, inverted path end) share the launches.
//   k_relz_sheds  pit / peak of every shed id (pond table or ddir == 5 cells), max elevation per fill_shed, DEM min / max
//   k_relz_cells  everything else is per-cell arithmetic in the reference's operation order (no FMA)
// drec is a global cell number, so the arrays are dense NR x NC (no pitch) on this path.
#include <vector>
#include "cellmath.h"
#include "litap_internal.h"

namespace {

constexpr int RELZ_NOUT = LITAP_RELZ_NOUT;
constexpr int RZ_THREADS = 256;
constexpr int RZ_MAX_ROUNDS = 64;
constexpr unsigned long long RZ_FINAL = 1ULL << 63;

__device__ __forceinline__ double rz_na() { return __longlong_as_double(0x7ff8000000000000LL); }
// order-preserving map double -> uint64 (for atomicMin / atomicMax)
__device__ __forceinline__ unsigned long long rz_enc(double x) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double rz_dec(unsigned long long u) {
  const unsigned long long b = (u >> 63) ? (u & 0x7fffffffffffffffULL) : ~u;
  return __longlong_as_double((long long)b);
}

struct Trace {
  const double* elev;
  const int32_t* drec;
  const int32_t* upslope;
  double val;
  int use_val;
  unsigned long long* pair;
};
struct Traces { Trace t[4]; const int32_t* sheds[4]; };   // sheds[y]: a shed-number array trace y also range-checks

__global__ void __launch_bounds__(RZ_THREADS) k_relz_init(Traces T, int64_t n, int* __restrict__ bad) {
  const int64_t i = (int64_t)blockIdx.x * RZ_THREADS + threadIdx.x;
  if (i >= n) return;
  const Trace& t = T.t[blockIdx.y];
  const int32_t d = t.drec[i];
  const double e = t.elev[i];
  bool stop = e != e || d == 0 || (int64_t)d - 1 == i;
  if (!stop && ((int64_t)d < 1 || (int64_t)d > n)) { *bad = 1; stop = true; }
  if (T.sheds[blockIdx.y]) { const int32_t s = T.sheds[blockIdx.y][i]; if (s < 0 || (int64_t)s > n) *bad = 2; }
  if (!stop && t.use_val) { const int32_t u = t.upslope[i]; stop = u != 0 && (double)u >= t.val; }
  t.pair[i] = stop ? (RZ_FINAL | (unsigned long long)(uint32_t)i) : ((1ULL << 32) | (uint32_t)(d - 1));
}

// bit 63 of a pair: its target is where the walk stops (nothing left to do for this cell).
// A thread advances RZ_U cells at once: the hops of one cell depend on each other, the cells do not, and the
// kernel is bound by the latency of the scattered loads.
constexpr int RZ_U = 4;
__global__ void __launch_bounds__(RZ_THREADS) k_relz_jump(Traces T, int64_t n, int* __restrict__ flags) {
  // plain (L1-cached) loads: a stale pair is still a true statement about that cell's path, it only delays the
  // jump to the next launch
  unsigned long long* pair = T.t[blockIdx.y].pair;
  const int64_t base = ((int64_t)blockIdx.x * RZ_THREADS + (threadIdx.x & ~31)) * RZ_U + (threadIdx.x & 31);
  unsigned long long p[RZ_U];
  bool act[RZ_U], moved[RZ_U];
#pragma unroll
  for (int u = 0; u < RZ_U; ++u) {
    const int64_t i = base + 32 * u;
    p[u] = i < n ? pair[i] : RZ_FINAL;
    act[u] = !(p[u] & RZ_FINAL);
    moved[u] = false;
  }
#pragma unroll 1
  for (int k = 0; k < 4; ++k) {   // a few hops per launch: fewer rounds, each ends with a host check
    unsigned long long q[RZ_U];
#pragma unroll
    for (int u = 0; u < RZ_U; ++u) q[u] = act[u] ? pair[(uint32_t)p[u]] : 0ULL;
#pragma unroll
    for (int u = 0; u < RZ_U; ++u) {
      if (!act[u]) continue;
      const unsigned long long steps = (p[u] >> 32) + ((q[u] & ~RZ_FINAL) >> 32);
      if (steps > (unsigned long long)n) { flags[1] = 3; act[u] = false; continue; }   // longer than the grid: a circle
      p[u] = (steps << 32) | (uint32_t)q[u] | (q[u] & RZ_FINAL);
      moved[u] = true;
      act[u] = !(p[u] & RZ_FINAL);
    }
  }
  bool open = false;
#pragma unroll
  for (int u = 0; u < RZ_U; ++u) {
    if (moved[u]) pair[base + 32 * u] = p[u];
    open = open || act[u];
  }
  if (open) flags[0] = 1;
}

struct ShedTab { int32_t* row; int32_t* col; double* elev; };   // by shed id; row 0 = no pit known

__global__ void __launch_bounds__(RZ_THREADS) k_relz_pond(int64_t npond, const int32_t* __restrict__ shedno,
                                                          const int32_t* __restrict__ row, const int32_t* __restrict__ col,
                                                          const double* __restrict__ elev, ShedTab tab) {
  const int64_t k = (int64_t)blockIdx.x * RZ_THREADS + threadIdx.x;
  if (k >= npond) return;
  const int32_t s = shedno[k];
  tab.row[s] = row[k]; tab.col[s] = col[k]; tab.elev[s] = elev[k];
}
// the ddir == 5 cell of every shed (calc_pit3 without a pond table; the peaks of the inverted DEM)
__global__ void __launch_bounds__(RZ_THREADS) k_relz_pitcells(int64_t n, int NC, const uint8_t* __restrict__ ddir,
                                                              const int32_t* __restrict__ shed, const double* __restrict__ elev,
                                                              ShedTab tab) {
  const int64_t i = (int64_t)blockIdx.x * RZ_THREADS + threadIdx.x;
  if (i >= n || ddir[i] != 5 || shed[i] == 0) return;
  const int32_t s = shed[i];
  tab.row[s] = (int32_t)(i / NC) + 1; tab.col[s] = (int32_t)(i % NC) + 1; tab.elev[s] = elev[i];
}
// mm[0] = min, mm[1] = max of the DEM (one atomic pair per block), shed_max[id] = max elevation of the fill_shed
// (encoded doubles; 0 = none)
__global__ void __launch_bounds__(RZ_THREADS) k_relz_sheds(int64_t n, const double* __restrict__ elev,
                                                           const int32_t* __restrict__ fill_shed,
                                                           unsigned long long* __restrict__ mm,
                                                           unsigned long long* __restrict__ shed_max) {
  __shared__ unsigned long long s_hi[RZ_THREADS / 32], s_lo[RZ_THREADS / 32];
  const int64_t i = (int64_t)blockIdx.x * RZ_THREADS + threadIdx.x;
  const double e = i < n ? elev[i] : rz_na();
  const bool ok = e == e;
  const int32_t s = ok ? fill_shed[i] : -1;
  unsigned long long hi = ok ? rz_enc(e) : 0ULL, lo = ok ? rz_enc(e) : ~0ULL;
  int same;
  __match_all_sync(0xffffffffu, s, &same);
  const unsigned long long shi = hi;
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const unsigned long long a = __shfl_xor_sync(0xffffffffu, hi, o), b = __shfl_xor_sync(0xffffffffu, lo, o);
    hi = a > hi ? a : hi;
    lo = b < lo ? b : lo;
  }
  if (same) { if ((threadIdx.x & 31) == 0 && ok) atomicMax(shed_max + s, hi); }   // the whole warp in one shed: one atomic
  else if (ok) atomicMax(shed_max + s, shi);
  if ((threadIdx.x & 31) == 0) { s_hi[threadIdx.x >> 5] = hi; s_lo[threadIdx.x >> 5] = lo; }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < RZ_THREADS / 32; ++w) { hi = s_hi[w] > hi ? s_hi[w] : hi; lo = s_lo[w] < lo ? s_lo[w] : lo; }
    if (hi) { atomicMax(mm + 1, hi); atomicMin(mm, lo); }
  }
}

struct RelzIn {
  const double *elev, *ielev;
  const int32_t *fill_shed, *local_shed, *inv_local_shed;
  const unsigned long long *p_st, *p_pit, *p_cr, *p_peak;
  ShedTab pits, peaks;
  const unsigned long long *mm, *shed_max;
};
struct RelzOut { double* o[RELZ_NOUT]; };

__device__ __forceinline__ double rz_sq(double x) { return x * x; }                       // R: x^2 is x * x
__device__ __forceinline__ double rz_pct(double a, double b) { return trunc((a / b) * 100); }
__device__ __forceinline__ double rz_len3(double r2, double c2, double z, double row, double col, double grid) {
  const double l = sqrt(rz_sq(c2 - col) + rz_sq(r2 - row)) * grid;
  return sqrt(rz_sq(l) + rz_sq(z * grid));
}

__global__ void __launch_bounds__(RZ_THREADS, 3) k_relz_cells(int64_t n, int NC, double grid, RelzIn I, RelzOut O) {
  const int64_t i = (int64_t)blockIdx.x * RZ_THREADS + threadIdx.x;
  if (i >= n) return;
  const double NA = rz_na();
  const double e = I.elev[i], ie = I.ielev[i];
  const double row = (double)(i / NC + 1), col = (double)(i % NC + 1);
  const double mn = rz_dec(I.mm[0]), mx = rz_dec(I.mm[1]);   // (an all-NA DEM never gets here with e valid)
  // every column is stored as soon as it is known (39 live doubles would cost the occupancy)
  auto put = [&](int k, double x) { if (O.o[k]) O.o[k][i] = x; };
  // streams / ridges: the cell the trace stopped at
  double st_row = NA, st_col = NA, st_elev = NA, z2st = 0, n2st = NA;   // z2st starts as 0 in the reference (:198)
  if (e == e) {
    const unsigned long long p = I.p_st[i];
    const int64_t t = (uint32_t)p;
    st_row = (double)(t / NC + 1); st_col = (double)(t % NC + 1); st_elev = I.elev[t];
    z2st = e - st_elev; if (z2st < 0) z2st = 0;
    n2st = (double)(uint32_t)((p & ~RZ_FINAL) >> 32);
  }
  double cr_row = NA, cr_col = NA, cr_ie = NA, z2cr = 0, n2cr = NA;
  if (ie == ie) {
    const unsigned long long p = I.p_cr[i];
    const int64_t t = (uint32_t)p;
    cr_row = (double)(t / NC + 1); cr_col = (double)(t % NC + 1); cr_ie = I.ielev[t];
    z2cr = ie - cr_ie; if (z2cr < 0) z2cr = 0;
    n2cr = (double)(uint32_t)((p & ~RZ_FINAL) >> 32);
  }
  if (e != e) { z2st = NA; z2cr = NA; }                                     // calc_relz :41
  // pits / peaks: by shed id
  double pit_row = NA, pit_col = NA, pit_elev = NA, z2pit = NA, n2pit = NA;
  const int32_t ls = I.local_shed[i];
  if (ls && I.pits.row[ls]) {
    pit_row = I.pits.row[ls]; pit_col = I.pits.col[ls]; pit_elev = I.pits.elev[ls];
    z2pit = e - pit_elev; if (z2pit < 0) z2pit = 0;
  }
  if (ls && e == e) n2pit = (double)(uint32_t)((I.p_pit[i] & ~RZ_FINAL) >> 32);
  double pk_row = NA, pk_col = NA, pk_ie = NA, z2peak = NA, n2peak = NA;
  const int32_t is = I.inv_local_shed[i];
  if (is && I.peaks.row[is]) {
    pk_row = I.peaks.row[is]; pk_col = I.peaks.col[is]; pk_ie = I.peaks.elev[is];
    z2peak = ie - pk_ie; if (z2peak < 0) z2peak = 0;
  }
  if (is && ie == ie) n2peak = (double)(uint32_t)((I.p_peak[i] & ~RZ_FINAL) >> 32);
  put(0, st_row); put(1, st_col); put(2, st_elev); put(3, z2st); put(4, n2st);
  put(5, cr_row); put(6, cr_col); put(7, mx - cr_ie); put(8, z2cr); put(9, n2cr);
  put(10, pit_row); put(11, pit_col); put(12, pit_elev); put(13, z2pit); put(14, n2pit);
  put(15, pk_row); put(16, pk_col); put(17, mx - pk_ie); put(18, z2peak); put(19, n2peak);
  // calc_relief
  double max_shed = NA;
  if (e == e) { const unsigned long long u = I.shed_max[I.fill_shed[i]]; if (u) max_shed = rz_dec(u); }
  const double zpit2peak = z2pit + z2peak, zcr2st = z2cr + z2st, zcr2pit = z2cr + z2pit;
  const double z2top = max_shed - e, ztop2pit = max_shed - pit_elev, ncr2st = n2cr + n2st;
  const double range = mx - mn;
  double pctz2top = trunc(100 - (z2top / ztop2pit) * 100), pctz2st = rz_pct(z2st, zcr2st);
  const double pctz2pit = rz_pct(z2pit, zpit2peak);
  double pctn2st = rz_pct(n2st, ncr2st), pmin2max = rz_pct(e - mn, range);
  if (ztop2pit == 0) pctz2top = 0;
  if (zcr2st == 0) pctz2st = 0;
  if (ncr2st == 0) pctn2st = 0;
  if (range == 0) pmin2max = 0;
  put(20, z2top); put(21, zpit2peak); put(22, zcr2st); put(23, zcr2pit); put(24, ztop2pit); put(25, ncr2st);
  put(26, pmin2max); put(27, pctz2st); put(28, pctz2pit); put(29, pctz2top); put(30, pctn2st);
  // calc_length
  const double l2pit = rz_len3(pit_row, pit_col, z2pit, row, col, grid), l2peak = rz_len3(pk_row, pk_col, z2peak, row, col, grid);
  const double l2str = rz_len3(st_row, st_col, z2st, row, col, grid), l2div = rz_len3(cr_row, cr_col, z2cr, row, col, grid);
  const double lpit2peak = l2pit + l2peak, lstr2div = l2str + l2div;
  double ppit2peakl = rz_pct(l2pit, lpit2peak), pstr2divl = rz_pct(l2str, lstr2div);
  const double r_lpit2peak = r_round1(lpit2peak), r_lstr2div = r_round1(lstr2div);
  if (r_lpit2peak == 0) ppit2peakl = 0;
  if (r_lstr2div == 0) pstr2divl = 0;
  put(31, r_round1(l2pit)); put(32, r_round1(l2peak)); put(33, r_lpit2peak); put(34, ppit2peakl);
  put(35, r_round1(l2str)); put(36, r_round1(l2div)); put(37, r_lstr2div); put(38, pstr2divl);
}

#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

template <typename T>
int to_dev(PoolScratch& S, T** dst, const T* src, size_t n) {
  *dst = S.get<T>(n);
  if (!*dst) return set_error(LITAP_ENOMEM, "device allocation failed (relz inputs)");
  // as rows of at most 32 MB so that pageable sources go through the staged upload
  const size_t bytes = n * sizeof(T), row = 32u << 20, full = bytes / row;
  if (full) { int rc = h2d_2d(*dst, row, src, row, row, full, ctx().stream); if (rc) return rc; }
  if (bytes > full * row) {
    int rc = h2d_2d((char*)*dst + full * row, bytes - full * row, (const char*)src + full * row, bytes - full * row, bytes - full * row, 1, ctx().stream);
    if (rc) return rc;
  }
  return LITAP_OK;
}

}  // namespace

extern "C" const char* litap_relz_names(void) {
  return "st_row,st_col,st_elev,z2st,n2st,cr_row,cr_col,cr_elev,z2cr,n2cr,pit_row,pit_col,pit_elev,z2pit,n2pit,"
         "peak_row,peak_col,peak_elev,z2peak,n2peak,z2top,zpit2peak,zcr2st,zcr2pit,ztop2pit,ncr2st,pmin2max,pctz2st,"
         "pctz2pit,pctz2top,pctn2st,l2pit,l2peak,lpit2peak,ppit2peakl,l2str,l2div,lstr2div,pstr2divl";
}

extern "C" int litap_relz(int64_t NR, int64_t NC, const double* elev, const uint8_t* ddir, const int32_t* drec,
                          const int32_t* upslope, const int32_t* fill_shed, const int32_t* local_shed,
                          const double* ielev, const uint8_t* iddir, const int32_t* idrec, const int32_t* iupslope,
                          const int32_t* inv_local_shed, int64_t npond, const int32_t* pond_shedno,
                          const int32_t* pond_row, const int32_t* pond_col, const double* pond_elev, double str_val,
                          double ridge_val, double grid, double* const* out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (NR < 3 || NC < 3 || NR * NC > 0x7fffffffLL) return set_error(LITAP_EARG, "litap_relz: grid must be 3 x 3 .. 2^31 - 1 cells");
  if (!elev || !drec || !upslope || !fill_shed || !local_shed || !ielev || !iddir || !idrec || !iupslope || !inv_local_shed || !out)
    return set_error(LITAP_EARG, "litap_relz: NULL input array");
  if (npond < 0 || (npond > 0 && (!pond_shedno || !pond_row || !pond_col || !pond_elev)))
    return set_error(LITAP_EARG, "litap_relz: pond table incomplete");
  if (npond == 0 && !ddir) return set_error(LITAP_EARG, "litap_relz: ddir is needed when there is no pond table");
  if (!(grid > 0)) return set_error(LITAP_EARG, "litap_relz: grid must be > 0");
  const int64_t n = NR * NC;
  for (int64_t k = 0; k < npond; ++k)
    if (pond_shedno[k] < 1 || pond_shedno[k] > n) return set_error(LITAP_EARG, "litap_relz: pond shedno out of range");
  RC(ensure_init());
  hostcopy_reset();
  Ctx& c = ctx();
  cudaStream_t st = c.stream;
  c.tm = litap_timings{};
  const int64_t l0 = c.launches;
  PoolScratch S;
  double *d_elev, *d_ielev;
  int32_t *d_drec, *d_ups, *d_fs, *d_ls, *d_idrec, *d_iups, *d_ils;
  uint8_t *d_ddir = nullptr, *d_iddir;
  CUDA_TRY(cudaEventRecord(c.ev[6], st));
  RC(to_dev(S, &d_elev, elev, n)); RC(to_dev(S, &d_drec, drec, n)); RC(to_dev(S, &d_ups, upslope, n));
  RC(to_dev(S, &d_fs, fill_shed, n)); RC(to_dev(S, &d_ls, local_shed, n));
  RC(to_dev(S, &d_ielev, ielev, n)); RC(to_dev(S, &d_idrec, idrec, n)); RC(to_dev(S, &d_iups, iupslope, n));
  RC(to_dev(S, &d_ils, inv_local_shed, n)); RC(to_dev(S, &d_iddir, iddir, n));
  if (npond == 0) RC(to_dev(S, &d_ddir, ddir, n));
  int32_t *d_ps = nullptr, *d_pr = nullptr, *d_pc = nullptr;
  double* d_pe = nullptr;
  if (npond > 0) {
    RC(to_dev(S, &d_ps, pond_shedno, npond)); RC(to_dev(S, &d_pr, pond_row, npond));
    RC(to_dev(S, &d_pc, pond_col, npond)); RC(to_dev(S, &d_pe, pond_elev, npond));
  }
  CUDA_TRY(cudaEventRecord(c.ev[7], st));
  // work arrays
  Traces T;
  unsigned long long* pairs = S.get<unsigned long long>(4 * (size_t)n);
  ShedTab pits{S.get<int32_t>(n + 1), S.get<int32_t>(n + 1), S.get<double>(n + 1)};
  ShedTab peaks{S.get<int32_t>(n + 1), S.get<int32_t>(n + 1), S.get<double>(n + 1)};
  unsigned long long* shed_max = S.get<unsigned long long>(n + 1);
  unsigned long long* mm = S.get<unsigned long long>(2);
  int* flags = S.get<int>(2);
  if (!pairs || !pits.row || !pits.col || !pits.elev || !peaks.row || !peaks.col || !peaks.elev || !shed_max || !mm || !flags)
    return set_error(LITAP_ENOMEM, "device allocation failed (relz work arrays)");
  T.t[0] = Trace{d_elev, d_drec, d_ups, str_val, 1, pairs};
  T.t[1] = Trace{d_elev, d_drec, d_ups, 0.0, 0, pairs + n};
  T.t[2] = Trace{d_ielev, d_idrec, d_iups, ridge_val, 1, pairs + 2 * n};
  T.t[3] = Trace{d_ielev, d_idrec, d_iups, 0.0, 0, pairs + 3 * n};
  T.sheds[0] = d_fs; T.sheds[1] = d_ls; T.sheds[2] = d_ils; T.sheds[3] = nullptr;
  CUDA_TRY(cudaMemsetAsync(pits.row, 0, (size_t)(n + 1) * sizeof(int32_t), st));
  CUDA_TRY(cudaMemsetAsync(peaks.row, 0, (size_t)(n + 1) * sizeof(int32_t), st));
  CUDA_TRY(cudaMemsetAsync(shed_max, 0, (size_t)(n + 1) * sizeof(unsigned long long), st));
  CUDA_TRY(cudaMemsetAsync(mm, 0xff, sizeof(unsigned long long), st));
  CUDA_TRY(cudaMemsetAsync(mm + 1, 0, sizeof(unsigned long long), st));
  CUDA_TRY(cudaMemsetAsync(flags, 0, 2 * sizeof(int), st));
  const unsigned nb = (unsigned)((n + RZ_THREADS - 1) / RZ_THREADS);
  KERNEL("k_relz_init", k_relz_init<<<dim3(nb, 4), RZ_THREADS, 0, st>>>(T, n, flags + 1));
  int rounds = 0;
  for (;; ++rounds) {
    if (rounds >= RZ_MAX_ROUNDS) return set_error(LITAP_EARG, "litap_relz: flow paths do not end (cycle in drec)");
    CUDA_TRY(cudaMemsetAsync(flags, 0, sizeof(int), st));
    KERNEL("k_relz_jump", k_relz_jump<<<dim3((nb + RZ_U - 1) / RZ_U, 4), RZ_THREADS, 0, st>>>(T, n, flags));
    int h[2];
    CUDA_TRY(cudaMemcpyAsync(h, flags, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (h[1] == 3) return set_error(LITAP_EARG, "litap_relz: flow paths do not end (cycle in drec)");
    if (h[1]) return set_error(LITAP_EARG, h[1] == 2 ? "litap_relz: shed number outside 0 .. NR * NC" : "litap_relz: drec outside 1 .. NR * NC");
    if (!h[0]) break;
  }
  c.tm.plateau_rounds = rounds + 1;
  if (npond > 0)
    KERNEL("k_relz_pond", k_relz_pond<<<(unsigned)((npond + RZ_THREADS - 1) / RZ_THREADS), RZ_THREADS, 0, st>>>(npond, d_ps, d_pr, d_pc, d_pe, pits));
  else
    KERNEL("k_relz_pitcells", k_relz_pitcells<<<nb, RZ_THREADS, 0, st>>>(n, (int)NC, d_ddir, d_ls, d_elev, pits));
  KERNEL("k_relz_pitcells", k_relz_pitcells<<<nb, RZ_THREADS, 0, st>>>(n, (int)NC, d_iddir, d_ils, d_ielev, peaks));
  KERNEL("k_relz_sheds", k_relz_sheds<<<nb, RZ_THREADS, 0, st>>>(n, d_elev, d_fs, mm, shed_max));
  RelzOut O;
  for (int k = 0; k < RELZ_NOUT; ++k) {
    O.o[k] = nullptr;
    if (out[k]) {
      O.o[k] = S.get<double>(n);
      if (!O.o[k]) return set_error(LITAP_ENOMEM, "device allocation failed (relz outputs)");
    }
  }
  RelzIn I{d_elev, d_ielev, d_fs, d_ls, d_ils, pairs, pairs + n, pairs + 2 * n, pairs + 3 * n, pits, peaks, mm, shed_max};
  KERNEL("k_relz_cells", k_relz_cells<<<nb, RZ_THREADS, 0, st>>>(n, (int)NC, grid, I, O));
  CUDA_TRY(cudaEventRecord(c.ev[4], st));
  for (int k = 0; k < RELZ_NOUT; ++k)   // pageable destinations (R vectors, numpy arrays) go through pinned staging buffers
    if (out[k]) RC(d2h_2d(out[k], (size_t)n * sizeof(double), O.o[k], (size_t)n * sizeof(double), (size_t)n * sizeof(double), 1, st));
  RC(hostcopy_flush());
  CUDA_TRY(cudaEventRecord(c.ev[5], st));
  CUDA_TRY(cudaStreamSynchronize(st));
  cudaEventElapsedTime(&c.tm.h2d_ms, c.ev[6], c.ev[7]);
  cudaEventElapsedTime(&c.tm.relz_ms, c.ev[7], c.ev[4]);
  cudaEventElapsedTime(&c.tm.d2h_ms, c.ev[4], c.ev[5]);
  c.tm.total_ms = c.tm.h2d_ms + c.tm.relz_ms + c.tm.d2h_ms;
  c.tm.kernel_launches = c.launches - l0;
  return LITAP_OK;
}
