// ddir.cu — calc_ddir2 on the device (R/flow_01_calc_ddir.R:2-118).
//   k_elev_max / k_invert : invert()            R/flow_06_invert.R:1-5
//   k_d8                  : finddir2            R/LITAP_functions.R:257-273  (3x3 stencil, 9 B/cell)
//   k_collect_flats       : worklist of ddir==5 cells
//   k_plateau_probe/apply : synchronous plateau rounds   R/flow_01_calc_ddir.R:12-41
//   k_patch_mark + host   : flat-pit patches + get_dir   R/flow_01_calc_ddir.R:48-107
#include <algorithm>
#include <unordered_map>
#include <vector>
#include "cellmath.h"
#include "litap_internal.h"

namespace {

__device__ __forceinline__ unsigned long long f64_orderable(double x) {
  unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double f64_from_orderable(unsigned long long k) {
  unsigned long long b = (k & 0x8000000000000000ull) ? (k & 0x7FFFFFFFFFFFFFFFull) : ~k;
  return __longlong_as_double((long long)b);
}

// ---- invert ------------------------------------------------------------------------------------
__global__ void k_elev_max(const double* __restrict__ elev, GridDims d, unsigned long long* out) {
  unsigned long long best = 0;  // below every orderable finite value
  for (int r = blockIdx.y; r < d.NR; r += gridDim.y)
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

// ---- D8 stencil --------------------------------------------------------------------------------
// One thread per column, a warp = 32 consecutive columns starting on an aligned interior column,
// each thread slides a 3-row register window down D8_ROWS rows: one coalesced 256 B row load per
// warp per row; left/right neighbours come from shuffles (+1 halo load on lanes 0 and 31).
constexpr int D8_ROWS = 32;
constexpr int D8_THREADS = 128;

struct Row3 { double l, m, r; };

__device__ __forceinline__ Row3 load_row3(const double* __restrict__ elev, const GridDims& d, int r, int c, int lane) {
  const double NA = __longlong_as_double(0x7ff8000000000000LL);
  Row3 w;
  const bool rv = (r >= 0 && r < d.NR);
  const double* row = elev + (size_t)(rv ? r : 0) * d.pitch;
  w.m = (rv && c < d.NC) ? __ldg(row + c) : NA;
  w.l = __shfl_up_sync(0xffffffffu, w.m, 1);
  w.r = __shfl_down_sync(0xffffffffu, w.m, 1);
  if (lane == 0) w.l = (rv && c - 1 >= 0 && c - 1 < d.NC) ? __ldg(row + c - 1) : NA;
  if (lane == 31) w.r = (rv && c + 1 < d.NC) ? __ldg(row + c + 1) : NA;
  return w;
}

__global__ void __launch_bounds__(D8_THREADS) k_d8(const double* __restrict__ elev, GridDims d,
                                                   uint8_t* __restrict__ ddir, uint8_t* __restrict__ flat,
                                                   unsigned long long* nflat) {
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * D8_THREADS + threadIdx.x + 1;  // column 1 is the aligned one
  const int r0 = blockIdx.y * D8_ROWS;
  Row3 p = load_row3(elev, d, r0 - 1, c, lane);
  Row3 q = load_row3(elev, d, r0, c, lane);
  unsigned cnt = 0;
  const int rend = min(r0 + D8_ROWS, d.NR);
  for (int r = r0; r < rend; ++r) {
    Row3 n = load_row3(elev, d, r + 1, c, lane);
    double nb[10];
    nb[7] = p.l; nb[8] = p.m; nb[9] = p.r;
    nb[4] = q.l; nb[5] = q.m; nb[6] = q.r;
    nb[1] = n.l; nb[2] = n.m; nb[3] = n.r;
    int dd = d8_dir(q.m, nb);
    if (c < d.NC) {
      size_t i = (size_t)r * d.pitch + c;
      ddir[i] = (uint8_t)dd;
      flat[i] = dd == 0 ? 2 : (dd == 5);
      cnt += (dd == 5);
      if (c == 1) { ddir[i - 1] = 0; flat[i - 1] = 2; }  // ring column 0 (always NA)
    }
    p = q;
    q = n;
  }
  for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if (lane == 0 && cnt) atomicAdd(nflat, (unsigned long long)cnt);
}

// ---- flat worklist -----------------------------------------------------------------------------
__global__ void k_collect_flats(const uint8_t* __restrict__ ddir, GridDims d, uint32_t* __restrict__ list,
                                unsigned long long* cursor) {
  const int lane = threadIdx.x & 31;
  const int c = blockIdx.x * blockDim.x + threadIdx.x + 1;
  const int r = blockIdx.y;
  bool f = (c < d.NC) && ddir[(size_t)r * d.pitch + c] == 5;
  unsigned m = __ballot_sync(0xffffffffu, f);
  if (!m) return;
  unsigned long long base = 0;
  if (lane == 0) base = atomicAdd(cursor, (unsigned long long)__popc(m));
  base = __shfl_sync(0xffffffffu, base, 0);
  if (f) list[base + __popc(m & ((1u << lane) - 1))] = (uint32_t)((size_t)r * d.pitch + c);
}

// ---- plateau rounds (Jacobi: probe reads the previous round's ddir, apply writes) ---------------
__global__ void k_plateau_probe(const uint32_t* __restrict__ list, unsigned n, const uint8_t* __restrict__ ddir,
                                const double* __restrict__ elev, GridDims d, uint8_t* __restrict__ cand,
                                unsigned* changed) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  bool any = false;
  if (i < n) {
    uint32_t id = list[i];
    uint8_t res = 0;
    if (ddir[id] == 5) {
      const double e = elev[id];
      const int r = id / d.pitch, c = id % d.pitch;
#pragma unroll
      for (int k = 1; k <= 9; ++k) {
        if (k == 5 || res) continue;
        int rr = r + nb_dr(k), cc = c + nb_dc(k);
        if (rr < 0 || rr >= d.NR || cc < 0 || cc >= d.NC) continue;
        size_t j = (size_t)rr * d.pitch + cc;
        double en = elev[j];
        if (is_nan(en) || !(e - en >= 0)) continue;
        uint8_t dn = ddir[j];
        if (dn == 5 || dn == 0) continue;
        res = (uint8_t)k;
      }
    }
    cand[i] = res;
    any = res != 0;
  }
  if (__syncthreads_or(any) && threadIdx.x == 0) atomicAdd(changed, 1u);
}
__global__ void k_plateau_apply(const uint32_t* __restrict__ list, unsigned n, const uint8_t* __restrict__ cand,
                                uint8_t* __restrict__ ddir) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && cand[i]) ddir[list[i]] = cand[i];
}

// ---- flat-pit patch cells: ddir==5 with an 8-neighbour ddir==5 ----------------------------------
__global__ void k_patch_mark(const uint32_t* __restrict__ list, unsigned n, const uint8_t* __restrict__ ddir,
                             GridDims d, uint32_t* __restrict__ plist, unsigned* pcount) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t id = list[i];
  if (ddir[id] != 5) return;
  const int r = id / d.pitch, c = id % d.pitch;
  bool hit = false;
#pragma unroll
  for (int k = 1; k <= 9; ++k) {
    if (k == 5) continue;
    int rr = r + nb_dr(k), cc = c + nb_dc(k);
    if (rr < 0 || rr >= d.NR || cc < 0 || cc >= d.NC) continue;
    hit |= ddir[(size_t)rr * d.pitch + cc] == 5;
  }
  if (hit) plist[atomicAdd(pcount, 1u)] = id;
}
__global__ void k_patch_apply(const uint32_t* __restrict__ ids, const uint8_t* __restrict__ dirs, unsigned n,
                              uint8_t* __restrict__ ddir) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ddir[ids[i]] = dirs[i];
}

// Host part of the patch step: components, median centre, get_dir.  The patch cell set is tiny
// (a few cells per 10^5) and the reference's rule is a per-patch sequential one
// (R/flow_01_calc_ddir.R:59-106); only the cell ids travel, never a raster.
void resolve_patches_host(std::vector<uint32_t>& P, int pitch, std::vector<uint8_t>& newdir) {
  std::sort(P.begin(), P.end());
  const size_t n = P.size();
  newdir.assign(n, 5);
  std::unordered_map<uint32_t, uint32_t> pos;
  pos.reserve(n * 2);
  for (size_t i = 0; i < n; ++i) pos[P[i]] = (uint32_t)i;
  std::vector<int> comp(n, -1);
  std::vector<uint32_t> stack, members;
  std::vector<long long> cols, rows;
  auto nbr = [&](uint32_t id, int k) -> long long {
    return (long long)id + (long long)nb_dr(k) * pitch + nb_dc(k);
  };
  for (size_t s = 0; s < n; ++s) {
    if (comp[s] >= 0) continue;
    members.clear();
    stack.assign(1, (uint32_t)s);
    comp[s] = (int)s;
    while (!stack.empty()) {
      uint32_t k = stack.back();
      stack.pop_back();
      members.push_back(k);
      for (int dn = 1; dn <= 9; ++dn) {
        if (dn == 5) continue;
        auto it = pos.find((uint32_t)nbr(P[k], dn));
        if (it != pos.end() && comp[it->second] < 0) { comp[it->second] = (int)s; stack.push_back(it->second); }
      }
    }
    std::sort(members.begin(), members.end());  // ascending cell id == ascending seqno
    cols.clear(); rows.clear();
    for (uint32_t k : members) { cols.push_back(P[k] % pitch + 1); rows.push_back(P[k] / pitch + 1); }
    auto ceil_median2 = [](std::vector<long long> v) {  // ceiling(median(v)) for integers
      std::sort(v.begin(), v.end());
      size_t m = v.size();
      if (m % 2) return v[m / 2];
      long long s2 = v[m / 2 - 1] + v[m / 2];  // median = s2/2; coordinates are positive
      return (s2 + 1) / 2;
    };
    long long cc = ceil_median2(cols), cr = ceil_median2(rows);
    long long best = -1;
    uint32_t centre = members[0];
    for (size_t t = 0; t < members.size(); ++t) {
      long long dd = (cols[t] - cc) * (cols[t] - cc) + (rows[t] - cr) * (rows[t] - cr);
      if (best < 0 || dd < best) { best = dd; centre = members[t]; }  // first minimum = lowest seqno
    }
    long long row_f = P[centre] / pitch + 1, col_f = P[centre] % pitch + 1;
    for (uint32_t k : members) {
      unsigned opts = 0;
      for (int dn = 1; dn <= 9; ++dn) {
        if (dn == 5) continue;
        if (pos.count((uint32_t)nbr(P[k], dn))) opts |= 1u << dn;
      }
      newdir[k] = (uint8_t)get_dir_int(P[k] / pitch + 1, P[k] % pitch + 1, row_f, col_f, opts);
    }
  }
}

}  // namespace

int run_ddir(litap_grid* g, int invert) {
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  if (!g->ddir) g->ddir = grid_alloc_array<uint8_t>(d, 0);
  if (!g->flat) g->flat = grid_alloc_array<uint8_t>(d, 2);
  if (!g->ddir || !g->flat) return set_error(LITAP_ENOMEM, "device allocation failed (ddir)");
  // scratch: [0] flat count, [1] cursor, [2] elev max, [3] changed/pcount (as 2 x u32)
  unsigned long long* scr = (unsigned long long*)pool_alloc(8 * sizeof(unsigned long long));
  if (!scr) return set_error(LITAP_ENOMEM, "device allocation failed (scratch)");
  struct Free { void* p; ~Free() { if (p) pool_free(p); } } f_scr{scr};
  CUDA_TRY(cudaMemsetAsync(scr, 0, 8 * sizeof(unsigned long long), st));

  if (invert) {
    dim3 gb((d.NC + 255) / 256 < 64 ? (d.NC + 255) / 256 : 64, d.NR < 1024 ? d.NR : 1024);
    KERNEL("k_elev_max", k_elev_max<<<gb, 256, 0, st>>>(g->elev, d, scr + 2));
    dim3 gi((d.NC + 31 + 255) / 256, d.NR);
    KERNEL("k_invert", k_invert<<<gi, 256, 0, st>>>(g->elev, d, scr + 2));
  }
  {
    dim3 grid((d.NC - 1 + D8_THREADS - 1) / D8_THREADS, (d.NR + D8_ROWS - 1) / D8_ROWS);
    KERNEL("k_d8", k_d8<<<grid, D8_THREADS, 0, st>>>(g->elev, d, g->ddir, g->flat, scr));
  }
  unsigned long long nflat = 0;
  CUDA_TRY(cudaMemcpyAsync(&nflat, scr, sizeof(nflat), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  c.tm.flat_cells = (int64_t)nflat;
  c.tm.plateau_rounds = 0;
  c.tm.patch_cells = 0;
  if (nflat == 0) return LITAP_OK;
  if (nflat > 0xFFFFFFF0ull) return set_error(LITAP_EARG, "too many flat cells");
  const unsigned n = (unsigned)nflat;
  uint32_t* list = (uint32_t*)pool_alloc((size_t)n * sizeof(uint32_t));
  uint32_t* plist = (uint32_t*)pool_alloc((size_t)n * sizeof(uint32_t));
  uint8_t* cand = (uint8_t*)pool_alloc((size_t)n);
  Free f1{list}, f2{plist}, f3{cand};
  if (!list || !plist || !cand) return set_error(LITAP_ENOMEM, "device allocation failed (flat worklist)");
  {
    dim3 grid((d.NC - 1 + 255) / 256, d.NR);
    KERNEL("k_collect_flats", k_collect_flats<<<grid, 256, 0, st>>>(g->ddir, d, list, scr + 1));
  }
  unsigned* changed = (unsigned*)(scr + 3);
  const unsigned nblk = (n + 255) / 256;
  // Every flat cell that can drain is at BFS distance <= nflat from a draining cell.
  for (unsigned round = 0; round <= n; ++round) {
    CUDA_TRY(cudaMemsetAsync(changed, 0, sizeof(unsigned), st));
    KERNEL("k_plateau_probe", k_plateau_probe<<<nblk, 256, 0, st>>>(list, n, g->ddir, g->elev, d, cand, changed));
    unsigned h = 0;
    CUDA_TRY(cudaMemcpyAsync(&h, changed, sizeof(h), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (!h) break;
    KERNEL("k_plateau_apply", k_plateau_apply<<<nblk, 256, 0, st>>>(list, n, cand, g->ddir));
    c.tm.plateau_rounds++;
  }
  unsigned* pcount = changed + 1;
  KERNEL("k_patch_mark", k_patch_mark<<<nblk, 256, 0, st>>>(list, n, g->ddir, d, plist, pcount));
  unsigned np = 0;
  CUDA_TRY(cudaMemcpyAsync(&np, pcount, sizeof(np), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  c.tm.patch_cells = np;
  if (np) {
    std::vector<uint32_t> P(np);
    std::vector<uint8_t> nd;
    CUDA_TRY(cudaMemcpyAsync(P.data(), plist, (size_t)np * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    resolve_patches_host(P, d.pitch, nd);
    // ids are relative to the XOFF-shifted pointer on both sides (plist was built from it)
    CUDA_TRY(cudaMemcpyAsync(plist, P.data(), (size_t)np * 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaMemcpyAsync(cand, nd.data(), (size_t)np, cudaMemcpyHostToDevice, st));
    KERNEL("k_patch_apply", k_patch_apply<<<(np + 255) / 256, 256, 0, st>>>(plist, cand, np, g->ddir));
    CUDA_TRY(cudaStreamSynchronize(st));  // P / nd are stack-owned
  }
  return LITAP_OK;
}
