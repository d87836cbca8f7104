// uced.cu — calc_upslopes(type = "uced") on the device: upslope cumulative elevation drop
// (R/flow_02_calc_shed.R:60-145; uced_values :118-130; called by pit removal, R/flow_04_remove_pit_fast.R:73,345).
//
// The reference visits the cells of a shed in the order (elev desc, seqno asc).  A cell that has no value yet takes
// its whole flow track down to the pit; the cells of the track that had no value (always a prefix of it: the SEGMENT
// the cell heads) each add   (double) sum_longdouble(elev[segment cells above] - elev[cell])   to every cell below
// them on the track, the additions to one cell happening in visiting order of the heads.  Neither the long double
// sums nor the order of the double additions can be changed without changing bits, so the device reproduces both:
//
//   k_uced_keys + cub radix sort    rank of every cell in the visiting order (one stable descending sort of the
//                                   order-preserving integer images of the elevations; ties stay in seqno order)
//   k_uced_mwalk                    m[c] = lowest rank in the upslope tree of c: every cell that precedes all its
//                                   donors walks down its track with atomicMin and stops where a lower rank has been
//   heads                           h with m[h] == rank[h]; its segment is the stretch of its track on which m == rank[h]
//   k_uced_count                    a thread per head walks its track: contributions per cell and in total
//   k_uced_fill                     the same walk with the arithmetic: per (head, cell below it) one long double sum
//                                   (cellmath.h lds_*: 64-bit significand, round to nearest even after every add) over the
//                                   segment's elevations, emitted as (cell << 32 | rank of the head, value)
//   cub radix sort + k_uced_sum     contributions of a cell in visiting order of their heads, added left to right
//
// The work is that of the reference: the sum over all cells of the cells upslope of them (~17 per cell on the
// bench DEM).  When the contributions of a DEM do not fit the memory budget the cells are processed in ranges
// (each range repeats the walk).  Output NaN = NA (cells without a shed; pits nothing flows into).
#include <cub/cub.cuh>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include "cellmath.h"
#include "litap_internal.h"

namespace uced {

constexpr uint32_t NONE = 0xFFFFFFFFu;
constexpr int UT = 256;
constexpr int KSEG = 24;   // elevations of a segment kept in registers / local memory; longer segments re-walk

struct Arrays {
  const double* elev;
  const uint8_t* ddir;
  const int32_t* shed;
  uint32_t* rank;   // visiting order, NONE for cells without a shed
  uint32_t* m;      // lowest rank in the upslope tree
  uint32_t* cnt;    // contributions per cell
  unsigned long long* total;   // [0] contributions, [1] emitted, [2] error flag (track longer than the DEM: cycle)
  int pitch;
  unsigned maxsteps;
};

__device__ __forceinline__ unsigned long long f64_orderable(double x) {
  unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ int nb_off(int n, int pitch) { return nb_dr(n) * pitch + nb_dc(n); }
// the next cell of the track, NONE at its end (pit, NA direction, or a direction that leaves the shed)
__device__ __forceinline__ uint32_t next_cell(const Arrays& A, uint32_t id) {
  const int dv = A.ddir[id];
  if (dv < 1 || dv > 9 || dv == 5) return NONE;
  const uint32_t n = id + nb_off(dv, A.pitch);
  return A.shed[n] == A.shed[id] ? n : NONE;
}

// sort keys of all pitched cells: the image of the elevation for cells with a shed, 0 (sorts last) for the others
__global__ void __launch_bounds__(UT) k_uced_keys(Arrays A, GridDims d, unsigned long long* __restrict__ keys,
                                                  uint32_t* __restrict__ ids, unsigned* nvalid) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)d.cells) return;
  const int r = (int)(i / d.pitch), c = (int)(i % d.pitch);
  bool valid = false;
  unsigned long long k = 0;
  if (c < d.NC && r > 0 && r < d.NR - 1 && c > 0 && c < d.NC - 1) {
    const double e = A.elev[i];
    if (A.shed[i] > 0 && !is_nan(e)) { valid = true; k = f64_orderable(e + 0.0); }
  }
  keys[i] = k;
  ids[i] = (uint32_t)i;
  A.rank[i] = NONE;
  A.cnt[i] = 0;
  const unsigned mv = __ballot_sync(0xFFFFFFFFu, valid);
  if ((threadIdx.x & 31) == 0 && mv) atomicAdd(nvalid, __popc(mv));
}
__global__ void __launch_bounds__(UT) k_uced_rank(Arrays A, const uint32_t* __restrict__ sorted_ids, unsigned nvalid) {
  const unsigned k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nvalid) return;
  const uint32_t id = sorted_ids[k];
  A.rank[id] = k;
  A.m[id] = k;
}

// cells that precede all their donors carry their rank down the track
__global__ void __launch_bounds__(UT) k_uced_mwalk(Arrays A, GridDims d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d.NC) return;
  const uint32_t id = (uint32_t)((size_t)r * d.pitch + c);
  const uint32_t rk = A.rank[id];
  if (rk == NONE) return;
#pragma unroll
  for (int n = 1; n <= 9; ++n) {
    if (n == 5) continue;
    const uint32_t nid = id + nb_off(n, d.pitch);
    if (A.ddir[nid] == 10 - n && A.shed[nid] == A.shed[id] && A.rank[nid] < rk) return;   // a donor comes first
  }
  uint32_t x = id;
  for (unsigned step = 0; step < A.maxsteps; ++step) {
    x = next_cell(A, x);
    if (x == NONE) return;
    if (atomicMin(&A.m[x], rk) < rk) return;
  }
  A.total[2] = 1;
}

// a thread per head: length of its track
__global__ void __launch_bounds__(UT) k_uced_count(Arrays A, GridDims d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  unsigned long long n = 0;
  if (c < d.NC) {
    const uint32_t id = (uint32_t)((size_t)r * d.pitch + c);
    const uint32_t rk = A.rank[id];
    if (rk != NONE && A.m[id] == rk) {
      uint32_t x = id;
      unsigned step = 0;
      for (; step < A.maxsteps; ++step) {
        x = next_cell(A, x);
        if (x == NONE) break;
        atomicAdd(&A.cnt[x], 1u);
        ++n;
      }
      if (step == A.maxsteps) A.total[2] = 1;
    }
  }
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xFFFFFFFFu, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(&A.total[0], n);
}

// cell ranges [bound[j], bound[j+1]) holding about `per` contributions each (pre = inclusive scan of cnt)
__global__ void k_uced_bounds(const unsigned long long* __restrict__ pre, size_t ncell, unsigned long long per, int nchunk,
                              unsigned* __restrict__ bound) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j > nchunk) return;
  if (j == nchunk) { bound[j] = (unsigned)ncell; return; }
  const unsigned long long want = per * (unsigned long long)j;
  size_t lo = 0, hi = ncell;   // first cell whose inclusive prefix exceeds want
  while (lo < hi) {
    const size_t mid = (lo + hi) >> 1;
    if (pre[mid] <= want) lo = mid + 1; else hi = mid;
  }
  bound[j] = (unsigned)lo;
}

// a thread per head: the contributions of its segment to the cells of its track that lie in [c_lo, c_hi)
__global__ void __launch_bounds__(UT) k_uced_fill(Arrays A, GridDims d, uint32_t c_lo, uint32_t c_hi,
                                                  unsigned long long* __restrict__ okeys, double* __restrict__ ovals,
                                                  unsigned long long cap) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  const int lane = threadIdx.x & 31;
  uint32_t head = NONE, rk = NONE;
  if (c < d.NC) {
    const uint32_t id = (uint32_t)((size_t)r * d.pitch + c);
    rk = A.rank[id];
    if (rk != NONE && A.m[id] == rk) head = id;
  }
  double seg[KSEG];      // elevations of the segment, head first
  int k = 0;             // cells of the segment seen so far
  bool inseg = true;     // the walk is still inside the segment
  uint32_t x = head;
  if (head != NONE) { seg[0] = A.elev[head]; k = 1; }
  bool alive = head != NONE;
  unsigned step = 0;
  while (__any_sync(0xFFFFFFFFu, alive)) {
    bool emit = false;
    unsigned long long key = 0;
    double val = 0;
    if (alive) {
      x = next_cell(A, x);
      if (x == NONE || ++step > A.maxsteps) alive = false;
      else {
        const double ec = A.elev[x];
        if (inseg && A.m[x] != rk) inseg = false;   // x had its value before this head was visited
        if (x >= c_lo && x < c_hi) {
          // sum(elev[segment cells above x] - elev[x]) with R's long double accumulator (:126)
          Lds s;
          s.m = 0; s.e = 0; s.neg = 0;
          const int kk = k < KSEG ? k : KSEG;
          for (int j = 0; j < kk; ++j) s = lds_add(s, lds_from_double(seg[j] - ec));
          if (k > KSEG) {   // a long segment: its tail is read again from the DEM
            uint32_t y = head;
            for (int j = 0; j < k; ++j) {
              if (j >= KSEG) s = lds_add(s, lds_from_double(A.elev[y] - ec));
              y = next_cell(A, y);
            }
          }
          emit = true;
          key = ((unsigned long long)x << 32) | rk;
          val = lds_to_double(s);
        }
        if (inseg) {
          if (k < KSEG) seg[k] = ec;
          ++k;
        }
      }
    }
    const unsigned em = __ballot_sync(0xFFFFFFFFu, emit);
    if (em) {
      unsigned long long base = 0;
      const int leader = __ffs(em) - 1;
      if (lane == leader) base = atomicAdd(&A.total[1], (unsigned long long)__popc(em));
      base = __shfl_sync(0xFFFFFFFFu, base, leader);
      if (emit) {
        const unsigned long long at = base + __popc(em & ((1u << lane) - 1));
        if (at < cap) { okeys[at] = key; ovals[at] = val; }
      }
    }
  }
}

__global__ void __launch_bounds__(UT) k_uced_init_out(Arrays A, GridDims d, double* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d.NC) return;
  const uint32_t id = (uint32_t)((size_t)r * d.pitch + c);
  double v = __longlong_as_double(0x7ff8000000000000LL);
  // a cell with a track of two or more cells gets 0 + its contributions (:98, :128-129); a pit only through the
  // tracks that end in it
  if (A.rank[id] != NONE && next_cell(A, id) != NONE) v = 0.0;
  out[id] = v;
}

// sorted contributions: the first of a cell's run adds the run left to right (visiting order of the heads)
__global__ void __launch_bounds__(UT) k_uced_sum(const unsigned long long* __restrict__ keys, const double* __restrict__ vals,
                                                 unsigned long long n, double* __restrict__ out) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t c = (uint32_t)(keys[i] >> 32);
  if (i && (uint32_t)(keys[i - 1] >> 32) == c) return;
  double acc = 0.0;
  for (unsigned long long j = i; j < n && (uint32_t)(keys[j] >> 32) == c; ++j) acc = acc + vals[j];
  out[c] = acc;
}

}  // namespace uced

namespace {
inline unsigned nblk(size_t n, unsigned t = 256) { return (unsigned)((n + t - 1) / t); }
}

int run_uced(litap_grid* g) {
  using namespace uced;
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  if (!g->uced) g->uced = grid_alloc_array<double>(d, -1);
  if (!g->uced) return set_error(LITAP_ENOMEM, "device allocation failed (uced)");
  const size_t ncell = (size_t)d.cells;
  PoolScratch S;
  Arrays A = {};
  A.elev = g->elev;
  A.ddir = g->ddir;
  A.shed = g->shed;
  A.rank = S.get<uint32_t>(ncell + 64);
  A.m = S.get<uint32_t>(ncell + 64);
  A.cnt = S.get<uint32_t>(ncell + 64);
  A.total = S.get<unsigned long long>(4);
  A.pitch = d.pitch;
  A.maxsteps = (unsigned)std::min<size_t>((size_t)d.NR * d.NC, 0x7FFFFFFFu);
  unsigned* nvalid_d = S.get<unsigned>(1);
  unsigned long long* keys_a = S.get<unsigned long long>(ncell);
  unsigned long long* keys_b = S.get<unsigned long long>(ncell);
  uint32_t* ids_a = S.get<uint32_t>(ncell);
  uint32_t* ids_b = S.get<uint32_t>(ncell);
  size_t t1 = 0;
  cub::DeviceRadixSort::SortPairsDescending(nullptr, t1, keys_a, keys_b, ids_a, ids_b, (int64_t)ncell, 0, 64, st);
  void* tmp1 = S.get<unsigned char>(t1 + 256);
  if (!A.rank || !A.m || !A.cnt || !A.total || !nvalid_d || !keys_a || !keys_b || !ids_a || !ids_b || !tmp1)
    return set_error(LITAP_ENOMEM, "device allocation failed (uced scratch)");
  CUDA_TRY(cudaMemsetAsync(A.total, 0, 4 * sizeof(unsigned long long), st));
  CUDA_TRY(cudaMemsetAsync(nvalid_d, 0, sizeof(unsigned), st));
  CUDA_TRY(cudaMemsetAsync(A.m, 0xFF, (ncell + 64) * sizeof(uint32_t), st));
  KERNEL("k_uced_keys", k_uced_keys<<<nblk(ncell, UT), UT, 0, st>>>(A, d, keys_a, ids_a, nvalid_d));
  CUDA_TRY(cub::DeviceRadixSort::SortPairsDescending(tmp1, t1, keys_a, keys_b, ids_a, ids_b, (int64_t)ncell, 0, 64, st));
  c.launches += 1;
  unsigned nvalid = 0;
  CUDA_TRY(cudaMemcpyAsync(&nvalid, nvalid_d, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  dim3 cgrid(nblk(d.NC, UT), d.NR);
  if (nvalid) {
    KERNEL("k_uced_rank", k_uced_rank<<<nblk(nvalid, UT), UT, 0, st>>>(A, ids_b, nvalid));
    KERNEL("k_uced_mwalk", k_uced_mwalk<<<cgrid, UT, 0, st>>>(A, d));
    KERNEL("k_uced_count", k_uced_count<<<cgrid, UT, 0, st>>>(A, d));
  }
  KERNEL("k_uced_init_out", k_uced_init_out<<<cgrid, UT, 0, st>>>(A, d, g->uced));
  unsigned long long tot[4] = {0, 0, 0, 0};
  CUDA_TRY(cudaMemcpyAsync(tot, A.total, sizeof(tot), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  if (tot[2]) return set_error(LITAP_EARG, "litap_uced: a flow track does not end (cycle in ddir)");
  const unsigned long long npairs = tot[0];
  c.tm.uced_pairs = (double)npairs;
  if (!npairs) return LITAP_OK;
  // the rank sort's buffers are free again: return them before the contributions are allocated
  // (PoolScratch frees at scope end; the big ones are released explicitly)
  for (void* p : {(void*)keys_a, (void*)keys_b, (void*)ids_a, (void*)ids_b, tmp1}) {
    S.ptrs.erase(std::find(S.ptrs.begin(), S.ptrs.end(), p));
    pool_free(p);
  }
  // memory budget of the contributions: keys + values, double buffered, + the sort's scratch (~ one more buffer)
  size_t free_b = 0, total_b = 0;
  CUDA_TRY(cudaMemGetInfo(&free_b, &total_b));
  unsigned long long budget = (unsigned long long)(free_b / 2) / 40;
  if (const char* e = getenv("LITAP_UCED_PAIRS")) budget = strtoull(e, nullptr, 10);   // tests: force several ranges
  budget = std::max<unsigned long long>(budget, 1024);
  const int nchunk = (int)((npairs + budget - 1) / budget);
  std::vector<unsigned> bound(nchunk + 1, 0);
  bound[nchunk] = (unsigned)ncell;
  unsigned long long cap = npairs;
  if (nchunk > 1) {
    unsigned long long* pre = S.get<unsigned long long>(ncell);
    unsigned* bound_d = S.get<unsigned>(nchunk + 1);
    size_t t2 = 0;
    cub::DeviceScan::InclusiveSum(nullptr, t2, A.cnt, pre, (int64_t)ncell, st);
    void* tmp2 = S.get<unsigned char>(t2 + 256);
    if (!pre || !bound_d || !tmp2) return set_error(LITAP_ENOMEM, "device allocation failed (uced ranges)");
    CUDA_TRY(cub::DeviceScan::InclusiveSum(tmp2, t2, A.cnt, pre, (int64_t)ncell, st));
    const unsigned long long per = (npairs + nchunk - 1) / nchunk;
    KERNEL("k_uced_bounds", k_uced_bounds<<<nblk(nchunk + 1, 64), 64, 0, st>>>(pre, ncell, per, nchunk, bound_d));
    CUDA_TRY(cudaMemcpyAsync(bound.data(), bound_d, (nchunk + 1) * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    // a range ends with a whole cell, so it can exceed `per` by one cell's contributions: size for the largest
    std::vector<unsigned long long> hp(1);
    cap = 0;
    for (int j = 0; j < nchunk; ++j) {
      unsigned long long a = 0, b = 0;
      if (bound[j] > 0) CUDA_TRY(cudaMemcpy(&a, pre + bound[j] - 1, 8, cudaMemcpyDeviceToHost));
      if (bound[j + 1] > 0) CUDA_TRY(cudaMemcpy(&b, pre + bound[j + 1] - 1, 8, cudaMemcpyDeviceToHost));
      cap = std::max(cap, b - a);
    }
  }
  unsigned long long* pk_a = S.get<unsigned long long>(cap);
  unsigned long long* pk_b = S.get<unsigned long long>(cap);
  double* pv_a = S.get<double>(cap);
  double* pv_b = S.get<double>(cap);
  size_t t3 = 0;
  int end_bit = 33;
  while (end_bit < 64 && (ncell >> (end_bit - 32)) != 0) ++end_bit;
  cub::DeviceRadixSort::SortPairs(nullptr, t3, pk_a, pk_b, pv_a, pv_b, (int64_t)cap, 0, end_bit, st);
  void* tmp3 = S.get<unsigned char>(t3 + 256);
  if (!pk_a || !pk_b || !pv_a || !pv_b || !tmp3)
    return set_error(LITAP_ENOMEM, "device allocation failed (uced contributions: " + std::to_string(cap) + " pairs)");
  for (int j = 0; j < nchunk; ++j) {
    if (bound[j + 1] <= bound[j]) continue;
    CUDA_TRY(cudaMemsetAsync(A.total + 1, 0, sizeof(unsigned long long), st));
    KERNEL("k_uced_fill", k_uced_fill<<<cgrid, UT, 0, st>>>(A, d, bound[j], bound[j + 1], pk_a, pv_a, cap));
    unsigned long long got = 0;
    CUDA_TRY(cudaMemcpyAsync(&got, A.total + 1, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (got > cap) return set_error(LITAP_ECUDA, "litap_uced: internal: more contributions than counted");
    if (!got) continue;
    size_t t4 = t3;
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp3, t4, pk_a, pk_b, pv_a, pv_b, (int64_t)got, 0, end_bit, st));
    c.launches += 1;
    KERNEL("k_uced_sum", k_uced_sum<<<nblk(got, UT), UT, 0, st>>>(pk_b, pv_b, got, g->uced));
  }
  CUDA_TRY(cudaStreamSynchronize(st));   // scratch goes back to the pool
  c.tm.uced_ranges = nchunk;
  return LITAP_OK;
}
