// weti.cu — calc_weti2 on the device (R/form_02b_calc_weti.R:3-110): Quinn-style multiple-flow-
// direction upslope area (qarea1/2) and wetness index (qweti1/2).
//
// The reference visits the cells in the order (elev desc, upslope asc, seqno asc) (:23-26) and
// PUSHES qc * tan_k into its lower neighbours (:69-77) or, for a cell without a lower neighbour
// that is not a pit, its whole qa into drec (:79-84).  fp64 addition is not associative and the
// outputs are truncated to 2 decimals, so every cell must receive its contributions in exactly that
// order.  The device version is a deterministic PULL over the same dependency DAG:
//
//   k_weti_cls      class of every cell (A: S != 0, B: S == 0 & ddir != 5, C: pit), S with R's long
//                   double accumulator (cellmath.h ld80_*), lower-neighbour mask, sort keys
//   cub radix sort  the reference's visiting order: stable LSD, upslope asc then elev desc, over
//                   cells enumerated by seqno -> rank[cell]
//   k_weti_donors   per cell: its donors that precede it (strictly higher neighbours; B neighbours
//                   whose drec is the cell and whose rank is lower), sorted by rank, packed in 32
//                   bits; pending counter; cells without donors seed wave 0
//   k_weti_waves    ONE cooperative launch: wave w = cells whose donors are all done; a ready cell
//                   adds its donors' contributions in rank order (no FP atomics), derives qc, and
//                   decrements its successors' counters (packed 8-bit counters, 32-bit atomics)
//   k_weti_flags + cub scan   the reference's qc1/qc2 are variables that survive between loop
//                   iterations (:70-71,:80 vs :87-88): a B cell reads the qc2, a pit the qc1 and
//                   qc2 of the last cell before it in the visiting order that assigned them ->
//                   inclusive max-scan of "last setter position" over the sorted order
//   k_weti_final    late contributions (B donors ranked after their receiver only change the
//                   final qa), qweti from the effective qc, trunc_dec(., 2), NA
//
// Algorithmic traffic 45 B/cell (elev 8 + ddir 1 + upslope 4 read, 4 x 8 written); the sort and the
// per-cell state make the real traffic several times that: this stage is latency/ordering-bound.
#include <cooperative_groups.h>
#include <cub/cub.cuh>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "cellmath.h"
#include "litap_internal.h"

namespace cg = cooperative_groups;

namespace weti {

enum : uint8_t { CLS_NA = 0, CLS_A = 1, CLS_B = 2, CLS_C = 3 };
constexpr uint8_t F_BEARLY = 0x10;  // Meta.cls flag: B cell that precedes its drec in the visiting order

struct __align__(16) Meta {
  double S;          // sum(tan) of the cell (A cells: != 0)
  uint32_t donors;   // up to 8 nibbles: direction codes (from the cell towards the donor) in rank order, 0 ends
  uint16_t lower;    // bit n: neighbour n is strictly lower (successors of an A cell)
  uint8_t cls;       // CLS_* | F_BEARLY
  uint8_t bmask;     // bit j: j-th donor is a B donor (contributes its whole qa)
};
struct __align__(32) State {
  double qc1, qc2;   // as assigned by the cell itself (NaN where the reference leaves the variable stale)
  double qa1, qa2;   // qa at processing time
};

struct Arrays {
  const double* elev;
  const uint8_t* ddir;
  const uint32_t* upslope;
  Meta* meta;
  State* state;
  uint8_t* cls;        // CLS_* only (neighbour lookups)
  uint32_t* rank;      // position in the reference's visiting order
  uint32_t* pend32;    // packed 8-bit pending-donor counters
  uint32_t* order;     // ready queue = a topological order; wave w is a contiguous segment
};

__device__ __forceinline__ unsigned long long f64_orderable(double x) {
  unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ int nb_off(int n, int pitch) { return nb_dr(n) * pitch + nb_dc(n); }

// ------------------------------------------------------------------------------------------------
// class, S, lower mask; sort input (dense seqno enumeration: i = r*NC + c)
__global__ void k_weti_cls(Arrays A, GridDims d, WetiK K, uint32_t* __restrict__ key_up,
                           uint32_t* __restrict__ val_id) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d.NC) return;
  const size_t id = (size_t)r * d.pitch + c;
  const size_t i = (size_t)r * d.NC + c;
  const double f = A.elev[id];
  const bool ring = r == 0 || c == 0 || r == d.NR - 1 || c == d.NC - 1;
  Meta m;
  m.S = 0; m.donors = 0; m.lower = 0; m.cls = CLS_NA; m.bmask = 0;
  if (!ring && !is_nan(f)) {
    double e[10];
#pragma unroll
    for (int n = 1; n <= 9; ++n) e[n] = A.elev[id + nb_off(n, d.pitch)];
    unsigned lower;
    m.S = weti_S(e, K, &lower);
    m.lower = (uint16_t)lower;
    m.cls = m.S != 0 ? CLS_A : (A.ddir[id] != 5 ? CLS_B : CLS_C);
  }
  A.meta[id] = m;
  A.cls[id] = m.cls;
  key_up[i] = m.cls == CLS_NA ? 0xFFFFFFFFu : A.upslope[id];
  val_id[i] = (uint32_t)id;
}
// second sort key: elevation descending (invalid cells last)
__global__ void k_weti_key_elev(const Arrays A, const uint32_t* __restrict__ ids, size_t n,
                                unsigned long long* __restrict__ key) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t id = ids[i];
  key[i] = A.cls[id] == CLS_NA ? ~0ull : ~f64_orderable(A.elev[id]);
}
__global__ void k_weti_rank(const uint32_t* __restrict__ ids, size_t n, uint32_t* __restrict__ rank) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rank[ids[i]] = (uint32_t)i;
}

// ------------------------------------------------------------------------------------------------
// block-aggregated append of newly ready cells to the queue: every thread contributes up to 8 ids
constexpr int WT = 256;            // threads per block of the queue kernels
constexpr int STAGE = WT * 8;

struct Stager {
  uint32_t* s_ids;     // STAGE entries
  unsigned* s_n;
  unsigned* s_base;
};
__device__ __forceinline__ void stage_push(const Stager& s, uint32_t id) {
  s.s_ids[atomicAdd(s.s_n, 1u)] = id;
}
// all threads of the block must call; `counter` counts appends relative to `base`
__device__ __forceinline__ void stage_flush(const Stager& s, uint32_t* order, unsigned base, unsigned* counter) {
  __syncthreads();
  const unsigned n = *s.s_n;
  if (threadIdx.x == 0 && n) *s.s_base = atomicAdd(counter, n);
  __syncthreads();
  if (n) {
    const unsigned b = base + *s.s_base;
    for (unsigned k = threadIdx.x; k < n; k += blockDim.x) order[b + k] = s.s_ids[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) *s.s_n = 0;
  __syncthreads();
}

// is neighbour n (code from the cell at `id`) a B cell whose drec is the cell at `id`?
__device__ __forceinline__ bool b_points_here(const Arrays& A, size_t nid, int n) {
  // the neighbour in direction n sees this cell in direction 10 - n
  return A.cls[nid] == CLS_B && A.ddir[nid] == 10 - n;
}

// donors that precede the cell, in rank order; pending counter; wave 0
__global__ void __launch_bounds__(WT) k_weti_donors(Arrays A, GridDims d, unsigned* counters) {
  __shared__ uint32_t s_ids[STAGE];
  __shared__ unsigned s_n, s_base;
  Stager st{s_ids, &s_n, &s_base};
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c < d.NC) {
    const size_t id = (size_t)r * d.pitch + c;
    Meta m = A.meta[id];
    if (m.cls != CLS_NA) {
      const double f = A.elev[id];
      const uint32_t rk = A.rank[id];
      uint32_t drank[8];
      uint8_t dcode[8];
      int nd = 0;
#pragma unroll
      for (int n = 1; n <= 9; ++n) {
        if (n == 5) continue;
        const size_t nid = id + nb_off(n, d.pitch);
        if (A.cls[nid] == CLS_NA) continue;
        const double en = A.elev[nid];
        bool donor = en > f;                       // an A cell that pushes qc*tan into this cell
        bool isb = false;
        if (!donor && b_points_here(A, nid, n)) {  // a B cell that pushes its qa into this cell
          isb = true;
          donor = A.rank[nid] < rk;                // ... before this cell is visited
        }
        if (!donor) continue;
        drank[nd] = A.rank[nid];
        dcode[nd] = (uint8_t)(n | (isb ? 16 : 0));
        ++nd;
      }
      // insertion sort by rank (nd <= 8)
      for (int a = 1; a < nd; ++a) {
        uint32_t kr = drank[a];
        uint8_t kc = dcode[a];
        int b = a - 1;
        while (b >= 0 && drank[b] > kr) { drank[b + 1] = drank[b]; dcode[b + 1] = dcode[b]; --b; }
        drank[b + 1] = kr;
        dcode[b + 1] = kc;
      }
      uint32_t donors = 0;
      uint8_t bmask = 0;
      for (int a = 0; a < nd; ++a) {
        donors |= (uint32_t)(dcode[a] & 15) << (4 * a);
        if (dcode[a] & 16) bmask |= (uint8_t)(1u << a);
      }
      m.donors = donors;
      m.bmask = bmask;
      if (m.cls == CLS_B) {
        const int dv = A.ddir[id];
        if (dv >= 1 && dv <= 9) {
          const size_t tid = id + nb_off(dv, d.pitch);
          if (A.cls[tid] != CLS_NA && rk < A.rank[tid]) m.cls |= F_BEARLY;
        }
      }
      A.meta[id] = m;
      reinterpret_cast<uint8_t*>(A.pend32)[id] = (uint8_t)nd;
      if (nd == 0) stage_push(st, (uint32_t)id);
    }
  }
  stage_flush(st, A.order, 0, &counters[0]);
}

// ------------------------------------------------------------------------------------------------
// counters[0..2]: cells appended for wave (w % 3); counters[3]: number of waves (reported);
__device__ __forceinline__ void notify(const Arrays& A, const Stager& st, size_t tid) {
  const unsigned sh = 8u * (unsigned)(tid & 3);
  // the successor will be visited soon: pull its descriptor towards L2 while the counter round trip is in flight
  asm volatile("prefetch.global.L2 [%0];" ::"l"(&A.meta[tid]));
  asm volatile("prefetch.global.L2 [%0];" ::"l"(&A.elev[tid]));
  const unsigned old = atomicSub(&A.pend32[tid >> 2], 1u << sh);
  if (((old >> sh) & 0xFFu) == 1u) stage_push(st, (uint32_t)tid);
}

// visit one ready cell: add the donors' contributions in the reference's order, derive qc, publish
// the state, release the successors (newly ready ones go to the block's staging buffer)
__device__ __forceinline__ void visit(const Arrays& A, const GridDims& d, const WetiK& K, const Stager& st, size_t id) {
  const Meta m = A.meta[id];
  const double f = A.elev[id];
  // all donor loads are issued before the (ordered) additions: one memory round trip, not eight
  double2 dv[8];
  double de[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int n = (m.donors >> (4 * j)) & 15;
    dv[j] = make_double2(0.0, 0.0);
    de[j] = 0.0;
    if (n) {
      const size_t nid = id + nb_off(n, d.pitch);
      const bool isb = (m.bmask >> j) & 1;
      dv[j] = __ldcg(reinterpret_cast<const double2*>(&A.state[nid]) + (isb ? 1 : 0));
      if (!isb) de[j] = A.elev[nid];
    }
  }
  double qa1 = 1.0, qa2 = K.g2;  // :35-36
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int n = (m.donors >> (4 * j)) & 15;
    if (!n) break;
    if ((m.bmask >> j) & 1) {  // :82-83, the donor's whole qa
      qa1 = qa1 + dv[j].x;
      qa2 = qa2 + dv[j].y;
    } else {                   // :74-75, qc * tan of the donor towards this cell
      const double t = weti_tan(de[j], f, n, K);
      qa1 = qa1 + (dv[j].x * t);
      qa2 = qa2 + (dv[j].y * t);
    }
  }
  const int cls = m.cls & 15;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  double2 qc;
  if (cls == CLS_A) { qc.x = qa1 / m.S; qc.y = qa2 / m.S; }       // :70-71
  else if (cls == CLS_B) { qc.x = qa1 / 0.0001; qc.y = nan; }      // :80 (qc2 keeps its old value)
  else { qc.x = nan; qc.y = nan; }
  double2* op = reinterpret_cast<double2*>(&A.state[id]);
  __stcg(op, qc);
  __stcg(op + 1, make_double2(qa1, qa2));
  __threadfence();
  if (cls == CLS_A) {
#pragma unroll
    for (int n = 1; n <= 9; ++n)
      if ((m.lower >> n) & 1) notify(A, st, id + nb_off(n, d.pitch));
  } else if (m.cls & F_BEARLY) {
    notify(A, st, id + nb_off(A.ddir[id], d.pitch));
  }
}

// Scheduling.  A cell in the queue is ready: all its donors are done and their state is visible
// (state store -> __threadfence -> counter decrement).  A grid-wide "wave" hands every block a chunk
// of the global queue; the successors a block releases go to its own shared-memory queue and the
// block keeps visiting them (LIFO, depth first) for up to `local_rounds` rounds before it appends
// what is left to the global queue for the next wave.  The DAG of the bench DEM is ~3900 levels deep
// with a broad spread of level sizes: chasing locally divides the number of grid barriers by the
// number of levels a block can advance on its own.
constexpr int QCAP = STAGE * 3;  // shared queue entries; a visiting round adds at most STAGE

__global__ void __launch_bounds__(WT) k_weti_waves(Arrays A, GridDims d, WetiK K, unsigned* counters,
                                                   unsigned* wavelog, unsigned wavelog_cap, int local_rounds) {
  __shared__ uint32_t s_ids[QCAP];
  __shared__ unsigned s_n, s_base;
  Stager st{s_ids, &s_n, &s_base};
  cg::grid_group grid = cg::this_grid();
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  unsigned begin = 0, end = *reinterpret_cast<volatile unsigned*>(&counters[0]);
  unsigned wave = 0;
  while (begin < end) {
    unsigned* next_counter = &counters[(wave + 1) % 3];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      counters[(wave + 2) % 3] = 0;  // untouched during this wave
      if (wavelog && wave < wavelog_cap) wavelog[wave] = end - begin;
    }
    // a wave that fits one block leaves every other block idle: let that block run on
    const int rounds = (end - begin <= WT) ? (1 << 30) : local_rounds;
    for (unsigned base = begin + blockIdx.x * WT; base < end; base += gridDim.x * WT) {
      const unsigned i = base + threadIdx.x;
      if (i < end) visit(A, d, K, st, __ldcg(&A.order[i]));
      __syncthreads();
      const unsigned filled = s_n;
      __syncthreads();  // nobody pushes before everybody has read
      if (filled > QCAP - STAGE) stage_flush(st, A.order, end, next_counter);
    }
    for (int round = 0; round < rounds; ++round) {
      __syncthreads();
      const unsigned n = s_n;
      if (n == 0 || n > QCAP - STAGE) break;
      const unsigned take = n < WT ? n : WT;
      const uint32_t my = threadIdx.x < take ? s_ids[n - take + threadIdx.x] : 0xFFFFFFFFu;
      __syncthreads();
      if (threadIdx.x == 0) s_n = n - take;
      __syncthreads();
      if (my != 0xFFFFFFFFu) visit(A, d, K, st, my);
    }
    stage_flush(st, A.order, end, next_counter);
    grid.sync();
    begin = end;
    end = begin + *reinterpret_cast<volatile unsigned*>(next_counter);
    ++wave;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) counters[3] = wave;
}

// ------------------------------------------------------------------------------------------------
// "last setter" positions over the visiting order: vA[i] = i if the i-th visited cell assigns qc2
// (class A), vAB[i] = i if it assigns qc1 (class A or B); -1 otherwise.  Inclusive max-scans turn
// them into the position of the last setter at or before i.
__global__ void k_weti_flags(const uint8_t* __restrict__ cls, const uint32_t* __restrict__ ids, size_t n,
                             int* __restrict__ vA, int* __restrict__ vAB) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t c = cls[ids[i]];
  vA[i] = c == CLS_A ? (int)i : -1;
  vAB[i] = (c == CLS_A || c == CLS_B) ? (int)i : -1;
}

__device__ __forceinline__ double trunc_dec2(double x) { return trunc(x * 100.0) / 100.0; }  // R/LITAP_utils.R:93-95

struct OutPtrs { double* o[4]; };

__global__ void k_weti_final(Arrays A, GridDims d, const uint32_t* __restrict__ ids, const int* __restrict__ pA,
                             const int* __restrict__ pAB, OutPtrs P) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d.NC) return;
  const size_t id = (size_t)r * d.pitch + c;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  const int cls = A.cls[id];
  if (cls == CLS_NA) {  // ring and missing cells (:98-101)
#pragma unroll
    for (int k = 0; k < 4; ++k) P.o[k][id] = nan;
    return;
  }
  const State s = A.state[id];
  double qa1 = s.qa1, qa2 = s.qa2;
  // B neighbours that drain into this cell but are visited after it: they only change the final qa
  {
    const uint32_t rk = A.rank[id];
    uint32_t lrank[8];
    uint8_t lcode[8];
    int nl = 0;
#pragma unroll
    for (int n = 1; n <= 9; ++n) {
      if (n == 5) continue;
      const size_t nid = id + nb_off(n, d.pitch);
      if (b_points_here(A, nid, n) && A.rank[nid] > rk) { lrank[nl] = A.rank[nid]; lcode[nl] = (uint8_t)n; ++nl; }
    }
    for (int a = 1; a < nl; ++a) {
      uint32_t kr = lrank[a];
      uint8_t kc = lcode[a];
      int b = a - 1;
      while (b >= 0 && lrank[b] > kr) { lrank[b + 1] = lrank[b]; lcode[b + 1] = lcode[b]; --b; }
      lrank[b + 1] = kr;
      lcode[b + 1] = kc;
    }
    for (int a = 0; a < nl; ++a) {
      const State t = A.state[id + nb_off(lcode[a], d.pitch)];
      qa1 = qa1 + t.qa1;
      qa2 = qa2 + t.qa2;
    }
  }
  // effective qc1 / qc2 at the time the reference evaluates :87-88
  double qc1 = s.qc1, qc2 = s.qc2;
  if (cls != CLS_A) {
    const uint32_t rk = A.rank[id];
    const int a = pA[rk];
    qc2 = a >= 0 ? A.state[ids[a]].qc2 : nan;
    if (cls == CLS_C) {
      const int b = pAB[rk];
      qc1 = b >= 0 ? A.state[ids[b]].qc1 : nan;
    }
  }
  double qw1 = (qc1 > 1) ? log(qc1) : log(1 + qc1);
  double qw2 = (qc2 > 1) ? log(qc2) : log(1 + qc2);
  if (is_nan(qw1)) qw1 = qa1 / 0.0001;  // :94 (the :95 line tests the refilled qw1, so qw2 stays)
  P.o[0][id] = trunc_dec2(qa1);
  P.o[1][id] = trunc_dec2(qa2);
  P.o[2][id] = trunc_dec2(qw1);
  P.o[3][id] = trunc_dec2(qw2);
}

}  // namespace weti

// ------------------------------------------------------------------------------------------------
namespace {
inline unsigned nblk(size_t n, unsigned t = 256) { return (unsigned)((n + t - 1) / t); }
}

int run_weti(litap_grid* g, double grid_m) {
  using namespace weti;
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  OutPtrs P;
  for (int k = 0; k < 4; ++k) {
    if (!g->weti[k]) g->weti[k] = grid_alloc_array<double>(d, -1);
    if (!g->weti[k]) return set_error(LITAP_ENOMEM, "device allocation failed (weti outputs)");
    P.o[k] = g->weti[k];
  }
  WetiK K;
  K.l1 = grid_m * 0.5;              // R/form_02b_calc_weti.R:5-9
  K.l2 = grid_m * 0.354;
  K.orth = grid_m;
  K.diag = grid_m * std::sqrt(2.0);
  K.g2 = grid_m * grid_m;

  const size_t ncell = (size_t)d.cells + 64;          // pitched ids
  const size_t nd = (size_t)d.NR * d.NC;               // dense (seqno) enumeration
  PoolScratch S;
  Arrays A;
  A.elev = g->elev;
  A.ddir = g->ddir;
  A.upslope = g->upslope;
  A.meta = S.get<Meta>(ncell);
  A.state = S.get<State>(ncell);
  A.cls = S.get<uint8_t>(ncell);
  A.rank = S.get<uint32_t>(ncell);
  A.pend32 = S.get<uint32_t>(ncell / 4 + 2);
  A.order = S.get<uint32_t>(nd);
  uint32_t* key_a = S.get<uint32_t>(nd);
  uint32_t* key_b = S.get<uint32_t>(nd);
  uint32_t* id_a = S.get<uint32_t>(nd);
  uint32_t* id_b = S.get<uint32_t>(nd);
  unsigned long long* k64_a = S.get<unsigned long long>(nd);
  unsigned long long* k64_b = S.get<unsigned long long>(nd);
  unsigned* counters = S.get<unsigned>(8);
  if (!A.meta || !A.state || !A.cls || !A.rank || !A.pend32 || !A.order || !key_a || !key_b || !id_a || !id_b ||
      !k64_a || !k64_b || !counters)
    return set_error(LITAP_ENOMEM, "device allocation failed (weti scratch)");
  CUDA_TRY(cudaMemsetAsync(counters, 0, 8 * sizeof(unsigned), st));
  CUDA_TRY(cudaMemsetAsync(A.cls, 0, ncell, st));  // pad columns read as NA

  dim3 cgrid(nblk(d.NC, WT), d.NR);
  KERNEL("k_weti_cls", k_weti_cls<<<cgrid, WT, 0, st>>>(A, d, K, key_a, id_a));
  // the reference's visiting order: stable LSD radix sort, upslope asc then elevation desc
  size_t t1 = 0, t2 = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, t1, key_a, key_b, id_a, id_b, (int)nd, 0, 32, st);
  cub::DeviceRadixSort::SortPairs(nullptr, t2, k64_a, k64_b, id_b, id_a, (int)nd, 0, 64, st);
  size_t t3 = 0;
  cub::DeviceScan::InclusiveScan(nullptr, t3, (int*)nullptr, (int*)nullptr, cub::Max(), (int)nd, st);
  const size_t tn = std::max(t1, std::max(t2, t3));
  void* tmp = S.get<unsigned char>(tn + 256);
  if (!tmp) return set_error(LITAP_ENOMEM, "device allocation failed (weti sort scratch)");
  size_t tcap = tn + 256;
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tcap, key_a, key_b, id_a, id_b, (int)nd, 0, 32, st));
  KERNEL("k_weti_key_elev", k_weti_key_elev<<<nblk(nd), 256, 0, st>>>(A, id_b, nd, k64_a));
  tcap = tn + 256;
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tcap, k64_a, k64_b, id_b, id_a, (int)nd, 0, 64, st));
  c.launches += 2;
  const uint32_t* sorted = id_a;  // sorted[i] = pitched id of the i-th visited cell
  KERNEL("k_weti_rank", k_weti_rank<<<nblk(nd), 256, 0, st>>>(sorted, nd, A.rank));
  KERNEL("k_weti_donors", k_weti_donors<<<cgrid, WT, 0, st>>>(A, d, counters));

  // LITAP_WETI_WAVELOG=1: print the wave-size histogram of the dependency DAG (tuning aid)
  unsigned* wavelog = nullptr;
  unsigned wavelog_cap = 0;
  if (getenv("LITAP_WETI_WAVELOG")) {
    wavelog_cap = 1u << 20;
    wavelog = S.get<unsigned>(wavelog_cap);
    if (wavelog) CUDA_TRY(cudaMemsetAsync(wavelog, 0, wavelog_cap * sizeof(unsigned), st));
  }
  int local_rounds = 16;
  if (const char* e = getenv("LITAP_WETI_ROUNDS")) local_rounds = atoi(e);  // tuning aid
  static int coop_blocks = 0;
  if (!coop_blocks) {
    int per_sm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_weti_waves, WT, 0));
    if (per_sm < 1) return set_error(LITAP_ECUDA, "k_weti_waves cannot be made resident");
    coop_blocks = c.sm_count * (per_sm < 4 ? per_sm : 4);
  }
  {
    GridDims dd = d;
    void* args[] = {(void*)&A, (void*)&dd, (void*)&K, (void*)&counters, (void*)&wavelog, (void*)&wavelog_cap, (void*)&local_rounds};
    ProfScope prof("k_weti_waves");
    CUDA_TRY(cudaLaunchCooperativeKernel((const void*)k_weti_waves, dim3(coop_blocks), dim3(WT), args, 0, st));
    c.launches++;
  }
  // stale qc1 / qc2: last setter before every visited position (key buffers are free again)
  int* vA = reinterpret_cast<int*>(key_a);
  int* vAB = reinterpret_cast<int*>(key_b);
  KERNEL("k_weti_flags", k_weti_flags<<<nblk(nd), 256, 0, st>>>(A.cls, sorted, nd, vA, vAB));
  tcap = tn + 256;
  CUDA_TRY(cub::DeviceScan::InclusiveScan(tmp, tcap, vA, vA, cub::Max(), (int)nd, st));
  tcap = tn + 256;
  CUDA_TRY(cub::DeviceScan::InclusiveScan(tmp, tcap, vAB, vAB, cub::Max(), (int)nd, st));
  c.launches += 2;
  KERNEL("k_weti_final", k_weti_final<<<cgrid, WT, 0, st>>>(A, d, sorted, vA, vAB, P));
  unsigned hc[8];
  CUDA_TRY(cudaMemcpyAsync(hc, counters, sizeof(hc), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));  // scratch goes back to the pool
  c.tm.weti_waves = (int)hc[3];
  if (wavelog) {
    std::vector<unsigned> h(std::min<unsigned>(hc[3], wavelog_cap));
    cudaMemcpy(h.data(), wavelog, h.size() * sizeof(unsigned), cudaMemcpyDeviceToHost);
    unsigned long long cnt[2][33] = {}, cells[2][33] = {};
    for (unsigned v : h) {
      int tail = 0;
      unsigned n = v & 0x7FFFFFFFu;
      int b = 0;
      while ((1u << b) < n) ++b;
      cnt[tail][b]++;
      cells[tail][b] += n;
    }
    fprintf(stderr, "weti wave-size histogram (%zu waves): bucket <=2^b : grid-mode waves/cells | tail-mode waves/cells\n", h.size());
    for (int b = 0; b <= 32; ++b)
      if (cnt[0][b] || cnt[1][b])
        fprintf(stderr, "  2^%-2d : %8llu %12llu | %8llu %12llu\n", b, cnt[0][b], cells[0][b], cnt[1][b], cells[1][b]);
  }
  return LITAP_OK;
}
