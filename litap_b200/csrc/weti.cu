// weti.cu — calc_weti2 on the device (R/form_02b_calc_weti.R:3-110): Quinn-style multiple-flow-
// direction upslope area (qarea1/2) and wetness index (qweti1/2).
//
// The reference visits the cells in the order (elev desc, upslope asc, seqno asc) (:23-26) and
// PUSHES qc * tan_k into its lower neighbours (:69-77) or, for a cell without a lower neighbour
// that is not a pit, its whole qa into drec (:79-84).  fp64 addition is not associative and the
// outputs are truncated to 2 decimals, so every cell must receive its contributions in exactly that
// order.  The device version is a deterministic PULL over the same dependency DAG, run as a
// barrier-free DATAFLOW (round 2; round 1 met at a grid barrier per wave and let one block finish
// the tail of the DAG alone: 62 ms at 8192^2):
//
//   k_weti_cls      class of every cell (A: S != 0, B: S == 0 & ddir != 5, C: pit); counts
//   k_weti_queries  the B / C cells as 128-bit visiting-order keys (they are the only cells that read
//                   the reference's stale qc1 / qc2 variables, see below); cub radix sort of those
//                   (~2 % of the cells) instead of round 1's two full-grid sorts
//   k_weti_table    bucket table over the sorted queries (elevation is mapped linearly to 2^20
//                   buckets) so that any cell finds its place among the queries in 1-3 probes
//   k_weti_nodes    per cell: S with R's long double accumulator (cellmath.h ld80_*), the donors that
//                   precede it (strictly higher neighbours; B neighbours whose drec is the cell and
//                   that are visited before it — decided LOCALLY by comparing (elev, upslope, seqno),
//                   no global rank), sorted in visiting order, packed in one 32-byte node; pending
//                   counter; cells without donors seed the ready queues; A cells record themselves as
//                   "last setter" candidate of the query interval they fall in (128-bit CAS max)
//   k_weti_flow     persistent kernel, no grid barrier: a lane visits a ready cell (adds its donors'
//                   contributions in the reference's order — no FP atomics —, derives qc, publishes
//                   its state, decrements its successors' pending counters) and KEEPS one of the cells
//                   it released for itself.  A hop of a dependency chain costs TWO memory round trips:
//                   (1) the donors' state, (2) the 64-bit atomic decrement of the successors' topology
//                   words, which hands the releasing lane the released cell's whole topology (donor
//                   codes, successor mask, class).  There is no fence between a cell's state stores and
//                   its decrements: states start as an all-ones pattern and a reader that still sees it
//                   loads again (data as flag).  Other released cells go to free lanes of the warp, the
//                   rest to one of 64 global ready queues that idle lanes wait on (each on its own slot);
//                   queue claims and polls are issued under the shadow of the visit's loads
//   k_weti_qflags + cub scans   the reference's qc1/qc2 are variables that survive between loop
//                   iterations (:70-71,:80 vs :87-88): a B cell reads the qc2, a pit the qc1 and qc2
//                   of the last cell before it in the visiting order that assigned them -> prefix
//                   maxima over the query intervals
//   k_weti_final    late contributions (B donors visited after their receiver only change the final
//                   qa), qweti from the effective qc, trunc_dec(., 2), NA
//
// Algorithmic traffic 45 B/cell (elev 8 + ddir 1 + upslope 4 read, 4 x 8 written).
#include <cub/cub.cuh>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "cellmath.h"
#include "litap_internal.h"

namespace weti {

enum : uint8_t { CLS_NA = 0, CLS_A = 1, CLS_B = 2, CLS_C = 3 };
constexpr uint32_t NONE = 0xFFFFFFFFu;   // "no cell" / empty queue slot
constexpr int NQ = 64;                   // global ready queues (spreads the head / tail atomics over L2 slices)
constexpr int NT = 1 << 20;              // buckets of the query table
constexpr int WT = 256;                  // threads per block
constexpr int WPB = WT / 32;
constexpr int WBUF = 32 * 7;             // cells a warp can release beyond the ones its lanes keep

// topology word of a cell, 64 bits.  The lane that drops the pending counter to zero gets the whole word back from
// its atomic decrement and needs no further load to start the visit:
//   [0,4) pending donors   [4,36) donor codes (direction from the cell towards the donor), 4 bits each, in visiting
//   order, 0 ends   [36,45) bit 36+n-1: neighbour n waits for this cell (lower neighbours of an A cell; drec of an
//   early B cell)   [45,47) class   [47,55) bit 47+j: the j-th donor is a B donor (contributes its whole qa)
constexpr int TP_DON = 4, TP_SUCC = 36, TP_CLS = 45, TP_BM = 47;
// state of a cell: 8 doubles = two 32-byte sectors.  A receiver reads one sector per donor.
//   [0] qc1 [1] qc2 (as assigned by the cell itself; NaN where the reference leaves the variable stale) [2] f [3] S
//   [4] qa1 [5] qa2 (qa at processing time) [6] RN(1/S) or 0 [7] -
// qc* / qa* hold the all-ones pattern until the cell has been visited
constexpr int ST_QC = 0, ST_F = 2, ST_S = 3, ST_QA = 4, ST_RS = 6, ST_N = 8;
constexpr unsigned long long UNSET = ~0ull;
// visiting-order key: later visited <=> (hi, lo) larger.  hi = ~orderable(elev), lo = upslope << 32 | cell id
struct __align__(16) K128 { unsigned long long lo, hi; };
struct __align__(128) QCount { unsigned v; unsigned pad[31]; };
struct Ctl {
  unsigned nvalid, nq, done, finished, abort, pad[3];
  unsigned long long st_pushed, st_popped, st_kept, st_shared, st_iters, st_idle, st_respin;
};

struct Arrays {
  const double* elev;
  const uint8_t* ddir;
  const uint32_t* upslope;
  unsigned long long* topo;
  double* st;          // ST_N doubles per cell
  uint8_t* cls;        // CLS_* only (neighbour lookups)
  ulonglong2* order;   // NQ ready queues of qcap entries {topology word, cell id}; all ones = not yet written
  QCount* qtail;       // [NQ]
  QCount* qhead;       // [NQ]
  unsigned qcap;
  Ctl* ctl;
  // stale-qc queries (B and C cells) sorted in visiting order, bucket table, last-setter slots
  const unsigned long long* qhi;
  const unsigned long long* qlo;
  const unsigned* table;  // [NT + 1]
  K128* slotA;            // [nq + 1]: latest A cell between query j-1 and query j (after the scan: before query j)
  const int* lastB;       // [nq]: latest B query at or before query j
  unsigned nq;
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
__device__ __forceinline__ K128 make_key(double e, uint32_t up, uint32_t id) {
  K128 k;
  k.hi = ~f64_orderable(e + 0.0);   // -0.0 and +0.0 tie in the reference's ordering
  k.lo = ((unsigned long long)up << 32) | id;
  return k;
}
__device__ __forceinline__ bool key_less(const K128& a, const K128& b) { return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo); }
// is cell a = (ea, id a) visited before cell b?  upslope is only read when the elevations tie
__device__ __forceinline__ bool visited_before(const Arrays& A, double ea, size_t ida, double eb, size_t idb) {
  if (ea > eb) return true;
  if (ea < eb) return false;
  const uint32_t ua = A.upslope[ida], ub = A.upslope[idb];
  return ua < ub || (ua == ub && ida < idb);
}

// ------------------------------------------------------------------------------------------------
// class of every cell; number of valid cells and of stale-qc queries (B and C cells)
__global__ void __launch_bounds__(WT) k_weti_cls(Arrays A, GridDims d, WetiK K) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  uint8_t cls = CLS_NA;
  if (c < d.NC) {
    const size_t id = (size_t)r * d.pitch + c;
    const double f = A.elev[id];
    const bool ring = r == 0 || c == 0 || r == d.NR - 1 || c == d.NC - 1;
    if (!ring && !is_nan(f)) {
      bool any = false;
#pragma unroll
      for (int n = 1; n <= 9; ++n) {
        if (n == 5) continue;
        const double t = weti_tan(f, A.elev[id + nb_off(n, d.pitch)], n, K);
        any |= t > 0;   // NaN compares false (na.rm = TRUE)
      }
      cls = any ? CLS_A : (A.ddir[id] != 5 ? CLS_B : CLS_C);
    }
    A.cls[id] = cls;
  }
  const unsigned mv = __ballot_sync(0xFFFFFFFFu, cls != CLS_NA);
  const unsigned mq = __ballot_sync(0xFFFFFFFFu, cls == CLS_B || cls == CLS_C);
  __shared__ unsigned s_v, s_q;
  if (threadIdx.x == 0) { s_v = 0; s_q = 0; }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) {
    if (mv) atomicAdd(&s_v, __popc(mv));
    if (mq) atomicAdd(&s_q, __popc(mq));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (s_v) atomicAdd(&A.ctl->nvalid, s_v);
    if (s_q) atomicAdd(&A.ctl->nq, s_q);
  }
}

// the B / C cells as visiting-order keys (unsorted)
__global__ void __launch_bounds__(WT) k_weti_queries(Arrays A, GridDims d, unsigned long long* __restrict__ qhi,
                                                     unsigned long long* __restrict__ qlo, unsigned* counter) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  bool q = false;
  size_t id = 0;
  if (c < d.NC) {
    id = (size_t)r * d.pitch + c;
    const uint8_t cls = A.cls[id];
    q = cls == CLS_B || cls == CLS_C;
  }
  const unsigned m = __ballot_sync(0xFFFFFFFFu, q);
  if (!m) return;
  const int lane = threadIdx.x & 31;
  unsigned base = 0;
  if (lane == __ffs(m) - 1) base = atomicAdd(counter, __popc(m));
  base = __shfl_sync(0xFFFFFFFFu, base, __ffs(m) - 1);
  if (q) {
    const unsigned i = base + __popc(m & ((1u << lane) - 1));
    const K128 k = make_key(A.elev[id], A.upslope[id], (uint32_t)id);
    qhi[i] = k.hi;
    qlo[i] = k.lo;
  }
}

// linear map elevation -> bucket, monotone: a cell in a lower bucket is visited before one in a higher bucket
struct BucketMap { double emax, scale; };
__device__ __forceinline__ BucketMap bucket_map_compute(const Arrays& A) {
  BucketMap m;
  m.emax = f64_from_orderable(~A.qhi[0]);
  const double emin = f64_from_orderable(~A.qhi[A.nq - 1]);
  const double span = m.emax - emin;
  m.scale = (span > 0 && span < 1.7e308) ? (double)(NT - 1) / span : 0.0;
  return m;
}
__device__ __forceinline__ int bucket_of(const BucketMap& m, double e) {
  if (!(e < m.emax)) return 0;
  const double t = (m.emax - e) * m.scale;
  if (!(t < (double)(NT - 1))) return NT - 1;
  return (int)t;
}
// the map is computed once (k_weti_table stores it behind the table) and loaded by the 67 M lookups
__device__ __forceinline__ BucketMap bucket_map(const Arrays& A) {
  const double2 v = __ldg(reinterpret_cast<const double2*>(A.table + NT + 4));
  BucketMap m;
  m.emax = v.x;
  m.scale = v.y;
  return m;
}
// table[b] = number of queries in buckets < b
__global__ void k_weti_table(Arrays A, unsigned* __restrict__ table) {
  const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > (unsigned)NT) return;
  const BucketMap m = bucket_map_compute(A);
  if (b == 0) *reinterpret_cast<double2*>(table + NT + 4) = make_double2(m.emax, m.scale);
  unsigned lo = 0, hi = A.nq;
  while (lo < hi) {
    const unsigned mid = (lo + hi) >> 1;
    if ((unsigned)bucket_of(m, f64_from_orderable(~A.qhi[mid])) < b) lo = mid + 1; else hi = mid;
  }
  table[b] = lo;
}
// number of queries visited before the key (= index of the key itself when it is a query)
__device__ __forceinline__ unsigned locate(const Arrays& A, const BucketMap& m, double e, const K128& k) {
  const int b = bucket_of(m, e);
  unsigned lo = A.table[b], hi = A.table[b + 1];
  while (lo < hi) {
    const unsigned mid = (lo + hi) >> 1;
    const unsigned long long h = A.qhi[mid];
    const bool less = h < k.hi || (h == k.hi && A.qlo[mid] < k.lo);
    if (less) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ K128 cas128(K128* p, const K128& cmp, const K128& val) {
  K128 old;
  asm volatile(
      "{\n .reg .b128 c, v, o;\n mov.b128 c, {%3, %4};\n mov.b128 v, {%5, %6};\n"
      " atom.global.relaxed.gpu.cas.b128 o, [%2], c, v;\n mov.b128 {%0, %1}, o;\n}"
      : "=l"(old.lo), "=l"(old.hi)
      : "l"(p), "l"(cmp.lo), "l"(cmp.hi), "l"(val.lo), "l"(val.hi)
      : "memory");
  return old;
}
__device__ __forceinline__ void slot_max(K128* p, const K128& k) {
  const ulonglong2 t = __ldcg(reinterpret_cast<const ulonglong2*>(p));
  K128 cur;
  cur.lo = t.x; cur.hi = t.y;
  while (key_less(cur, k)) {
    const K128 old = cas128(p, cur, k);
    if (old.lo == cur.lo && old.hi == cur.hi) break;
    cur = old;
  }
}

// ------------------------------------------------------------------------------------------------
// memory accesses of the dataflow, all at L2 (the data is shared between SMs), in program order (asm volatile)
__device__ __forceinline__ double2 ld_cg_f64x2(const double* p) {
  double2 v;
  asm volatile("ld.global.cg.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double ld_cg_f64(const double* p) {
  double v;
  asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_cg_f64x2(double* p, double x, double y) {
  asm volatile("st.global.cg.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(x), "d"(y) : "memory");
}
__device__ __forceinline__ ulonglong2 ld_entry(const ulonglong2* p) {
  ulonglong2 v;
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_entry(ulonglong2* p, unsigned long long tp, uint32_t id) {
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(tp), "l"((unsigned long long)id) : "memory");
}
__device__ __forceinline__ unsigned long long atom_dec_u64(unsigned long long* p) {
  unsigned long long old;
  asm volatile("atom.global.relaxed.gpu.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(p), "l"(~0ull) : "memory");
  return old;
}
__device__ __forceinline__ void st_volatile(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }
__device__ __forceinline__ uint32_t ld_volatile(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }
__device__ __forceinline__ bool is_unset(double x) { return (unsigned long long)__double_as_longlong(x) == UNSET; }

// ready queues.  Queue q owns order[q*qcap, (q+1)*qcap); an entry is written once (all ones -> {topology, cell})
// and read by the one lane that claimed its index from qhead.  Items that do not fit a queue go to the next
// one; a claim beyond qcap is void and sends the claiming warp on to the next queue.  Total capacity
// exceeds cells + waiting lanes, so neither side can run out.
// the whole warp pushes n items (shared memory)
__device__ __forceinline__ void warp_push(const Arrays& A, unsigned& pushq, const uint32_t* ids, const unsigned long long* tps,
                                          unsigned n, int lane) {
  while (n) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(&A.qtail[pushq].v, n);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    const unsigned fit = base >= A.qcap ? 0u : min(n, A.qcap - base);
    ulonglong2* dst = A.order + (size_t)pushq * A.qcap + base;
    for (unsigned k = lane; k < fit; k += 32) st_entry(dst + k, tps[k], ids[k]);
    ids += fit;
    tps += fit;
    n -= fit;
    pushq = (pushq + 1) % NQ;
  }
}

// ------------------------------------------------------------------------------------------------
// topology word and state of every cell: S, donors in visiting order, successors, pending counter; seeds;
// last-setter slots
__global__ void __launch_bounds__(WT, 4) k_weti_nodes(Arrays A, GridDims d, WetiK K) {
  bool seed = false;          // a cell without donors: ready from the start
  uint32_t seed_id = 0;
  unsigned long long seed_tp = 0;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c < d.NC) {
    const size_t id = (size_t)r * d.pitch + c;
    const uint8_t cls = A.cls[id];
    unsigned long long tp = 0;
    if (cls != CLS_NA) {
      double e[10];
#pragma unroll
      for (int n = 1; n <= 9; ++n) e[n] = A.elev[id + nb_off(n, d.pitch)];
      const double f = e[5];
      double S = 0;
      unsigned succ = 0;
      if (cls == CLS_A) S = weti_S(e, K, &succ);
      const int dv = A.ddir[id];
      if (cls == CLS_B && dv >= 1 && dv <= 9) {   // an early B cell pushes its qa into drec before drec is visited
        const size_t tid = id + nb_off(dv, d.pitch);
        if (A.cls[tid] != CLS_NA && visited_before(A, f, id, e[dv], tid)) succ = 1u << dv;
      }
      // donors that precede the cell: strictly higher neighbours (they are A cells and push qc*tan, :74-75) and B
      // neighbours whose drec is this cell and that are visited first (they push their whole qa, :82-83).
      // (Loops over the neighbours / the donors, not unrolled: code size, see weti_S.)
      unsigned dmask = 0, bdon = 0;
      unsigned clsw = 0;   // classes of the eight neighbours, 2 bits each: eight independent loads in flight
#pragma unroll
      for (int n = 1; n <= 9; ++n)
        if (n != 5) clsw |= (unsigned)A.cls[id + nb_off(n, d.pitch)] << (2 * n);
#pragma unroll 1
      for (int n = 1; n <= 9; ++n) {
        if (n == 5) continue;
        const size_t nid = id + nb_off(n, d.pitch);
        const uint8_t nc = (uint8_t)((clsw >> (2 * n)) & 3u);
        if (nc == CLS_A) {
          // the same test that sets the donor's succ bit: tan > 0.  A difference above 1e-280 cannot underflow on its way
          // through / d * l (cell sizes are within 1e-30 .. 1e30 or the reciprocals are off), so only tiny ones are computed
          const double df = e[n] - f;
          if (df > 1e-280 ? K.rorth != 0.0 || weti_tan(e[n], f, 10 - n, K) > 0 : weti_tan(e[n], f, 10 - n, K) > 0) dmask |= 1u << n;
        } else if (nc == CLS_B && A.ddir[nid] == 10 - n) {
          if (visited_before(A, e[n], nid, f, id)) { dmask |= 1u << n; bdon |= 1u << n; }
        }
      }
      // position of a donor in the visiting order = number of donors visited before it, 4 bits per neighbour code
      unsigned long long posw = 0;
      for (unsigned ma = dmask; ma; ma &= ma - 1) {
        const int a = __ffs(ma) - 1;
        for (unsigned mb = ma & (ma - 1); mb; mb &= mb - 1) {
          const int b = __ffs(mb) - 1;
          const bool ab = visited_before(A, e[a], id + nb_off(a, d.pitch), e[b], id + nb_off(b, d.pitch));
          posw += 1ull << (4 * (ab ? b : a));
        }
      }
      const int nd = __popc(dmask);
      for (unsigned m = dmask; m; m &= m - 1) {
        const int n = __ffs(m) - 1;
        const int pos = (int)(posw >> (4 * n)) & 15;
        tp |= (unsigned long long)n << (TP_DON + 4 * pos);
        if ((bdon >> n) & 1) tp |= 1ull << (TP_BM + pos);
      }
      tp |= (unsigned long long)nd | ((unsigned long long)succ << (TP_SUCC - 1)) | ((unsigned long long)cls << TP_CLS);
      if (nd == 0) { seed = true; seed_id = (uint32_t)id; seed_tp = tp; }
      const double unset = __longlong_as_double((long long)UNSET);
      double4* sp = reinterpret_cast<double4*>(A.st + id * ST_N);
      sp[0] = make_double4(unset, unset, f, S);
      sp[1] = make_double4(unset, unset, cls == CLS_A ? div_rcp_prepare(S) : 0.0, 0.0);
      // last-setter candidate: the A cell records itself in the interval between two queries it falls in
      if (cls == CLS_A && A.nq) {
        const K128 k = make_key(f, A.upslope[id], (uint32_t)id);
        const BucketMap bm = bucket_map(A);
        slot_max(&A.slotA[locate(A, bm, f, k)], k);
      }
    }
    A.topo[id] = tp;
  }
  // seeds of this warp -> one ready queue, one atomic per warp (an entry that does not fit goes on to the next queue)
  unsigned q = (blockIdx.y * gridDim.x + blockIdx.x + (threadIdx.x >> 5)) % NQ;
  const int lane = threadIdx.x & 31;
  for (int tries = 0; tries < NQ; ++tries) {
    const unsigned m = __ballot_sync(0xFFFFFFFFu, seed);
    if (!m) break;
    const int leader = __ffs(m) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(&A.qtail[q].v, (unsigned)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    if (seed) {
      const unsigned idx = base + __popc(m & ((1u << lane) - 1));
      if (idx < A.qcap) {
        A.order[(size_t)q * A.qcap + idx] = make_ulonglong2(seed_tp, (unsigned long long)seed_id);
        seed = false;
      }
    }
    q = (q + 1) % NQ;
  }
}

// ------------------------------------------------------------------------------------------------
// The dataflow.  A lane owns at most one ready cell (cur, with its topology word tp).  Per iteration of a warp:
//   issue   the queue claim of the lanes that have nothing, and the loads of the lanes that visit: own f / S / 1/S and the
//           donors' states (NB at a time)
//   wait    for the claim, issue the poll of the claimed entries
//   visit   add the donors' contributions in the reference's order, derive qc, store the state, decrement the successors
//   wait    for the decrements: cells whose counter dropped to zero are released; the lane keeps the first one, free
//           lanes of the warp take the next ones, the rest goes to a ready queue; polled entries that were ready are taken
template <int NB>
__global__ void __launch_bounds__(WT, NB == 8 ? 2 : (NB == 6 ? 3 : 4)) k_weti_flow(Arrays A, GridDims d, WetiK K, unsigned idle_limit) {
  __shared__ uint32_t s_id[WPB][WBUF];
  __shared__ unsigned long long s_tp[WPB][WBUF];
  __shared__ int s_off[16];   // neighbour code -> cell offset (the donors' codes are data, not loop indices)
  if (threadIdx.x < 16) s_off[threadIdx.x] = (threadIdx.x >= 1 && threadIdx.x <= 9) ? nb_off((int)threadIdx.x, d.pitch) : 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  const unsigned gw = blockIdx.x * WPB + wib;   // global warp number
  uint32_t* wid = s_id[wib];
  unsigned long long* wtp = s_tp[wib];
  const unsigned lt = (1u << lane) - 1;
  const int pitch = d.pitch;
  unsigned home = gw % NQ, pushq = (gw * 7 + 3) % NQ;
  uint32_t cur = NONE, claim = NONE;
  unsigned long long tp = 0;
  unsigned visited = 0, idle = 0;
  unsigned n_pushed = 0, n_popped = 0, n_kept = 0, n_shared = 0, n_iters = 0, n_idle = 0, n_respin = 0;
  const unsigned nvalid = A.ctl->nvalid;
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  for (;;) {
    // ---- issue: queue claim
    const unsigned need = __ballot_sync(0xFFFFFFFFu, cur == NONE && claim == NONE);
    const int leader = __ffs(need) - 1;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(&A.qhead[home].v, (unsigned)__popc(need));
    // ---- issue: loads of the visit
    const bool go = cur != NONE;
    const size_t id = go ? cur : 0;
    const double* sp = A.st + id * ST_N;
    double2 fs = make_double2(0.0, 0.0), dv[NB];
    double rs = 0.0, de[NB];
    if (go) {
      fs = ld_cg_f64x2(sp + ST_F);
      rs = ld_cg_f64(sp + ST_RS);
#pragma unroll
      for (int j = 0; j < NB; ++j) {
        const int n = (int)(tp >> (TP_DON + 4 * j)) & 15;
        dv[j] = make_double2(0.0, 0.0);
        de[j] = 0.0;
        if (n) {
          const double* dp = A.st + (id + s_off[n]) * ST_N;
          if ((tp >> (TP_BM + j)) & 1) dv[j] = ld_cg_f64x2(dp + ST_QA);
          else { dv[j] = ld_cg_f64x2(dp + ST_QC); de[j] = ld_cg_f64(dp + ST_F); }
        }
      }
    }
    // ---- the claim comes back; poll the claimed entries
    if (need) {
      base = __shfl_sync(0xFFFFFFFFu, base, leader);
      const unsigned n = __popc(need);
      if ((need >> lane) & 1) {
        const unsigned idx = base + __popc(need & lt);
        if (idx < A.qcap) claim = home * A.qcap + idx;
      }
      if (base + n > A.qcap) home = (home + 1) % NQ;   // that queue is used up
    }
    ulonglong2 polled = make_ulonglong2(UNSET, UNSET);
    if (!go && claim != NONE) polled = ld_entry(&A.order[claim]);
    ++n_iters;
    // ---- visit
    unsigned rel = 0;
    unsigned long long old[10];
    if (go) {
      const double f = fs.x;
      double qa1 = 1.0, qa2 = K.g2;  // :35-36
#pragma unroll
      for (int h = 0; h < 8; h += NB) {
        if (h) {   // donors NB .. 7 (few cells have that many)
          if (!((tp >> (TP_DON + 4 * h)) & 15)) break;
#pragma unroll
          for (int j = 0; j < NB; ++j) {
            const int n = h + j < 8 ? (int)(tp >> (TP_DON + 4 * (h + j))) & 15 : 0;
            if (n) {
              const double* dp = A.st + (id + s_off[n]) * ST_N;
              if ((tp >> (TP_BM + h + j)) & 1) dv[j] = ld_cg_f64x2(dp + ST_QA);
              else { dv[j] = ld_cg_f64x2(dp + ST_QC); de[j] = ld_cg_f64(dp + ST_F); }
            }
          }
        }
#pragma unroll
        for (int j = 0; j < NB; ++j) {
          const int n = h + j < 8 ? (int)(tp >> (TP_DON + 4 * (h + j))) & 15 : 0;
          if (!n) break;
          const bool isb = (tp >> (TP_BM + h + j)) & 1;
          while (is_unset(dv[j].x) || is_unset(dv[j].y)) {   // the donor's decrement overtook its state stores
            dv[j] = ld_cg_f64x2(A.st + (id + s_off[n]) * ST_N + (isb ? ST_QA : ST_QC));
            ++n_respin;
          }
          if (isb) {  // :82-83, the donor's whole qa
            qa1 = qa1 + dv[j].x;
            qa2 = qa2 + dv[j].y;
          } else {    // :74-75, qc * tan of the donor towards this cell
            const double t = weti_tan(de[j], f, n, K);
            qa1 = qa1 + (dv[j].x * t);
            qa2 = qa2 + (dv[j].y * t);
          }
        }
      }
      const int cls = (int)(tp >> TP_CLS) & 3;
      double qc1 = nan, qc2 = nan;
      if (cls == CLS_A) { qc1 = div_rcp(qa1, fs.y, rs); qc2 = div_rcp(qa2, fs.y, rs); }   // :70-71
      else if (cls == CLS_B) qc1 = div_rcp(qa1, 0.0001, K.r1e4);                           // :80 (qc2 keeps its old value)
      double* op = A.st + id * ST_N;
      st_cg_f64x2(op + ST_QC, qc1, qc2);
      st_cg_f64x2(op + ST_QA, qa1, qa2);
      const unsigned succ = (unsigned)(tp >> (TP_SUCC - 1)) & 0x3FEu;
#pragma unroll
      for (int n = 1; n <= 9; ++n) {
        old[n] = 0;
        if ((succ >> n) & 1) old[n] = atom_dec_u64(&A.topo[id + nb_off(n, pitch)]);
      }
#pragma unroll
      for (int n = 1; n <= 9; ++n)
        if ((old[n] & 15ull) == 1ull) rel |= 1u << n;
    }
    // ---- hand out the released cells: the lane keeps the first one, free lanes of the warp take the next ones
    const unsigned active = __ballot_sync(0xFFFFFFFFu, go);
    visited += __popc(active);
    if (go) {
      cur = NONE;
      tp = 0;
      if (rel) {
        const int n0 = __ffs(rel) - 1;
#pragma unroll
        for (int n = 1; n <= 9; ++n)
          if (n == n0) tp = old[n] - 1;
        cur = (uint32_t)(id + s_off[n0]);
        rel &= rel - 1;
        ++n_kept;
      }
    }
    // a polled entry that was ready
    if (polled.x != UNSET && polled.y != UNSET) {
      cur = (uint32_t)polled.y;
      tp = polled.x;
      claim = NONE;
      ++n_popped;
    }
    const unsigned nextra = __popc(rel);
    if (__ballot_sync(0xFFFFFFFFu, nextra != 0)) {
      unsigned x = nextra;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= o) x += y;
      }
      const unsigned total = __shfl_sync(0xFFFFFFFFu, x, 31);
      unsigned at = x - nextra;
#pragma unroll
      for (int n = 1; n <= 9; ++n) {
        if ((rel >> n) & 1) {
          wid[at] = (uint32_t)(id + nb_off(n, pitch));
          wtp[at] = old[n] - 1;
          ++at;
        }
      }
      __syncwarp();
      const unsigned freem = __ballot_sync(0xFFFFFFFFu, cur == NONE && claim == NONE);
      const unsigned give = min(total, (unsigned)__popc(freem));
      if ((freem >> lane) & 1) {
        const unsigned rk = __popc(freem & lt);
        if (rk < give) { cur = wid[rk]; tp = wtp[rk]; ++n_shared; }
      }
      __syncwarp();
      if (total > give) {
        warp_push(A, pushq, wid + give, wtp + give, total - give, lane);
        if (lane == 0) n_pushed += total - give;
      }
      __syncwarp();
    }
    if (active) { idle = 0; continue; }
    // ---- nothing to visit in this warp
    ++n_idle;
    if (__ballot_sync(0xFFFFFFFFu, cur != NONE)) continue;   // a poll just delivered
    if (visited) {
      if (lane == 0) {
        const unsigned before = atomicAdd(&A.ctl->done, visited);
        if (before + visited == nvalid) { __threadfence(); st_volatile(&A.ctl->finished, 1u); }
      }
      visited = 0;
    }
    unsigned fin = 0;
    if (lane == 0) fin = ld_volatile(&A.ctl->finished);
    fin = __shfl_sync(0xFFFFFFFFu, fin, 0);
    if (fin) break;
    if (++idle > idle_limit) {   // watchdog: a stalled dataflow must not hang the device
      if (lane == 0) { st_volatile(&A.ctl->abort, 1u); __threadfence(); st_volatile(&A.ctl->finished, 1u); }
      break;
    }
    if (idle > 2) __nanosleep(idle > 64 ? 400 : 100);   // idle warps must not take issue slots from the visiting ones
  }
  // statistics (tuning aid)
  for (int o = 16; o; o >>= 1) {
    n_popped += __shfl_xor_sync(0xFFFFFFFFu, n_popped, o);
    n_kept += __shfl_xor_sync(0xFFFFFFFFu, n_kept, o);
    n_shared += __shfl_xor_sync(0xFFFFFFFFu, n_shared, o);
    n_respin += __shfl_xor_sync(0xFFFFFFFFu, n_respin, o);
  }
  if (lane == 0) {
    atomicAdd(&A.ctl->st_pushed, (unsigned long long)n_pushed);
    atomicAdd(&A.ctl->st_popped, (unsigned long long)n_popped);
    atomicAdd(&A.ctl->st_kept, (unsigned long long)n_kept);
    atomicAdd(&A.ctl->st_shared, (unsigned long long)n_shared);
    atomicAdd(&A.ctl->st_iters, (unsigned long long)n_iters);
    atomicAdd(&A.ctl->st_idle, (unsigned long long)n_idle);
    atomicAdd(&A.ctl->st_respin, (unsigned long long)n_respin);
  }
}

// ------------------------------------------------------------------------------------------------
// queries in visiting order: index of the query itself where it is a B cell (it assigns qc1), else -1
__global__ void k_weti_qflags(Arrays A, int* __restrict__ bflag) {
  const unsigned j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A.nq) return;
  const uint32_t id = (uint32_t)(A.qlo[j] & 0xFFFFFFFFu);
  bflag[j] = A.cls[id] == CLS_B ? (int)j : -1;
}
struct K128Max {
  __device__ __forceinline__ K128 operator()(const K128& a, const K128& b) const { return key_less(a, b) ? b : a; }
};

__device__ __forceinline__ double trunc_dec2(double x) { return trunc(x * 100.0) / 100.0; }  // R/LITAP_utils.R:93-95

struct OutPtrs { double* o[4]; };

__global__ void __launch_bounds__(WT) k_weti_final(Arrays A, GridDims d, OutPtrs P) {
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
  const double* sp = A.st + id * ST_N;
  const double2 sqc = *reinterpret_cast<const double2*>(sp + ST_QC);
  const double2 sqa = *reinterpret_cast<const double2*>(sp + ST_QA);
  double qa1 = sqa.x, qa2 = sqa.y;
  // B neighbours that drain into this cell but are visited after it: they only change the final qa
  {
    uint8_t lcode[8];
    int nl = 0;
    double f = 0;
#pragma unroll
    for (int n = 1; n <= 9; ++n) {
      if (n == 5) continue;
      const size_t nid = id + nb_off(n, d.pitch);
      if (A.cls[nid] == CLS_B && A.ddir[nid] == 10 - n) {
        if (nl == 0) f = A.elev[id];
        const double en = A.elev[nid];
        if (visited_before(A, en, nid, f, id)) continue;   // an early donor: already in qa
        int b = nl - 1;
        while (b >= 0) {
          const int nb = lcode[b];
          const size_t bid = id + nb_off(nb, d.pitch);
          if (visited_before(A, A.elev[bid], bid, en, nid)) break;
          lcode[b + 1] = lcode[b];
          --b;
        }
        lcode[b + 1] = (uint8_t)n;
        ++nl;
      }
    }
    for (int a = 0; a < nl; ++a) {
      const double2 t = *reinterpret_cast<const double2*>(A.st + (id + nb_off(lcode[a], d.pitch)) * ST_N + ST_QA);
      qa1 = qa1 + t.x;
      qa2 = qa2 + t.y;
    }
  }
  // effective qc1 / qc2 at the time the reference evaluates :87-88
  double qc1 = sqc.x, qc2 = sqc.y;
  if (cls != CLS_A) {
    const double f = A.elev[id];
    const K128 k = make_key(f, A.upslope[id], (uint32_t)id);
    const BucketMap bm = bucket_map(A);
    const unsigned j = locate(A, bm, f, k);       // this cell's own place among the queries
    const K128 a = A.slotA[j];                     // latest A cell before it
    const bool havea = (a.lo | a.hi) != 0;
    qc2 = havea ? A.st[(size_t)(a.lo & 0xFFFFFFFFu) * ST_N + ST_QC + 1] : nan;
    if (cls == CLS_C) {
      const int b = A.lastB[j];                    // latest B cell before it
      uint32_t src = havea ? (uint32_t)(a.lo & 0xFFFFFFFFu) : NONE;
      if (b >= 0) {
        K128 kb;
        kb.hi = A.qhi[b]; kb.lo = A.qlo[b];
        if (!havea || key_less(a, kb)) src = (uint32_t)(kb.lo & 0xFFFFFFFFu);
      }
      qc1 = src != NONE ? A.st[(size_t)src * ST_N + ST_QC] : nan;
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
  K.rorth = div_rcp_prepare(K.orth);
  K.rdiag = div_rcp_prepare(K.diag);
  K.r1e4 = div_rcp_prepare(0.0001);

  // launch shape of the dataflow kernel: as many blocks as are resident (waiting lanes must never keep a block from running)
  static int flow_blocks = 0, flow_nb = 8;
  if (!flow_blocks) {
    if (const char* e = getenv("LITAP_WETI_NB")) flow_nb = atoi(e);  // tuning aid: donors loaded per batch (8, 6 or 4)
    if (flow_nb != 6 && flow_nb != 4) flow_nb = 8;
    int per_sm = 0;
    if (flow_nb == 8) CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_weti_flow<8>, WT, 0));
    else if (flow_nb == 6) CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_weti_flow<6>, WT, 0));
    else CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_weti_flow<4>, WT, 0));
    if (per_sm < 1) return set_error(LITAP_ECUDA, "k_weti_flow cannot be made resident");
    if (const char* e = getenv("LITAP_WETI_BPS")) per_sm = std::max(1, std::min(per_sm, atoi(e)));  // tuning aid
    flow_blocks = c.sm_count * per_sm;
  }

  const size_t ncell = (size_t)d.cells + 64;          // pitched ids
  const size_t nd = (size_t)d.NR * d.NC;
  const size_t lanes = (size_t)flow_blocks * WT;
  const unsigned qcap = (unsigned)((nd + lanes) / NQ + 4096);
  PoolScratch S;
  Arrays A = {};
  A.elev = g->elev;
  A.ddir = g->ddir;
  A.upslope = g->upslope;
  A.topo = S.get<unsigned long long>(ncell);
  A.st = S.get<double>(ncell * ST_N);
  A.cls = S.get<uint8_t>(ncell);
  A.order = S.get<ulonglong2>((size_t)NQ * qcap);
  A.qcap = qcap;
  unsigned char* ctlmem = S.get<unsigned char>(sizeof(Ctl) + 2 * NQ * sizeof(QCount) + 256);
  unsigned* table = S.get<unsigned>(NT + 4 + 4);   // [NT + 1] counts, then (16-byte aligned) the bucket map
  if (!A.topo || !A.st || !A.cls || !A.order || !ctlmem || !table)
    return set_error(LITAP_ENOMEM, "device allocation failed (weti scratch)");
  A.qtail = reinterpret_cast<QCount*>(ctlmem);
  A.qhead = A.qtail + NQ;
  A.ctl = reinterpret_cast<Ctl*>(A.qhead + NQ);
  A.table = table;
  CUDA_TRY(cudaMemsetAsync(ctlmem, 0, sizeof(Ctl) + 2 * NQ * sizeof(QCount), st));
  CUDA_TRY(cudaMemsetAsync(A.cls, 0, ncell, st));  // pad columns read as NA
  CUDA_TRY(cudaMemsetAsync(A.order, 0xFF, (size_t)NQ * qcap * sizeof(ulonglong2), st));

  dim3 cgrid(nblk(d.NC, WT), d.NR);
  KERNEL("k_weti_cls", k_weti_cls<<<cgrid, WT, 0, st>>>(A, d, K));
  Ctl hctl;
  CUDA_TRY(cudaMemcpyAsync(&hctl, A.ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));   // the number of B / C cells sizes the query arrays
  const unsigned nq = hctl.nq;
  A.nq = nq;
  if (hctl.nvalid) {
    int* lastB = nullptr;
    if (nq) {
      unsigned long long* hi_a = S.get<unsigned long long>(nq);
      unsigned long long* hi_b = S.get<unsigned long long>(nq);
      unsigned long long* lo_a = S.get<unsigned long long>(nq);
      unsigned long long* lo_b = S.get<unsigned long long>(nq);
      A.slotA = S.get<K128>((size_t)nq + 1);
      lastB = S.get<int>(nq);
      unsigned* qcount = S.get<unsigned>(1);
      size_t t1 = 0, t2 = 0, t3 = 0;
      cub::DeviceRadixSort::SortPairs(nullptr, t1, lo_a, lo_b, hi_a, hi_b, (int)nq, 0, 64, st);
      cub::DeviceScan::InclusiveScan(nullptr, t2, (K128*)nullptr, (K128*)nullptr, K128Max(), (int)nq + 1, st);
      cub::DeviceScan::InclusiveScan(nullptr, t3, (int*)nullptr, (int*)nullptr, cub::Max(), (int)nq, st);
      const size_t tn = std::max(t1, std::max(t2, t3)) + 256;
      void* tmp = S.get<unsigned char>(tn);
      if (!hi_a || !hi_b || !lo_a || !lo_b || !A.slotA || !lastB || !qcount || !tmp)
        return set_error(LITAP_ENOMEM, "device allocation failed (weti query scratch)");
      CUDA_TRY(cudaMemsetAsync(qcount, 0, sizeof(unsigned), st));
      CUDA_TRY(cudaMemsetAsync(A.slotA, 0, ((size_t)nq + 1) * sizeof(K128), st));
      KERNEL("k_weti_queries", k_weti_queries<<<cgrid, WT, 0, st>>>(A, d, hi_a, lo_a, qcount));
      // visiting order of the queries: stable LSD passes, (upslope, id) then elevation
      size_t tcap = tn;
      CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tcap, lo_a, lo_b, hi_a, hi_b, (int)nq, 0, 64, st));
      tcap = tn;
      CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tcap, hi_b, hi_a, lo_b, lo_a, (int)nq, 0, 64, st));
      c.launches += 2;
      A.qhi = hi_a;
      A.qlo = lo_a;
      KERNEL("k_weti_table", k_weti_table<<<nblk(NT + 1), 256, 0, st>>>(A, table));
      KERNEL("k_weti_qflags", k_weti_qflags<<<nblk(nq), 256, 0, st>>>(A, lastB));
      tcap = tn;
      CUDA_TRY(cub::DeviceScan::InclusiveScan(tmp, tcap, lastB, lastB, cub::Max(), (int)nq, st));
      c.launches += 1;
      A.lastB = lastB;
      KERNEL("k_weti_nodes", k_weti_nodes<<<cgrid, WT, 0, st>>>(A, d, K));
      tcap = tn;
      CUDA_TRY(cub::DeviceScan::InclusiveScan(tmp, tcap, A.slotA, A.slotA, K128Max(), (int)nq + 1, st));
      c.launches += 1;
    } else {
      KERNEL("k_weti_nodes", k_weti_nodes<<<cgrid, WT, 0, st>>>(A, d, K));
    }
    unsigned idle_limit = 1u << 22;   // ~ 2 s of polling without any progress anywhere
    if (flow_nb == 8) KERNEL("k_weti_flow", k_weti_flow<8><<<flow_blocks, WT, 0, st>>>(A, d, K, idle_limit));
    else if (flow_nb == 6) KERNEL("k_weti_flow", k_weti_flow<6><<<flow_blocks, WT, 0, st>>>(A, d, K, idle_limit));
    else KERNEL("k_weti_flow", k_weti_flow<4><<<flow_blocks, WT, 0, st>>>(A, d, K, idle_limit));
  }
  KERNEL("k_weti_final", k_weti_final<<<cgrid, WT, 0, st>>>(A, d, P));
  CUDA_TRY(cudaMemcpyAsync(&hctl, A.ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));  // scratch goes back to the pool
  if (hctl.abort) return set_error(LITAP_ECUDA, "calc_weti2: the dataflow stalled (dependency cycle in the inputs?)");
  c.tm.weti_queued = (int)std::min<unsigned long long>(hctl.st_pushed, 2147483647ull);
  if (getenv("LITAP_WETI_STATS"))
    fprintf(stderr,
            "weti: cells %u queries %u | kept by the releasing lane %llu, shared in the warp %llu, queued %llu (popped %llu) | "
            "warp iterations %llu (idle %llu), state re-reads %llu, %d blocks, %d donors per batch\n",
            hctl.nvalid, hctl.nq, hctl.st_kept, hctl.st_shared, hctl.st_pushed, hctl.st_popped, hctl.st_iters, hctl.st_idle,
            hctl.st_respin, flow_blocks, flow_nb);
  return LITAP_OK;
}
