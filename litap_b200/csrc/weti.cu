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
//                   it released for itself, so a dependency chain costs one atomic + two memory round
//                   trips per hop; other released cells go to free lanes of the warp, the rest to one
//                   of 64 global ready queues that idle lanes wait on (each waits on its own slot)
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

struct __align__(32) Node {   // everything a visit needs about the cell itself: one 32-byte sector
  double S;          // sum(tan) of the cell (A cells: != 0)
  double f;          // its elevation
  uint32_t donors;   // up to 8 nibbles: direction codes (from the cell towards the donor) in visiting order, 0 ends
  uint16_t succ;     // bit n: neighbour n waits for this cell (lower neighbours of an A cell; drec of an early B cell)
  uint8_t cls;       // CLS_*
  uint8_t bmask;     // bit j: j-th donor is a B donor (contributes its whole qa)
  uint32_t pad0, pad1;
};
struct __align__(32) State {
  double qc1, qc2;   // as assigned by the cell itself (NaN where the reference leaves the variable stale)
  double qa1, qa2;   // qa at processing time
};
// visiting-order key: later visited <=> (hi, lo) larger.  hi = ~orderable(elev), lo = upslope << 32 | cell id
struct __align__(16) K128 { unsigned long long lo, hi; };
struct __align__(128) QCount { unsigned v; unsigned pad[31]; };
struct Ctl {
  unsigned nvalid, nq, done, finished, abort, pad[3];
  unsigned long long st_pushed, st_popped, st_kept, st_shared, st_iters, st_idle;
};

struct Arrays {
  const double* elev;
  const uint8_t* ddir;
  const uint32_t* upslope;
  Node* node;
  State* state;
  uint8_t* cls;        // CLS_* only (neighbour lookups)
  uint32_t* pend32;    // packed 8-bit pending-donor counters
  uint32_t* order;     // NQ ready queues of qcap slots each, NONE = not yet written
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
__device__ __forceinline__ BucketMap bucket_map(const Arrays& A) {
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
// table[b] = number of queries in buckets < b
__global__ void k_weti_table(Arrays A, unsigned* __restrict__ table) {
  const unsigned b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b > (unsigned)NT) return;
  const BucketMap m = bucket_map(A);
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
// ready queues.  Queue q owns order[q*qcap, (q+1)*qcap); a slot is written once (NONE -> cell id) and
// read by the one lane that claimed its index from qhead.  Items that do not fit a queue go to the next
// one; a claim beyond qcap is void and sends the claiming warp on to the next queue.  Total capacity
// exceeds cells + waiting lanes, so neither side can run out.
__device__ __forceinline__ void st_volatile(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }
__device__ __forceinline__ uint32_t ld_volatile(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }

// the whole warp pushes items[0, n) (shared memory)
__device__ __forceinline__ void warp_push(const Arrays& A, unsigned& pushq, const uint32_t* items, unsigned n, int lane) {
  while (n) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(&A.qtail[pushq].v, n);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    const unsigned fit = base >= A.qcap ? 0u : min(n, A.qcap - base);
    uint32_t* dst = A.order + (size_t)pushq * A.qcap + base;
    for (unsigned k = lane; k < fit; k += 32) st_volatile(dst + k, items[k]);
    items += fit;
    n -= fit;
    pushq = (pushq + 1) % NQ;
  }
}

// ------------------------------------------------------------------------------------------------
// node of every cell: S, donors in visiting order, successors, pending counter; seeds; last-setter slots
__global__ void __launch_bounds__(WT) k_weti_nodes(Arrays A, GridDims d, WetiK K) {
  __shared__ uint32_t s_ids[WT];
  __shared__ unsigned s_n;
  if (threadIdx.x == 0) s_n = 0;
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c < d.NC) {
    const size_t id = (size_t)r * d.pitch + c;
    const uint8_t cls = A.cls[id];
    Node m;
    m.S = 0; m.f = 0; m.donors = 0; m.succ = 0; m.cls = cls; m.bmask = 0; m.pad0 = 0; m.pad1 = 0;
    if (cls != CLS_NA) {
      double e[10];
#pragma unroll
      for (int n = 1; n <= 9; ++n) e[n] = A.elev[id + nb_off(n, d.pitch)];
      const double f = e[5];
      m.f = f;
      unsigned lower = 0;
      if (cls == CLS_A) m.S = weti_S(e, K, &lower);
      m.succ = (uint16_t)lower;
      const int dv = A.ddir[id];
      if (cls == CLS_B && dv >= 1 && dv <= 9) {   // an early B cell pushes its qa into drec before drec is visited
        const size_t tid = id + nb_off(dv, d.pitch);
        if (A.cls[tid] != CLS_NA && visited_before(A, f, id, e[dv], tid)) m.succ = (uint16_t)(1u << dv);
      }
      // donors that precede the cell: strictly higher neighbours (they are A cells and push qc*tan, :74-75) and B
      // neighbours whose drec is this cell and that are visited first (they push their whole qa, :82-83)
      unsigned dmask = 0, bdon = 0;
#pragma unroll
      for (int n = 1; n <= 9; ++n) {
        if (n == 5) continue;
        const size_t nid = id + nb_off(n, d.pitch);
        const uint8_t nc = A.cls[nid];
        if (nc == CLS_A) {
          if (weti_tan(e[n], f, 10 - n, K) > 0) dmask |= 1u << n;   // the same test that sets the donor's succ bit
        } else if (nc == CLS_B && A.ddir[nid] == 10 - n) {
          if (visited_before(A, e[n], nid, f, id)) { dmask |= 1u << n; bdon |= 1u << n; }
        }
      }
      // position of a donor in the visiting order = number of donors visited before it (static indices: registers)
      int pos[10];
#pragma unroll
      for (int n = 1; n <= 9; ++n) pos[n] = 0;
#pragma unroll
      for (int a = 1; a <= 9; ++a) {
        if (a == 5) continue;
#pragma unroll
        for (int b = a + 1; b <= 9; ++b) {
          if (b == 5) continue;
          if (((dmask >> a) & 1) && ((dmask >> b) & 1)) {
            if (visited_before(A, e[a], id + nb_off(a, d.pitch), e[b], id + nb_off(b, d.pitch))) ++pos[b]; else ++pos[a];
          }
        }
      }
      const int nd = __popc(dmask);
#pragma unroll
      for (int n = 1; n <= 9; ++n) {
        if (n == 5 || !((dmask >> n) & 1)) continue;
        m.donors |= (uint32_t)n << (4 * pos[n]);
        if ((bdon >> n) & 1) m.bmask |= (uint8_t)(1u << pos[n]);
      }
      reinterpret_cast<uint8_t*>(A.pend32)[id] = (uint8_t)nd;
      if (nd == 0) s_ids[atomicAdd(&s_n, 1u)] = (uint32_t)id;
      // last-setter candidate: the A cell records itself in the interval between two queries it falls in
      if (cls == CLS_A && A.nq) {
        const K128 k = make_key(f, A.upslope[id], (uint32_t)id);
        const BucketMap bm = bucket_map(A);
        slot_max(&A.slotA[locate(A, bm, f, k)], k);
      }
    }
    *reinterpret_cast<uint4*>(&A.node[id]) = *reinterpret_cast<const uint4*>(&m);
    *(reinterpret_cast<uint4*>(&A.node[id]) + 1) = *(reinterpret_cast<const uint4*>(&m) + 1);
  }
  __syncthreads();
  // seeds of this block -> one ready queue (overflow goes on to the next)
  unsigned n = s_n;
  if (!n) return;
  __shared__ unsigned s_base;
  unsigned q = (blockIdx.y * gridDim.x + blockIdx.x) % NQ;
  const uint32_t* items = s_ids;
  while (n) {
    if (threadIdx.x == 0) s_base = atomicAdd(&A.qtail[q].v, n);
    __syncthreads();
    const unsigned base = s_base;
    const unsigned fit = base >= A.qcap ? 0u : min(n, A.qcap - base);
    for (unsigned k = threadIdx.x; k < fit; k += blockDim.x) A.order[(size_t)q * A.qcap + base + k] = items[k];
    items += fit;
    n -= fit;
    q = (q + 1) % NQ;
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// visit one ready cell: add the donors' contributions in the reference's order, derive qc, publish the
// state, release the successors.  Returns the mask of neighbour codes whose cells became ready.
__device__ __forceinline__ unsigned visit(const Arrays& A, int pitch, const WetiK& K, size_t id, int flags) {
  Node m;
  {
    const uint4* p = reinterpret_cast<const uint4*>(&A.node[id]);
    const uint4 a = __ldg(p), b = __ldg(p + 1);
    *reinterpret_cast<uint4*>(&m) = a;
    *(reinterpret_cast<uint4*>(&m) + 1) = b;
  }
  const double f = m.f;
  // all donor loads are issued before the (ordered) additions: one memory round trip, not eight
  double2 dv[8];
  double de[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int n = (m.donors >> (4 * j)) & 15;
    dv[j] = make_double2(0.0, 0.0);
    de[j] = 0.0;
    if (n) {
      const size_t nid = id + nb_off(n, pitch);
      const bool isb = (m.bmask >> j) & 1;
      dv[j] = __ldcg(reinterpret_cast<const double2*>(&A.state[nid]) + (isb ? 1 : 0));
      if (!isb) de[j] = __ldg(&A.elev[nid]);
    }
  }
  if (flags & 1) {   // pull the successors' nodes towards L2 while the donors' state is in flight
#pragma unroll
    for (int n = 1; n <= 9; ++n)
      if ((m.succ >> n) & 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(&A.node[id + nb_off(n, pitch)]));
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
  const double nan = __longlong_as_double(0x7ff8000000000000LL);
  double2 qc;
  if (m.cls == CLS_A) { qc.x = qa1 / m.S; qc.y = qa2 / m.S; }       // :70-71
  else if (m.cls == CLS_B) { qc.x = qa1 / 0.0001; qc.y = nan; }      // :80 (qc2 keeps its old value)
  else { qc.x = nan; qc.y = nan; }
  double2* op = reinterpret_cast<double2*>(&A.state[id]);
  __stcg(op, qc);
  __stcg(op + 1, make_double2(qa1, qa2));
  unsigned rel = 0;
  if (m.succ) {
    __threadfence();
    unsigned old[10];
#pragma unroll
    for (int n = 1; n <= 9; ++n) {
      old[n] = 0;
      if ((m.succ >> n) & 1) {
        const size_t tid = id + nb_off(n, pitch);
        old[n] = atomicSub(&A.pend32[tid >> 2], 1u << (8u * (unsigned)(tid & 3)));
      }
    }
#pragma unroll
    for (int n = 1; n <= 9; ++n) {
      if ((m.succ >> n) & 1) {
        const size_t tid = id + nb_off(n, pitch);
        if (((old[n] >> (8u * (unsigned)(tid & 3))) & 0xFFu) == 1u) rel |= 1u << n;
      }
    }
  }
  return rel;
}

__global__ void __launch_bounds__(WT) k_weti_flow(Arrays A, GridDims d, WetiK K, int flags, unsigned idle_limit) {
  __shared__ uint32_t s_buf[WPB][WBUF];
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  const unsigned gw = blockIdx.x * WPB + wib;   // global warp number
  uint32_t* wbuf = s_buf[wib];
  const unsigned lt = (1u << lane) - 1;
  unsigned home = gw % NQ, pushq = (gw * 7 + 3) % NQ;
  uint32_t cur = NONE, claim = NONE;
  unsigned visited = 0, idle = 0;
  unsigned long long n_pushed = 0, n_popped = 0, n_kept = 0, n_shared = 0, n_iters = 0, n_idle = 0;
  const unsigned nvalid = A.ctl->nvalid;
  for (;;) {
    // lanes without a cell and without a claim take the next index of the warp's home queue
    const unsigned need = __ballot_sync(0xFFFFFFFFu, cur == NONE && claim == NONE);
    if (need) {
      const int leader = __ffs(need) - 1;
      const unsigned n = __popc(need);
      unsigned base = 0;
      if (lane == leader) base = atomicAdd(&A.qhead[home].v, n);
      base = __shfl_sync(0xFFFFFFFFu, base, leader);
      if ((need >> lane) & 1) {
        const unsigned idx = base + __popc(need & lt);
        if (idx < A.qcap) claim = home * A.qcap + idx;
      }
      if (base + n > A.qcap) home = (home + 1) % NQ;   // that queue is used up
    }
    if (cur == NONE && claim != NONE) {
      const uint32_t v = ld_volatile(&A.order[claim]);
      if (v != NONE) { cur = v; claim = NONE; ++n_popped; }
    }
    const unsigned active = __ballot_sync(0xFFFFFFFFu, cur != NONE);
    ++n_iters;
    if (!active) {
      ++n_idle;
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
      if (idle > 16) __nanosleep(idle > 256 ? 400 : 100);
      continue;
    }
    idle = 0;
    unsigned rel = 0;
    size_t id = cur;
    if (cur != NONE) rel = visit(A, d.pitch, K, id, flags);
    __syncwarp();
    visited += __popc(active);
    // the lane keeps the first cell it released; the others go to free lanes of the warp, then to a queue
    uint32_t next = NONE;
    if (rel) {
      const int n0 = __ffs(rel) - 1;
      next = (uint32_t)(id + nb_off(n0, d.pitch));
      rel &= rel - 1;
      ++n_kept;
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
      while (rel) {
        const int n = __ffs(rel) - 1;
        rel &= rel - 1;
        wbuf[at++] = (uint32_t)(id + nb_off(n, d.pitch));
      }
      __syncwarp();
      if (cur != NONE) cur = next;
      const unsigned freem = __ballot_sync(0xFFFFFFFFu, cur == NONE && claim == NONE);
      const unsigned nfree = __popc(freem);
      const unsigned give = min(total, nfree);
      if ((freem >> lane) & 1) {
        const unsigned rk = __popc(freem & lt);
        if (rk < give) { cur = wbuf[rk]; ++n_shared; }
      }
      __syncwarp();
      if (total > give) {
        warp_push(A, pushq, wbuf + give, total - give, lane);
        if (lane == 0) n_pushed += total - give;
      }
      __syncwarp();
    } else {
      if (cur != NONE) cur = next;
    }
  }
  // statistics (tuning aid)
  for (int o = 16; o; o >>= 1) {
    n_popped += __shfl_xor_sync(0xFFFFFFFFu, n_popped, o);
    n_kept += __shfl_xor_sync(0xFFFFFFFFu, n_kept, o);
    n_shared += __shfl_xor_sync(0xFFFFFFFFu, n_shared, o);
  }
  if (lane == 0) {
    atomicAdd(&A.ctl->st_pushed, n_pushed);
    atomicAdd(&A.ctl->st_popped, n_popped);
    atomicAdd(&A.ctl->st_kept, n_kept);
    atomicAdd(&A.ctl->st_shared, n_shared);
    atomicAdd(&A.ctl->st_iters, n_iters);
    atomicAdd(&A.ctl->st_idle, n_idle);
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
  const State s = A.state[id];
  double qa1 = s.qa1, qa2 = s.qa2;
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
      const State t = A.state[id + nb_off(lcode[a], d.pitch)];
      qa1 = qa1 + t.qa1;
      qa2 = qa2 + t.qa2;
    }
  }
  // effective qc1 / qc2 at the time the reference evaluates :87-88
  double qc1 = s.qc1, qc2 = s.qc2;
  if (cls != CLS_A) {
    const double f = A.elev[id];
    const K128 k = make_key(f, A.upslope[id], (uint32_t)id);
    const BucketMap bm = bucket_map(A);
    const unsigned j = locate(A, bm, f, k);       // this cell's own place among the queries
    const K128 a = A.slotA[j];                     // latest A cell before it
    const bool havea = (a.lo | a.hi) != 0;
    qc2 = havea ? A.state[a.lo & 0xFFFFFFFFu].qc2 : nan;
    if (cls == CLS_C) {
      const int b = A.lastB[j];                    // latest B cell before it
      uint32_t src = havea ? (uint32_t)(a.lo & 0xFFFFFFFFu) : NONE;
      if (b >= 0) {
        K128 kb;
        kb.hi = A.qhi[b]; kb.lo = A.qlo[b];
        if (!havea || key_less(a, kb)) src = (uint32_t)(kb.lo & 0xFFFFFFFFu);
      }
      qc1 = src != NONE ? A.state[src].qc1 : nan;
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

  // launch shape of the dataflow kernel: as many blocks as are resident (waiting lanes must never keep a block from running)
  static int flow_blocks = 0;
  if (!flow_blocks) {
    int per_sm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_weti_flow, WT, 0));
    if (per_sm < 1) return set_error(LITAP_ECUDA, "k_weti_flow cannot be made resident");
    if (const char* e = getenv("LITAP_WETI_BPS")) per_sm = std::max(1, std::min(per_sm, atoi(e)));  // tuning aid
    flow_blocks = c.sm_count * per_sm;
  }
  int flags = 1;
  if (const char* e = getenv("LITAP_WETI_FLAGS")) flags = atoi(e);  // tuning aid: bit 0 = prefetch successor nodes

  const size_t ncell = (size_t)d.cells + 64;          // pitched ids
  const size_t nd = (size_t)d.NR * d.NC;
  const size_t lanes = (size_t)flow_blocks * WT;
  const unsigned qcap = (unsigned)((nd + lanes) / NQ + 4096);
  PoolScratch S;
  Arrays A = {};
  A.elev = g->elev;
  A.ddir = g->ddir;
  A.upslope = g->upslope;
  A.node = S.get<Node>(ncell);
  A.state = S.get<State>(ncell);
  A.cls = S.get<uint8_t>(ncell);
  A.pend32 = S.get<uint32_t>(ncell / 4 + 2);
  A.order = S.get<uint32_t>((size_t)NQ * qcap);
  A.qcap = qcap;
  unsigned char* ctlmem = S.get<unsigned char>(sizeof(Ctl) + 2 * NQ * sizeof(QCount) + 256);
  unsigned* table = S.get<unsigned>(NT + 1);
  if (!A.node || !A.state || !A.cls || !A.pend32 || !A.order || !ctlmem || !table)
    return set_error(LITAP_ENOMEM, "device allocation failed (weti scratch)");
  A.qtail = reinterpret_cast<QCount*>(ctlmem);
  A.qhead = A.qtail + NQ;
  A.ctl = reinterpret_cast<Ctl*>(A.qhead + NQ);
  A.table = table;
  CUDA_TRY(cudaMemsetAsync(ctlmem, 0, sizeof(Ctl) + 2 * NQ * sizeof(QCount), st));
  CUDA_TRY(cudaMemsetAsync(A.cls, 0, ncell, st));  // pad columns read as NA
  CUDA_TRY(cudaMemsetAsync(A.order, 0xFF, (size_t)NQ * qcap * sizeof(uint32_t), st));

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
    KERNEL("k_weti_flow", k_weti_flow<<<flow_blocks, WT, 0, st>>>(A, d, K, flags, idle_limit));
  }
  KERNEL("k_weti_final", k_weti_final<<<cgrid, WT, 0, st>>>(A, d, P));
  CUDA_TRY(cudaMemcpyAsync(&hctl, A.ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));  // scratch goes back to the pool
  if (hctl.abort) return set_error(LITAP_ECUDA, "calc_weti2: the dataflow stalled (dependency cycle in the inputs?)");
  c.tm.weti_queued = (int)std::min<unsigned long long>(hctl.st_pushed, 2147483647ull);
  if (getenv("LITAP_WETI_STATS"))
    fprintf(stderr,
            "weti: cells %u queries %u | kept by the releasing lane %llu, shared in the warp %llu, queued %llu (popped %llu) | "
            "warp iterations %llu (idle %llu), %d blocks\n",
            hctl.nvalid, hctl.nq, hctl.st_kept, hctl.st_shared, hctl.st_pushed, hctl.st_popped, hctl.st_iters, hctl.st_idle,
            flow_blocks);
  return LITAP_OK;
}
