// rowring.cuh — a ring of DEM row segments in shared memory, filled by the TMA unit.
//
// A block of a stencil kernel owns RR_COLS consecutive columns and walks down its rows.  Row
// segments [cb-2, cb+RR_COLS+2) (one halo column each side, widened to 16-byte alignment) are
// fetched with 1-D bulk async copies (cp.async.bulk ... mbarrier::complete_tx::bytes): the copy
// engine writes shared memory and signals an mbarrier per ring slot, so loads cost no registers,
// no issue slots and run RR_STAGES - RR_GROUP .. RR_STAGES rows ahead of the arithmetic.
// One thread issues; slots are recycled in groups of RR_GROUP rows behind a __syncthreads().
//
// Layout facts used (litap_internal.h): column 1 of every row is 256-byte aligned, so column
// cb - 2 with cb = 1 + k*RR_COLS is 16-byte aligned; the allocation extends 32 elements past the
// last row (reads past a row's pitch land in the next row's pads or in that tail).
#pragma once
#include <stdint.h>
#include "litap_internal.h"

template <typename T, int COLS, int STAGES>
struct RowRing {
  static constexpr int SEG = COLS + 4;              // elements per slot
  alignas(16) T row[STAGES][SEG];
  alignas(8) unsigned long long bar[STAGES];
};

__device__ __forceinline__ unsigned rr_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void rr_bar_init(unsigned long long* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(rr_smem(bar)) : "memory");
}
__device__ __forceinline__ void rr_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// one thread: arm the slot's barrier with the byte count and start the copy
__device__ __forceinline__ void rr_issue(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  const unsigned b = rr_smem(bar);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(rr_smem(dst)),
               "l"(src), "r"(bytes), "r"(b)
               : "memory");
}
// one thread: complete the slot's phase without data (a row outside the grid), so that the slot's
// phase parity keeps following k / STAGES
__device__ __forceinline__ void rr_skip(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rr_smem(bar)) : "memory");
}
__device__ __forceinline__ void rr_wait(unsigned long long* bar, unsigned parity) {
  const unsigned b = rr_smem(bar);
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "RR_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra RR_DONE;\n"
      "bra RR_WAIT;\n"
      "RR_DONE:\n"
      "}\n" ::"r"(b),
      "r"(parity)
      : "memory");
}
