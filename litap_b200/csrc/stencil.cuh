// stencil.cuh — the row-sliding 3x3 window shared by the D8 (ddir.cu) and form (form.cu) kernels.
//
// A thread owns SW_CPT consecutive columns; a warp therefore reads 32*SW_CPT*8 contiguous bytes of
// a row with 16-byte vector loads (the first interior column is 256-byte aligned, LITAP_XOFF).
// The two window columns a thread does not own come from its lane neighbours by shuffle (lanes 0 and
// 31 fetch them from memory).  Rows slide through registers: the loads of row r+2 are issued before
// the cells of row r are computed, so every warp keeps two row requests in flight.
//
// Columns NC..pitch-1 of every row are readable pad (NaN for elev), so no column bound checks are
// needed on loads; rows outside [0, NR) read as NaN.
#pragma once
#include "litap_internal.h"

template <int CPT>
struct RowWin {          // v[0] = left neighbour column, v[1..CPT] = own columns, v[CPT+1] = right neighbour
  double v[CPT + 2];
};

__device__ __forceinline__ double sw_nan() { return __longlong_as_double(0x7ff8000000000000LL); }

// issue the loads of one row (no use of the values yet)
template <int CPT>
__device__ __forceinline__ void sw_issue(RowWin<CPT>& w, const double* __restrict__ elev, const GridDims& d, int r,
                                         int c0, int lane) {
  static_assert(CPT == 2 || CPT == 4, "vector loads of 2 doubles");
  const bool rv = (r >= 0 && r < d.NR);
  if (rv) {
    const double* row = elev + (size_t)r * d.pitch;
#pragma unroll
    for (int k = 0; k < CPT; k += 2) {
      double2 t = __ldg(reinterpret_cast<const double2*>(row + c0 + k));
      w.v[1 + k] = t.x;
      w.v[2 + k] = t.y;
    }
    w.v[0] = w.v[CPT + 1] = 0.0;
    if (lane == 0) w.v[0] = __ldg(row + c0 - 1);
    if (lane == 31) w.v[CPT + 1] = __ldg(row + c0 + CPT);
  } else {
#pragma unroll
    for (int k = 0; k < CPT + 2; ++k) w.v[k] = sw_nan();
  }
}
// complete the row: neighbour columns from the adjacent lanes
template <int CPT>
__device__ __forceinline__ void sw_finish(RowWin<CPT>& w, int lane) {
  double l = __shfl_up_sync(0xffffffffu, w.v[CPT], 1);
  double r = __shfl_down_sync(0xffffffffu, w.v[1], 1);
  if (lane != 0) w.v[0] = l;
  if (lane != 31) w.v[CPT + 1] = r;
}
