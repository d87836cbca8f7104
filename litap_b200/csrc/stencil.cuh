// stencil.cuh — the 3x3 row window shared by the D8 (ddir.cu) and form (form.cu) kernels.
//
// A thread owns CPT consecutive columns of a block's column range and keeps, for each of the rows r-1, r, r+1, its
// own columns plus the column either side in registers (RowWin).  The rows reach the block through the TMA row ring
// (rowring.cuh): a thread reads its window columns from the ring slot of the row, so there are no neighbour-lane
// shuffles and no edge-lane loads; rows slide through the three windows as the block walks down.
// (Round 1 prefetched the rows through registers, one row ahead per warp: ~18 KB in flight per SM - latency-bound
// once the arithmetic was cheap enough.)
//
// Columns NC..pitch-1 of every row are readable pad (NaN for elev); rows outside [0, NR) read as NaN.
#pragma once
#include "litap_internal.h"

template <int CPT>
struct RowWin {          // v[0] = left neighbour column, v[1..CPT] = own columns, v[CPT+1] = right neighbour
  double v[CPT + 2];
};

__device__ __forceinline__ double sw_nan() { return __longlong_as_double(0x7ff8000000000000LL); }
