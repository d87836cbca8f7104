// shed_tile.cuh — tile-local building blocks of calc_shed4 on the device (included by shed.cu).
//
// The drec forest is processed in tiles of TW x TH interior cells staged in shared memory with a
// 1-cell ring (local grid LW x LH, 16-bit local indices).  Inside a tile every cell is resolved
// to its local terminal by pointer jumping in shared memory:
//     pit   (ddir == 5)                       -> a root of the forest
//     exit  (ddir points out of the tile)     -> continues at the entry cell of the next tile
//     NA    (ddir NA, chain leaves the grid, or a cycle)
// and flow is accumulated inside the tile by chasing from the source cells with shared-memory
// atomics (each cell forwards its count once all its in-tile donors have arrived).  Only the tile
// border cells (4 x 63 = 252 "slots" per tile) take part in the global, level-2 graph.
#pragma once
#include "cellmath.h"
#include "litap_internal.h"

namespace shed {

constexpr int NT = 512;            // threads per tile CTA
constexpr int CPT = TW * TH / NT;  // interior cells per thread (8)
static_assert(TW * TH % NT == 0, "tile must divide evenly over the CTA");

struct TileGeom {
  GridDims d;
  int tilesX, tilesY;
};

__device__ __forceinline__ int loc_off(int dir) { return nb_dr(dir) * LW + nb_dc(dir); }
__device__ __forceinline__ bool loc_is_ring(int li) {
  int ly = li / LW, lx = li - ly * LW;
  return lx == 0 || lx == LW - 1 || ly == 0 || ly == LH - 1;
}
// k-th interior cell owned by thread t: consecutive threads take consecutive columns of one row
__device__ __forceinline__ int owned_local(int t, int k) {
  int ii = k * NT + t;  // 0..TW*TH-1
  return (ii / TW + 1) * LW + (ii % TW + 1);
}
__device__ __forceinline__ int border_slot_local(int lx, int ly) {
  if (ly == 1) return lx - 1;
  if (ly == TH) return TW + lx - 1;
  if (lx == 1) return 2 * TW + (ly - 2);
  if (lx == TW) return 2 * TW + (TH - 2) + (ly - 2);
  return -1;
}
// global slot index of a cell (r >= 1, c >= 1 required); -1 if the cell is not on a tile border
__device__ __forceinline__ long long slot_of_rc(int r, int c, int tilesX) {
  int tx = (c - 1) / TW, ty = (r - 1) / TH;
  int s = border_slot_local((c - 1) % TW + 1, (r - 1) % TH + 1);
  return s < 0 ? -1 : ((long long)(ty * tilesX + tx) * BSLOTS + s);
}
__device__ __forceinline__ long long slot_of_id(uint32_t id, const TileGeom& g) {
  int r = id / g.d.pitch, c = id - r * g.d.pitch;
  return slot_of_rc(r, c, g.tilesX);
}

// terminal kinds
enum { T_NA = 0, T_PIT = 1, T_EXIT = 2 };

struct TileCtx {
  int tx, ty, r_base, c_base;  // global (r, c) of local (0, 0)
};
__device__ __forceinline__ TileCtx tile_ctx() {
  TileCtx t;
  t.tx = blockIdx.x;
  t.ty = blockIdx.y;
  t.r_base = t.ty * TH;
  t.c_base = t.tx * TW;
  return t;
}
__device__ __forceinline__ bool in_grid(const TileGeom& g, int r, int c) { return r < g.d.NR && c < g.d.NC; }

// Phase 0: stage ddir of the local grid (0 outside the grid and on buffered row 0 / column 0, which
// belong to no tile: they are the NA ring).  Cells whose ddir points outside the grid or onto
// row 0 / column 0 become NA themselves (their chain has an NA drec).
__device__ __forceinline__ void tile_load_dir(const TileGeom& g, const TileCtx& t, const uint8_t* __restrict__ ddir,
                                              uint8_t* s_dir) {
  for (int li = threadIdx.x; li < LC; li += NT) {
    int ly = li / LW, lx = li - ly * LW;
    int r = t.r_base + ly, c = t.c_base + lx;
    uint8_t dv = 0;
    if (r >= 1 && c >= 1 && in_grid(g, r, c)) {
      dv = ddir[(size_t)r * g.d.pitch + c];
      if (dv > 9) dv = 0;
      if (dv != 0 && dv != 5) {
        int rr = r + nb_dr(dv), cc = c + nb_dc(dv);
        if (rr < 1 || cc < 1 || !in_grid(g, rr, cc)) dv = 0;
      }
    }
    s_dir[li] = dv;
  }
}

// Phase 1: local terminals by pointer jumping.  On return s_ptr[i] (interior i) is self-pointing
// for terminals; use term_kind() to classify.  Cells caught in a cycle are classified T_NA.
__device__ __forceinline__ void tile_pointer_jump(const uint8_t* s_dir, uint16_t* s_ptr) {
  for (int li = threadIdx.x; li < LC; li += NT) {
    int p = li;
    if (!loc_is_ring(li)) {
      int dv = s_dir[li];
      if (dv != 0 && dv != 5) {
        int t = li + loc_off(dv);
        if (!loc_is_ring(t)) p = t;  // else: exit cell, stays self-pointing
      }
    }
    s_ptr[li] = (uint16_t)p;
  }
  __syncthreads();
  for (int round = 0; round < 14; ++round) {  // 2^13 > LC: chains converge in <= 13 rounds
    bool changed = false;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      int i = owned_local(threadIdx.x, k);
      int p = s_ptr[i];
      int pp = s_ptr[p];
      if (pp != p) { s_ptr[i] = (uint16_t)pp; changed = true; }
    }
    if (!__syncthreads_or(changed)) break;
  }
}
__device__ __forceinline__ int term_kind(const uint8_t* s_dir, const uint16_t* s_ptr, int t) {
  if (s_ptr[t] != t) return T_NA;  // never converged: on / behind a cycle
  int dv = s_dir[t];
  if (dv == 0) return T_NA;
  if (dv == 5) return T_PIT;
  return loc_is_ring(t + loc_off(dv)) ? T_EXIT : T_NA;  // self-pointing non-exit = 2-cycle
}

// In-tile donors of each interior cell (pull, no atomics) and the source mask of the owned cells.
__device__ __forceinline__ unsigned tile_in_degree(const uint8_t* s_dir, uint8_t* s_deg) {
  unsigned srcmask = 0;
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    int i = owned_local(threadIdx.x, k);
    int deg = 0;
#pragma unroll
    for (int n = 1; n <= 9; ++n) {
      if (n == 5) continue;
      int j = i + loc_off(n);
      if (loc_is_ring(j)) continue;
      int dj = s_dir[j];
      // neighbour n of i flows into i iff its direction is the opposite code 10 - n
      deg += (dj == 10 - n);
    }
    s_deg[i] = (uint8_t)deg;
    if (deg == 0 && s_dir[i] != 0) srcmask |= 1u << k;
  }
  return srcmask;
}

// Chase: every source cell pushes its count downstream; a cell continues the walk only when it
// was the last donor to arrive (its in-degree byte drops from 1 to 0).  Integer adds: exact in
// any order.  s_cnt must hold the initial weights; s_deg the in-degrees (4 per 32-bit word).
__device__ __forceinline__ void tile_chase(const uint8_t* s_dir, unsigned* s_cnt, uint8_t* s_deg, unsigned srcmask) {
  unsigned* deg32 = reinterpret_cast<unsigned*>(s_deg);
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    if (!((srcmask >> k) & 1u)) continue;
    int cur = owned_local(threadIdx.x, k);
    unsigned carry = s_cnt[cur];
    while (true) {
      int dv = s_dir[cur];
      if (dv == 5 || dv == 0) break;
      int t = cur + loc_off(dv);
      if (loc_is_ring(t)) break;  // exit cell: its count leaves the tile
      atomicAdd(&s_cnt[t], carry);
      __threadfence_block();
      unsigned sh = 8u * (t & 3);
      unsigned old = atomicSub(&deg32[t >> 2], 1u << sh);
      if (((old >> sh) & 0xFFu) != 1u) break;
      __threadfence_block();
      carry = *reinterpret_cast<volatile unsigned*>(&s_cnt[t]);
      cur = t;
    }
  }
}

}  // namespace shed
