// shed_tile.cuh — tile-local building blocks of calc_shed4 on the device (included by shed.cu).
//
// The drec forest is processed in tiles of TW x TH interior cells staged in shared memory with a
// 1-cell ring (local grid LW x LH, 16-bit local indices).  Inside a tile every cell is resolved
// to its local terminal by pointer jumping in shared memory:
//     pit   (ddir == 5)                       -> a root of the forest
//     exit  (ddir points out of the tile)     -> continues at the entry cell of the next tile
//     NA    (ddir NA, chain leaves the grid, or a cycle)
// and flow is accumulated inside the tile in the same pointer-doubling loop.  Only the tile
// border cells (BUSED "slots" per tile) take part in the global, level-2 graph.
#pragma once
#include "cellmath.h"
#include "litap_internal.h"

namespace shed {

constexpr int NT = TW * TH / 8;    // threads per tile CTA: 8 interior cells per thread
constexpr int CPT = TW * TH / NT;  // interior cells per thread (8)
constexpr int RPK = NT / TW;       // tile rows covered by one pass of the CTA
static_assert(NT % TW == 0 && (TW & (TW - 1)) == 0, "row-wise ownership needs a power-of-two tile width");
constexpr uint8_t RING = 0x80;     // s_dir flag of the local ring cells

struct TileGeom {
  GridDims d;
  int tilesX, tilesY;
};

// local index step of direction code 1..9 (0 for 5): rows grow southward, R/LITAP_functions.R:212-220
__device__ __forceinline__ int loc_off(int dir) {
  int dr = (dir < 4) - (dir > 6);
  int dc = (int)((0x92490u >> (2 * dir)) & 3u) - 1;  // 2-bit table of (dir-1)%3 for dir 0..9 (dir 0 -> unused)
  return dr * LW + dc;
}
// k-th interior cell owned by this thread: consecutive threads take consecutive columns of a row
__device__ __forceinline__ int owned_ly(int k) { return k * RPK + (int)(threadIdx.x / TW) + 1; }
__device__ __forceinline__ int owned_lx() { return (int)(threadIdx.x % TW) + 1; }
__device__ __forceinline__ int owned_local(int k) { return owned_ly(k) * LW + owned_lx(); }

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

struct L2 {                    // level-2 arrays, one entry per border slot
  uint32_t* jump0;             // original edge: entry cell id | ID_ROOT|pid | ID_NA
  uint32_t* jres;              // resolved (compressed) pointer
  unsigned long long* pkmax;   // exit cells: max orderable elevation draining through it inside its tile (0 = none)
  uint32_t* pkidx;             //             lowest cell id attaining pkmax
  uint32_t* ext;               // external inflow into the border cell (starts as base, accumulated in place)
  uint32_t* indeg;
  uint8_t* src;
};
struct Pits {                  // one entry per pit id
  unsigned long long* rootmax; // max orderable elevation of the shed
  unsigned long long* pitkmax; // the pit tile's own partial
  uint32_t* pitkidx;
  uint32_t* rootidx;           // lowest cell id attaining rootmax
  uint32_t* rank;              // final initial_shed number
  uint32_t* cell;              // cell id of the pit
  uint32_t* cnt;               // cells draining into the pit inside its own tile (itself included)
};

// row band: external inflow of the entry cells (first / last own row), indexed by column; nullptr = none
struct BandSeeds {
  const uint32_t* x_top = nullptr;
  const uint32_t* x_bot = nullptr;
  int row_top = -1, row_bot = -1;
};

// everything one calc_shed4 run keeps between its phases (the row-band driver interleaves exchanges)
struct ShedWork {
  PoolScratch S;
  TileGeom tg;
  size_t nslots = 0;
  unsigned npits = 0;
  L2 l2 = {};
  Pits pits = {};
  unsigned long long* counters = nullptr;  // [0] npits, [1] pit cursor, [2] flags (2 x u32)
};

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

// Phase 0: stage ddir of the interior (0 outside the grid; buffered row 0 / column 0 belong to no
// tile: they are the NA ring) and flag the local ring.  A cell whose ddir points outside the grid
// or onto row 0 / column 0 becomes NA itself (its chain has an NA drec).
__device__ __forceinline__ void tile_load_dir(const TileGeom& g, const TileCtx& t, const uint8_t* __restrict__ ddir,
                                              uint8_t* s_dir) {
  for (int q = threadIdx.x; q < 2 * LW + 2 * TH; q += NT) {
    int li;
    if (q < LW) li = q;
    else if (q < 2 * LW) li = (LH - 1) * LW + (q - LW);
    else if (q < 2 * LW + TH) li = (q - 2 * LW + 1) * LW;
    else li = (q - 2 * LW - TH + 1) * LW + LW - 1;
    s_dir[li] = RING;
  }
  const int lx = owned_lx(), c = t.c_base + lx;
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    int ly = owned_ly(k), r = t.r_base + ly;
    uint8_t dv = 0;
    if (in_grid(g, r, c)) {
      dv = ddir[(size_t)r * g.d.pitch + c];
      if (dv > 9) dv = 0;
      if (dv != 0 && dv != 5) {
        int rr = r + nb_dr(dv), cc = c + nb_dc(dv);
        if (rr < 1 || cc < 1 || !in_grid(g, rr, cc)) dv = 0;
      }
    }
    s_dir[ly * LW + lx] = dv;
  }
}

// Phase 1: local terminals AND in-tile flow accumulation in one pointer-doubling loop, no
// divergent walks and no return-value atomics.  p_k(c) is the cell 2^k steps downstream of c inside
// the tile, flagged OVER once the walk would pass the local terminal; S_k(c) is the weight of all
// cells at distance < 2^k upstream of c (S_0 = own weight).  One round does
//     S_{k+1}(p_k(y)) += S_k(y)   for every y whose p_k is not OVER      (32-bit shared RED)
//     p_{k+1}(y) = p_k(p_k(y))
// because the cells at distance 2^k .. 2^(k+1)-1 above c are exactly the S_k sets of the cells y
// sitting 2^k steps above c.  ceil(log2(longest in-tile path)) rounds, two barriers per round.
// On return, for every interior cell i: s_cnt[i] = weight draining through i inside the tile
// (itself included) and s_ptr[i] = local index of its valid terminal (a pit or an exit cell), or
// 0 (the ring corner, never a terminal) when the chain is NA: NA ddir, leaves the grid, or a cycle.
constexpr int OVER = 0x8000;
constexpr int MAX_ROUNDS = LC > 2048 ? 14 : 12;  // 2^(MAX_ROUNDS-1) > LC: only cycles are still live after that

__device__ __forceinline__ void tile_jump_accumulate(const uint8_t* s_dir, uint16_t* s_ptr, unsigned* s_cnt,
                                                     const unsigned* w) {
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    int i = owned_local(k);
    int dv = s_dir[i], p = i | OVER;
    if (dv != 0 && dv != 5) {
      int t = i + loc_off(dv);
      if (!(s_dir[t] & RING)) p = t;  // else: exit cell, a terminal
    }
    s_ptr[i] = (uint16_t)p;
    s_cnt[i] = w[k];
  }
  __syncthreads();
  for (int round = 0; round < MAX_ROUNDS; ++round) {
    int a[CPT], b[CPT];
    unsigned sv[CPT];
    bool any = false;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      int i = owned_local(k);
      a[k] = s_ptr[i];
      if (!(a[k] & OVER)) {
        b[k] = s_ptr[a[k]];
        sv[k] = s_cnt[i];
        any = true;
      }
    }
    if (!__syncthreads_or(any)) break;  // also separates this round's reads from its writes
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      if (a[k] & OVER) continue;
      s_ptr[owned_local(k)] = (uint16_t)b[k];
      atomicAdd(&s_cnt[a[k]], sv[k]);
    }
    __syncthreads();
  }
  // classify: a valid terminal was reached (OVER) and is a pit or an exit
  int res[CPT];
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    int pv = s_ptr[owned_local(k)];
    int tt = pv & (OVER - 1);
    int dv = s_dir[tt];
    bool ok = (pv & OVER) && dv != 0 && (dv == 5 || (s_dir[tt + loc_off(dv)] & RING));
    res[k] = ok ? tt : 0;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < CPT; ++k) s_ptr[owned_local(k)] = (uint16_t)res[k];
  __syncthreads();
}

}  // namespace shed
