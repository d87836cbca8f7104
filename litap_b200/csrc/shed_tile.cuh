// shed_tile.cuh — tile-local building blocks of calc_shed4 on the device (included by shed.cu).
//
// The drec forest is processed in tiles of TW x TH interior cells staged in shared memory with a
// 1-cell ring (local grid LW x LH, row pitch LP, 16-bit local indices).  Inside a tile every cell is resolved
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
constexpr int CPT = 8;             // interior cells per thread: 8 consecutive columns of one tile row, so that
                                   // every global access of a thread is one aligned vector (8 B of ddir ... 64 B of elev)
constexpr int GX = TW / CPT;       // threads per tile row
static_assert(TW % CPT == 0 && NT == GX * TH, "a thread owns CPT consecutive columns of one tile row");
// local pointer word (16 bit): low 11 bits = local index, OVER = the walk has reached / passed the
// terminal whose index is in the low bits, T_PIT / T_EXIT = kind of that terminal (neither = NA)
constexpr int OVER = 0x8000, T_PIT = 0x4000, T_EXIT = 0x2000, PIDX = 0x07FF;
// the pointer words of a tile in shared memory: 32-bit slots (the words themselves stay 16-bit, as stored per cell for
// k_shed_final).  With 16-bit slots two threads' own-cell stores shared a bank word (1.7 wavefronts per store); 32-bit
// slots with the odd row pitch are conflict-free
#ifndef SHED_PTR16
typedef uint32_t ptr_t;
#else
typedef uint16_t ptr_t;
#endif
static_assert(LC <= PIDX + 1, "local indices must fit 11 bits");

struct TileGeom {
  GridDims d;
  int tilesX, tilesY;
};

// local index step of direction code 1..9 (0 for 5): rows grow southward, R/LITAP_functions.R:212-220
__device__ __forceinline__ int loc_off(int dir) {
  int dr = (dir < 4) - (dir > 6);
  int dc = (int)((0x92490u >> (2 * dir)) & 3u) - 1;  // 2-bit table of (dir-1)%3 for dir 0..9 (dir 0 -> unused)
  return dr * LP + dc;
}
// k-th interior cell owned by this thread: local row owned_ly(), local column owned_lx0() + k.  Column 1
// of every grid row is 32-element aligned (LITAP_XOFF) and tiles are 32 columns wide, so the thread's
// 8 cells start at an 8-element boundary of every per-cell array.
__device__ __forceinline__ int owned_ly() { return (int)(threadIdx.x / GX) + 1; }
__device__ __forceinline__ int owned_lx0() { return (int)(threadIdx.x % GX) * CPT + 1; }
__device__ __forceinline__ int owned_local(int k) { return owned_ly() * LP + owned_lx0() + k; }

// local coordinates of border slot s (0 <= s < BUSED): inverse of border_slot_local
__device__ __forceinline__ void slot_to_local(int s, int& lx, int& ly) {
  if (s < TW) { ly = 1; lx = s + 1; }
  else if (s < 2 * TW) { ly = TH; lx = s - TW + 1; }
  else if (s < 2 * TW + TH - 2) { lx = 1; ly = s - 2 * TW + 2; }
  else { lx = TW; ly = s - (2 * TW + TH - 2) + 2; }
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

struct L2 {                    // level-2 arrays, one entry per border slot
  uint32_t* jump0;             // original edge: entry cell id | ID_ROOT|pid | ID_NA
  uint32_t* jres;              // resolved (compressed) pointer
  // exit cells, compacted (k_shed_local appends one record per border cell that is an exit itself; the cursor is
  // counters[3]): x = max orderable elevation draining through it inside its tile, y = slot << 32 | lowest cell id
  // attaining that maximum
  ulonglong2* xrec;
  // external inflow of the border cell and the state of the level-2 chase in ONE word: inflow known so far << 32 (starts
  // as the counts handed over by the exits of the neighbouring tiles) | inflows that have arrived << 16 | in-edges
  // (border slots whose flow continues at this cell; < 2^16: at most 3 tiles x BUSED).  k_shed_local builds it with
  // 64-bit atomic adds; only the chase changes the middle field.
  unsigned long long* xi;
  uint32_t* slab;              // final shed number of the slot (k_final_labels)
};
__device__ __forceinline__ uint32_t xi_inflow(unsigned long long w) { return (uint32_t)(w >> 32); }
constexpr unsigned long long XI_ARRIVED = 1ull << 16, XI_EDGES = 0xFFFFull;
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
  uint16_t* tword = nullptr;               // per cell: final local pointer word of k_shed_local (terminal + kind)
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

// 8 bytes of a u8 array at the thread's owned cells (0 outside the grid rows)
__device__ __forceinline__ unsigned long long load_owned_u8(const TileGeom& g, const TileCtx& t, const uint8_t* __restrict__ a) {
  const int r = t.r_base + owned_ly();
  if (r >= g.d.NR) return 0ull;
  return __ldg(reinterpret_cast<const unsigned long long*>(a + (size_t)r * g.d.pitch + t.c_base + owned_lx0()));
}
// Phase 0: stage ddir of the owned cells and their initial pointer words.  Cells outside the grid,
// with NA ddir, or whose ddir leaves the grid / points onto buffered row 0 or column 0 (which belong
// to no tile: the NA ring) are NA terminals.  Returns the 8 direction codes packed 4 bits each; p[k]
// gets the pointer word of the k-th owned cell.
__device__ __forceinline__ unsigned tile_setup(const TileGeom& g, const TileCtx& t, const uint8_t* __restrict__ ddir,
                                               ptr_t* s_ptr, uint8_t* s_dir, int* p) {
  const int ly = owned_ly(), r = t.r_base + ly, lx0 = owned_lx0();
  const unsigned long long raw = load_owned_u8(g, t, ddir);
  unsigned pack = 0;
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    const int lx = lx0 + k, c = t.c_base + lx, i = ly * LP + lx;
    int dv = (int)((raw >> (8 * k)) & 0xFFull);
    if (c >= g.d.NC || dv > 9) dv = 0;
    int pw = i | OVER;  // NA terminal unless shown otherwise
    if (dv == 5) pw |= T_PIT;
    else if (dv != 0) {
      const int dr = (dv < 4) - (dv > 6);
      const int dc = (int)((0x92490u >> (2 * dv)) & 3u) - 1;  // (dv-1)%3 - 1
      const int rr = r + dr, cc = c + dc;
      if (rr < 1 || cc < 1 || !in_grid(g, rr, cc)) dv = 0;
      else {
        const bool ex = (dc < 0 && lx == 1) || (dc > 0 && lx == TW) || (dr < 0 && ly == 1) || (dr > 0 && ly == TH);
        pw = ex ? (i | OVER | T_EXIT) : (i + dr * LP + dc);
      }
    }
    s_ptr[i] = (ptr_t)pw;
    s_dir[i] = (uint8_t)dv;
    p[k] = pw;
    pack |= (unsigned)dv << (4 * k);
  }
  return pack;
}

// Phase 1: local terminals AND in-tile flow accumulation in one pointer-doubling loop, no
// divergent walks and no return-value atomics.  p_k(c) is the cell 2^k steps downstream of c inside
// the tile, flagged OVER (with the terminal's index and kind) once the walk would pass the local
// terminal; S_k(c) is the weight of all cells at distance < 2^k upstream of c (S_0 = own weight).
// One round does
//     S_{k+1}(p_k(y)) += S_k(y)   for every y whose p_k is not OVER      (32-bit shared RED)
//     p_{k+1}(y) = p_k(p_k(y))
// because the cells at distance 2^k .. 2^(k+1)-1 above c are exactly the S_k sets of the cells y
// sitting 2^k steps above c.  ceil(log2(longest in-tile path)) rounds, two barriers per round.
// The caller has stored the weights in s_cnt and synchronised.  On return p[k] is the final pointer
// word of the k-th owned cell (a cycle never gets OVER: treat as NA) and s_cnt[i] the weight draining
// through cell i inside the tile, itself included.
constexpr int MAX_ROUNDS = LC > 2048 ? 14 : 12;  // 2^(MAX_ROUNDS-1) > LC: only cycles are still live after that

// 32-bit shared-window addresses + explicit ld/st/red.shared: with ordinary pointers the compiler
// rebuilt the window base (S2R CgaCtaId, MOV, VIADD, LEA) in front of every data-dependent access of
// the hot loop — 4 extra instructions per access.  The empty asm makes the base opaque, so it stays
// in a register.
__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("" : "+r"(a));
  return a;
}
__device__ __forceinline__ unsigned lds_u16(uint32_t a) {
  unsigned v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ unsigned lds_u32(uint32_t a) {
  unsigned v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t a, unsigned v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_u16(uint32_t a, unsigned v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_u32(uint32_t a, unsigned v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void red_max_u32(uint32_t a, unsigned v) {
  asm volatile("red.shared.max.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void red_min_u32(uint32_t a, unsigned v) {
  asm volatile("red.shared.min.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

__device__ __forceinline__ void tile_jump_accumulate(ptr_t* s_ptr, unsigned* s_cnt, int* p) {
  const int base = owned_local(0);
  const uint32_t a_ptr = smem_addr(s_ptr), a_cnt = smem_addr(s_cnt);
  const uint32_t own_ptr = a_ptr + (uint32_t)sizeof(ptr_t) * base, own_cnt = a_cnt + 4 * base;
  for (int round = 0; round < MAX_ROUNDS; ++round) {
    unsigned b[CPT], sv[CPT];
    bool any = false;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      if (!(p[k] & OVER)) {
        b[k] = sizeof(ptr_t) == 2 ? lds_u16(a_ptr + 2 * p[k]) : lds_u32(a_ptr + 4 * p[k]);
        sv[k] = lds_u32(own_cnt + 4 * k);
        any = true;
      }
    }
    if (!__syncthreads_or(any)) break;  // also separates this round's reads from its writes
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      if (p[k] & OVER) continue;
      if (sizeof(ptr_t) == 2) sts_u16(own_ptr + 2 * k, b[k]); else sts_u32(own_ptr + 4 * k, b[k]);
      red_add_u32(a_cnt + 4 * p[k], sv[k]);
      p[k] = (int)b[k];
    }
    __syncthreads();
  }
}

}  // namespace shed
