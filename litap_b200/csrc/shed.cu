// shed.cu — calc_shed4 + calc_upslopes("upslope") on the device (R/flow_02_calc_shed.R:10-145).
//
//   k_count_pits    pits = cells with ddir == 5                                   (:12-20)
//   k_shed_local    per tile: local terminals (pointer jumping), per-terminal max-elevation key,
//                   local flow counts; writes the level-2 edges of the 124 border cells, counts the
//                   in-edges / base inflows of the entry cells they lead to (L2::xi) and appends
//                   one key record per exit cell
//   k_l2_*          level-2 graph over border slots: resolve every border cell to its pit
//                   (pointer jumping with path compression), accumulate external inflow
//                   (chase: one 64-bit atomic per hop), merge the per-shed keys (exit records) (:25-29)
//   cub radix sort  rank pits by (max elev desc, seqno of that cell asc)                  (:32-38)
//   k_final_labels  final shed number of every border slot and pit cell
//   k_shed_final    per tile: final shed label, upslope count (local accumulation seeded with
//                   the external inflow) and ridge flag                                 (:42-57)
//
// Algorithmic traffic: ddir is read twice (1 B), elev once (8 B), shed + upslope + ridge written
// once (9 B) => 19 B/cell; the level-2 arrays are ~1.9 B/cell and stay in L2.
#include <algorithm>
#include <vector>
#include <cub/cub.cuh>
#include "shed_tile.cuh"

namespace shed {

__device__ __forceinline__ unsigned long long f64_orderable(double x) {
  unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}

__global__ void k_count_pits(const uint8_t* __restrict__ ddir, GridDims d, unsigned long long* count) {
  unsigned n = 0;
  const int c = blockIdx.x * blockDim.x + threadIdx.x + 1;
  if (c < d.NC)
    for (int r = 1 + blockIdx.y; r < d.NR; r += gridDim.y) n += ddir[(size_t)r * d.pitch + c] == 5;
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(count, (unsigned long long)n);
}

// ------------------------------------------------------------------------------------------------
// shared memory of k_shed_local: ptr, dir, cnt, key hi / lo / idx (no aliasing: fewer barriers)
constexpr size_t SMEM_A = (size_t)LC * (sizeof(ptr_t) + 1 + 4 + 4 + 4 + 4) + 64;

#ifndef SHED_MINB
#define SHED_MINB 8
#endif
__global__ void __launch_bounds__(NT, SHED_MINB) k_shed_local(TileGeom g, const uint8_t* __restrict__ ddir,
                                                      const double* __restrict__ elev, uint32_t* __restrict__ lab,
                                                      uint16_t* __restrict__ tword, uint32_t* __restrict__ lcnt,
                                                      L2 l2, Pits pits, unsigned long long* pit_cursor) {
  // static shared arrays: the compiler addresses them with immediate offsets (with dynamic shared
  // memory it rebuilt the shared-window base, 4 instructions, in front of most accesses)
  __shared__ unsigned s_cnt[LC];
  __shared__ uint32_t s_khi[LC];   // key, high word
  __shared__ uint32_t s_klo[LC];   // key, low word
  __shared__ uint32_t s_kid[LC];   // lowest cell id attaining the key; later the pit id
  __shared__ ptr_t s_ptr[LC];
  __shared__ uint8_t s_dir[LC];
  __shared__ unsigned s_npit, s_nx, s_xcur;   // pits / exit cells of the tile, cursor of its exit records
  __shared__ unsigned long long s_pitbase, s_xbase;

  const TileCtx t = tile_ctx();
  const int r = t.r_base + owned_ly(), c0 = t.c_base + owned_lx0();
  const bool row_ok = r < g.d.NR;
  const int base = owned_local(0);
  const size_t tile_slot0 = (size_t)(t.ty * g.tilesX + t.tx) * BSLOTS;
  const size_t cell0 = (size_t)r * g.d.pitch + c0;  // k-th owned cell: cell0 + k (an 8-element boundary)
  if (threadIdx.x == 0) s_npit = s_nx = s_xcur = 0;
  int p[CPT];
  const unsigned dvp = tile_setup(g, t, ddir, s_ptr, s_dir, p);
  // elevations of the owned cells (64 contiguous bytes): issued now, consumed after the pointer-doubling loop
  double ev[CPT];
  if (row_ok) {
#pragma unroll
    for (int k = 0; k < CPT; k += 2) {
      const double2 v = __ldg(reinterpret_cast<const double2*>(elev + cell0 + k));
      ev[k] = v.x;
      ev[k + 1] = v.y;
    }
  }
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    const bool valid = ((dvp >> (4 * k)) & 15u) != 0;
    if (!valid) ev[k] = 0.0;
    const int i = base + k;
    s_cnt[i] = valid ? 1u : 0u;  // out-of-grid and NA cells carry no weight
    s_khi[i] = 0u;
    s_klo[i] = 0u;
    s_kid[i] = 0xFFFFFFFFu;
  }
  __syncthreads();
  const unsigned pit_mask_src = dvp;  // dv == 5 <-> pit
  tile_jump_accumulate(s_ptr, s_cnt, p);
  // here: all REDs into s_cnt are complete (the loop ends with a barrier or with the all-done vote)

  // ---- exits hand their count to the entry cell's slot (base of the level-2 sum); pits keep theirs
  unsigned pcnt[CPT], nx = 0;
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    const int dv = (dvp >> (4 * k)) & 15u;
    const int i = base + k;
    pcnt[k] = s_cnt[i];
    // own initial pointer word had T_EXIT <=> the final one has it and points at the cell itself
    if ((p[k] & (OVER | T_EXIT)) == (OVER | T_EXIT) && (p[k] & PIDX) == i) {
      const long long es = slot_of_rc(r + nb_dr(dv), c0 + k + nb_dc(dv), g.tilesX);
      if (es >= 0) atomicAdd(&l2.xi[es], (unsigned long long)pcnt[k] << 32);
      ++nx;
    }
  }
  if (nx) atomicAdd(&s_nx, nx);   // exit records of the tile: their space is claimed together with the pit ids
  if (row_ok) {  // kept for k_shed_final: local terminal word and in-tile flow count of every cell (pads included)
    uint4 w;
    w.x = (unsigned)p[0] & 0xFFFFu | (unsigned)p[1] << 16;
    w.y = (unsigned)p[2] & 0xFFFFu | (unsigned)p[3] << 16;
    w.z = (unsigned)p[4] & 0xFFFFu | (unsigned)p[5] << 16;
    w.w = (unsigned)p[6] & 0xFFFFu | (unsigned)p[7] << 16;
    *reinterpret_cast<uint4*>(tword + cell0) = w;
    *reinterpret_cast<uint4*>(lcnt + cell0) = make_uint4(pcnt[0], pcnt[1], pcnt[2], pcnt[3]);
    *reinterpret_cast<uint4*>(lcnt + cell0 + 4) = make_uint4(pcnt[4], pcnt[5], pcnt[6], pcnt[7]);
  }

  // ---- per-terminal key: max elevation, then lowest cell id attaining it; three rounds of native
  // 32-bit shared REDs (high word, low word among the high-word winners, id among the winners)
  uint32_t khi[CPT], klo[CPT];
  int term[CPT];  // 4 * local index of the terminal; 0 = NA (the ring corner is never a terminal)
  const uint32_t a_khi = smem_addr(s_khi), a_klo = smem_addr(s_klo), a_kid = smem_addr(s_kid);
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    const bool ok = (p[k] & OVER) && (p[k] & (T_PIT | T_EXIT));
    term[k] = ok ? 4 * (p[k] & PIDX) : 0;
    const unsigned long long key = f64_orderable(ev[k]);
    khi[k] = (uint32_t)(key >> 32);
    klo[k] = (uint32_t)key;
    if (term[k]) red_max_u32(a_khi + term[k], khi[k]);
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    if (term[k] && khi[k] == lds_u32(a_khi + term[k])) red_max_u32(a_klo + term[k], klo[k]);
    else term[k] = 0;  // out of the race
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < CPT; ++k)
    if (term[k] && klo[k] == lds_u32(a_klo + term[k])) red_min_u32(a_kid + term[k], (uint32_t)(cell0 + k));

  // ---- pit ids (order of ids is irrelevant: the canonical numbering comes from the sort)
  unsigned rk[CPT];
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    rk[k] = 0xFFFFFFFFu;
    if (((pit_mask_src >> (4 * k)) & 15u) == 5u) rk[k] = atomicAdd(&s_npit, 1u);
  }
  __syncthreads();  // keys complete, s_npit complete
  if (threadIdx.x == 0 && (s_npit | s_nx)) {   // one atomic per tile: pits in the low half of the cursor, exit records in the high half
    const unsigned long long old = atomicAdd(pit_cursor, (unsigned long long)s_npit | ((unsigned long long)s_nx << 32));
    s_pitbase = old & 0xFFFFFFFFull;
    s_xbase = old >> 32;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    if (rk[k] == 0xFFFFFFFFu) continue;
    const int i = base + k;
    const uint32_t pid = (uint32_t)(s_pitbase + rk[k]);
    const unsigned long long key = ((unsigned long long)s_khi[i] << 32) | s_klo[i];
    pits.rootmax[pid] = key;
    pits.pitkmax[pid] = key;
    pits.pitkidx[pid] = s_kid[i];
    pits.rootidx[pid] = 0xFFFFFFFFu;
    pits.cell[pid] = (uint32_t)(cell0 + k);
    pits.cnt[pid] = pcnt[k];
    lab[cell0 + k] = pid;  // read back by k_shed_final of this tile
    s_cnt[i] = pid;        // from here on s_cnt[pit] is its pid (counts are consumed)
  }
  __syncthreads();

  // ---- level-2 edges and key partials: one thread per border slot
  unsigned long long pk = 0;   // this thread's border cell is an exit itself: its key partial
  uint32_t pi = 0xFFFFFFFFu;
  bool is_exit = false;
  if (threadIdx.x < BUSED) {
    int bx, by;
    slot_to_local((int)threadIdx.x, bx, by);
    const int i = by * LP + bx;
    const int pw = s_ptr[i];
    uint32_t edge = ID_NA;
    if ((pw & OVER) && (pw & (T_PIT | T_EXIT))) {
      const int tt = pw & PIDX;
      if (pw & T_PIT) edge = ID_ROOT | s_cnt[tt];
      else {  // exit: continue at the entry cell of the neighbouring tile
        const int dv = s_dir[tt];
        const int ty_ = tt / LP, tx_ = tt - ty_ * LP;
        const int er = t.r_base + ty_ + nb_dr(dv), ec = t.c_base + tx_ + nb_dc(dv);
        edge = (uint32_t)((size_t)er * g.d.pitch + ec);
        const long long es = slot_of_rc(er, ec, g.tilesX);   // one more in-edge of the entry cell's slot
        if (es >= 0) atomicAdd(&l2.xi[es], 1ull);
        if (tt == i) {  // this border cell is the exit itself
          pk = ((unsigned long long)s_khi[i] << 32) | s_klo[i];
          pi = s_kid[i];
          is_exit = true;
        }
      }
    }
    l2.jump0[tile_slot0 + threadIdx.x] = edge;
    l2.jres[tile_slot0 + threadIdx.x] = edge;
  }
  // exit records: the key passes run over the exits only, not over every slot
  if (is_exit) l2.xrec[s_xbase + atomicAdd(&s_xcur, 1u)] = make_ulonglong2(pk, ((unsigned long long)(tile_slot0 + threadIdx.x) << 32) | pi);
}

// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool is_cell(uint32_t v) { return !(v & ID_ROOT); }  // ID_NA has the flag bit too

__global__ void k_l2_resolve(TileGeom g, uint32_t* __restrict__ jres, size_t nslots, unsigned* unresolved) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool open = false;
  if (s < nslots) {
    uint32_t v = jres[s];
    if (is_cell(v)) {
      for (int hop = 0; hop < 8 && is_cell(v); ++hop) {
        long long s2 = slot_of_id(v, g);
        v = s2 < 0 ? ID_NA : jres[s2];
      }
      jres[s] = v;
      open = is_cell(v);
    }
  }
  if (__syncthreads_or(open) && threadIdx.x == 0) atomicAdd(unresolved, 1u);
}
__global__ void k_l2_close_cycles(uint32_t* __restrict__ jres, size_t nslots) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < nslots && is_cell(jres[s])) jres[s] = ID_NA;
}
// External inflow of every border cell.  A thread per source (a slot without in-edges) carries its inflow down the
// chain of border slots; at every slot the LAST inflow to arrive carries the sum on.  One 64-bit atomic per hop adds the
// carry to the slot's inflow field and counts the arrival; the value it returns tells the thread whether it was the
// last and what had arrived before it (atomics on one address are ordered: no fence), and the next edge is loaded
// under the shadow of the atomic: one memory round trip per hop.  The in-edge counts and the base inflows come
// from k_shed_local in the same word (L2::xi), so there is no in-degree pass and no source pass over the slots.
__global__ void k_l2_chase(TileGeom g, const uint32_t* __restrict__ jump0, size_t nslots, unsigned long long* xi) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots) return;
  const unsigned long long w = xi[s];
  if (w & XI_EDGES) return;          // not a source (the in-edge field does not change during the chase)
  uint32_t carry = xi_inflow(w);     // nothing is ever added to a source
  uint32_t v = jump0[s];
  for (size_t hop = 0; hop < nslots; ++hop) {
    if (!is_cell(v)) break;
    const long long t = slot_of_id(v, g);
    if (t < 0) break;
    const uint32_t vn = jump0[t];
    const unsigned long long old = atomicAdd(&xi[t], ((unsigned long long)carry << 32) + XI_ARRIVED);
    if (((old >> 16) & 0xFFFFull) + 1ull != (old & XI_EDGES)) break;   // other inflows are still to come
    carry += xi_inflow(old);
    v = vn;
  }
}
// shed keys: exits' partials merged into their root; then the index among the maxima (grid-stride over the exit
// records; their number stays on the device)
__global__ void k_keys_max(const L2 l2, const unsigned long long* __restrict__ nrec, Pits pits) {
  const size_t n = (size_t)(*nrec >> 32);   // (the pit cursor: exit records in its high half)
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const ulonglong2 rec = l2.xrec[i];
    const uint32_t root = l2.jres[rec.y >> 32];
    if (rec.x && root != ID_NA && (root & ID_ROOT)) atomicMax(&pits.rootmax[root & ~ID_ROOT], rec.x);
  }
}
__global__ void k_keys_idx(const L2 l2, const unsigned long long* __restrict__ nrec, Pits pits, unsigned npits) {
  const size_t n = (size_t)(*nrec >> 32), stride = (size_t)gridDim.x * blockDim.x, t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (size_t i = t0; i < n; i += stride) {
    const ulonglong2 rec = l2.xrec[i];
    const uint32_t root = l2.jres[rec.y >> 32];
    if (rec.x && root != ID_NA && (root & ID_ROOT)) {
      const uint32_t pid = root & ~ID_ROOT;
      if (rec.x == pits.rootmax[pid]) atomicMin(&pits.rootidx[pid], (uint32_t)rec.y);
    }
  }
  for (size_t s = t0; s < npits; s += stride)
    if (pits.pitkmax[s] == pits.rootmax[s]) atomicMin(&pits.rootidx[s], pits.pitkidx[s]);
}
__global__ void k_iota(uint32_t* a, unsigned n) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = i;
}
__global__ void k_gather_u64(const unsigned long long* __restrict__ src, const uint32_t* __restrict__ idx,
                             unsigned long long* __restrict__ dst, unsigned n) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[idx[i]];
}
__global__ void k_scatter_rank(const uint32_t* __restrict__ order, uint32_t* __restrict__ rank, unsigned n) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) rank[order[i]] = i + 1;
}

// ------------------------------------------------------------------------------------------------
// Final shed number of every border slot (the rank of the pit its level-2 chain ends in, 0 = none) and, in the label
// array, of every pit cell (k_shed_local left the pit id there): k_shed_final then needs ONE load per border / ring /
// pit cell where it used to chase slot -> pit id -> rank, two dependent L2 round trips in front of its barriers.
__global__ void k_final_labels(const uint32_t* __restrict__ jres, size_t nslots, const uint32_t* __restrict__ rank,
                               const uint32_t* __restrict__ pitcell, unsigned npits, uint32_t* __restrict__ slab,
                               uint32_t* __restrict__ lab) {
  const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < nslots) {
    const uint32_t v = jres[s];
    slab[s] = (v != ID_NA && (v & ID_ROOT)) ? rank[v & ~ID_ROOT] : 0u;
  }
  if (s < npits) lab[pitcell[s]] = rank[s];
}

// ------------------------------------------------------------------------------------------------
// k_shed_final: labels, upslope and ridge of one tile from what k_shed_local left behind — the
// local terminal of every cell (tword) and its in-tile flow count (upslope array) — plus the level-2
// results.  Flow entering the tile (external inflow of a border cell; flow entering a row band at its
// first / last own row) is walked down its in-tile path to the local terminal: at most BUSED + 2 TW
// short serial walks per tile instead of a second pointer-doubling pass over all its cells.
__device__ __forceinline__ bool tw_labelled(unsigned w) { return (w & OVER) && (w & (T_PIT | T_EXIT)); }

// add x to every cell from local cell i down to its local terminal (inclusive)
// s_td: per local cell, terminal word (low half) | direction code << 16 - one 32-bit slot, so that the own-cell stores
// are conflict-free (separate u16 / u8 arrays took 2 and 3 wavefronts per store) and a walk step is one load
__device__ __forceinline__ void tile_walk_add(uint32_t a_cnt, const uint32_t* s_td, int i, unsigned x) {
  const unsigned w = s_td[i] & 0xFFFFu;
  if (!x || !tw_labelled(w)) return;  // NA chains and cycles carry no upslope
  const int term = w & PIDX;
  for (int step = 0; step < LC; ++step) {
    red_add_u32(a_cnt + 4 * i, x);
    if (i == term) break;
    i += loc_off((int)(s_td[i] >> 16));
  }
}

#ifndef SHEDF_MINB
#define SHEDF_MINB 8
#endif
__global__ void __launch_bounds__(NT, SHEDF_MINB) k_shed_final(TileGeom g, const uint8_t* __restrict__ ddir, const L2 l2,
                                                      const uint32_t* __restrict__ slab,
                                                      const uint16_t* __restrict__ tword, int32_t* shed_out,
                                                      uint32_t* upslope, uint8_t* __restrict__ ridge, BandSeeds seeds) {
  __shared__ unsigned s_cnt[LC];
  __shared__ uint32_t s_tlab[LC];  // label of a terminal
  __shared__ uint32_t s_lab[LC];   // label of every local cell
  __shared__ uint32_t s_td[LC];
  const uint32_t* lab_in = reinterpret_cast<const uint32_t*>(shed_out);      // pit cells hold their final label (k_final_labels)

  const TileCtx t = tile_ctx();
  const int r = t.r_base + owned_ly(), c0 = t.c_base + owned_lx0();
  const bool row_ok = r < g.d.NR;
  const int base = owned_local(0);
  const size_t tile_slot0 = (size_t)(t.ty * g.tilesX + t.tx) * BSLOTS;
  const size_t cell0 = (size_t)r * g.d.pitch + c0;  // k-th owned cell: cell0 + k (an 8-element boundary)
  unsigned tw[CPT];
  {
    unsigned long long draw = 0;
    uint4 wraw = make_uint4(0, 0, 0, 0), n0 = wraw, n1 = wraw;
    if (row_ok) {
      draw = __ldg(reinterpret_cast<const unsigned long long*>(ddir + cell0));
      wraw = __ldg(reinterpret_cast<const uint4*>(tword + cell0));
      n0 = *reinterpret_cast<const uint4*>(upslope + cell0);
      n1 = *reinterpret_cast<const uint4*>(upslope + cell0 + 4);
    }
    const unsigned wv[4] = {wraw.x, wraw.y, wraw.z, wraw.w};
    const unsigned nv[8] = {n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n1.z, n1.w};
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
      const int i = base + k;
      const bool in = c0 + k < g.d.NC;  // (row_ok is folded into the zero loads)
      const unsigned w = in ? (wv[k >> 1] >> (16 * (k & 1))) & 0xFFFFu : 0u;
      tw[k] = w;
      s_td[i] = w | (in ? (unsigned)((draw >> (8 * k)) & 0xFFull) << 16 : 0u);
      s_cnt[i] = in ? nv[k] : 0u;
    }
  }
  // ---- labels of the terminals (pits: their own cell of the label array; exits: their border slot) and of the ring
  // cells (border slots of the neighbouring tiles, for the ridge test): one independent load each, all issued here,
  // in front of the first barrier, and stored after the walks
  uint32_t pitlab[CPT];
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    pitlab[k] = 0xFFFFFFFFu;
    if ((tw[k] & (OVER | T_PIT)) == (OVER | T_PIT) && (int)(tw[k] & PIDX) == base + k) pitlab[k] = lab_in[cell0 + k];
  }
  int bi = -1;  // this thread's border slot as a local index
  uint32_t bext = 0, blab = 0;
  if (threadIdx.x < BUSED) {
    int bx, by;
    slot_to_local((int)threadIdx.x, bx, by);
    bi = by * LP + bx;
    bext = xi_inflow(l2.xi[tile_slot0 + threadIdx.x]);
    blab = slab[tile_slot0 + threadIdx.x];
  }
  constexpr int NRING = 2 * LW + 2 * TH, RPT = (NRING + NT - 1) / NT;
  uint32_t rlab[RPT];
  int rloc[RPT];
#pragma unroll
  for (int j = 0; j < RPT; ++j) {
    const int q = (int)threadIdx.x + j * NT;
    rloc[j] = -1;
    rlab[j] = 0;
    if (q < NRING) {
      int ry, rx;
      if (q < LW) { ry = 0; rx = q; }
      else if (q < 2 * LW) { ry = LH - 1; rx = q - LW; }
      else if (q < 2 * LW + TH) { ry = q - 2 * LW + 1; rx = 0; }
      else { ry = q - 2 * LW - TH + 1; rx = LW - 1; }
      rloc[j] = ry * LP + rx;
      const int rr = t.r_base + ry, cc = t.c_base + rx;
      if (rr >= 1 && cc >= 1 && in_grid(g, rr, cc)) {
        const long long sl = slot_of_rc(rr, cc, g.tilesX);
        if (sl >= 0) rlab[j] = slab[sl];
      }
    }
  }
  __syncthreads();
  const uint32_t a_cnt = smem_addr(s_cnt);
  // ---- flow entering the tile, walked down to the local terminal
  if (bi >= 0) tile_walk_add(a_cnt, s_td, bi, bext);
  if (seeds.x_top || seeds.x_bot) {
    const int lt = seeds.row_top - t.r_base, lb = seeds.row_bot - t.r_base;
    if (threadIdx.x < TW && t.c_base + (int)threadIdx.x + 1 < g.d.NC) {
      const int cc = t.c_base + (int)threadIdx.x + 1;
      if (seeds.x_top && lt >= 1 && lt <= TH) tile_walk_add(a_cnt, s_td, lt * LP + (int)threadIdx.x + 1, seeds.x_top[cc]);
      if (seeds.x_bot && lb >= 1 && lb <= TH) tile_walk_add(a_cnt, s_td, lb * LP + (int)threadIdx.x + 1, seeds.x_bot[cc]);
    }
  }
#pragma unroll
  for (int k = 0; k < CPT; ++k)
    if (pitlab[k] != 0xFFFFFFFFu) s_tlab[base + k] = pitlab[k];
  if (bi >= 0) {
    const int pw = (int)(s_td[bi] & 0xFFFFu);
    if ((pw & (OVER | T_EXIT)) == (OVER | T_EXIT) && (pw & PIDX) == bi) s_tlab[bi] = blab;
  }
#pragma unroll
  for (int j = 0; j < RPT; ++j)
    if (rloc[j] >= 0) s_lab[rloc[j]] = rlab[j];
  __syncthreads();
  uint32_t mylab[CPT];
  const uint32_t a_tlab = smem_addr(s_tlab);
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    mylab[k] = tw_labelled(tw[k]) ? lds_u32(a_tlab + 4 * (tw[k] & PIDX)) : 0u;
    s_lab[base + k] = mylab[k];
  }
  __syncthreads();

  // ---- outputs.  ridge = more than one distinct non-NA label in the 3 x 3 window
  // (R/flow_02_calc_shed.R:48-57) <=> max label != min non-zero label; with 0 -> 0xFFFFFFFF by
  // "label - 1" the minimum skips NA cells.  Column triples (rows above / at / below) of the 10
  // window columns of the thread's 8 cells, then 3 adjacent columns per cell.
  if (!row_ok) return;
  uint32_t cmx[CPT + 2], cmn[CPT + 2];
#pragma unroll
  for (int j = 0; j < CPT + 2; ++j) {
    const uint32_t a = s_lab[base - LP - 1 + j], b = s_lab[base - 1 + j], cc = s_lab[base + LP - 1 + j];
    cmx[j] = __vimax3_u32(a, b, cc);
    cmn[j] = __vimin3_u32(a - 1u, b - 1u, cc - 1u);
  }
  uint32_t up[CPT];
  unsigned long long rg8 = 0;
  const bool row_in = r < g.d.NR - 1;
#pragma unroll
  for (int k = 0; k < CPT; ++k) {
    const uint32_t lb = mylab[k];
    unsigned rg = 2;  // NA on the ring
    if (row_in && c0 + k < g.d.NC - 1) {
      const uint32_t mx = __vimax3_u32(cmx[k], cmx[k + 1], cmx[k + 2]), mn = __vimin3_u32(cmn[k], cmn[k + 1], cmn[k + 2]);
      rg = (mx != 0u) & (mn + 1u != mx);
    }
    up[k] = lb ? s_cnt[base + k] : 0u;
    rg8 |= (unsigned long long)rg << (8 * k);
  }
  // pad columns (>= NC) get label 0 / upslope 0 / ridge NA: one aligned vector per array
  *reinterpret_cast<uint4*>(shed_out + cell0) = make_uint4(mylab[0], mylab[1], mylab[2], mylab[3]);
  *reinterpret_cast<uint4*>(shed_out + cell0 + 4) = make_uint4(mylab[4], mylab[5], mylab[6], mylab[7]);
  *reinterpret_cast<uint4*>(upslope + cell0) = make_uint4(up[0], up[1], up[2], up[3]);
  *reinterpret_cast<uint4*>(upslope + cell0 + 4) = make_uint4(up[4], up[5], up[6], up[7]);
  *reinterpret_cast<unsigned long long*>(ridge + cell0) = rg8;
}

// ring row 0 / column 0 (never inside a tile) and, for uniformity, the other two ring lines
__global__ void k_ring_fill(GridDims d, int32_t* __restrict__ shed_out, uint32_t* __restrict__ upslope,
                            uint8_t* __restrict__ ridge, int keep_last_row) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  int line = blockIdx.y;
  if (line == 1 && keep_last_row) return;  // south halo of a row band: its pit cells still hold their ids
  int r, c;
  if (line == 0) { r = 0; c = t; }
  else if (line == 1) { r = d.NR - 1; c = t; }
  else if (line == 2) { r = t; c = 0; }
  else { r = t; c = d.NC - 1; }
  if (r >= d.NR || c >= d.NC) return;
  size_t o = (size_t)r * d.pitch + c;
  shed_out[o] = 0;
  upslope[o] = 0;
  ridge[o] = 2;
}

}  // namespace shed

// ------------------------------------------------------------------------------------------------
// host side: calc_shed4 in phases (the row-band driver in bands.cu runs exchanges between them)
//   shed_phase_local   tiles + level-2 graph: every border cell resolved to a pit id, external
//                      inflow of every border cell, per-pit key (max elevation, lowest cell)
//   shed_phase_rank    single device: canonical numbering of the pits (two radix sorts)
//   shed_phase_final   labels, upslope, ridge
int shed_phase_local(litap_grid* g, shed::ShedWork& W) {
  using namespace shed;
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  if (!g->shed) g->shed = grid_alloc_array<int32_t>(d, -1);
  if (!g->upslope) g->upslope = grid_alloc_array<uint32_t>(d, -1);
  if (!g->ridge) g->ridge = grid_alloc_array<uint8_t>(d, -1);
  if (!g->shed || !g->upslope || !g->ridge) return set_error(LITAP_ENOMEM, "device allocation failed (shed outputs)");

  static_assert(SMEM_A <= 48 * 1024, "tile arrays are static shared memory");
  TileGeom& tg = W.tg;
  tg.d = d;
  tg.tilesX = (d.NC - 1 + TW - 1) / TW;
  tg.tilesY = (d.NR - 1 + TH - 1) / TH;
  const size_t nslots = W.nslots = (size_t)tg.tilesX * tg.tilesY * BSLOTS;

  PoolScratch& S = W.S;
  unsigned long long* counters = W.counters = S.get<unsigned long long>(4);
  if (!counters) return set_error(LITAP_ENOMEM, "device allocation failed (shed scratch)");
  CUDA_TRY(cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned long long), st));
  // capacity of the pit arrays: the flat-cell count of the D8 stencil when calc_ddir2 ran on this grid
  // (pits are a subset; a row band adds its portals), else an exact count (one more pass over ddir and a host sync)
  size_t np = 1;
  if (g->pit_bound >= 0) np = g->pit_bound ? (size_t)g->pit_bound : 1;
  else {
    dim3 grid((d.NC - 1 + 255) / 256, d.NR - 1 < 256 ? d.NR - 1 : 256);
    KERNEL("k_count_pits", k_count_pits<<<grid, 256, 0, st>>>(g->ddir, d, counters));
    unsigned long long npits64 = 0;
    CUDA_TRY(cudaMemcpyAsync(&npits64, counters, 8, cudaMemcpyDeviceToHost, st));
    STREAM_SYNC(st);
    np = npits64 ? (size_t)npits64 : 1;
  }

  L2& l2 = W.l2;
  l2.jump0 = S.get<uint32_t>(nslots);
  l2.jres = S.get<uint32_t>(nslots);
  l2.xrec = S.get<ulonglong2>(nslots);
  l2.xi = S.get<unsigned long long>(nslots);
  l2.slab = S.get<uint32_t>(nslots);
  Pits& pits = W.pits;
  pits.rootmax = S.get<unsigned long long>(np);
  pits.pitkmax = S.get<unsigned long long>(np);
  pits.pitkidx = S.get<uint32_t>(np);
  pits.rootidx = S.get<uint32_t>(np);
  pits.rank = S.get<uint32_t>(np);
  pits.cell = S.get<uint32_t>(np);
  pits.cnt = S.get<uint32_t>(np);
  if (!l2.jump0 || !l2.jres || !l2.xrec || !l2.xi || !l2.slab || !pits.rootmax ||
      !pits.pitkmax || !pits.pitkidx || !pits.rootidx || !pits.rank || !pits.cell || !pits.cnt)
    return set_error(LITAP_ENOMEM, "device allocation failed (shed level-2 arrays)");
  CUDA_TRY(cudaMemsetAsync(l2.xi, 0, nslots * 8, st));
  // slots BUSED.. of every tile are never written by k_shed_local: keep them NA / empty
  CUDA_TRY(cudaMemsetAsync(l2.jump0, 0xFF, nslots * 4, st));
  CUDA_TRY(cudaMemsetAsync(l2.jres, 0xFF, nslots * 4, st));

  dim3 tgrid(tg.tilesX, tg.tilesY);
  W.tword = S.get<uint16_t>((size_t)d.cells + 64);  // pitched and offset like the grid arrays (aligned 16-byte vectors)
  if (!W.tword) return set_error(LITAP_ENOMEM, "device allocation failed (shed terminal words)");
  W.tword += LITAP_XOFF;
  KERNEL("k_shed_local", k_shed_local<<<tgrid, NT, 0, st>>>(tg, g->ddir, g->elev, reinterpret_cast<uint32_t*>(g->shed), W.tword,
                                         g->upslope, l2, pits, counters + 1));

  const unsigned sblk = (unsigned)((nslots + 255) / 256);
  // ---- resolve border cells to pits; a drec cycle across tiles never resolves -> NA after the cap
  // (two rounds per host check; the same read-back brings the pit count = the pit cursor of k_shed_local)
  unsigned* flag = reinterpret_cast<unsigned*>(counters + 2);
  bool open = true;
  unsigned long long hb[2] = {0, 0};
  for (int round = 0; round < 64 && open; round += 2) {
    for (int k = 0; k < 2; ++k) {
      CUDA_TRY(cudaMemsetAsync(flag, 0, 4, st));
      KERNEL("k_l2_resolve", k_l2_resolve<<<sblk, 256, 0, st>>>(tg, l2.jres, nslots, flag));
    }
    CUDA_TRY(cudaMemcpyAsync(hb, counters + 1, 16, cudaMemcpyDeviceToHost, st));
    STREAM_SYNC(st);
    open = (unsigned)(hb[1] & 0xFFFFFFFFull) != 0;
  }
  const unsigned npits = W.npits = (unsigned)hb[0];
  if (open) {
    KERNEL("k_l2_close_cycles", k_l2_close_cycles<<<sblk, 256, 0, st>>>(l2.jres, nslots));
  }
  // ---- external inflow of every border cell
  KERNEL("k_l2_chase", k_l2_chase<<<sblk, 256, 0, st>>>(tg, l2.jump0, nslots, l2.xi));
  // ---- per-pit keys
  if (npits) {
    const unsigned kblk = (unsigned)ctx().sm_count * 16;
    KERNEL("k_keys_max", k_keys_max<<<kblk, 256, 0, st>>>(l2, counters + 1, pits));
    KERNEL("k_keys_idx", k_keys_idx<<<kblk, 256, 0, st>>>(l2, counters + 1, pits, npits));
  }
  return LITAP_OK;
}

int shed_phase_rank(shed::ShedWork& W) {
  using namespace shed;
  Ctx& c = ctx();
  cudaStream_t st = c.stream;
  const unsigned npits = W.npits;
  if (!npits) return LITAP_OK;
  PoolScratch& S = W.S;
  Pits& pits = W.pits;
  uint32_t* ord_a = S.get<uint32_t>(npits);
  uint32_t* ord_b = S.get<uint32_t>(npits);
  uint32_t* key32_b = S.get<uint32_t>(npits);
  unsigned long long* key64_a = S.get<unsigned long long>(npits);
  unsigned long long* key64_b = S.get<unsigned long long>(npits);
  if (!ord_a || !ord_b || !key32_b || !key64_a || !key64_b)
    return set_error(LITAP_ENOMEM, "device allocation failed (shed sort buffers)");
  const unsigned pblk = (npits + 255) / 256;
  KERNEL("k_iota", k_iota<<<pblk, 256, 0, st>>>(ord_a, npits));
  size_t tmp1 = 0, tmp2 = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp1, pits.rootidx, key32_b, ord_a, ord_b, (int)npits, 0, 32, st);
  cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp2, key64_a, key64_b, ord_b, ord_a, (int)npits, 0, 64, st);
  size_t tmpn = tmp1 > tmp2 ? tmp1 : tmp2;
  void* tmp = S.get<unsigned char>(tmpn ? tmpn : 256);
  if (!tmp) return set_error(LITAP_ENOMEM, "device allocation failed (sort scratch)");
  // stable LSD: first by the tie-break (cell id asc), then by max elevation desc.  Cell ids need few bits: only those
  // are sorted (one more, so that the all-ones "none" still sorts last)
  int idx_bits = 1;
  while (idx_bits < 32 && ((unsigned long long)W.tg.d.cells >> idx_bits) != 0) ++idx_bits;
  idx_bits = idx_bits < 31 ? idx_bits + 1 : 32;
  { ProfScope _p("cub_sort_pits_idx"); CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tmpn, pits.rootidx, key32_b, ord_a, ord_b, (int)npits, 0, idx_bits, st)); }
  KERNEL("k_gather_u64", k_gather_u64<<<pblk, 256, 0, st>>>(pits.rootmax, ord_b, key64_a, npits));
  { ProfScope _p("cub_sort_pits_key"); CUDA_TRY(cub::DeviceRadixSort::SortPairsDescending(tmp, tmpn, key64_a, key64_b, ord_b, ord_a, (int)npits, 0, 64, st)); }
  KERNEL("k_scatter_rank", k_scatter_rank<<<pblk, 256, 0, st>>>(ord_a, pits.rank, npits));
  c.launches += 4;  // cub passes (at least)
  return LITAP_OK;
}

int shed_phase_final(litap_grid* g, shed::ShedWork& W, const shed::BandSeeds& seeds) {
  using namespace shed;
  Ctx& c = ctx();
  const GridDims d = g->d;
  cudaStream_t st = c.stream;
  {
    int m = d.NR > d.NC ? d.NR : d.NC;
    dim3 grid((m + 255) / 256, 4);
    KERNEL("k_ring_fill", k_ring_fill<<<grid, 256, 0, st>>>(d, g->shed, g->upslope, g->ridge, g->halo_bot ? 1 : 0));
  }
  uint32_t* slab = W.l2.slab;
  {
    const size_t n = std::max(W.nslots, (size_t)W.npits);
    KERNEL("k_final_labels", k_final_labels<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(
        W.l2.jres, W.nslots, W.pits.rank, W.pits.cell, W.npits, slab, reinterpret_cast<uint32_t*>(g->shed)));
  }
  dim3 tgrid(W.tg.tilesX, W.tg.tilesY);
  KERNEL("k_shed_final", k_shed_final<<<tgrid, NT, 0, st>>>(W.tg, g->ddir, W.l2, slab, W.tword, g->shed, g->upslope,
                                                                g->ridge, seeds));
  return LITAP_OK;
}

int run_shed(litap_grid* g) {
  shed::ShedWork W;
  int rc = shed_phase_local(g, W);
  if (rc) return rc;
  g->npits = (int64_t)W.npits;
  ctx().tm.npits = g->npits;
  rc = shed_phase_rank(W);
  if (rc) return rc;
  rc = shed_phase_final(g, W, shed::BandSeeds());
  if (rc) return rc;
  STREAM_SYNC(ctx().stream);  // scratch is returned to the pool when W dies
  return LITAP_OK;
}
