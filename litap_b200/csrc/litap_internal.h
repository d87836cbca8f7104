// litap_internal.h — shared declarations of liblitap_cuda.so (not installed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/litap.h"

// ---- HBM layout -------------------------------------------------------------------------------
// Every per-cell array of a grid uses one pitched layout: buffered cell (r, c), 0 <= r < NR,
// 0 <= c < NC, lives at element r*pitch + c of a pointer that is XOFF elements past a
// 256-byte-aligned allocation; pitch is a multiple of 32 elements.  So the first interior
// column (c = 1) of every row is 32-element aligned for every dtype (32 B for u8, 128 B for
// u32, 256 B for f64): warp-wide row accesses are sector-aligned and rows are TMA-legal.
// A "cell id" on the device is that pitched linear index (uint32, < 2^31).
constexpr int LITAP_XOFF = 31;

struct GridDims {
  int NR, NC;      // buffered rows / cols
  int pitch;       // elements per row (multiple of 32)
  int64_t cells;   // NR * pitch (allocation is cells + 64 elements)
};

inline GridDims make_dims(int64_t NR, int64_t NC) {
  GridDims d;
  d.NR = (int)NR;
  d.NC = (int)NC;
  d.pitch = (int)(((NC + LITAP_XOFF + 31) / 32) * 32);
  d.cells = (int64_t)d.NR * d.pitch;
  return d;
}

// ---- shed / upslope tiling --------------------------------------------------------------------
#ifndef LITAP_TILE
#define LITAP_TILE 32
#endif
constexpr int TW = LITAP_TILE, TH = LITAP_TILE;  // interior cells per tile
constexpr int LW = TW + 2, LH = TH + 2;    // local grid incl. the 1-cell ring
#ifndef LITAP_LPAD
#define LITAP_LPAD 1
#endif
// row pitch of the local grid in shared memory: odd (35), so that the 32 threads of a warp - 8 tile rows x 4 groups of
// 8 columns - touch 32 different banks when each accesses the k-th of its own cells (with a pitch of 34 they met in
// pairs: every own-cell access of the tile kernels took two wavefronts, and k_shed_local is bound by the LSU pipe)
constexpr int LP = LW + LITAP_LPAD;
constexpr int LC = LP * LH;                // local cells (1190 for 32 x 32)
constexpr int BUSED = 2 * TW + 2 * (TH - 2);        // border cells per tile (124)
constexpr int BSLOTS = (BUSED + 31) / 32 * 32;      // border slots per tile (128)
constexpr uint32_t ID_NA = 0xFFFFFFFFu;    // "no shed" (NA drec chain / cycle)
constexpr uint32_t ID_ROOT = 0x80000000u;  // flag: low 31 bits are a pit id (pid)

struct litap_grid {
  GridDims d;
  // per-cell arrays (pitched, pointers already offset by XOFF); nullptr until first needed
  double* elev = nullptr;
  uint8_t* ddir = nullptr;
  uint8_t* flat = nullptr;
  int32_t* shed = nullptr;
  uint32_t* upslope = nullptr;
  uint8_t* ridge = nullptr;
  double* form[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  double* weti[4] = {nullptr, nullptr, nullptr, nullptr};
  double* uced = nullptr;
  // slope_gc: sgre, sgr, sgcn, sgc, scr, scc / hill_r_dir, hill_c_dir / hill_r_n, hill_r_cell, hill_c_n, hill_c_cell
  double* gc_f64[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  uint8_t* gc_u8[2] = {nullptr, nullptr};
  int32_t* gc_i32[4] = {nullptr, nullptr, nullptr, nullptr};
  int32_t* gc_sum = nullptr;   // slope_gc: per (column, row segment) label summaries (gc.cu)
  int64_t npits = 0;
  bool have_ddir = false, have_shed = false;
  // row-band geometry: rows [own_lo, own_hi) are this grid's own cells; row 0 / NR-1 are the NA ring
  // of the DEM or halo rows of the neighbouring bands; row_off = DEM row of local row 0
  int own_lo = 0, own_hi = 0;
  long long row_off = 0;
  bool halo_top = false, halo_bot = false;
  // row band of a larger DEM (bands.cu): nbands == 0 for a whole DEM
  int band = 0, nbands = 0;
  long long gnr = 0;                        // rows of the whole DEM
  int64_t pit_bound_all = -1;               // the same bound for the band with the most flat cells (all bands agree on it)
  int64_t pit_bound = -1;                   // >= number of pits of the ddir on the grid (flat cells of the D8 stencil), -1 unknown
  int edge_top_row = -1, edge_bot_row = -1; // local row holding the first / last DEM row, -1 if not in this grid
  // calc_ddir2 worklists (ddir.cu)
  uint32_t* dd_list = nullptr;
  uint8_t* dd_cand = nullptr;
  unsigned long long* dd_keys = nullptr;
  void* dd_ctr = nullptr;
  size_t dd_cap = 0;
};

// ---- context ----------------------------------------------------------------------------------
struct Ctx {
  bool ready = false;
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[8] = {};
  litap_timings tm = {};
  int64_t launches = 0;
};
Ctx& ctx();
int set_error(int code, const std::string& msg);
// The library keeps process-wide state (pool, staging buffers, timings, the stream swap of *_flow_form) and is driven
// by one host thread at a time; callers that release a global lock around foreign calls (ctypes does) could still
// enter twice, so every extern "C" entry point takes this (recursive: entry points call each other) mutex.
struct ApiLock {
  ApiLock();
  ~ApiLock();
};
int ensure_init();

// litap_*_flow_form: the flow stages run on a high-priority stream of the library beside calc_form on the
// caller's stream; ctx().stream is what every stage driver enqueues on, so the flow stages are redirected
// by swapping it for their duration
int flow_streams(cudaStream_t* hi, cudaEvent_t* fork, cudaEvent_t* join);  // api.cu (created on first use)
struct StreamSwap {
  cudaStream_t keep;
  bool joined = false;   // set once the join event is recorded; a scope left early (an error) drains the swapped-in
                         // stream so that nothing is still running on it when scratch goes back to the pool
  explicit StreamSwap(cudaStream_t s) : keep(ctx().stream) { ctx().stream = s; }
  ~StreamSwap() {
    if (!joined) cudaStreamSynchronize(ctx().stream);
    ctx().stream = keep;
  }
};

// pooled device allocations (freed by litap_shutdown); zero = cudaMemsetAsync to 0
void* pool_alloc(size_t bytes);
void pool_free(void* p);
void pool_release_all();
// hostcopy.cu: device -> caller memory; pageable destinations go through pinned staging buffers + host threads
int d2h_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows, cudaStream_t st);
int d2h_2d_i32_expand(double* dst, size_t dpitch, const int32_t* src, size_t spitch, size_t ncol, size_t rows, int kind, cudaStream_t st);
int h2d_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows, cudaStream_t st);
bool hostcopy_is_pageable(const void* p);   // would d2h_2d stage a copy into p?
int hostcopy_flush();    // pageable destinations are complete after this
void hostcopy_reset();   // start of a call: forget chunks a failed call left pending
void hostcopy_forget_staging();
void rasterio_forget();   // rasterio.cu: writer lanes (litap_shutdown)
int hostcopy_staging(int which, void* buf[2], cudaEvent_t ev[2], size_t* bytes);   // pinned pairs for file I/O (rasterio.cu)

template <typename T>
inline T* grid_alloc_array(const GridDims& d, int fill_byte) {
  size_t n = (size_t)d.cells + 64;
  T* base = (T*)pool_alloc(n * sizeof(T));
  if (!base) return nullptr;
  if (fill_byte >= 0) cudaMemsetAsync(base, fill_byte, n * sizeof(T), ctx().stream);
  return base + LITAP_XOFF;
}
template <typename T>
inline void grid_free_array(T*& p) {
  if (p) pool_free((void*)(p - LITAP_XOFF));
  p = nullptr;
}

// scope-owned pool allocations (returned to the pool when the scope dies; sync the stream first)
#include <vector>
struct PoolScratch {
  std::vector<void*> ptrs;
  ~PoolScratch() { for (void* p : ptrs) pool_free(p); }
  template <typename T> T* get(size_t n) {
    T* p = (T*)pool_alloc((n ? n : 1) * sizeof(T));
    if (p) ptrs.push_back(p);
    return p;
  }
};

#define CUDA_TRY(expr)                                                                      \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return set_error(LITAP_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));    \
  } while (0)
// a host wait on the stream inside a stage (counted: litap_timings.host_syncs)
#define STREAM_SYNC(st)                        \
  do {                                         \
    CUDA_TRY(cudaStreamSynchronize(st));       \
    ++ctx().tm.host_syncs;                     \
  } while (0)
#define LAUNCH_CHECK()                                                                      \
  do {                                                                                      \
    ctx().launches++;                                                                       \
    cudaError_t _e = cudaGetLastError();                                                    \
    if (_e != cudaSuccess)                                                                  \
      return set_error(LITAP_ECUDA, std::string("kernel launch: ") + cudaGetErrorString(_e) + \
                                        " at " + __FILE__ + ":" + std::to_string(__LINE__)); \
  } while (0)

// Optional per-kernel timing (litap_profile_enable): CUDA events around every launch, summed
// per kernel name.  Off by default; the bench turns it on for a separate, untimed pass.
struct ProfScope {
  int idx;
  explicit ProfScope(const char* name);
  ~ProfScope();
};
#define KERNEL(name, ...)   \
  do {                      \
    ProfScope _prof(name);  \
    __VA_ARGS__;            \
    LAUNCH_CHECK();         \
  } while (0)

struct ArrRef { const void* p; int esize; };
ArrRef find_array_impl(litap_grid* g, const char* name);   // api.cu: result array of a grid by name
ArrRef grid_find_array(litap_grid* g, const char* name);

// ---- stage drivers (each enqueues on ctx().stream; host syncs only where stated) --------------
int run_ddir(litap_grid* g, int invert);               // ddir.cu
int run_shed(litap_grid* g);                           // shed.cu
int run_form(litap_grid* g, double grid_m);            // form.cu
int run_form_rows(litap_grid* g, double grid_m, int row_lo, int row_hi);  // form.cu: a row range (pipelined upload)
int run_form_pack(litap_grid* g, int32_t* const* out, int row_lo, int row_hi, int* flags);  // form.cu: outputs as int32 (packed downloads)
void form_forget_trig();                               // form.cu (pool-owned table dropped at shutdown)
int run_weti(litap_grid* g, double grid_m);            // weti.cu
int run_uced(litap_grid* g);                           // uced.cu
int run_pit_stat(litap_grid* g, const uint8_t* edge, int method, int64_t cap, double* const* out, int64_t* nrows);  // pitstat.cu
int run_slope_gc(litap_grid* g, double grid_m);        // gc.cu
int run_synth(litap_grid* g, int64_t row0, int64_t gnr, uint32_t seed, double quant, int na_permille);  // synth.cu
int run_checksum(const void* arr, int elem_size, const GridDims& d, uint64_t* out);                     // synth.cu
int run_checksum_rows(const void* arr, int elem_size, const GridDims& d, int row_lo, int row_hi, long long row_off, uint64_t* out);
