// bands.cu — row-band sharding of one DEM over several B200s.
//
// The DEM is cut into `nbands` contiguous row bands.  A band is an ordinary litap_grid whose local
// rows are
//     row 0            unused (NA)
//     row 1            north halo: last own row of the band above, or the NA ring of the DEM (band 0)
//     rows 2 .. NR-2   own rows
//     row NR-1         south halo: first own row of the band below, or the NA ring (last band)
// so that the single-device kernels run on it unchanged (tiles of calc_shed4 start at row 1).
//
// Two transports carry the (small) exchanges between bands:
//   local  all bands live in this process on the current device: device-to-device copies.  This is
//          the "multi-shard on one device" mode the tests use to prove that the sharded algorithm
//          gives the bits of the unsharded one.
//   nccl   one band per process / GPU (torchrun): ncclSend/ncclRecv for halo rows, ncclAllGather
//          for the boundary tables.  NCCL is loaded with dlopen at run time (the copy PyTorch has
//          already loaded when there is one), so the library itself has no link dependency on it.
//
// What crosses a band boundary:
//   calc_ddir2   one row of ddir (1 B/col) per plateau round + an "any change" flag; max(elev) for
//                invert(); the keys of flat-pit patch cells (all bands resolve all patches)
//   calc_form    nothing (halo rows of elev are generated / uploaded with the band)
//   calc_shed4   per band 48 B/col: for the entry cells of its first and last own row the band-local
//                root (a pit of the band, or the entry cell of the neighbouring band the path leaves
//                through), and for the cells of its halo rows ("portals") the flow and the
//                max-elevation key that leave through them; plus the per-pit keys for the global
//                canonical numbering.  Every band solves the same small boundary graph.
#include <dlfcn.h>
#include <algorithm>
#include <cstring>
#include <vector>
#include <cub/cub.cuh>
#include "cellmath.h"
#include "shed_tile.cuh"

// phases of the single-device stages (ddir.cu, shed.cu, form.cu)
int shed_phase_local(litap_grid* g, shed::ShedWork& W);
int shed_phase_final(litap_grid* g, shed::ShedWork& W, const shed::BandSeeds& seeds);
int ddir_band_begin(litap_grid* g);                                   // counters + D8 stencil
int ddir_band_counts(litap_grid* g, unsigned long long* nflat, unsigned* npatch);
int ddir_band_probe(litap_grid* g, unsigned nflat, unsigned* changed_dev);
int ddir_band_apply(litap_grid* g, unsigned nflat);
int ddir_band_mark(litap_grid* g);                                    // patch cells of the band -> g->dd_keys
int ddir_band_patch_apply(litap_grid* g, const unsigned long long* K, const uint8_t* nd, unsigned np);
int ddir_patch_resolve(const unsigned long long* keys, unsigned np, PoolScratch& S, unsigned long long** K_out,
                       uint8_t** dirs_out);
int ddir_band_elev_max(litap_grid* g, unsigned long long* out_dev);
int ddir_band_invert(litap_grid* g, const unsigned long long* max_dev);
int ddir_band_grow(litap_grid* g, size_t cap);

// ------------------------------------------------------------------------------------------------
// NCCL through dlopen
namespace {
typedef struct ncclComm* ncclComm_t;
struct NcclUniqueId { char internal[128]; };
struct NcclApi {
  void* h = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, NcclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
ncclComm_t g_comm = nullptr;
int g_rank = 0, g_nranks = 1;
constexpr int NCCL_CHAR = 0;  // ncclInt8 / ncclChar

int nccl_load() {
  if (g_nccl.h) return LITAP_OK;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // PyTorch's copy, if loaded
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return set_error(LITAP_ECOMM, std::string("cannot load libnccl.so.2: ") + dlerror());
#define SYM(field, name)                                                                       \
  *(void**)(&g_nccl.field) = dlsym(h, name);                                                   \
  if (!g_nccl.field) return set_error(LITAP_ECOMM, std::string("libnccl lacks ") + name)
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(Send, "ncclSend");
  SYM(Recv, "ncclRecv");
  SYM(AllGather, "ncclAllGather");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  g_nccl.h = h;
  return LITAP_OK;
}
#define NCCL_TRY(expr)                                                                          \
  do {                                                                                          \
    int _r = (expr);                                                                            \
    if (_r != 0) return set_error(LITAP_ECOMM, std::string(#expr) + ": " + g_nccl.GetErrorString(_r)); \
  } while (0)
}  // namespace

extern "C" int litap_comm_unique_id(void* id128) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!id128) return set_error(LITAP_EARG, "litap_comm_unique_id: NULL");
  int rc = nccl_load();
  if (rc) return rc;
  NcclUniqueId id;
  NCCL_TRY(g_nccl.GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return LITAP_OK;
}
extern "C" int litap_comm_init(int rank, int nranks, const void* id128) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!id128 || nranks < 1 || rank < 0 || rank >= nranks) return set_error(LITAP_EARG, "litap_comm_init: bad arguments");
  int rc = ensure_init();
  if (rc) return rc;
  rc = nccl_load();
  if (rc) return rc;
  if (g_comm) { g_nccl.CommDestroy(g_comm); g_comm = nullptr; }
  NcclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  NCCL_TRY(g_nccl.CommInitRank(&g_comm, nranks, id, rank));
  g_rank = rank;
  g_nranks = nranks;
  return LITAP_OK;
}
extern "C" void litap_comm_destroy(void) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (g_comm) g_nccl.CommDestroy(g_comm);
  g_comm = nullptr;
  g_rank = 0;
  g_nranks = 1;
}

// ------------------------------------------------------------------------------------------------
// band geometry and transport
namespace {
struct BandGeom { int band, nbands; long long r0, r1; };  // own DEM rows [r0, r1)

inline BandGeom band_rows(long long gnr, int band, int nbands) {
  BandGeom b;
  b.band = band;
  b.nbands = nbands;
  b.r0 = gnr * band / nbands;
  b.r1 = gnr * (band + 1) / nbands;
  return b;
}

// The caller passes the bands this process owns, in band order.  Local transport: all nbands bands.
// NCCL transport: exactly one band, band index == rank.
int check_bands(litap_grid** B, int n) {
  if (!B || n < 1) return set_error(LITAP_EARG, "bands: empty band list");
  for (int i = 0; i < n; ++i) {
    if (!B[i] || B[i]->nbands < 1) return set_error(LITAP_EARG, "bands: not a band grid (use litap_band_create)");
    if (B[i]->nbands != B[0]->nbands || B[i]->d.NC != B[0]->d.NC || B[i]->band != B[0]->band + i)
      return set_error(LITAP_EARG, "bands: the list must hold consecutive bands of one DEM");
  }
  if (g_comm) {
    if (n != 1 || B[0]->band != g_rank || B[0]->nbands != g_nranks)
      return set_error(LITAP_EARG, "bands: with a communicator every process owns exactly the band of its rank");
  } else if (n != B[0]->nbands || B[0]->band != 0) {
    return set_error(LITAP_EARG, "bands: without a communicator all bands of the DEM must be passed");
  }
  return LITAP_OK;
}

// own boundary rows -> halo rows of the neighbouring bands, for one per-cell array (esize bytes/cell)
template <typename T>
int halo_exchange(litap_grid** B, int n, T* litap_grid::*arr) {
  cudaStream_t st = ctx().stream;
  if (g_comm) {
    litap_grid* g = B[0];
    T* a = g->*arr;
    const size_t bytes = (size_t)g->d.NC * sizeof(T);
    NCCL_TRY(g_nccl.GroupStart());
    if (g->halo_top) {
      NCCL_TRY(g_nccl.Send(a + (size_t)g->own_lo * g->d.pitch, bytes, NCCL_CHAR, g_rank - 1, g_comm, st));
      NCCL_TRY(g_nccl.Recv(a + (size_t)(g->own_lo - 1) * g->d.pitch, bytes, NCCL_CHAR, g_rank - 1, g_comm, st));
    }
    if (g->halo_bot) {
      NCCL_TRY(g_nccl.Send(a + (size_t)(g->own_hi - 1) * g->d.pitch, bytes, NCCL_CHAR, g_rank + 1, g_comm, st));
      NCCL_TRY(g_nccl.Recv(a + (size_t)g->own_hi * g->d.pitch, bytes, NCCL_CHAR, g_rank + 1, g_comm, st));
    }
    NCCL_TRY(g_nccl.GroupEnd());
    return LITAP_OK;
  }
  for (int i = 0; i + 1 < n; ++i) {
    litap_grid *u = B[i], *l = B[i + 1];
    const size_t bytes = (size_t)u->d.NC * sizeof(T);
    CUDA_TRY(cudaMemcpyAsync((l->*arr) + (size_t)(l->own_lo - 1) * l->d.pitch, (u->*arr) + (size_t)(u->own_hi - 1) * u->d.pitch,
                             bytes, cudaMemcpyDeviceToDevice, st));
    CUDA_TRY(cudaMemcpyAsync((u->*arr) + (size_t)u->own_hi * u->d.pitch, (l->*arr) + (size_t)l->own_lo * l->d.pitch, bytes,
                             cudaMemcpyDeviceToDevice, st));
  }
  return LITAP_OK;
}

// every band contributes `bytes` from send[i]; afterwards recv[i] holds nbands * bytes, band order
int all_gather(litap_grid** B, int n, void* const* send, void* const* recv, size_t bytes) {
  cudaStream_t st = ctx().stream;
  if (g_comm) {
    NCCL_TRY(g_nccl.AllGather(send[0], recv[0], bytes, NCCL_CHAR, g_comm, st));
    return LITAP_OK;
  }
  (void)B;
  for (int dst = 0; dst < n; ++dst)
    for (int src = 0; src < n; ++src)
      CUDA_TRY(cudaMemcpyAsync((char*)recv[dst] + (size_t)src * bytes, send[src], bytes, cudaMemcpyDeviceToDevice, st));
  return LITAP_OK;
}

// a few 64-bit words per band, gathered and brought to the host (this synchronises the stream)
int gather_words_host(litap_grid** B, int n, const std::vector<std::vector<unsigned long long>>& mine, int nwords,
                      std::vector<unsigned long long>& all) {
  const int P = B[0]->nbands;
  PoolScratch S;
  std::vector<void*> send(n), recv(n);
  for (int i = 0; i < n; ++i) {
    send[i] = S.get<unsigned long long>(nwords);
    recv[i] = S.get<unsigned long long>((size_t)nwords * P);
    if (!send[i] || !recv[i]) return set_error(LITAP_ENOMEM, "device allocation failed (gather words)");
    CUDA_TRY(cudaMemcpyAsync(send[i], mine[i].data(), nwords * 8, cudaMemcpyHostToDevice, ctx().stream));
  }
  int rc = all_gather(B, n, send.data(), recv.data(), (size_t)nwords * 8);
  if (rc) return rc;
  all.resize((size_t)nwords * P);
  CUDA_TRY(cudaMemcpyAsync(all.data(), recv[0], all.size() * 8, cudaMemcpyDeviceToHost, ctx().stream));
  STREAM_SYNC(ctx().stream);
  return LITAP_OK;
}
}  // namespace

extern "C" int litap_band_create(int64_t gnr, int64_t nc, int band, int nbands, litap_grid** out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!out) return set_error(LITAP_EARG, "litap_band_create: NULL out");
  *out = nullptr;
  if (nbands < 1 || band < 0 || band >= nbands) return set_error(LITAP_EARG, "litap_band_create: bad band index");
  BandGeom bg = band_rows(gnr, band, nbands);
  if (bg.r1 - bg.r0 < 3) return set_error(LITAP_EARG, "litap_band_create: every band needs at least 3 DEM rows");
  int rc = litap_grid_create(bg.r1 - bg.r0 + 3, nc + 2, out);
  if (rc) return rc;
  litap_grid* g = *out;
  g->band = band;
  g->nbands = nbands;
  g->gnr = gnr;
  g->own_lo = 2;
  g->own_hi = g->d.NR - 1;
  g->row_off = bg.r0 - 1;  // buffered DEM row of local row 0
  g->halo_top = band > 0;
  g->halo_bot = band < nbands - 1;
  g->edge_top_row = band == 0 ? 2 : -1;
  g->edge_bot_row = band == nbands - 1 ? g->d.NR - 2 : -1;
  return LITAP_OK;
}

// synthetic DEM of the band, halo rows included (bit-identical with the unsharded generator)
extern "C" int litap_band_synth(litap_grid* g, uint32_t seed, double quant, int na_permille) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || g->nbands < 1) return set_error(LITAP_EARG, "litap_band_synth: not a band grid");
  // run_synth maps local row R to DEM row row0 + R - 1; local row 2 is DEM row r0
  int rc = run_synth(g, g->row_off, g->gnr, seed, quant, na_permille);
  if (rc) return rc;
  CUDA_TRY(cudaMemsetAsync(g->elev, 0xFF, (size_t)g->d.pitch * sizeof(double), ctx().stream));  // row 0 = NA
  g->have_ddir = g->have_shed = false;
  STREAM_SYNC(ctx().stream);
  return LITAP_OK;
}

// the band's own copy of its rows: local rows 1 .. NR-1 (halo rows included), NC wide, dense
extern "C" int litap_band_upload_rows(litap_grid* g, const double* rows) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || g->nbands < 1 || !rows) return set_error(LITAP_EARG, "litap_band_upload_rows: bad arguments");
  const GridDims d = g->d;
  CUDA_TRY(cudaMemcpy2DAsync(g->elev + d.pitch, (size_t)d.pitch * 8, rows, (size_t)d.NC * 8, (size_t)d.NC * 8, d.NR - 1,
                             cudaMemcpyHostToDevice, ctx().stream));
  CUDA_TRY(cudaMemsetAsync(g->elev, 0xFF, (size_t)d.pitch * sizeof(double), ctx().stream));
  g->have_ddir = g->have_shed = false;
  STREAM_SYNC(ctx().stream);
  return LITAP_OK;
}

// rows of the full buffered DEM (gnr+2 rows, NC wide) that belong to this band, halo rows included
extern "C" int litap_band_upload_elev(litap_grid* g, const double* elev_full) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || g->nbands < 1 || !elev_full) return set_error(LITAP_EARG, "litap_band_upload_elev: bad arguments");
  const GridDims d = g->d;
  // local rows 1 .. NR-1 are buffered DEM rows row_off+1 .. row_off+NR-1
  const double* src = elev_full + (size_t)(g->row_off + 1) * d.NC;
  CUDA_TRY(cudaMemcpy2DAsync(g->elev + d.pitch, (size_t)d.pitch * 8, src, (size_t)d.NC * 8, (size_t)d.NC * 8, d.NR - 1,
                             cudaMemcpyHostToDevice, ctx().stream));
  CUDA_TRY(cudaMemsetAsync(g->elev, 0xFF, (size_t)d.pitch * sizeof(double), ctx().stream));
  g->have_ddir = g->have_shed = false;
  STREAM_SYNC(ctx().stream);
  return LITAP_OK;
}

// ------------------------------------------------------------------------------------------------
// calc_ddir2 over row bands (R/flow_01_calc_ddir.R:2-118)
namespace {
__global__ void k_max_words(const unsigned long long* __restrict__ w, int n, unsigned long long* out) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long m = 0;
    for (int i = 0; i < n; ++i) m = w[i] > m ? w[i] : m;
    *out = m;
  }
}
}  // namespace

extern "C" int litap_bands_ddir(litap_grid** B, int n, int invert) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  int rc = check_bands(B, n);
  if (rc) return rc;
  Ctx& c = ctx();
  cudaStream_t st = c.stream;
  const int P = B[0]->nbands;
  const int64_t l0 = c.launches;
  cudaEventRecord(c.ev[0], st);
  PoolScratch S;

  if (invert) {  // invert(): elev <- max(elev) - elev with the maximum of the whole DEM
    std::vector<void*> send(n), recv(n);
    for (int i = 0; i < n; ++i) {
      send[i] = S.get<unsigned long long>(1);
      recv[i] = S.get<unsigned long long>(P + 1);
      if (!send[i] || !recv[i]) return set_error(LITAP_ENOMEM, "device allocation failed (invert)");
      CUDA_TRY(cudaMemsetAsync(send[i], 0, 8, st));
      if ((rc = ddir_band_elev_max(B[i], (unsigned long long*)send[i]))) return rc;
    }
    if ((rc = all_gather(B, n, send.data(), recv.data(), 8))) return rc;
    for (int i = 0; i < n; ++i) {
      unsigned long long* r = (unsigned long long*)recv[i];
      KERNEL("k_max_words", k_max_words<<<1, 32, 0, st>>>(r, P, r + P));
      if ((rc = ddir_band_invert(B[i], r + P))) return rc;
    }
  }

  std::vector<unsigned long long> nflat(n);
  for (int attempt = 0;; ++attempt) {
    for (int i = 0; i < n; ++i)
      if ((rc = ddir_band_begin(B[i]))) return rc;
    bool overflow = false;
    for (int i = 0; i < n; ++i) {
      if ((rc = ddir_band_counts(B[i], &nflat[i], nullptr))) return rc;
      if (nflat[i] > B[i]->dd_cap) {  // unusually flat band: regrow its worklist and redo the stencil
        overflow = true;
        if ((rc = ddir_band_grow(B[i], (size_t)nflat[i] + 4096))) return rc;
      }
    }
    if (!overflow) break;
    if (attempt) return set_error(LITAP_ECUDA, "internal: flat worklist overflow after regrowth");
  }
  if ((rc = halo_exchange(B, n, &litap_grid::ddir))) return rc;

  // plateau rounds: probe everywhere (reads the previous round), apply everywhere, exchange the
  // boundary rows, stop when no band changed anything
  // The rounds keep their Jacobi semantics (every round ends with the halo rows exchanged, in stream order), but the
  // "did anything change" flags are only looked at every KR rounds: one all-gather + host read-back per batch instead
  // of per round.  A round after the fixed point changes nothing, so running up to KR - 1 of them is harmless.
  constexpr int KR = 3;
  std::vector<void*> fsend(n), frecv(n);
  for (int i = 0; i < n; ++i) {
    fsend[i] = S.get<unsigned long long>(KR);
    frecv[i] = S.get<unsigned long long>((size_t)KR * P);
    if (!fsend[i] || !frecv[i]) return set_error(LITAP_ENOMEM, "device allocation failed (round flags)");
  }
  int rounds = 0;
  std::vector<unsigned long long> flags((size_t)KR * P);
  for (bool open = true; open;) {
    for (int i = 0; i < n; ++i) CUDA_TRY(cudaMemsetAsync(fsend[i], 0, 8 * KR, st));
    for (int k = 0; k < KR; ++k) {
      for (int i = 0; i < n; ++i)
        if ((rc = ddir_band_probe(B[i], (unsigned)nflat[i], (unsigned*)((unsigned long long*)fsend[i] + k)))) return rc;
      for (int i = 0; i < n; ++i)
        if ((rc = ddir_band_apply(B[i], (unsigned)nflat[i]))) return rc;
      if ((rc = halo_exchange(B, n, &litap_grid::ddir))) return rc;
    }
    if ((rc = all_gather(B, n, fsend.data(), frecv.data(), 8 * KR))) return rc;
    CUDA_TRY(cudaMemcpyAsync(flags.data(), frecv[0], (size_t)KR * P * 8, cudaMemcpyDeviceToHost, st));
    STREAM_SYNC(st);
    for (int k = 0; k < KR && open; ++k) {
      bool any = false;
      for (int b = 0; b < P; ++b) any |= flags[(size_t)b * KR + k] != 0;
      if (any) ++rounds; else open = false;
    }
  }

  // flat-pit patches: every band marks its own patch cells; all bands resolve all patches
  std::vector<std::vector<unsigned long long>> mine(n, std::vector<unsigned long long>(2));
  for (int i = 0; i < n; ++i)
    if ((rc = ddir_band_mark(B[i]))) return rc;
  for (int i = 0; i < n; ++i) {
    unsigned np = 0;
    if ((rc = ddir_band_counts(B[i], nullptr, &np))) return rc;
    mine[i][0] = np;
    mine[i][1] = nflat[i];   // travels along: calc_shed4 sizes the global pit table from it without a gather of its own
  }
  std::vector<unsigned long long> both, counts(P);
  if ((rc = gather_words_host(B, n, mine, 2, both))) return rc;
  unsigned long long total = 0, mx = 0, flat_total = 0, flat_max = 0;
  for (int b = 0; b < P; ++b) {
    counts[b] = both[2 * b];
    total += counts[b];
    mx = std::max(mx, counts[b]);
    flat_max = std::max(flat_max, both[2 * b + 1]);
  }
  for (int i = 0; i < n; ++i) {
    flat_total += nflat[i];
    // pits of the band in calc_shed4: own cells that stay flat + one portal per halo cell
    B[i]->pit_bound = (int64_t)nflat[i] + 2 * (int64_t)B[i]->d.NC;
    B[i]->pit_bound_all = (int64_t)flat_max + 2 * (int64_t)B[i]->d.NC;   // the same bound for the largest band
  }
  if (total) {
    std::vector<void*> ksend(n), krecv(n);
    for (int i = 0; i < n; ++i) {
      ksend[i] = S.get<unsigned long long>(mx);
      krecv[i] = S.get<unsigned long long>(mx * P);
      if (!ksend[i] || !krecv[i]) return set_error(LITAP_ENOMEM, "device allocation failed (patch keys)");
      if (mine[i][0])
        CUDA_TRY(cudaMemcpyAsync(ksend[i], B[i]->dd_keys, mine[i][0] * 8, cudaMemcpyDeviceToDevice, st));
    }
    if ((rc = all_gather(B, n, ksend.data(), krecv.data(), mx * 8))) return rc;
    for (int i = 0; i < n; ++i) {
      unsigned long long* all = S.get<unsigned long long>(total);
      if (!all) return set_error(LITAP_ENOMEM, "device allocation failed (patch keys)");
      size_t off = 0;
      for (int b = 0; b < P; ++b) {  // compact the padded gather
        if (counts[b])
          CUDA_TRY(cudaMemcpyAsync(all + off, (unsigned long long*)krecv[i] + (size_t)b * mx, counts[b] * 8,
                                   cudaMemcpyDeviceToDevice, st));
        off += counts[b];
      }
      unsigned long long* K = nullptr;
      uint8_t* nd = nullptr;
      if ((rc = ddir_patch_resolve(all, (unsigned)total, S, &K, &nd))) return rc;
      if ((rc = ddir_band_patch_apply(B[i], K, nd, (unsigned)total))) return rc;
    }
  }
  cudaEventRecord(c.ev[1], st);
  STREAM_SYNC(st);
  cudaEventElapsedTime(&c.tm.ddir_ms, c.ev[0], c.ev[1]);
  c.tm.plateau_rounds = rounds;
  c.tm.flat_cells = (int64_t)flat_total;  // of the bands of this process
  c.tm.patch_cells = (int64_t)total;
  c.tm.kernel_launches = c.launches - l0;
  for (int i = 0; i < n; ++i) B[i]->have_ddir = true;
  return LITAP_OK;
}

extern "C" int litap_bands_form(litap_grid** B, int n, double grid_m) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  int rc = check_bands(B, n);
  if (rc) return rc;
  if (!(grid_m > 0)) return set_error(LITAP_EARG, "litap_bands_form: grid must be > 0");
  Ctx& c = ctx();
  const int64_t l0 = c.launches;
  cudaEventRecord(c.ev[4], c.stream);
  for (int i = 0; i < n; ++i)
    if ((rc = run_form(B[i], grid_m))) return rc;
  cudaEventRecord(c.ev[5], c.stream);
  STREAM_SYNC(c.stream);
  cudaEventElapsedTime(&c.tm.form_ms, c.ev[4], c.ev[5]);
  c.tm.kernel_launches = c.launches - l0;
  return LITAP_OK;
}

// ------------------------------------------------------------------------------------------------
// calc_shed4 over row bands (R/flow_02_calc_shed.R:10-145)
extern "C" int litap_bands_shed(litap_grid** B, int n);
// form + ddir + shed of the local bands as one call: calc_form on the caller's stream, the flow stages (with
// their exchanges) on the library's high-priority stream, like litap_grid_flow_form
extern "C" int litap_bands_flow_form(litap_grid** B, int n, double grid_m, int invert) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  int rc = check_bands(B, n);
  if (rc) return rc;
  if (!(grid_m > 0)) return set_error(LITAP_EARG, "litap_bands_flow_form: grid must be > 0");
  Ctx& c = ctx();
  cudaStream_t hi;
  cudaEvent_t fork, join;
  if ((rc = flow_streams(&hi, &fork, &join))) return rc;
  cudaStream_t user = c.stream;
  const int64_t l0 = c.launches;
  c.tm.host_syncs = 0;
  CUDA_TRY(cudaEventRecord(fork, user));
  CUDA_TRY(cudaStreamWaitEvent(hi, fork, 0));
  CUDA_TRY(cudaEventRecord(c.ev[4], user));
  for (int i = 0; i < n; ++i)
    if ((rc = run_form(B[i], grid_m))) return rc;
  CUDA_TRY(cudaEventRecord(c.ev[5], user));
  if (invert) {  // invert() rewrites elev in place: the flow stages then wait for calc_form (no overlap)
    CUDA_TRY(cudaEventRecord(fork, user));
    CUDA_TRY(cudaStreamWaitEvent(hi, fork, 0));
  }
  {
    StreamSwap sw(hi);
    if ((rc = litap_bands_ddir(B, n, invert))) return rc;
    if ((rc = litap_bands_shed(B, n))) return rc;
    CUDA_TRY(cudaEventRecord(join, hi));
    sw.joined = true;
  }
  CUDA_TRY(cudaStreamWaitEvent(user, join, 0));
  STREAM_SYNC(user);
  cudaEventElapsedTime(&c.tm.form_ms, c.ev[4], c.ev[5]);
  c.tm.kernel_launches = c.launches - l0;
  return LITAP_OK;
}

namespace bg {
using namespace shed;

__device__ __forceinline__ bool is_cell(uint32_t v) { return !(v & ID_ROOT); }  // ID_NA carries the flag bit too
// node of the boundary graph: entry cell (band, side, col); side 0 = first own row, 1 = last own row
__device__ __forceinline__ uint32_t node_of(int band, int side, int c, int NC) {
  return (uint32_t)((band * 2 + side) * NC + c);
}

// halo rows become portals: pseudo-pits that collect what leaves the band through them
__global__ void k_band_portals(uint8_t* __restrict__ ddir, const double* __restrict__ elev, GridDims d, int top, int bot) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.NC) return;
  const size_t a = (size_t)1 * d.pitch + c, b = (size_t)(d.NR - 1) * d.pitch + c;
  ddir[a] = (top && !is_nan(elev[a])) ? 5 : 0;
  ddir[b] = (bot && !is_nan(elev[b])) ? 5 : 0;
}

// cells draining into every pit: inside its tile + through the border cells of its tile
__global__ void k_pit_totals(const L2 l2, size_t nslots, uint32_t* __restrict__ tot) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nslots) return;
  uint32_t v = l2.jump0[s];
  if (v != ID_NA && (v & ID_ROOT)) {
    uint32_t e = xi_inflow(l2.xi[s]);
    if (e) atomicAdd(&tot[v & ~ID_ROOT], e);
  }
}

// band-local root (ID_ROOT|pid or ID_NA) of a cell that is not necessarily a tile-border cell
__device__ uint32_t local_root(const TileGeom& g, const uint8_t* __restrict__ ddir, const uint32_t* __restrict__ lab,
                               const uint32_t* __restrict__ jres, int r, int c) {
  const int ty = (r - 1) / TH, tx = (c - 1) / TW;
  for (int step = 0; step <= LC; ++step) {
    const int dv = ddir[(size_t)r * g.d.pitch + c];
    if (dv == 0 || dv > 9) return ID_NA;
    if (dv == 5) return ID_ROOT | lab[(size_t)r * g.d.pitch + c];
    const int rr = r + nb_dr(dv), cc = c + nb_dc(dv);
    if (rr < 1 || cc < 1 || rr >= g.d.NR || cc >= g.d.NC) return ID_NA;
    if ((rr - 1) / TH != ty || (cc - 1) / TW != tx) {
      const long long s = slot_of_rc(rr, cc, g.tilesX);
      return s < 0 ? ID_NA : jres[s];
    }
    r = rr;
    c = cc;
  }
  return ID_NA;  // a cycle inside the tile
}

struct Record {  // what one band publishes, all arrays [2][NC] (side-major)
  unsigned long long* pkmax;
  unsigned long long* pkidx;
  uint32_t* nxt;
  uint32_t* L;
};
__host__ __device__ inline Record record_at(void* base, int NC) {
  Record r;
  r.pkmax = (unsigned long long*)base;
  r.pkidx = r.pkmax + 2 * (size_t)NC;
  r.nxt = (uint32_t*)(r.pkidx + 2 * (size_t)NC);
  r.L = r.nxt + 2 * (size_t)NC;
  return r;
}
inline size_t record_bytes(int NC) { return (size_t)NC * 48; }

struct BandInfo { int band, nbands, own_lo, own_hi, halo_top, halo_bot; long long row_off; unsigned pit_stride; };

__device__ __forceinline__ unsigned long long gseq(const BandInfo& bi, const GridDims& d, uint32_t cell) {
  const int r = cell / d.pitch, c = cell - r * d.pitch;
  return (unsigned long long)(bi.row_off + r) * (unsigned long long)d.NC + (unsigned long long)c;
}
// is this pit id a portal?  returns the boundary-graph node its flow continues at, or ID_NA for a real pit
__device__ __forceinline__ uint32_t portal_node(const BandInfo& bi, const GridDims& d, uint32_t cell) {
  const int r = cell / d.pitch, c = cell - r * d.pitch;
  if (bi.halo_top && r == 1) return node_of(bi.band - 1, 1, c, d.NC);
  if (bi.halo_bot && r == d.NR - 1) return node_of(bi.band + 1, 0, c, d.NC);
  return ID_NA;
}

__global__ void k_band_tables(TileGeom g, BandInfo bi, const uint8_t* __restrict__ ddir, const uint32_t* __restrict__ lab,
                              const L2 l2, const Pits pits, const uint32_t* __restrict__ tot, Record out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int side = blockIdx.y;
  if (c >= g.d.NC) return;
  const int NC = g.d.NC;
  // entry cell of this band
  {
    const int r = side == 0 ? bi.own_lo : bi.own_hi - 1;
    uint32_t v = c >= 1 ? local_root(g, ddir, lab, l2.jres, r, c) : ID_NA;
    if (v != ID_NA) {
      const uint32_t pid = v & ~ID_ROOT;
      const uint32_t pn = portal_node(bi, g.d, pits.cell[pid]);
      v = pn != ID_NA ? pn : (ID_ROOT | ((uint32_t)bi.band * bi.pit_stride + pid));
    }
    out.nxt[(size_t)side * NC + c] = v;
  }
  // portal of this band (halo row): what leaves through it
  {
    const int r = side == 0 ? 1 : g.d.NR - 1;
    const bool on = side == 0 ? bi.halo_top : bi.halo_bot;
    uint32_t L = 0;
    unsigned long long km = 0, ki = ~0ull;
    if (on && ddir[(size_t)r * g.d.pitch + c] == 5) {
      const uint32_t pid = lab[(size_t)r * g.d.pitch + c];
      L = tot[pid] - 1;  // the portal cell itself belongs to the other band
      km = pits.rootmax[pid];
      const uint32_t ci = pits.rootidx[pid];
      ki = ci == 0xFFFFFFFFu ? ~0ull : gseq(bi, g.d, ci);
    }
    out.L[(size_t)side * NC + c] = L;
    out.pkmax[(size_t)side * NC + c] = km;
    out.pkidx[(size_t)side * NC + c] = ki;
  }
}

// keys of the real pits of the band, padded to pit_stride entries (0 = no pit here)
__global__ void k_band_pitkeys(BandInfo bi, GridDims d, const Pits pits, unsigned npits, unsigned long long* __restrict__ kmax,
                               unsigned long long* __restrict__ kidx) {
  unsigned pid = blockIdx.x * blockDim.x + threadIdx.x;
  if (pid >= bi.pit_stride) return;
  unsigned long long km = 0, ki = ~0ull;
  if (pid < npits && portal_node(bi, d, pits.cell[pid]) == ID_NA) {
    km = pits.rootmax[pid];
    const uint32_t ci = pits.rootidx[pid];
    ki = ci == 0xFFFFFFFFu ? ~0ull : gseq(bi, d, ci);
  }
  kmax[pid] = km;
  kidx[pid] = ki;
}

// ---- the boundary graph (2 * nbands * NC nodes), solved identically by every band ----------------
struct Graph {
  uint32_t* nxt;    // original edge
  uint32_t* jres;   // resolved
  uint32_t* X;      // flow arriving at the entry cell from the other band (starts as L, accumulated in place)
  uint32_t* indeg;
  uint8_t* src;
  unsigned long long* pkmax;
  unsigned long long* pkidx;
};
__global__ void k_bg_build(const char* __restrict__ records, size_t rec_bytes, int P, int NC, Graph G) {
  const size_t n = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= (size_t)2 * P * NC) return;
  const int c = (int)(n % NC), bs = (int)(n / NC), b = bs >> 1, s = bs & 1;
  const Record mine = record_at((void*)(records + (size_t)b * rec_bytes), NC);
  const uint32_t v = mine.nxt[(size_t)s * NC + c];
  G.nxt[n] = v;
  G.jres[n] = v;
  const int nb = s == 0 ? b - 1 : b + 1;  // the band whose portal row is the image of this entry row
  uint32_t L = 0;
  unsigned long long km = 0, ki = ~0ull;
  if (nb >= 0 && nb < P) {
    const Record o = record_at((void*)(records + (size_t)nb * rec_bytes), NC);
    const size_t k = (size_t)(1 - s) * NC + c;
    L = o.L[k];
    km = o.pkmax[k];
    ki = o.pkidx[k];
  }
  G.X[n] = L;
  G.indeg[n] = 0;
  G.pkmax[n] = km;
  G.pkidx[n] = ki;
}
__global__ void k_bg_resolve(uint32_t* __restrict__ jres, size_t n, unsigned* unresolved) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool open = false;
  if (s < n) {
    uint32_t v = jres[s];
    if (is_cell(v)) {
      for (int hop = 0; hop < 8 && is_cell(v); ++hop) v = jres[v];
      jres[s] = v;
      open = is_cell(v);
    }
  }
  if (__syncthreads_or(open) && threadIdx.x == 0) atomicAdd(unresolved, 1u);
}
__global__ void k_bg_close(uint32_t* __restrict__ jres, size_t n) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n && is_cell(jres[s])) jres[s] = ID_NA;
}
__global__ void k_bg_indeg(Graph G, size_t n) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n && is_cell(G.nxt[s])) atomicAdd(&G.indeg[G.nxt[s]], 1u);
}
__global__ void k_bg_src(Graph G, size_t n) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) G.src[s] = G.indeg[s] == 0;
}
__global__ void k_bg_chase(Graph G, size_t n) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n || !G.src[s]) return;
  uint32_t carry = *reinterpret_cast<volatile uint32_t*>(&G.X[s]);
  while (true) {
    const uint32_t t = G.nxt[s];
    if (!is_cell(t)) break;
    if (carry) atomicAdd(&G.X[t], carry);
    __threadfence();
    if (atomicSub(&G.indeg[t], 1u) != 1u) break;
    __threadfence();
    carry = atomicAdd(&G.X[t], 0u);
    s = t;
  }
}
// merge the partial keys that travel through portals into the key of the pit they end in
__global__ void k_bg_keys_max(Graph G, size_t n, unsigned long long* __restrict__ kfin) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const unsigned long long pk = G.pkmax[s];
  const uint32_t root = G.jres[s];
  if (pk && root != ID_NA && (root & ID_ROOT)) atomicMax(&kfin[root & ~ID_ROOT], pk);
}
__global__ void k_bg_keys_own(const unsigned long long* __restrict__ kmax, const unsigned long long* __restrict__ kfin,
                              unsigned long long* __restrict__ kidx, size_t m) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m && kmax[i] != kfin[i]) kidx[i] = ~0ull;  // the shed's highest cell lies in another band
}
__global__ void k_bg_keys_idx(Graph G, size_t n, const unsigned long long* __restrict__ kfin,
                              unsigned long long* __restrict__ kidx) {
  size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const unsigned long long pk = G.pkmax[s];
  const uint32_t root = G.jres[s];
  if (pk && root != ID_NA && (root & ID_ROOT) && pk == kfin[root & ~ID_ROOT]) atomicMin(&kidx[root & ~ID_ROOT], G.pkidx[s]);
}
__global__ void k_iota32(uint32_t* a, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = (uint32_t)i;
}
__global__ void k_gather64(const unsigned long long* __restrict__ src, const uint32_t* __restrict__ idx,
                           unsigned long long* __restrict__ dst, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[idx[i]];
}
// lists = P sorted key lists of `stride` entries each, list b = [kfin desc (0 = empty slot, last) | the seqno keys in
// the same order].  Number of entries of list b that come before the key (f, x): kfin greater, or equal with a lower seqno
__device__ __forceinline__ unsigned keys_before(const unsigned long long* __restrict__ lists, int b, unsigned stride,
                                                unsigned long long f, unsigned long long x) {
  const unsigned long long* kf = lists + (size_t)b * 2 * stride;
  const unsigned long long* kx = kf + stride;
  unsigned lo = 0, hi = stride;
  while (lo < hi) {
    const unsigned mid = (lo + hi) >> 1;
    const unsigned long long mf = kf[mid];
    const bool before = mf > f || (mf == f && kx[mid] < x);
    if (before) lo = mid + 1; else hi = mid;
  }
  return lo;
}
// global rank (1-based, 0 = empty slot) of the band's own slots: place in the own sorted list + keys before it elsewhere
__global__ void k_bg_rank_own(const unsigned long long* __restrict__ lists, int P, int own, unsigned stride,
                              const uint32_t* __restrict__ order, uint32_t* __restrict__ rank_own) {
  const unsigned j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= stride) return;
  const unsigned long long* mf = lists + (size_t)own * 2 * stride;
  const unsigned long long f = mf[j], x = mf[stride + j];
  uint32_t r = 0;
  if (f != 0) {
    r = j + 1;
    for (int b = 0; b < P; ++b)
      if (b != own) r += keys_before(lists, b, stride, f, x);
  }
  rank_own[order[j]] = r;
}
// pits of the whole DEM = non-empty entries of all lists
__global__ void k_bg_count_real(const unsigned long long* __restrict__ lists, int P, unsigned stride, unsigned long long* count) {
  for (int b = threadIdx.x; b < P; b += blockDim.x) {
    const unsigned long long* kf = lists + (size_t)b * 2 * stride;
    unsigned lo = 0, hi = stride;
    while (lo < hi) {
      const unsigned mid = (lo + hi) >> 1;
      if (kf[mid] != 0) lo = mid + 1; else hi = mid;
    }
    atomicAdd(count, (unsigned long long)lo);
  }
}
// final shed number of every local pit id: real pits look up their global rank, portals the rank of the
// pit their flow ends in; and the seeds (flow entering at the entry cells of this band)
__global__ void k_band_fill_rank(BandInfo bi, GridDims d, const Pits pits, unsigned npits, const uint32_t* __restrict__ grank,
                                 const uint32_t* __restrict__ jres) {
  unsigned pid = blockIdx.x * blockDim.x + threadIdx.x;
  if (pid >= npits) return;
  const uint32_t pn = portal_node(bi, d, pits.cell[pid]);
  uint32_t rk;
  if (pn == ID_NA) rk = grank[(size_t)bi.band * bi.pit_stride + pid];
  else {
    const uint32_t v = jres[pn];
    rk = (v != ID_NA && (v & ID_ROOT)) ? grank[v & ~ID_ROOT] : 0;
  }
  pits.rank[pid] = rk;
}
// add the flow X entering at an entry cell to the external inflow of every tile-border cell on its way
__global__ void k_band_seed_walk(TileGeom g, BandInfo bi, const uint8_t* __restrict__ ddir, const L2 l2, size_t nslots,
                                 const uint32_t* __restrict__ X) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int side = blockIdx.y;
  if (c < 1 || c >= g.d.NC) return;
  const uint32_t x = X[node_of(bi.band, side, c, g.d.NC)];
  if (!x) return;
  int r = side == 0 ? bi.own_lo : bi.own_hi - 1, cc = c;
  const int ty = (r - 1) / TH, tx = (cc - 1) / TW;
  long long s = -1;
  for (int step = 0; step <= LC; ++step) {  // inside the entry cell's own tile: k_shed_final adds x itself
    const int dv = ddir[(size_t)r * g.d.pitch + cc];
    if (dv == 0 || dv == 5 || dv > 9) return;
    const int rr = r + nb_dr(dv), c2 = cc + nb_dc(dv);
    if (rr < 1 || c2 < 1 || rr >= g.d.NR || c2 >= g.d.NC) return;
    if ((rr - 1) / TH != ty || (c2 - 1) / TW != tx) { s = slot_of_rc(rr, c2, g.tilesX); break; }
    r = rr;
    cc = c2;
  }
  for (size_t hop = 0; s >= 0 && hop < nslots; ++hop) {
    atomicAdd(&l2.xi[s], (unsigned long long)x << 32);
    const uint32_t v = l2.jump0[s];
    if (!is_cell(v)) return;
    s = slot_of_id(v, g);
  }
}
__global__ void k_band_seeds_extract(BandInfo bi, int NC, const uint32_t* __restrict__ X, uint32_t* __restrict__ x_top,
                                     uint32_t* __restrict__ x_bot) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= NC) return;
  x_top[c] = X[node_of(bi.band, 0, c, NC)];
  x_bot[c] = X[node_of(bi.band, 1, c, NC)];
}
// the NA ring row of the DEM sits inside the tile grid of the first band: restore its NA outputs
__global__ void k_band_ring_row(GridDims d, int row, int32_t* __restrict__ shed_out, uint32_t* __restrict__ upslope,
                                uint8_t* __restrict__ ridge) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.NC) return;
  const size_t o = (size_t)row * d.pitch + c;
  shed_out[o] = 0;
  upslope[o] = 0;
  ridge[o] = 2;
}
}  // namespace bg

// cells with a shed number and the upslope counts of the pits (ddir 5) among them, rows [row_lo, row_hi): every cell drains
// into exactly one pit, so over a whole DEM the two numbers are equal -- unless a 32-bit count wrapped
__global__ void k_conservation(const int32_t* __restrict__ shed, const uint8_t* __restrict__ ddir, const uint32_t* __restrict__ up,
                               GridDims d, int row_lo, int row_hi, unsigned long long* __restrict__ out) {
  unsigned long long cells = 0, pits = 0;
  for (int r = row_lo + blockIdx.y; r < row_hi; r += gridDim.y)
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < d.NC; c += gridDim.x * blockDim.x) {
      const size_t id = (size_t)r * d.pitch + c;
      if (shed[id] > 0) {
        ++cells;
        if (ddir[id] == 5) pits += up[id];
      }
    }
  for (int o = 16; o; o >>= 1) {
    cells += __shfl_xor_sync(0xffffffffu, cells, o);
    pits += __shfl_xor_sync(0xffffffffu, pits, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (cells) atomicAdd(&out[0], cells);
    if (pits) atomicAdd(&out[1], pits);
  }
}
static int conservation_enqueue(litap_grid* g, unsigned long long* out2) {
  const int lo = g->nbands ? g->own_lo : 1, hi = g->nbands ? g->own_hi : g->d.NR - 1;
  CUDA_TRY(cudaMemsetAsync(out2, 0, 16, ctx().stream));
  dim3 grid(std::min((g->d.NC + 255) / 256, 32), std::min(std::max(hi - lo, 1), 2048));
  KERNEL("k_conservation", k_conservation<<<grid, 256, 0, ctx().stream>>>(g->shed, g->ddir, g->upslope, g->d, lo, hi, out2));
  return LITAP_OK;
}
// (cells with a shed, sum of upslope over the pits) of a grid or of a band's own rows
extern "C" int litap_grid_conservation(litap_grid* g, uint64_t* cells, uint64_t* pit_upslope) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || !cells || !pit_upslope) return set_error(LITAP_EARG, "litap_grid_conservation: NULL");
  if (!g->shed || !g->ddir || !g->upslope) return set_error(LITAP_EARG, "litap_grid_conservation: needs ddir, shed and upslope on the grid");
  PoolScratch S;
  unsigned long long* o = S.get<unsigned long long>(2);
  if (!o) return set_error(LITAP_ENOMEM, "device allocation failed (conservation)");
  int rc = conservation_enqueue(g, o);
  if (rc) return rc;
  unsigned long long h[2];
  CUDA_TRY(cudaMemcpyAsync(h, o, 16, cudaMemcpyDeviceToHost, ctx().stream));
  STREAM_SYNC(ctx().stream);
  *cells = h[0];
  *pit_upslope = h[1];
  return LITAP_OK;
}

extern "C" int litap_bands_shed(litap_grid** B, int n) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  using namespace bg;
  int rc = check_bands(B, n);
  if (rc) return rc;
  Ctx& c = ctx();
  cudaStream_t st = c.stream;
  const int P = B[0]->nbands, NC = B[0]->d.NC;
  const int64_t l0 = c.launches;
  cudaEventRecord(c.ev[2], st);
  auto blk = [](size_t m) { return (unsigned)((m + 255) / 256); };
  const unsigned cblk = blk(NC);

  // ---- phase A: every band on its own, halo rows as portals
  std::vector<ShedWork> W(n);
  for (int i = 0; i < n; ++i) {
    litap_grid* g = B[i];
    if (!g->ddir) return set_error(LITAP_EARG, "litap_bands_shed: no ddir on a band (run litap_bands_ddir first)");
    KERNEL("k_band_portals", k_band_portals<<<cblk, 256, 0, st>>>(g->ddir, g->elev, g->d, g->halo_top, g->halo_bot));
    if ((rc = shed_phase_local(g, W[i]))) return rc;
  }
  // slots per band in the global pit table: the largest pit count of any band -- or, when calc_ddir2 ran on these
  // bands, its bound on that (flat cells of the D8 stencil + portals), which saves a gather + host round trip
  unsigned long long stride64 = 1;
  bool bounded = true;
  for (int i = 0; i < n; ++i) bounded = bounded && B[i]->pit_bound_all > 0;   // the same answer on every rank
  if (bounded) {
    stride64 = (unsigned long long)B[0]->pit_bound_all;
  } else {
    std::vector<std::vector<unsigned long long>> mine(n, std::vector<unsigned long long>(1));
    for (int i = 0; i < n; ++i) mine[i][0] = W[i].npits;
    std::vector<unsigned long long> np_all;
    if ((rc = gather_words_host(B, n, mine, 1, np_all))) return rc;
    for (int b = 0; b < P; ++b) stride64 = std::max(stride64, np_all[b]);
  }
  if (stride64 * (unsigned long long)P >= 0x7FFFFFFFull)
    return set_error(LITAP_EARG, "litap_bands_shed: too many pits for 31-bit pit numbers");
  const unsigned stride = (unsigned)stride64;
  const size_t M = (size_t)stride * P;          // global pit slots: band * stride + pid
  const size_t Nn = (size_t)2 * P * NC;         // boundary-graph nodes
  const size_t rec = record_bytes(NC);

  PoolScratch S;
  std::vector<void*> rsend(n), rrecv(n), ksend(n), krecv(n), isend(n), irecv(n);
  std::vector<uint32_t*> tot(n);
  std::vector<BandInfo> BI(n);
  for (int i = 0; i < n; ++i) {
    litap_grid* g = B[i];
    BandInfo& bi = BI[i];
    bi.band = g->band; bi.nbands = P; bi.own_lo = g->own_lo; bi.own_hi = g->own_hi;
    bi.halo_top = g->halo_top; bi.halo_bot = g->halo_bot; bi.row_off = g->row_off; bi.pit_stride = stride;
    rsend[i] = S.get<unsigned char>(rec);
    rrecv[i] = S.get<unsigned char>(rec * P);
    ksend[i] = S.get<unsigned long long>(stride);
    krecv[i] = S.get<unsigned long long>(M);
    isend[i] = S.get<unsigned long long>(stride);
    irecv[i] = S.get<unsigned long long>(M);
    tot[i] = S.get<uint32_t>(W[i].npits ? W[i].npits : 1);
    if (!rsend[i] || !rrecv[i] || !ksend[i] || !krecv[i] || !isend[i] || !irecv[i] || !tot[i])
      return set_error(LITAP_ENOMEM, "device allocation failed (band tables)");
    if (W[i].npits) {
      CUDA_TRY(cudaMemcpyAsync(tot[i], W[i].pits.cnt, (size_t)W[i].npits * 4, cudaMemcpyDeviceToDevice, st));
      KERNEL("k_pit_totals", k_pit_totals<<<blk(W[i].nslots), 256, 0, st>>>(W[i].l2, W[i].nslots, tot[i]));
    }
    KERNEL("k_band_tables", k_band_tables<<<dim3(cblk, 2), 256, 0, st>>>(W[i].tg, bi, g->ddir, reinterpret_cast<uint32_t*>(g->shed),
                                                                       W[i].l2, W[i].pits, tot[i], record_at(rsend[i], NC)));
    KERNEL("k_band_pitkeys", k_band_pitkeys<<<blk(stride), 256, 0, st>>>(bi, g->d, W[i].pits, W[i].npits,
                                                                       (unsigned long long*)ksend[i],
                                                                       (unsigned long long*)isend[i]));
  }
  // ---- exchange: 48 B per column and band, 16 B per pit
  if ((rc = all_gather(B, n, rsend.data(), rrecv.data(), rec))) return rc;
  if ((rc = all_gather(B, n, ksend.data(), krecv.data(), (size_t)stride * 8))) return rc;
  if ((rc = all_gather(B, n, isend.data(), irecv.data(), (size_t)stride * 8))) return rc;

  // ---- phase B: the same boundary graph on every band; the global pit ranking is shared out: a band sorts the keys
  // of ITS pits, the sorted lists are all-gathered, and a pit's rank is its place in its own list plus, by binary
  // search, the number of keys before it in every other list (round 1: every band sorted all M keys, 1.6 ms of an
  // 8-band step at 32768^2)
  unsigned long long npits_global = 0;
  std::vector<uint32_t*> xt(n), xb(n);
  struct PhaseB {
    Graph G;
    unsigned long long *kfin, *cnt;
    uint32_t *ord, *grank;
  };
  std::vector<PhaseB> Q(n);
  std::vector<void*> lsend(n), lrecv(n), gsend(n), grecv(n);
  int idx_bits = 1;   // seqnos of the whole DEM need few bits (one more keeps the all-ones "none" last)
  while (idx_bits < 63 && (((unsigned long long)(B[0]->gnr + 2) * (unsigned long long)NC) >> idx_bits) != 0) ++idx_bits;
  ++idx_bits;
  for (int i = 0; i < n; ++i) {
    litap_grid* g = B[i];
    Graph& G = Q[i].G;
    G.nxt = S.get<uint32_t>(Nn);
    G.jres = S.get<uint32_t>(Nn);
    G.X = S.get<uint32_t>(Nn);
    G.indeg = S.get<uint32_t>(Nn);
    G.src = S.get<uint8_t>(Nn);
    G.pkmax = S.get<unsigned long long>(Nn);
    G.pkidx = S.get<unsigned long long>(Nn);
    unsigned long long* kmax = (unsigned long long*)krecv[i];
    unsigned long long* kidx = (unsigned long long*)irecv[i];
    unsigned long long* kfin = Q[i].kfin = S.get<unsigned long long>(M);
    unsigned long long* k64b = S.get<unsigned long long>(stride);
    unsigned long long* k64c = S.get<unsigned long long>(stride);
    uint32_t* ord_a = Q[i].ord = S.get<uint32_t>(stride);
    uint32_t* ord_b = S.get<uint32_t>(stride);
    Q[i].grank = S.get<uint32_t>(M);
    unsigned long long* cnt = Q[i].cnt = S.get<unsigned long long>(2);
    lsend[i] = S.get<unsigned long long>((size_t)2 * stride);
    lrecv[i] = S.get<unsigned long long>((size_t)2 * M);
    gsend[i] = S.get<uint32_t>(stride);
    grecv[i] = Q[i].grank;
    xt[i] = S.get<uint32_t>(NC);
    xb[i] = S.get<uint32_t>(NC);
    size_t t1 = 0, t2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, t1, kidx, k64b, ord_a, ord_b, (int)stride, 0, 64, st);
    cub::DeviceRadixSort::SortPairsDescending(nullptr, t2, k64b, k64c, ord_b, ord_a, (int)stride, 0, 64, st);
    size_t tn = std::max(t1, t2) + 256;
    void* tmp = S.get<unsigned char>(tn);
    if (!G.nxt || !G.jres || !G.X || !G.indeg || !G.src || !G.pkmax || !G.pkidx || !kfin || !k64b || !k64c || !ord_a ||
        !ord_b || !Q[i].grank || !cnt || !lsend[i] || !lrecv[i] || !gsend[i] || !xt[i] || !xb[i] || !tmp)
      return set_error(LITAP_ENOMEM, "device allocation failed (boundary graph)");
    CUDA_TRY(cudaMemsetAsync(cnt, 0, 16, st));
    KERNEL("k_bg_build", k_bg_build<<<blk(Nn), 256, 0, st>>>((const char*)rrecv[i], rec, P, NC, G));
    unsigned* flag = reinterpret_cast<unsigned*>(cnt + 1);
    bool open = true;
    for (int round = 0; round < 64 && open; round += 4) {   // four doubling rounds per look at the flag (a round on a resolved graph is a no-op)
      for (int k = 0; k < 4; ++k) {
        if (k == 3) CUDA_TRY(cudaMemsetAsync(flag, 0, 4, st));
        KERNEL("k_bg_resolve", k_bg_resolve<<<blk(Nn), 256, 0, st>>>(G.jres, Nn, flag));
      }
      unsigned h = 0;
      CUDA_TRY(cudaMemcpyAsync(&h, flag, 4, cudaMemcpyDeviceToHost, st));
      STREAM_SYNC(st);
      open = h != 0;
    }
    if (open) KERNEL("k_bg_close", k_bg_close<<<blk(Nn), 256, 0, st>>>(G.jres, Nn));
    KERNEL("k_bg_indeg", k_bg_indeg<<<blk(Nn), 256, 0, st>>>(G, Nn));
    KERNEL("k_bg_src", k_bg_src<<<blk(Nn), 256, 0, st>>>(G, Nn));
    KERNEL("k_bg_chase", k_bg_chase<<<blk(Nn), 256, 0, st>>>(G, Nn));
    // keys: the highest cell of a shed may lie in any band its flow crosses
    CUDA_TRY(cudaMemcpyAsync(kfin, kmax, M * 8, cudaMemcpyDeviceToDevice, st));
    KERNEL("k_bg_keys_max", k_bg_keys_max<<<blk(Nn), 256, 0, st>>>(G, Nn, kfin));
    KERNEL("k_bg_keys_own", k_bg_keys_own<<<blk(M), 256, 0, st>>>(kmax, kfin, kidx, M));
    KERNEL("k_bg_keys_idx", k_bg_keys_idx<<<blk(Nn), 256, 0, st>>>(G, Nn, kfin, kidx));
    // canonical numbering (R/flow_02_calc_shed.R:32-38) of this band's slots: stable LSD, seqno asc then max elevation desc
    const size_t base = (size_t)g->band * stride;
    unsigned long long* ls = (unsigned long long*)lsend[i];
    KERNEL("k_iota32", k_iota32<<<blk(stride), 256, 0, st>>>(ord_a, stride));
    size_t tcap = tn;
    { ProfScope _p("cub_sort_pits_idx"); CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tcap, kidx + base, k64b, ord_a, ord_b, (int)stride, 0, idx_bits, st)); }
    KERNEL("k_gather64", k_gather64<<<blk(stride), 256, 0, st>>>(kfin + base, ord_b, k64b, stride));
    tcap = tn;
    { ProfScope _p("cub_sort_pits_key"); CUDA_TRY(cub::DeviceRadixSort::SortPairsDescending(tmp, tcap, k64b, ls, ord_b, ord_a, (int)stride, 0, 64, st)); }
    KERNEL("k_gather64", k_gather64<<<blk(stride), 256, 0, st>>>(kidx + base, ord_a, ls + stride, stride));
    c.launches += 4;
  }
  if ((rc = all_gather(B, n, lsend.data(), lrecv.data(), (size_t)2 * stride * 8))) return rc;
  for (int i = 0; i < n; ++i) {
    KERNEL("k_bg_rank_own", k_bg_rank_own<<<blk(stride), 256, 0, st>>>((const unsigned long long*)lrecv[i], P, B[i]->band, stride,
                                                                     Q[i].ord, (uint32_t*)gsend[i]));
    KERNEL("k_bg_count_real", k_bg_count_real<<<1, 32, 0, st>>>((const unsigned long long*)lrecv[i], P, stride, Q[i].cnt));
  }
  if ((rc = all_gather(B, n, gsend.data(), grecv.data(), (size_t)stride * 4))) return rc;
  for (int i = 0; i < n; ++i) {
    litap_grid* g = B[i];
    Graph& G = Q[i].G;
    if (W[i].npits)
      KERNEL("k_band_fill_rank", k_band_fill_rank<<<blk(W[i].npits), 256, 0, st>>>(BI[i], g->d, W[i].pits, W[i].npits, Q[i].grank, G.jres));
    // flow entering this band
    KERNEL("k_band_seeds_extract", k_band_seeds_extract<<<cblk, 256, 0, st>>>(BI[i], NC, G.X, xt[i], xb[i]));
    KERNEL("k_band_seed_walk", k_band_seed_walk<<<dim3(cblk, 2), 256, 0, st>>>(W[i].tg, BI[i], g->ddir, W[i].l2, W[i].nslots, G.X));
    CUDA_TRY(cudaMemcpyAsync(&npits_global, Q[i].cnt, 8, cudaMemcpyDeviceToHost, st));
    // ---- phase C: labels, upslope, ridge
    BandSeeds seeds;
    seeds.x_top = g->halo_top ? xt[i] : nullptr;
    seeds.x_bot = g->halo_bot ? xb[i] : nullptr;
    seeds.row_top = g->own_lo;
    seeds.row_bot = g->own_hi - 1;
    if ((rc = shed_phase_final(g, W[i], seeds))) return rc;
    if (!g->halo_top)
      KERNEL("k_band_ring_row", k_band_ring_row<<<cblk, 256, 0, st>>>(g->d, g->own_lo - 1, g->shed, g->upslope, g->ridge));
  }
  cudaEventRecord(c.ev[3], st);
  STREAM_SYNC(st);
  cudaEventElapsedTime(&c.tm.shed_ms, c.ev[2], c.ev[3]);
  // upslope is a 32-bit count (include/litap.h): a DEM of 2^32 or more cells (65536 x 65536 is exactly 2^32) could hold a
  // basin whose count wraps.  A wrapped count also wraps at the basin's pit, and then the pit counts no longer add up to
  // the number of cells: detect that and fail instead of returning wrapped numbers.
  if ((double)B[0]->gnr * (double)(NC - 2) >= 4294967296.0) {
    std::vector<std::vector<unsigned long long>> cons(n, std::vector<unsigned long long>(2));
    for (int i = 0; i < n; ++i) {
      uint64_t a = 0, b = 0;
      if ((rc = litap_grid_conservation(B[i], &a, &b))) return rc;
      cons[i][0] = a; cons[i][1] = b;
    }
    std::vector<unsigned long long> all;
    if ((rc = gather_words_host(B, n, cons, 2, all))) return rc;
    unsigned long long cells = 0, pits = 0;
    for (int b = 0; b < P; ++b) { cells += all[2 * b]; pits += all[2 * b + 1]; }
    if (cells != pits)
      return set_error(LITAP_EARG, "litap_bands_shed: a basin of this DEM has 2^32 or more cells: upslope (32-bit) would wrap");
  }
  for (int i = 0; i < n; ++i) {
    B[i]->npits = (int64_t)npits_global;
    B[i]->have_shed = true;
  }
  c.tm.npits = (int64_t)npits_global;
  c.tm.kernel_launches = c.launches - l0;
  return LITAP_OK;
}

// rows of the full buffered DEM this band is responsible for: its own rows, plus the NA ring row of
// the DEM for the first / last band.  dst receives those rows densely (NC wide); *row0 / *nrows
// (optional) tell where they go in the (gnr+2)-row buffered DEM.
extern "C" int litap_band_rows(const litap_grid* g, int64_t* row0, int64_t* nrows) {
  if (!g || g->nbands < 1) return set_error(LITAP_EARG, "litap_band_rows: not a band grid");
  const int lo = g->halo_top ? g->own_lo : g->own_lo - 1, hi = g->halo_bot ? g->own_hi : g->own_hi + 1;
  if (row0) *row0 = g->row_off + lo;
  if (nrows) *nrows = hi - lo;
  return LITAP_OK;
}
