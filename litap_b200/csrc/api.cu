// api.cu — C ABI of liblitap_cuda.so: context, device-memory pool, host-buffer and
// device-resident entry points.  See include/litap.h for the contract.
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include "litap_internal.h"

// ------------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static std::recursive_mutex g_api_mutex;
ApiLock::ApiLock() { g_api_mutex.lock(); }
ApiLock::~ApiLock() { g_api_mutex.unlock(); }
static Ctx g_ctx;
static bool g_own_stream = true;  // false after litap_set_stream(non-NULL)
static cudaStream_t g_hi_stream = nullptr;  // litap_grid_flow_form: the flow stages beside calc_form
static cudaEvent_t g_ev_fork = nullptr, g_ev_join = nullptr;
Ctx& ctx() { return g_ctx; }

int set_error(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
extern "C" const char* litap_last_error(void) { return g_err.c_str(); }
extern "C" const char* litap_version(void) { return "litap_b200 0.1 (sm_100a)"; }

extern "C" int litap_init(int device) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  Ctx& c = ctx();
  if (c.ready && (device < 0 || device == c.device)) return LITAP_OK;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return set_error(LITAP_ECUDA, std::string("no CUDA device (there is no CPU fallback): ") +
                                      (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0"));
  if (device < 0) {
    const char* s = getenv("LITAP_CUDA_DEVICE");
    if (!s) s = getenv("LOCAL_RANK");
    device = s ? atoi(s) % n : 0;
  }
  if (device >= n) return set_error(LITAP_EARG, "litap_init: device index out of range");
  if (c.ready) litap_shutdown();
  CUDA_TRY(cudaSetDevice(device));
  cudaDeviceProp p;
  CUDA_TRY(cudaGetDeviceProperties(&p, device));
  if (p.major < 10)
    return set_error(LITAP_ECUDA, std::string("device ") + p.name + " is not sm_100-class; this library targets B200 only");
  c.device = device;
  c.sm_count = p.multiProcessorCount;
  CUDA_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
  for (auto& ev : c.ev) CUDA_TRY(cudaEventCreate(&ev));
  c.ready = true;
  return LITAP_OK;
}

int ensure_init() { return ctx().ready ? LITAP_OK : litap_init(-1); }

// ------------------------------------------------------------------------------------------------
// pool: exact-ish size reuse; DEM-sized buffers recur with identical sizes call after call
struct Block { void* p; size_t bytes; bool used; };
static std::vector<Block> g_pool;

void* pool_alloc(size_t bytes) {
  if (bytes == 0) bytes = 256;
  int best = -1;
  for (size_t i = 0; i < g_pool.size(); ++i)
    if (!g_pool[i].used && g_pool[i].bytes >= bytes && g_pool[i].bytes <= bytes + bytes / 4 + 4096)
      if (best < 0 || g_pool[i].bytes < g_pool[best].bytes) best = (int)i;
  if (best >= 0) { g_pool[best].used = true; return g_pool[best].p; }
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) {
    cudaGetLastError();
    // drop unused cached blocks and retry once
    for (size_t i = 0; i < g_pool.size();)
      if (!g_pool[i].used) { cudaFree(g_pool[i].p); g_pool.erase(g_pool.begin() + i); } else ++i;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  }
  g_pool.push_back({p, bytes, true});
  return p;
}
void pool_free(void* p) {
  for (auto& b : g_pool)
    if (b.p == p) { b.used = false; return; }
}
// give unused cached blocks back to the driver (between phases with very different working sets)
extern "C" void litap_pool_trim(void) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (ctx().ready) cudaStreamSynchronize(ctx().stream);
  for (size_t i = 0; i < g_pool.size();)
    if (!g_pool[i].used) { cudaFree(g_pool[i].p); g_pool.erase(g_pool.begin() + i); } else ++i;
}
void pool_release_all() {
  for (auto& b : g_pool) cudaFree(b.p);
  g_pool.clear();
}

extern "C" void litap_shutdown(void) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  Ctx& c = ctx();
  if (!c.ready) return;
  cudaStreamSynchronize(c.stream);
  pool_release_all();
  form_forget_trig();
  hostcopy_forget_staging();
  rasterio_forget();
  for (auto& ev : c.ev) if (ev) { cudaEventDestroy(ev); ev = nullptr; }
  if (g_own_stream) cudaStreamDestroy(c.stream);
  if (g_hi_stream) { cudaStreamDestroy(g_hi_stream); cudaEventDestroy(g_ev_fork); cudaEventDestroy(g_ev_join); g_hi_stream = nullptr; }
  g_own_stream = true;
  c.stream = nullptr;
  c.ready = false;
}

extern "C" int litap_get_timings(litap_timings* out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!out) return set_error(LITAP_EARG, "litap_get_timings: NULL");
  *out = ctx().tm;
  return LITAP_OK;
}

// ------------------------------------------------------------------------------------------------
extern "C" int litap_grid_create(int64_t NR, int64_t NC, litap_grid** out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!out) return set_error(LITAP_EARG, "litap_grid_create: NULL out");
  *out = nullptr;
  if (NR < 3 || NC < 3) return set_error(LITAP_EARG, "grid must be at least 3 x 3 buffered cells (1 x 1 DEM + NA ring)");
  GridDims d = make_dims(NR, NC);
  if ((double)NR * d.pitch >= 2147483648.0 || NR > 65535)   // cell ids are 31-bit; rows are the y dimension of most launches
    return set_error(LITAP_EARG, "grid too large for one device: NR*pitch must be < 2^31 cells and NR <= 65535; shard into row bands");
  int rc = ensure_init();
  if (rc) return rc;
  litap_grid* g = new litap_grid();
  g->d = d;
  g->own_lo = 0;
  g->own_hi = d.NR;
  g->edge_top_row = 1;
  g->edge_bot_row = d.NR - 2;
  g->gnr = d.NR - 2;
  g->elev = grid_alloc_array<double>(d, 0xFF);  // pads = NaN
  if (!g->elev) { delete g; return set_error(LITAP_ENOMEM, "device allocation failed (elev)"); }
  *out = g;
  return LITAP_OK;
}

extern "C" void litap_grid_destroy(litap_grid* g) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return;
  if (ctx().ready) cudaStreamSynchronize(ctx().stream);
  grid_free_array(g->elev);
  grid_free_array(g->ddir);
  grid_free_array(g->flat);
  grid_free_array(g->shed);
  grid_free_array(g->upslope);
  grid_free_array(g->ridge);
  for (auto& p : g->form) grid_free_array(p);
  for (auto& p : g->weti) grid_free_array(p);
  grid_free_array(g->uced);
  for (auto& p : g->gc_f64) grid_free_array(p);
  for (auto& p : g->gc_u8) grid_free_array(p);
  for (auto& p : g->gc_i32) grid_free_array(p);
  if (g->gc_sum) { pool_free(g->gc_sum); g->gc_sum = nullptr; }
  if (g->dd_list) pool_free(g->dd_list);
  if (g->dd_cand) pool_free(g->dd_cand);
  if (g->dd_keys) pool_free(g->dd_keys);
  if (g->dd_ctr) pool_free(g->dd_ctr);
  delete g;
}

template <typename T>
static int upload2d(T* dst, const GridDims& d, const T* src) {
  CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)d.pitch * sizeof(T), src, (size_t)d.NC * sizeof(T),
                             (size_t)d.NC * sizeof(T), d.NR, cudaMemcpyHostToDevice, ctx().stream));
  return LITAP_OK;
}
template <typename T>
static int download2d(T* dst, const GridDims& d, const T* src) {
  CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)d.NC * sizeof(T), src, (size_t)d.pitch * sizeof(T),
                             (size_t)d.NC * sizeof(T), d.NR, cudaMemcpyDeviceToHost, ctx().stream));
  return LITAP_OK;
}

extern "C" int litap_grid_upload_elev(litap_grid* g, const double* elev) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || !elev) return set_error(LITAP_EARG, "litap_grid_upload_elev: NULL");
  int rc = upload2d(g->elev, g->d, elev);
  if (rc) return rc;
  g->have_ddir = g->have_shed = false;
  CUDA_TRY(cudaStreamSynchronize(ctx().stream));
  return LITAP_OK;
}

extern "C" int litap_grid_synth(litap_grid* g, int64_t row0, int64_t gnr, uint32_t seed, double quant, int na_permille) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_synth: NULL");
  g->have_ddir = g->have_shed = false;
  int rc = run_synth(g, row0, gnr, seed, quant, na_permille);
  if (rc) return rc;
  CUDA_TRY(cudaStreamSynchronize(ctx().stream));
  return LITAP_OK;
}

struct StageTimer {
  float* slot;
  cudaEvent_t a, b;
  StageTimer(float* s, int k) : slot(s), a(ctx().ev[k]), b(ctx().ev[k + 1]) { cudaEventRecord(a, ctx().stream); }
  int finish() {
    cudaEventRecord(b, ctx().stream);
    cudaError_t e = cudaEventSynchronize(b);
    if (e != cudaSuccess) return set_error(LITAP_ECUDA, std::string("stage failed: ") + cudaGetErrorString(e));
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    *slot = ms;
    return LITAP_OK;
  }
};

extern "C" int litap_grid_ddir(litap_grid* g, int invert) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_ddir: NULL");
  int64_t l0 = ctx().launches;
  StageTimer t(&ctx().tm.ddir_ms, 0);
  int rc = run_ddir(g, invert);
  if (rc) return rc;
  rc = t.finish();
  ctx().tm.kernel_launches = ctx().launches - l0;
  if (!rc) g->have_ddir = true;
  return rc;
}
extern "C" int litap_grid_shed(litap_grid* g) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_shed: NULL");
  if (!g->ddir) return set_error(LITAP_EARG, "litap_grid_shed: no ddir on the grid (run litap_grid_ddir first)");
  int64_t l0 = ctx().launches;
  StageTimer t(&ctx().tm.shed_ms, 2);
  int rc = run_shed(g);
  if (rc) return rc;
  rc = t.finish();
  ctx().tm.kernel_launches = ctx().launches - l0;
  if (!rc) g->have_shed = true;
  return rc;
}
extern "C" int litap_grid_form(litap_grid* g, double grid_m) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_form: NULL");
  if (!(grid_m > 0)) return set_error(LITAP_EARG, "litap_grid_form: grid must be > 0");
  int64_t l0 = ctx().launches;
  StageTimer t(&ctx().tm.form_ms, 4);
  int rc = run_form(g, grid_m);
  if (rc) return rc;
  rc = t.finish();
  ctx().tm.kernel_launches = ctx().launches - l0;
  return rc;
}
extern "C" int litap_grid_weti(litap_grid* g, double grid_m) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_weti: NULL");
  if (!(grid_m > 0)) return set_error(LITAP_EARG, "litap_grid_weti: grid must be > 0");
  if (g->nbands > 0)   // the visiting order couples all bands: a band on its own would give a band-local (wrong) answer
    return set_error(LITAP_EARG, "litap_grid_weti: not available on a row band (run it on a whole-DEM grid)");
  if (!g->ddir || !g->upslope) return set_error(LITAP_EARG, "litap_grid_weti: needs ddir and upslope on the grid");
  int64_t l0 = ctx().launches;
  StageTimer t(&ctx().tm.weti_ms, 6);
  int rc = run_weti(g, grid_m);
  if (rc) return rc;
  rc = t.finish();
  ctx().tm.kernel_launches = ctx().launches - l0;
  return rc;
}

extern "C" int litap_grid_uced(litap_grid* g) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_uced: NULL");
  if (g->nbands > 0)   // a track crosses band edges; not sharded
    return set_error(LITAP_EARG, "litap_grid_uced: not available on a row band (run it on a whole-DEM grid)");
  if (!g->ddir || !g->shed) return set_error(LITAP_EARG, "litap_grid_uced: needs ddir and shed on the grid");
  int64_t l0 = ctx().launches;
  StageTimer t(&ctx().tm.uced_ms, 6);
  int rc = run_uced(g);
  if (rc) return rc;
  rc = t.finish();
  ctx().tm.kernel_launches = ctx().launches - l0;
  return rc;
}

// calc_form + calc_ddir2 + calc_shed4 of one resident DEM as ONE call: calc_form does not depend on the
// flow stages, so it runs on the caller's stream while ddir / shed run on a high-priority stream of the
// library.  The flow stages are chains of small kernels, host read-backs and one cooperative launch; the
// form blocks (fp64-bound) fill the SMs those leave idle, and priority keeps them out of the way when a
// flow kernel has blocks to place.  The call returns after both streams have finished.
int flow_streams(cudaStream_t* hi, cudaEvent_t* fork, cudaEvent_t* join) {
  if (!g_hi_stream) {
    int lo = 0, top = 0;
    CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &top));
    CUDA_TRY(cudaStreamCreateWithPriority(&g_hi_stream, cudaStreamNonBlocking, top));
    CUDA_TRY(cudaEventCreateWithFlags(&g_ev_fork, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&g_ev_join, cudaEventDisableTiming));
  }
  *hi = g_hi_stream;
  *fork = g_ev_fork;
  *join = g_ev_join;
  return LITAP_OK;
}
extern "C" int litap_grid_flow_form(litap_grid* g, double grid_m, int invert) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_flow_form: NULL");
  if (!(grid_m > 0)) return set_error(LITAP_EARG, "litap_grid_flow_form: grid must be > 0");
  if (g->nbands > 0) return set_error(LITAP_EARG, "litap_grid_flow_form: row bands use litap_bands_*");
  Ctx& c = ctx();
  {
    cudaStream_t hs; cudaEvent_t ef, ej;
    int rc0 = flow_streams(&hs, &ef, &ej);
    if (rc0) return rc0;
  }
  const int64_t l0 = c.launches;
  c.tm.host_syncs = 0;
  cudaStream_t user = c.stream;
  CUDA_TRY(cudaEventRecord(g_ev_fork, user));            // elev (upload / synth) is ordered on the caller's stream
  CUDA_TRY(cudaStreamWaitEvent(g_hi_stream, g_ev_fork, 0));
  // form is queued first and returns at once; the flow kernels then take the SMs as its blocks retire
  CUDA_TRY(cudaEventRecord(c.ev[4], user));
  int rc = run_form(g, grid_m);
  if (rc) return rc;
  CUDA_TRY(cudaEventRecord(c.ev[5], user));
  if (invert) {  // invert() rewrites elev in place: the flow stages then wait for calc_form (no overlap)
    CUDA_TRY(cudaEventRecord(g_ev_fork, user));
    CUDA_TRY(cudaStreamWaitEvent(g_hi_stream, g_ev_fork, 0));
  }
  {
    StreamSwap sw(g_hi_stream);
    CUDA_TRY(cudaEventRecord(c.ev[0], g_hi_stream));
    rc = run_ddir(g, invert);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(c.ev[1], g_hi_stream));
    CUDA_TRY(cudaEventRecord(c.ev[2], g_hi_stream));
    rc = run_shed(g);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(c.ev[3], g_hi_stream));
    CUDA_TRY(cudaEventRecord(g_ev_join, g_hi_stream));
    sw.joined = true;
  }
  CUDA_TRY(cudaStreamWaitEvent(user, g_ev_join, 0));
  CUDA_TRY(cudaStreamSynchronize(user));
  cudaEventElapsedTime(&c.tm.ddir_ms, c.ev[0], c.ev[1]);
  cudaEventElapsedTime(&c.tm.shed_ms, c.ev[2], c.ev[3]);
  cudaEventElapsedTime(&c.tm.form_ms, c.ev[4], c.ev[5]);  // while the flow stages share the device
  g->have_ddir = g->have_shed = true;
  c.tm.npits = g->npits;
  c.tm.kernel_launches = c.launches - l0;
  return LITAP_OK;
}

extern "C" int litap_grid_slope_gc(litap_grid* g, double grid_m) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g) return set_error(LITAP_EARG, "litap_grid_slope_gc: NULL");
  if (!(grid_m > 0)) return set_error(LITAP_EARG, "litap_grid_slope_gc: grid must be > 0");
  if (g->nbands > 0) return set_error(LITAP_EARG, "litap_grid_slope_gc: not available on a row band (columns span all bands)");
  int64_t l0 = ctx().launches;
  StageTimer t(&ctx().tm.gc_ms, 4);
  int rc = run_slope_gc(g, grid_m);
  if (rc) return rc;
  rc = t.finish();
  ctx().tm.kernel_launches = ctx().launches - l0;
  return rc;
}

extern "C" int64_t litap_grid_npits(const litap_grid* g) { return g ? g->npits : -1; }

ArrRef grid_find_array(litap_grid* g, const char* name) { return find_array_impl(g, name); }
static ArrRef find_array(litap_grid* g, const char* name) { return find_array_impl(g, name); }
ArrRef find_array_impl(litap_grid* g, const char* name) {
  static const char* fn[5] = {"slope_pct", "slope_deg", "aspect", "prof", "plan"};
  static const char* wn[4] = {"qarea1", "qarea2", "qweti1", "qweti2"};
  if (!strcmp(name, "elev")) return {g->elev, 8};
  if (!strcmp(name, "ddir")) return {g->ddir, 1};
  if (!strcmp(name, "flatcell")) return {g->flat, 1};
  if (!strcmp(name, "shed")) return {g->shed, 4};
  if (!strcmp(name, "upslope")) return {g->upslope, 4};
  if (!strcmp(name, "ridge")) return {g->ridge, 1};
  for (int i = 0; i < 5; ++i) if (!strcmp(name, fn[i])) return {g->form[i], 8};
  for (int i = 0; i < 4; ++i) if (!strcmp(name, wn[i])) return {g->weti[i], 8};
  if (!strcmp(name, "uced")) return {g->uced, 8};
  static const char* gf[6] = {"sgre", "sgr", "sgcn", "sgc", "scr", "scc"};
  static const char* gu[2] = {"hill_r_dir", "hill_c_dir"};
  static const char* gi[4] = {"hill_r_n", "hill_r_cell", "hill_c_n", "hill_c_cell"};
  for (int i = 0; i < 6; ++i) if (!strcmp(name, gf[i])) return {g->gc_f64[i], 8};
  for (int i = 0; i < 2; ++i) if (!strcmp(name, gu[i])) return {g->gc_u8[i], 1};
  for (int i = 0; i < 4; ++i) if (!strcmp(name, gi[i])) return {g->gc_i32[i], 4};
  return {nullptr, 0};
}

extern "C" int litap_grid_download(litap_grid* g, const char* name, void* dst) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || !name || !dst) return set_error(LITAP_EARG, "litap_grid_download: NULL");
  ArrRef a = find_array(g, name);
  if (!a.p) return set_error(LITAP_EARG, std::string("litap_grid_download: no such result on the grid: ") + name);
  int rc;
  if (a.esize == 8) rc = download2d((double*)dst, g->d, (const double*)a.p);
  else if (a.esize == 4) rc = download2d((uint32_t*)dst, g->d, (const uint32_t*)a.p);
  else rc = download2d((uint8_t*)dst, g->d, (const uint8_t*)a.p);
  if (rc) return rc;
  CUDA_TRY(cudaStreamSynchronize(ctx().stream));
  return LITAP_OK;
}

// a band's share of the full buffered DEM (see litap_band_rows), copied densely into dst
extern "C" int litap_band_download(litap_grid* g, const char* name, void* dst) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || !name || !dst || g->nbands < 1) return set_error(LITAP_EARG, "litap_band_download: bad arguments");
  ArrRef a = find_array(g, name);
  if (!a.p) return set_error(LITAP_EARG, std::string("litap_band_download: no such result on the band: ") + name);
  const int lo = g->halo_top ? g->own_lo : g->own_lo - 1, hi = g->halo_bot ? g->own_hi : g->own_hi + 1;
  const GridDims& d = g->d;
  const char* src = (const char*)a.p + (size_t)lo * d.pitch * a.esize;
  CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)d.NC * a.esize, src, (size_t)d.pitch * a.esize, (size_t)d.NC * a.esize, hi - lo,
                             cudaMemcpyDeviceToHost, ctx().stream));
  CUDA_TRY(cudaStreamSynchronize(ctx().stream));
  return LITAP_OK;
}

extern "C" int litap_grid_checksum(litap_grid* g, const char* name, uint64_t* out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || !name || !out) return set_error(LITAP_EARG, "litap_grid_checksum: NULL");
  ArrRef a = find_array(g, name);
  if (!a.p) return set_error(LITAP_EARG, std::string("litap_grid_checksum: no such result on the grid: ") + name);
  return run_checksum(a.p, a.esize, g->d, out);
}

// a band's share of litap_grid_checksum of the whole DEM: the rows litap_band_rows names, hashed with their cell
// numbers in the whole buffered DEM.  The (wrapping) sum over all bands equals the checksum of the unsharded result.
extern "C" int litap_band_checksum(litap_grid* g, const char* name, uint64_t* out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || !name || !out || g->nbands < 1) return set_error(LITAP_EARG, "litap_band_checksum: bad arguments");
  ArrRef a = find_array(g, name);
  if (!a.p) return set_error(LITAP_EARG, std::string("litap_band_checksum: no such result on the band: ") + name);
  const int lo = g->halo_top ? g->own_lo : g->own_lo - 1, hi = g->halo_bot ? g->own_hi : g->own_hi + 1;
  return run_checksum_rows(a.p, a.esize, g->d, lo, hi, g->row_off, out);
}

// ------------------------------------------------------------------------------------------------
// caller-provided stream (e.g. the bench passes torch's current stream so that its CUDA events
// bracket this library's work) and the optional per-kernel profiler

extern "C" int litap_set_stream(void* cuda_stream) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  int rc = ensure_init();
  if (rc) return rc;
  Ctx& c = ctx();
  CUDA_TRY(cudaStreamSynchronize(c.stream));
  if (g_own_stream && c.stream) cudaStreamDestroy(c.stream);
  if (cuda_stream) {
    c.stream = (cudaStream_t)cuda_stream;
    g_own_stream = false;
  } else {
    CUDA_TRY(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    g_own_stream = true;
  }
  return LITAP_OK;
}

struct ProfRec { const char* name; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::vector<cudaEvent_t> g_prof_free;

ProfScope::ProfScope(const char* name) : idx(-1) {
  if (!g_prof_on) return;
  ProfRec r;
  r.name = name;
  for (cudaEvent_t* e : {&r.a, &r.b}) {
    if (!g_prof_free.empty()) { *e = g_prof_free.back(); g_prof_free.pop_back(); }
    else cudaEventCreate(e);
  }
  cudaEventRecord(r.a, ctx().stream);
  idx = (int)g_prof.size();
  g_prof.push_back(r);
}
ProfScope::~ProfScope() {
  if (idx >= 0) cudaEventRecord(g_prof[idx].b, ctx().stream);
}

extern "C" int litap_profile_enable(int on) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  g_prof_on = on != 0;
  return LITAP_OK;
}
// CSV "kernel,launches,total_ms\n" of everything recorded since the last report; clears the log
extern "C" int litap_profile_report(char* buf, int64_t cap) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!buf || cap < 1) return set_error(LITAP_EARG, "litap_profile_report: bad buffer");
  if (ctx().ready) cudaStreamSynchronize(ctx().stream);
  std::vector<std::string> names;
  std::vector<double> ms;
  std::vector<long> cnt;
  for (auto& r : g_prof) {
    float t = 0;
    cudaEventSynchronize(r.b);
    cudaEventElapsedTime(&t, r.a, r.b);
    size_t k = 0;
    for (; k < names.size(); ++k) if (names[k] == r.name) break;
    if (k == names.size()) { names.push_back(r.name); ms.push_back(0); cnt.push_back(0); }
    ms[k] += t;
    cnt[k]++;
    g_prof_free.push_back(r.a);
    g_prof_free.push_back(r.b);
  }
  g_prof.clear();
  std::string out;
  for (size_t k = 0; k < names.size(); ++k)
    out += names[k] + "," + std::to_string(cnt[k]) + "," + std::to_string(ms[k]) + "\n";
  if ((int64_t)out.size() + 1 > cap) return set_error(LITAP_EARG, "litap_profile_report: buffer too small");
  memcpy(buf, out.c_str(), out.size() + 1);
  return LITAP_OK;
}
