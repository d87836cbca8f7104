// api_host.cu — host-buffer entry points: what the R `.Call` shim binds (include/litap.h).
// Each call uploads its inputs, runs the stage(s) on the device and downloads the requested
// outputs.  No result is ever computed on the host.
#include <chrono>
#include "litap_internal.h"

namespace {
struct GridGuard {
  litap_grid* g = nullptr;
  ~GridGuard() { litap_grid_destroy(g); }
};
struct Wall {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  float ms() const { return std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
};

template <typename T>
int up(T* dst, const GridDims& d, const T* src) {
  CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)d.pitch * sizeof(T), src, (size_t)d.NC * sizeof(T),
                             (size_t)d.NC * sizeof(T), d.NR, cudaMemcpyHostToDevice, ctx().stream));
  return LITAP_OK;
}
template <typename T>
int down(T* dst, const GridDims& d, const T* src) {
  if (!dst) return LITAP_OK;
  if (!src) return set_error(LITAP_ECUDA, "internal: result array missing on the device");
  CUDA_TRY(cudaMemcpy2DAsync(dst, (size_t)d.NC * sizeof(T), src, (size_t)d.pitch * sizeof(T),
                             (size_t)d.NC * sizeof(T), d.NR, cudaMemcpyDeviceToHost, ctx().stream));
  return LITAP_OK;
}
int timed_sync(float* slot, cudaEvent_t a, cudaEvent_t b) {
  CUDA_TRY(cudaEventRecord(b, ctx().stream));
  CUDA_TRY(cudaEventSynchronize(b));
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  *slot += ms;
  return LITAP_OK;
}
#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

int begin_call(int64_t NR, int64_t NC, const void* must, GridGuard& gg) {
  if (!must) return set_error(LITAP_EARG, "NULL input array");
  RC(ensure_init());
  litap_timings keep = ctx().tm;
  (void)keep;
  ctx().tm = litap_timings{};
  RC(litap_grid_create(NR, NC, &gg.g));
  return LITAP_OK;
}
}  // namespace

extern "C" int litap_ddir(const double* elev, int64_t NR, int64_t NC, int invert, uint8_t* ddir, uint8_t* flatcell) {
  return litap_flow_form(elev, NR, NC, 0.0, invert, ddir, flatcell, nullptr, nullptr, nullptr, nullptr, nullptr,
                         nullptr, nullptr, nullptr, nullptr);
}

extern "C" int litap_flow_form(const double* elev, int64_t NR, int64_t NC, double grid, int invert, uint8_t* ddir,
                               uint8_t* flatcell, int32_t* shed, uint32_t* upslope, uint8_t* ridge, int64_t* npits,
                               double* slope_pct, double* slope_deg, double* aspect, double* prof, double* plan) {
  Wall w;
  GridGuard gg;
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  const bool want_shed = shed || upslope || ridge || npits;
  const bool want_form = slope_pct || slope_deg || aspect || prof || plan;
  if (want_form && !(grid > 0)) return set_error(LITAP_EARG, "grid must be > 0");
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(up(g->elev, g->d, elev));
  RC(timed_sync(&c.tm.h2d_ms, c.ev[6], c.ev[7]));
  // form is computed on the elevations as given (form_mapper never sees the inverted DEM)
  if (want_form) { RC(litap_grid_form(g, grid)); }
  RC(litap_grid_ddir(g, invert));
  if (want_shed) RC(litap_grid_shed(g));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(down(ddir, g->d, g->ddir));
  RC(down(flatcell, g->d, g->flat));
  if (want_shed) {
    RC(down(shed, g->d, g->shed));
    RC(down(upslope, g->d, g->upslope));
    RC(down(ridge, g->d, g->ridge));
    if (npits) *npits = g->npits;
  }
  if (want_form) {
    double* outs[5] = {slope_pct, slope_deg, aspect, prof, plan};
    for (int i = 0; i < 5; ++i) RC(down(outs[i], g->d, g->form[i]));
  }
  RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}

extern "C" int litap_shed(const double* elev, const uint8_t* ddir, int64_t NR, int64_t NC, int32_t* shed,
                          uint32_t* upslope, uint8_t* ridge, int64_t* npits) {
  Wall w;
  GridGuard gg;
  if (!ddir) return set_error(LITAP_EARG, "litap_shed: NULL ddir");
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  g->ddir = grid_alloc_array<uint8_t>(g->d, 0);
  if (!g->ddir) return set_error(LITAP_ENOMEM, "device allocation failed (ddir)");
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(up(g->elev, g->d, elev));
  RC(up(g->ddir, g->d, ddir));
  RC(timed_sync(&c.tm.h2d_ms, c.ev[6], c.ev[7]));
  RC(litap_grid_shed(g));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(down(shed, g->d, g->shed));
  RC(down(upslope, g->d, g->upslope));
  RC(down(ridge, g->d, g->ridge));
  RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));
  if (npits) *npits = g->npits;
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}

extern "C" int litap_form(const double* elev, int64_t NR, int64_t NC, double grid, double* slope_pct,
                          double* slope_deg, double* aspect, double* prof, double* plan) {
  Wall w;
  GridGuard gg;
  if (!(grid > 0)) return set_error(LITAP_EARG, "litap_form: grid must be > 0");
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(up(g->elev, g->d, elev));
  RC(timed_sync(&c.tm.h2d_ms, c.ev[6], c.ev[7]));
  RC(litap_grid_form(g, grid));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  double* outs[5] = {slope_pct, slope_deg, aspect, prof, plan};
  for (int i = 0; i < 5; ++i) RC(down(outs[i], g->d, g->form[i]));
  RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}

extern "C" int litap_weti(const double* elev, const uint8_t* ddir, const uint32_t* upslope, int64_t NR, int64_t NC,
                          double grid, double* qarea1, double* qarea2, double* qweti1, double* qweti2) {
  Wall w;
  GridGuard gg;
  if (!ddir || !upslope) return set_error(LITAP_EARG, "litap_weti: NULL ddir/upslope");
  if (!(grid > 0)) return set_error(LITAP_EARG, "litap_weti: grid must be > 0");
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  g->ddir = grid_alloc_array<uint8_t>(g->d, 0);
  g->upslope = grid_alloc_array<uint32_t>(g->d, 0);
  if (!g->ddir || !g->upslope) return set_error(LITAP_ENOMEM, "device allocation failed (ddir/upslope)");
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(up(g->elev, g->d, elev));
  RC(up(g->ddir, g->d, ddir));
  RC(up(g->upslope, g->d, upslope));
  RC(timed_sync(&c.tm.h2d_ms, c.ev[6], c.ev[7]));
  RC(litap_grid_weti(g, grid));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  double* outs[4] = {qarea1, qarea2, qweti1, qweti2};
  for (int i = 0; i < 4; ++i) RC(down(outs[i], g->d, g->weti[i]));
  RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}
