// api_host.cu — host-buffer entry points: what the R `.Call` shim binds (include/litap.h).
// Each call uploads its inputs, runs the stage(s) on the device and downloads the requested
// outputs.  No result is ever computed on the host.
#include <chrono>
#include <vector>
#include <condition_variable>
#include <deque>
#include <algorithm>
#include <mutex>
#include <string>
#include <thread>
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
  return h2d_2d(dst, (size_t)d.pitch * sizeof(T), src, (size_t)d.NC * sizeof(T), (size_t)d.NC * sizeof(T), d.NR, ctx().stream);
}
template <typename T>
int down(T* dst, const GridDims& d, const T* src) {
  if (!dst) return LITAP_OK;
  if (!src) return set_error(LITAP_ECUDA, "internal: result array missing on the device");
  return d2h_2d(dst, (size_t)d.NC * sizeof(T), src, (size_t)d.pitch * sizeof(T), (size_t)d.NC * sizeof(T), d.NR, ctx().stream);
}
int timed_sync(float* slot, cudaEvent_t a, cudaEvent_t b) {
  { int rc = hostcopy_flush(); if (rc) return rc; }   // staged downloads into pageable memory end here
  CUDA_TRY(cudaEventRecord(b, ctx().stream));
  CUDA_TRY(cudaEventSynchronize(b));
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  *slot += ms;
  return LITAP_OK;
}
#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

// second stream + event list of one pipelined call (destroyed with it; a failed call leaves nothing queued
// that refers to caller memory because the destructor drains the stream first)
struct Pipe {
  cudaStream_t copy = nullptr;
  std::vector<cudaEvent_t> ev, aux;
  int init() {
    CUDA_TRY(cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking));
    return LITAP_OK;
  }
  int mark(cudaStream_t st) {
    cudaEvent_t e;
    CUDA_TRY(cudaEventCreate(&e));
    ev.push_back(e);
    CUDA_TRY(cudaEventRecord(e, st));
    return LITAP_OK;
  }
  ~Pipe() {
    if (copy) { cudaStreamSynchronize(copy); cudaStreamDestroy(copy); }
    for (cudaEvent_t e : ev) cudaEventDestroy(e);
    for (cudaEvent_t e : aux) cudaEventDestroy(e);
  }
};

int begin_call(int64_t NR, int64_t NC, const void* must, GridGuard& gg) {
  if (!must) return set_error(LITAP_EARG, "NULL input array");
  RC(ensure_init());
  hostcopy_reset();
  litap_timings keep = ctx().tm;
  (void)keep;
  ctx().tm = litap_timings{};
  RC(litap_grid_create(NR, NC, &gg.g));
  return LITAP_OK;
}
}  // namespace

extern "C" int litap_ddir(const double* elev, int64_t NR, int64_t NC, int invert, uint8_t* ddir, uint8_t* flatcell) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  return litap_flow_form(elev, NR, NC, 0.0, invert, ddir, flatcell, nullptr, nullptr, nullptr, nullptr, nullptr,
                         nullptr, nullptr, nullptr, nullptr);
}

extern "C" int litap_flow_form(const double* elev, int64_t NR, int64_t NC, double grid, int invert, uint8_t* ddir,
                               uint8_t* flatcell, int32_t* shed, uint32_t* upslope, uint8_t* ridge, int64_t* npits,
                               double* slope_pct, double* slope_deg, double* aspect, double* prof, double* plan) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  Wall w;
  GridGuard gg;
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  const bool want_shed = shed || upslope || ridge || npits;
  const bool want_form = slope_pct || slope_deg || aspect || prof || plan;
  if (want_form && !(grid > 0)) return set_error(LITAP_EARG, "grid must be > 0");
  // The DEM goes up in row chunks; calc_form runs on each chunk as soon as its rows (and the row
  // below) are resident, and its five outputs go back on a second stream while the next chunk is
  // still uploading and while calc_ddir2 / calc_shed4 run: both PCIe directions stay busy.  That holds for
  // pinned / registered destinations (plain asynchronous copies).  Pageable destinations are staged by THIS
  // thread (hostcopy.cu): the transfers of a chunk's outputs still overlap the copy engine's next upload and the
  // kernels already queued, but later stages are enqueued only once the thread gets back from its host copies.
  // Nothing waits for the device in the middle of the call: the host syncs once, at the end.
  // form is computed on the elevations as given (form_mapper never sees the inverted DEM).
  const GridDims d = g->d;
  double* fouts[5] = {slope_pct, slope_deg, aspect, prof, plan};
  Pipe P;
  RC(P.init());
  // the form outputs are rounded numbers: they travel as int32 (20 instead of 40 B/cell) and are widened by the host
  // threads of hostcopy.cu while they are copied out of the staging buffers (bit-identical: form.cu k_form_pack checks
  // every value, hostexpand.cpp reproduces the quotient); LITAP_FORM_PACK=0 sends the doubles themselves
  const char* pack_env = getenv("LITAP_FORM_PACK");   // read per call (the tests compare both ways)
  const bool pack_off = pack_env && pack_env[0] == '0';
  const bool pack = want_form && !pack_off && (size_t)d.NC * 4 <= (64u << 20);
  int32_t* packed[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  int* pack_flags = nullptr;
  struct PackFree {
    int32_t** p; int** f;
    ~PackFree() { for (int i = 0; i < 5; ++i) grid_free_array(p[i]); if (*f) pool_free(*f); }
  } pack_free{packed, &pack_flags};
  if (pack) {
    for (int i = 0; i < 5; ++i)
      if (fouts[i] && !(packed[i] = grid_alloc_array<int32_t>(d, -1))) return set_error(LITAP_ENOMEM, "device allocation failed (packed form outputs)");
    pack_flags = (int*)pool_alloc(5 * sizeof(int));
    if (!pack_flags) return set_error(LITAP_ENOMEM, "device allocation failed (pack flags)");
    CUDA_TRY(cudaMemsetAsync(pack_flags, 0, 5 * sizeof(int), c.stream));
  }
  const int nchunk = d.NR >= 8 * 256 ? 8 : 1;
  const int crows = ((d.NR + nchunk - 1) / nchunk + 31) / 32 * 32;
  cudaEvent_t up0 = nullptr, up1 = nullptr;   // first / last upload (destroyed by the Pipe)
  RC(P.mark(c.stream));
  up0 = P.ev.back();
  const size_t ev_base = P.ev.size();
  // The downloads of the form outputs are driven by a helper thread: staged copies (pageable destinations, packed
  // outputs) keep a host thread busy with the copying / widening, and this thread has the uploads, calc_ddir2 and
  // calc_shed4 (with their host read-backs) to enqueue meanwhile.  The helper owns the download staging buffers of
  // hostcopy.cu until it is joined; this thread only uses the upload pair and plain asynchronous copies before that.
  struct Chunk { int lo, hi; cudaEvent_t ready; };
  std::vector<Chunk> chunks;
  chunks.reserve(nchunk + 1);
  std::mutex mu;
  std::condition_variable cv;
  size_t produced = 0;
  bool closed = false;
  // the small results (ddir, flatcell, shed, upslope, ridge) that go into pinned memory travel in pieces of <= 32 MB which
  // the helper slips between its staged chunks: one long copy queued by this thread would hold the D2H engine for
  // milliseconds and stall the staged downloads behind it (20 of 58 ms at 8192^2 before)
  struct Piece { void* dst; const void* src; size_t dpitch, spitch, width, rows; };
  std::deque<Piece> pieces;
  bool small_published = false;
  cudaEvent_t flow_done = nullptr;
  int helper_rc = LITAP_OK;
  std::string helper_err;
  const int device = c.device;
  cudaStream_t copy_stream = P.copy;
  std::thread helper;
  struct Joiner {   // whatever way the call ends, the helper is told to stop and joined before the staging state is touched
    std::thread& t; std::mutex& mu; std::condition_variable& cv; bool& closed; bool& published;
    ~Joiner() {
      if (!t.joinable()) return;
      { std::lock_guard<std::mutex> lk(mu); closed = true; published = true; }
      cv.notify_all();
      t.join();
    }
  } joiner{helper, mu, cv, closed, small_published};
  if (want_form)
    helper = std::thread([&] {
      cudaSetDevice(device);
      bool waited = false;
      auto send_pieces = [&](size_t maxn) -> int {
        std::vector<Piece> take;
        {
          std::lock_guard<std::mutex> lk(mu);
          if (!small_published) return LITAP_OK;
          while (take.size() < maxn && !pieces.empty()) { take.push_back(pieces.front()); pieces.pop_front(); }
        }
        if (take.empty()) return LITAP_OK;
        if (!waited) {
          if (cudaStreamWaitEvent(copy_stream, flow_done, 0) != cudaSuccess) return set_error(LITAP_ECUDA, "cudaStreamWaitEvent failed");
          waited = true;
        }
        for (const Piece& q : take)
          if (cudaMemcpy2DAsync(q.dst, q.dpitch, q.src, q.spitch, q.width, q.rows, cudaMemcpyDeviceToHost, copy_stream) != cudaSuccess)
            return set_error(LITAP_ECUDA, "cudaMemcpy2DAsync failed (small results)");
        return LITAP_OK;
      };
      for (size_t j = 0;; ++j) {
        Chunk ch;
        {
          std::unique_lock<std::mutex> lk(mu);
          cv.wait(lk, [&] { return produced > j || closed; });
          if (produced <= j) break;
          ch = chunks[j];
        }
        int rc = LITAP_OK;
        if (cudaStreamWaitEvent(copy_stream, ch.ready, 0) != cudaSuccess) rc = set_error(LITAP_ECUDA, "cudaStreamWaitEvent failed");
        for (int i = 0; i < 5 && !rc; ++i) {
          if (!fouts[i]) continue;
          if (pack)
            rc = d2h_2d_i32_expand(fouts[i] + (size_t)ch.lo * d.NC, (size_t)d.NC * 8, packed[i] + (size_t)ch.lo * d.pitch,
                                   (size_t)d.pitch * 4, (size_t)d.NC, ch.hi - ch.lo, i == 2 ? 2 : 1, copy_stream);
          else
            rc = d2h_2d(fouts[i] + (size_t)ch.lo * d.NC, (size_t)d.NC * 8, g->form[i] + (size_t)ch.lo * d.pitch,
                        (size_t)d.pitch * 8, (size_t)d.NC * 8, ch.hi - ch.lo, copy_stream);
          if (!rc) rc = send_pieces(1);
        }
        if (rc) { helper_rc = rc; helper_err = litap_last_error(); return; }
      }
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return small_published; });
      }
      int rc = send_pieces((size_t)-1);
      if (!rc) rc = hostcopy_flush();
      if (rc) { helper_rc = rc; helper_err = litap_last_error(); }
    });
  int form_done = 0;  // rows [0, form_done) of the form outputs are computed
  for (int a = 0; a < d.NR; a += crows) {
    const int b = a + crows < d.NR ? a + crows : d.NR;
    RC(h2d_2d(g->elev + (size_t)a * d.pitch, (size_t)d.pitch * 8, elev + (size_t)a * d.NC, (size_t)d.NC * 8,
              (size_t)d.NC * 8, b - a, c.stream));
    if (!want_form) continue;
    const int hi = b == d.NR ? d.NR : b - 1;  // row hi is needed as the window row below hi - 1
    if (hi <= form_done) continue;
    RC(P.mark(c.stream));
    RC(run_form_rows(g, grid, form_done, hi));
    RC(P.mark(c.stream));
    if (pack) RC(run_form_pack(g, packed, form_done, hi, pack_flags));
    cudaEvent_t ready;
    CUDA_TRY(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming));
    P.aux.push_back(ready);
    CUDA_TRY(cudaEventRecord(ready, c.stream));
    {
      std::lock_guard<std::mutex> lk(mu);
      chunks.push_back({form_done, hi, ready});
      produced = chunks.size();
    }
    cv.notify_all();
    form_done = hi;
  }
  { std::lock_guard<std::mutex> lk(mu); closed = true; }
  cv.notify_all();
  const size_t ev_end = P.ev.size();
  RC(P.mark(c.stream));
  up1 = P.ev.back();
  RC(litap_grid_ddir(g, invert));
  if (want_shed) RC(litap_grid_shed(g));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  // results that go into pinned memory are plain asynchronous copies: they can start while the helper is still busy;
  // pageable destinations need the staging buffers the helper holds
  struct Small { void* dst; const void* src; size_t esize; };
  std::vector<Small> small;
  small.push_back({ddir, g->ddir, 1});
  small.push_back({flatcell, g->flat, 1});
  if (want_shed) {
    small.push_back({shed, g->shed, 4});
    small.push_back({upslope, g->upslope, 4});
    small.push_back({ridge, g->ridge, 1});
    if (npits) *npits = g->npits;
  }
  auto send_small = [&](const Small& x) -> int {
    if (!x.dst) return LITAP_OK;
    if (!x.src) return set_error(LITAP_ECUDA, "internal: result array missing on the device");
    return d2h_2d(x.dst, (size_t)d.NC * x.esize, x.src, (size_t)d.pitch * x.esize, (size_t)d.NC * x.esize, d.NR, c.stream);
  };
  if (helper.joinable()) {   // pinned destinations: in pieces, through the helper
    CUDA_TRY(cudaEventCreateWithFlags(&flow_done, cudaEventDisableTiming));
    P.aux.push_back(flow_done);
    CUDA_TRY(cudaEventRecord(flow_done, c.stream));
    std::deque<Piece> mine;
    for (const Small& x : small) {
      if (!x.dst || hostcopy_is_pageable(x.dst)) continue;
      if (!x.src) return set_error(LITAP_ECUDA, "internal: result array missing on the device");
      const size_t rowb = (size_t)d.NC * x.esize;
      const size_t per = std::max<size_t>(1, ((size_t)32 << 20) / rowb);
      for (size_t r = 0; r < (size_t)d.NR; r += per) {
        const size_t nr = std::min(per, (size_t)d.NR - r);
        mine.push_back({(char*)x.dst + r * rowb, (const char*)x.src + r * (size_t)d.pitch * x.esize, rowb, (size_t)d.pitch * x.esize, rowb, nr});
      }
    }
    { std::lock_guard<std::mutex> lk(mu); pieces.swap(mine); small_published = true; }
    cv.notify_all();
  } else {
    for (const Small& x : small)
      if (x.dst && !hostcopy_is_pageable(x.dst)) RC(send_small(x));
  }
  if (helper.joinable()) helper.join();
  if (helper_rc) return set_error(helper_rc, helper_err);
  for (const Small& x : small)
    if (x.dst && hostcopy_is_pageable(x.dst)) RC(send_small(x));
  CUDA_TRY(cudaStreamSynchronize(P.copy));
  RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));  // what is left of the downloads after the last stage
  if (pack) {   // an output with a value its integer does not reproduce (|k| >= 2^31) goes again, as doubles
    int hf[5] = {0, 0, 0, 0, 0};
    CUDA_TRY(cudaMemcpy(hf, pack_flags, sizeof(hf), cudaMemcpyDeviceToHost));
    bool again = false;
    for (int i = 0; i < 5; ++i)
      if (fouts[i] && hf[i]) { RC(down(fouts[i], g->d, g->form[i])); again = true; }
    if (again) RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));
  }
  {
    float ms = 0;
    cudaEventElapsedTime(&ms, up0, up1);   // first upload .. last upload, with the form kernels between them
    c.tm.h2d_ms = ms;
    for (size_t i = ev_base; i + 1 < ev_end; i += 2) {
      cudaEventElapsedTime(&ms, P.ev[i], P.ev[i + 1]);
      c.tm.form_ms += ms;
    }
  }
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}

extern "C" int litap_shed(const double* elev, const uint8_t* ddir, int64_t NR, int64_t NC, int32_t* shed,
                          uint32_t* upslope, uint8_t* ridge, int64_t* npits) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
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
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
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
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
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

extern "C" int litap_uced(const double* elev, const uint8_t* ddir, const int32_t* shed, int64_t NR, int64_t NC, double* uced) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  Wall w;
  GridGuard gg;
  if (!ddir || !shed || !uced) return set_error(LITAP_EARG, "litap_uced: NULL ddir/shed/uced");
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  g->ddir = grid_alloc_array<uint8_t>(g->d, 0);
  g->shed = grid_alloc_array<int32_t>(g->d, 0);
  if (!g->ddir || !g->shed) return set_error(LITAP_ENOMEM, "device allocation failed (ddir/shed)");
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(up(g->elev, g->d, elev));
  RC(up(g->ddir, g->d, ddir));
  RC(up(g->shed, g->d, shed));
  RC(timed_sync(&c.tm.h2d_ms, c.ev[6], c.ev[7]));
  RC(litap_grid_uced(g));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(down(uced, g->d, g->uced));
  RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}

extern "C" int litap_pit_stat(const double* elev, const uint8_t* ddir, const int32_t* shed, const uint32_t* upslope,
                              const uint8_t* ridge, const uint8_t* edge_map, int64_t NR, int64_t NC, int method, int64_t cap,
                              double* const* out, int64_t* nrows) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  Wall w;
  GridGuard gg;
  if (!ddir || !shed || !upslope || !ridge || !edge_map || !out || !nrows) return set_error(LITAP_EARG, "litap_pit_stat: NULL argument");
  if (method != 0 && method != 1) return set_error(LITAP_EARG, "litap_pit_stat: method must be 0 (litap) or 1 (landmapr)");
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  g->ddir = grid_alloc_array<uint8_t>(g->d, 0);
  g->shed = grid_alloc_array<int32_t>(g->d, 0);
  g->upslope = grid_alloc_array<uint32_t>(g->d, 0);
  g->ridge = grid_alloc_array<uint8_t>(g->d, 0);
  uint8_t* edge = grid_alloc_array<uint8_t>(g->d, 0);
  struct EdgeFree { uint8_t* p; ~EdgeFree() { grid_free_array(p); } } ef{edge};
  if (!g->ddir || !g->shed || !g->upslope || !g->ridge || !edge) return set_error(LITAP_ENOMEM, "device allocation failed (pit_stat inputs)");
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(up(g->elev, g->d, elev));
  RC(up(g->ddir, g->d, ddir));
  RC(up(g->shed, g->d, shed));
  RC(up(g->upslope, g->d, upslope));
  RC(up(g->ridge, g->d, ridge));
  RC(up(edge, g->d, edge_map));
  RC(timed_sync(&c.tm.h2d_ms, c.ev[6], c.ev[7]));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(run_pit_stat(g, edge, method, cap, out, nrows));
  CUDA_TRY(cudaEventRecord(c.ev[7], c.stream));
  CUDA_TRY(cudaEventSynchronize(c.ev[7]));
  cudaEventElapsedTime(&c.tm.pitstat_ms, c.ev[6], c.ev[7]);
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}

extern "C" int litap_slope_gc(const double* elev, int64_t NR, int64_t NC, double grid, double* sgre, double* sgr,
                              double* sgcn, double* sgc, double* scr, double* scc, uint8_t* hill_r_dir,
                              uint8_t* hill_c_dir, int32_t* hill_r_n, int32_t* hill_r_cell, int32_t* hill_c_n,
                              int32_t* hill_c_cell) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  Wall w;
  GridGuard gg;
  if (!(grid > 0)) return set_error(LITAP_EARG, "litap_slope_gc: grid must be > 0");
  RC(begin_call(NR, NC, elev, gg));
  litap_grid* g = gg.g;
  Ctx& c = ctx();
  int64_t l0 = c.launches;
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  RC(up(g->elev, g->d, elev));
  RC(timed_sync(&c.tm.h2d_ms, c.ev[6], c.ev[7]));
  RC(litap_grid_slope_gc(g, grid));
  CUDA_TRY(cudaEventRecord(c.ev[6], c.stream));
  double* f[6] = {sgre, sgr, sgcn, sgc, scr, scc};
  uint8_t* u[2] = {hill_r_dir, hill_c_dir};
  int32_t* n[4] = {hill_r_n, hill_r_cell, hill_c_n, hill_c_cell};
  for (int i = 0; i < 6; ++i) RC(down(f[i], g->d, g->gc_f64[i]));
  for (int i = 0; i < 2; ++i) RC(down(u[i], g->d, g->gc_u8[i]));
  for (int i = 0; i < 4; ++i) RC(down(n[i], g->d, g->gc_i32[i]));
  RC(timed_sync(&c.tm.d2h_ms, c.ev[6], c.ev[7]));
  c.tm.kernel_launches = c.launches - l0;
  c.tm.total_ms = w.ms();
  return LITAP_OK;
}
