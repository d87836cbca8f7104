// hostcopy.cu — device results into host memory the caller owns.
//
// The callers of the host-buffer entry points hand over vectors they allocated with malloc (R vectors,
// numpy arrays): pageable and usually fresh pages.  cudaMemcpy into such memory is staged by the driver
// on one thread and ends up bound by page faults (~4 GB/s measured on the B200 box).  d2h_2d() looks at the
// destination: pinned / registered memory gets the plain asynchronous copy; pageable memory is filled
// through four pinned staging buffers in turn, a few host threads copying chunk j out while chunk j + 1
// crosses PCIe (3-4 x faster for the 39 columns of litap_relz).  The staged path returns when the data
// is in place (it is synchronous for the host, like a pageable cudaMemcpyAsync already is).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sys/mman.h>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <unistd.h>
#include <vector>
#include "litap_internal.h"

extern "C" void litap_host_expand_i32(const int32_t* src, double* dst, size_t n, int kind);   // hostexpand.cpp

namespace {
#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)
constexpr size_t HC_STAGE = 64u << 20;
constexpr int NSTAGE = 4;   // download staging buffers: the copy engine runs up to three chunks ahead of the host threads
void* g_stage[NSTAGE] = {};
cudaEvent_t g_stage_ev[NSTAGE] = {};

int stage_init() {
  for (int b = 0; b < NSTAGE; ++b) {
    if (!g_stage[b]) CUDA_TRY(cudaHostAlloc(&g_stage[b], HC_STAGE, cudaHostAllocDefault));
    if (!g_stage_ev[b]) CUDA_TRY(cudaEventCreateWithFlags(&g_stage_ev[b], cudaEventDisableTiming));
  }
  return LITAP_OK;
}
int copy_threads() {
  const char* e = getenv("LITAP_COPY_THREADS");
  int t = e ? atoi(e) : 12;   // 8 -> 12 threads: 244 -> 213 ms for the pageable 8192^2 litap_flow_form on the 16-core box; 16 is no faster
  const int hw = (int)std::thread::hardware_concurrency();
  if (hw > 0) t = std::min(t, hw);
  return std::max(t, 1);
}
// A small pool of host threads that stays alive between chunks: a staged chunk is a few tens of MB, i.e. ~1 ms of
// copying / widening for 12 threads, and creating and joining them for every chunk cost a quarter of that.
// run(n, f) calls f(0) .. f(n - 1), f(0) on the calling thread, and returns when all are done.  One caller at a time
// (the library is driven by one thread, or by the download helper of litap_flow_form while that thread enqueues).
class CopyPool {
 public:
  void run(int n, const std::function<void(int)>& f) {
    if (n <= 1) { if (n == 1) f(0); return; }
    std::lock_guard<std::mutex> one(user_);
    grow(n - 1);
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = &f; parts_ = n; next_ = 1; left_ = n - 1; ++gen_;
    }
    cv_.notify_all();
    f(0);
    std::unique_lock<std::mutex> lk(mu_);
    done_.wait(lk, [&] { return left_ == 0; });
    fn_ = nullptr;
  }
  void stop() {
    { std::lock_guard<std::mutex> lk(mu_); quit_ = true; }
    cv_.notify_all();
    for (auto& t : th_) t.join();
    th_.clear();
    quit_ = false;
  }
  ~CopyPool() { stop(); }

 private:
  void grow(int n) {
    while ((int)th_.size() < n) th_.emplace_back([this] { work(); });
  }
  void work() {
    unsigned long long seen = 0;
    for (;;) {
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [&] { return quit_ || (gen_ != seen && next_ < parts_); });
      if (quit_) return;
      if (gen_ != seen && next_ >= parts_) { seen = gen_; continue; }
      const int i = next_++;
      if (next_ >= parts_) seen = gen_;
      const std::function<void(int)>* f = fn_;
      lk.unlock();
      (*f)(i);
      lk.lock();
      if (--left_ == 0) done_.notify_all();
    }
  }
  std::mutex user_, mu_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> th_;
  const std::function<void(int)>* fn_ = nullptr;
  int parts_ = 0, next_ = 0, left_ = 0;
  unsigned long long gen_ = 0;
  bool quit_ = false;
};
CopyPool g_pool_threads;

void host_copy(char* dst, const char* src, size_t bytes, int nt) {
  if (nt <= 1 || bytes < (4u << 20)) { memcpy(dst, src, bytes); return; }
  const size_t per = ((bytes + nt - 1) / nt + 4095) & ~(size_t)4095;
  g_pool_threads.run(nt, [=](int t) {
    const size_t lo = std::min(bytes, per * t), hi = std::min(bytes, per * (t + 1));
    if (hi > lo) memcpy(dst + lo, src + lo, hi - lo);
  });
}
// Fresh anonymous memory is faulted in page by page while it is written; asking for transparent huge pages on the
// 2 MB-aligned inside of the destination makes that 512 x fewer faults where the system allows it
// (transparent_hugepage = madvise or always).  Advice only: it does not change what the caller's memory holds.
// The memory belongs to the caller, so this is OPT-IN: LITAP_D2H_HUGEPAGE=1 (the pageable figures of the bench say
// which way they were taken).
void advise_hugepages(void* dst, size_t bytes) {
  static const bool on = [] { const char* e = getenv("LITAP_D2H_HUGEPAGE"); return e && e[0] == '1'; }();
  if (!on) return;
  const uintptr_t H = (uintptr_t)2 << 20;
  const uintptr_t lo = ((uintptr_t)dst + H - 1) & ~(H - 1), hi = ((uintptr_t)dst + bytes) & ~(H - 1);
  if (hi > lo) (void)madvise((void*)lo, hi - lo, MADV_HUGEPAGE);
}
bool pageable(const void* p) {
  static const bool off = [] { const char* e = getenv("LITAP_STAGED_D2H"); return e && e[0] == '0'; }();   // A/B switch
  if (off) return false;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return true; }
  return at.type == cudaMemoryTypeUnregistered;
}
// A chunk that is on its way into a staging buffer: the host copy happens when the buffer is needed again or at
// hostcopy_flush(), so that the copy of one chunk overlaps the transfer of the next one across calls too.
struct Pending { bool on = false; char* dst = nullptr; size_t dpitch = 0, width = 0, nrows = 0, bytes = 0; int kind = 0; };   // kind 1 / 2: int32 -> fp64
Pending g_pend[NSTAGE];
int g_next = 0;
double g_wait_ms = 0, g_copy_ms = 0, g_bytes = 0;

int complete(int b) {
  Pending& p = g_pend[b];
  if (!p.on) return LITAP_OK;
  p.on = false;
  const auto t0 = std::chrono::steady_clock::now();
  CUDA_TRY(cudaEventSynchronize(g_stage_ev[b]));
  const auto t1 = std::chrono::steady_clock::now();
  const int nt = copy_threads();
  const char* s = (const char*)g_stage[b];
  if (p.kind) {   // packed download: width int32 per row in the staging buffer -> width doubles per row of the destination
    const size_t ncol = p.width, total = ncol * p.nrows;
    const int32_t* src = (const int32_t*)s;
    auto piece = [&](size_t lo, size_t hi) {   // elements [lo, hi) of the dense chunk, row by row where the destination is pitched
      if (p.dpitch == ncol * 8) { litap_host_expand_i32(src + lo, (double*)p.dst + lo, hi - lo, p.kind); return; }
      while (lo < hi) {
        const size_t r = lo / ncol, c = lo % ncol, n = std::min(hi - lo, ncol - c);
        litap_host_expand_i32(src + lo, (double*)(p.dst + r * p.dpitch) + c, n, p.kind);
        lo += n;
      }
    };
    if (nt <= 1 || total < (1u << 18)) piece(0, total);
    else {
      const size_t per = ((total + nt - 1) / nt + 15) & ~(size_t)15;
      g_pool_threads.run(nt, [&](int t) {
        const size_t lo = std::min(total, per * t), hi = std::min(total, per * (t + 1));
        if (hi > lo) piece(lo, hi);
      });
    }
  }
  else if (!p.nrows) host_copy(p.dst, s, p.bytes, nt);
  else if (p.dpitch == p.width) host_copy(p.dst, s, p.nrows * p.width, nt);
  else for (size_t r = 0; r < p.nrows; ++r) memcpy(p.dst + r * p.dpitch, s + r * p.width, p.width);
  const auto t2 = std::chrono::steady_clock::now();
  g_wait_ms += std::chrono::duration<double, std::milli>(t1 - t0).count();
  g_copy_ms += std::chrono::duration<double, std::milli>(t2 - t1).count();
  g_bytes += p.kind ? (double)p.nrows * p.width * 4 : p.nrows ? (double)p.nrows * p.width : (double)p.bytes;
  return LITAP_OK;
}
}  // namespace

// rows x width bytes from pitched device memory into dense-or-pitched host memory.  Pageable destinations are
// complete after hostcopy_flush() (every entry point calls it before it returns).
int d2h_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows, cudaStream_t st) {
  if (!rows || !width) return LITAP_OK;
  if (!pageable(dst)) {
    CUDA_TRY(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyDeviceToHost, st));
    return LITAP_OK;
  }
  RC(stage_init());
  advise_hugepages(dst, (rows - 1) * dpitch + width);
  // chunk size: the staging buffer, or less when LITAP_STAGE_KB says so (read per call: the tests use it to push
  // small grids through the multi-chunk paths)
  size_t stage = HC_STAGE;
  if (const char* e = getenv("LITAP_STAGE_KB")) stage = std::min(HC_STAGE, std::max<size_t>(4096, (size_t)atol(e) << 10));
  // chunks: whole rows while a row fits the staging buffer, else pieces of the one long row
  auto send = [&](size_t row, size_t off, size_t nrows, size_t bytes) -> int {
    const int b = g_next;
    RC(complete(b));                               // the buffer's previous chunk goes out while the other buffer fills
    const char* s = (const char*)src + row * spitch + off;
    if (nrows) CUDA_TRY(cudaMemcpy2DAsync(g_stage[b], width, s, spitch, width, nrows, cudaMemcpyDeviceToHost, st));
    else CUDA_TRY(cudaMemcpyAsync(g_stage[b], s, bytes, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaEventRecord(g_stage_ev[b], st));
    Pending& p = g_pend[b];
    p.on = true; p.dst = (char*)dst + row * dpitch + off; p.dpitch = dpitch; p.width = width; p.nrows = nrows; p.bytes = bytes; p.kind = 0;
    g_next = (g_next + 1) % NSTAGE;
    return LITAP_OK;
  };
  if (width <= stage) {
    const size_t per = std::max<size_t>(1, stage / width);
    for (size_t r = 0; r < rows; r += per) RC(send(r, 0, std::min(per, rows - r), 0));
  } else {
    for (size_t r = 0; r < rows; ++r)
      for (size_t o = 0; o < width; o += stage) RC(send(r, o, 0, std::min(stage, width - o)));
  }
  return LITAP_OK;
}

// rows x ncol int32 from pitched device memory, widened to fp64 (kind 1: k / 1000, kind 2: k; INT32_MIN = NA) on their way
// into dense-or-pitched host memory.  Always staged (pinned destinations too): the host threads do the widening.
int d2h_2d_i32_expand(double* dst, size_t dpitch, const int32_t* src, size_t spitch, size_t ncol, size_t rows, int kind, cudaStream_t st) {
  if (!rows || !ncol) return LITAP_OK;
  RC(stage_init());
  const size_t width = ncol * 4;
  if (width > HC_STAGE) return set_error(LITAP_EARG, "packed download: a row does not fit the staging buffer");
  if (pageable(dst)) advise_hugepages(dst, (rows - 1) * dpitch + ncol * 8);
  size_t stage = HC_STAGE;
  if (const char* e = getenv("LITAP_STAGE_KB")) stage = std::min(HC_STAGE, std::max<size_t>(width, (size_t)atol(e) << 10));
  const size_t per = std::max<size_t>(1, stage / width);
  for (size_t r = 0; r < rows; r += per) {
    const size_t nr = std::min(per, rows - r);
    const int b = g_next;
    RC(complete(b));
    CUDA_TRY(cudaMemcpy2DAsync(g_stage[b], width, (const char*)src + r * spitch, spitch, width, nr, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaEventRecord(g_stage_ev[b], st));
    Pending& p = g_pend[b];
    p.on = true; p.dst = (char*)dst + r * dpitch; p.dpitch = dpitch; p.width = ncol; p.nrows = nr; p.bytes = 0; p.kind = kind;
    g_next = (g_next + 1) % NSTAGE;
  }
  return LITAP_OK;
}

// ---- uploads from pageable caller memory ------------------------------------------------------------------
// Same idea the other way round (own pair of staging buffers): host threads fill buffer b while the copy engine
// drains buffer b ^ 1.  Nothing refers to the caller's memory any more when the call returns; the last chunks
// are still on their way from the staging buffers, in stream order before whatever the caller enqueues next.
namespace {
void* g_up[2] = {nullptr, nullptr};
cudaEvent_t g_up_ev[2] = {nullptr, nullptr};
int g_up_next = 0;
int up_init() {
  for (int b = 0; b < 2; ++b) {
    if (!g_up[b]) CUDA_TRY(cudaHostAlloc(&g_up[b], HC_STAGE, cudaHostAllocDefault));
    if (!g_up_ev[b]) CUDA_TRY(cudaEventCreateWithFlags(&g_up_ev[b], cudaEventDisableTiming));
  }
  return LITAP_OK;
}
}  // namespace

int h2d_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows, cudaStream_t st) {
  if (!rows || !width) return LITAP_OK;
  if (!pageable(src) || width > HC_STAGE || spitch != width) {
    CUDA_TRY(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyHostToDevice, st));
    return LITAP_OK;
  }
  RC(up_init());
  const int nt = copy_threads();
  size_t stage = HC_STAGE;
  if (const char* e = getenv("LITAP_STAGE_KB")) stage = std::min(HC_STAGE, std::max<size_t>(4096, (size_t)atol(e) << 10));
  const size_t per = std::max<size_t>(1, stage / width);
  if (width > stage) {   // (only with a test-sized LITAP_STAGE_KB)
    CUDA_TRY(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyHostToDevice, st));
    return LITAP_OK;
  }
  for (size_t r = 0; r < rows; r += per) {
    const size_t nr = std::min(per, rows - r);
    const int b = g_up_next;
    CUDA_TRY(cudaEventSynchronize(g_up_ev[b]));     // the copy engine is done with this buffer (no-op the first time)
    host_copy((char*)g_up[b], (const char*)src + r * spitch, nr * width, nt);
    CUDA_TRY(cudaMemcpy2DAsync((char*)dst + r * dpitch, dpitch, g_up[b], width, width, nr, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaEventRecord(g_up_ev[b], st));
    g_up_next ^= 1;
  }
  return LITAP_OK;
}

// the pinned staging pairs for code that moves files (rasterio.cu): which = 0 the upload pair, 1 the download pair
// (pending downloads are completed first)
int hostcopy_staging(int which, void* buf[2], cudaEvent_t ev[2], size_t* bytes) {
  if (which == 0) {
    RC(up_init());
    for (int b = 0; b < 2; ++b) { buf[b] = g_up[b]; ev[b] = g_up_ev[b]; }
  } else {
    RC(stage_init());
    for (int k = 0; k < NSTAGE; ++k) RC(complete((g_next + k) % NSTAGE));
    for (int b = 0; b < 2; ++b) { buf[b] = g_stage[b]; ev[b] = g_stage_ev[b]; }
  }
  *bytes = HC_STAGE;
  return LITAP_OK;
}

bool hostcopy_is_pageable(const void* p) { return pageable(p); }

// completes the staged downloads (older buffer first)
int hostcopy_flush() {
  for (int k = 0; k < NSTAGE; ++k) RC(complete((g_next + k) % NSTAGE));   // oldest first
  static const bool debug = getenv("LITAP_COPY_DEBUG") != nullptr;
  if (debug && g_bytes > 0)
    fprintf(stderr, "[hostcopy] cumulative: %.0f MB, waiting for the device %.1f ms, host copies %.1f ms\n", g_bytes / 1e6, g_wait_ms, g_copy_ms);
  return LITAP_OK;
}
// a call that failed half way may have left chunks pending: their destination belongs to that call, drop them
void hostcopy_reset() {
  bool any = false;
  for (int b = 0; b < NSTAGE; ++b) any = any || g_pend[b].on;
  if (any) cudaDeviceSynchronize();
  for (int b = 0; b < NSTAGE; ++b) g_pend[b].on = false;
}

void hostcopy_forget_staging() {   // litap_shutdown
  g_pool_threads.stop();
  for (int b = 0; b < NSTAGE; ++b) {
    if (g_stage[b]) { cudaFreeHost(g_stage[b]); g_stage[b] = nullptr; }
    if (g_stage_ev[b]) { cudaEventDestroy(g_stage_ev[b]); g_stage_ev[b] = nullptr; }
  }
  for (int b = 0; b < 2; ++b) {
    if (g_up[b]) { cudaFreeHost(g_up[b]); g_up[b] = nullptr; }
    if (g_up_ev[b]) { cudaEventDestroy(g_up_ev[b]); g_up_ev[b] = nullptr; }
  }
}
