// hostcopy.cu — device results into host memory the caller owns.
//
// The callers of the host-buffer entry points hand over vectors they allocated with malloc (R vectors,
// numpy arrays): pageable and usually fresh pages.  cudaMemcpy into such memory is staged by the driver
// on one thread and ends up bound by page faults (~4 GB/s measured on the B200 box).  d2h_2d() looks at the
// destination: pinned / registered memory gets the plain asynchronous copy; pageable memory is filled
// through two pinned staging buffers in turn, a few host threads copying chunk j out while chunk j + 1
// crosses PCIe (3-4 x faster for the 39 columns of litap_relz).  The staged path returns when the data
// is in place (it is synchronous for the host, like a pageable cudaMemcpyAsync already is).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sys/mman.h>
#include <thread>
#include <unistd.h>
#include <vector>
#include "litap_internal.h"

namespace {
#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)
constexpr size_t HC_STAGE = 64u << 20;
void* g_stage[2] = {nullptr, nullptr};
cudaEvent_t g_stage_ev[2] = {nullptr, nullptr};

int stage_init() {
  for (int b = 0; b < 2; ++b) {
    if (!g_stage[b]) CUDA_TRY(cudaHostAlloc(&g_stage[b], HC_STAGE, cudaHostAllocDefault));
    if (!g_stage_ev[b]) CUDA_TRY(cudaEventCreateWithFlags(&g_stage_ev[b], cudaEventDisableTiming));
  }
  return LITAP_OK;
}
int copy_threads() {
  const char* e = getenv("LITAP_COPY_THREADS");
  int t = e ? atoi(e) : 8;
  const int hw = (int)std::thread::hardware_concurrency();
  if (hw > 0) t = std::min(t, hw);
  return std::max(t, 1);
}
void host_copy(char* dst, const char* src, size_t bytes, int nt) {
  if (nt <= 1 || bytes < (4u << 20)) { memcpy(dst, src, bytes); return; }
  std::vector<std::thread> th;
  const size_t per = ((bytes + nt - 1) / nt + 4095) & ~(size_t)4095;
  for (int t = 1; t < nt; ++t) {
    const size_t lo = std::min(bytes, per * t), hi = std::min(bytes, per * (t + 1));
    if (hi > lo) th.emplace_back([=] { memcpy(dst + lo, src + lo, hi - lo); });
  }
  memcpy(dst, src, std::min(bytes, per));
  for (auto& x : th) x.join();
}
// Fresh anonymous memory is faulted in page by page while it is written; asking for transparent huge pages on the
// 2 MB-aligned inside of the destination makes that 512 x fewer faults where the system allows it
// (transparent_hugepage = madvise or always).  Advice only: it does not change what the caller's memory holds.
// LITAP_D2H_HUGEPAGE=0 turns it off.
void advise_hugepages(void* dst, size_t bytes) {
  static const bool off = [] { const char* e = getenv("LITAP_D2H_HUGEPAGE"); return e && e[0] == '0'; }();
  if (off) return;
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
}  // namespace

// rows x width bytes from pitched device memory into dense-or-pitched host memory
int d2h_2d(void* dst, size_t dpitch, const void* src, size_t spitch, size_t width, size_t rows, cudaStream_t st) {
  if (!rows || !width) return LITAP_OK;
  if (!pageable(dst)) {
    CUDA_TRY(cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, rows, cudaMemcpyDeviceToHost, st));
    return LITAP_OK;
  }
  RC(stage_init());
  const int nt = copy_threads();
  advise_hugepages(dst, (rows - 1) * dpitch + width);
  // chunk size: the staging buffer, or less when LITAP_STAGE_KB says so (read per call: the tests use it to push
  // small grids through the multi-chunk paths)
  size_t stage = HC_STAGE;
  if (const char* e = getenv("LITAP_STAGE_KB")) stage = std::min(HC_STAGE, std::max<size_t>(4096, (size_t)atol(e) << 10));
  // chunks: whole rows while a row fits the staging buffer, else pieces of the one long row
  struct Chunk { size_t row, off, nrows, bytes; };
  std::vector<Chunk> ch;
  if (width <= stage) {
    const size_t per = std::max<size_t>(1, stage / width);
    for (size_t r = 0; r < rows; r += per) ch.push_back({r, 0, std::min(per, rows - r), 0});
  } else {
    for (size_t r = 0; r < rows; ++r)
      for (size_t o = 0; o < width; o += stage) ch.push_back({r, o, 0, std::min(stage, width - o)});
  }
  auto issue = [&](size_t j) -> int {
    const Chunk& c = ch[j];
    const char* s = (const char*)src + c.row * spitch + c.off;
    if (c.nrows) CUDA_TRY(cudaMemcpy2DAsync(g_stage[j & 1], width, s, spitch, width, c.nrows, cudaMemcpyDeviceToHost, st));
    else CUDA_TRY(cudaMemcpyAsync(g_stage[j & 1], s, c.bytes, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaEventRecord(g_stage_ev[j & 1], st));
    return LITAP_OK;
  };
  static const bool debug = getenv("LITAP_COPY_DEBUG") != nullptr;
  static double t_wait = 0, t_copy = 0, t_bytes = 0;
  auto now = [] { return std::chrono::steady_clock::now(); };
  RC(issue(0));
  for (size_t j = 0; j < ch.size(); ++j) {
    if (j + 1 < ch.size()) RC(issue(j + 1));      // buffer (j + 1) & 1 was drained by the host copy of chunk j - 1
    const auto t0 = now();
    CUDA_TRY(cudaEventSynchronize(g_stage_ev[j & 1]));
    const auto t1 = now();
    struct Acc { decltype(t1) a; double& slot; ~Acc() { slot += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - a).count(); } } acc{t1, t_copy};
    t_wait += std::chrono::duration<double, std::milli>(t1 - t0).count();
    const Chunk& c = ch[j];
    t_bytes += c.nrows ? (double)c.nrows * width : (double)c.bytes;
    char* d = (char*)dst + c.row * dpitch + c.off;
    const char* s = (const char*)g_stage[j & 1];
    if (!c.nrows) host_copy(d, s, c.bytes, nt);
    else if (dpitch == width) host_copy(d, s, c.nrows * width, nt);
    else for (size_t r = 0; r < c.nrows; ++r) memcpy(d + r * dpitch, s + r * width, width);
  }
  if (debug) fprintf(stderr, "[hostcopy] cumulative: %.0f MB, waiting for the device %.1f ms, host copies %.1f ms\n", t_bytes / 1e6, t_wait, t_copy);
  return LITAP_OK;
}

void hostcopy_forget_staging() {   // litap_shutdown
  for (int b = 0; b < 2; ++b) {
    if (g_stage[b]) { cudaFreeHost(g_stage[b]); g_stage[b] = nullptr; }
    if (g_stage_ev[b]) { cudaEventDestroy(g_stage_ev[b]); g_stage_ev[b] = nullptr; }
  }
}
