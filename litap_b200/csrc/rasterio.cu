// rasterio.cu — big DEMs from and to disk without a tibble (first half).
//
// The reference loads rasters through raster::raster(file) %>% as.data.frame(xy = TRUE) (R/LITAP_load.R:187-192, the
// ".flt" branch of load_file, :77-79), turns the missing value into NA (db_format, :194-213), wraps the grid in a ring of
// NA (add_buffer, :273-332) and saves every stage as an .rds / .csv of the whole tibble (R/LITAP_read_write.R:8-66).
// At 8192^2 and above those tibbles are what limits the size and the speed, so this file moves the raster the direct way:
//
//   litap_flt_info        parses the ESRI header (.hdr: ncols, nrows, xllcorner / xllcenter, yll*, cellsize,
//                         NODATA_value, byteorder)
//   litap_grid_read_flt   .flt (float32, row-major, north row first) -> a litap_grid: the file is read in chunks of whole
//                         rows into two pinned buffers in turn, chunk j + 1 is read from disk while chunk j crosses PCIe
//                         and k_flt_rows widens it to fp64 in place in the pitched elev array (NODATA / missing_value ->
//                         NaN, optional byte swap); the NA ring is the grid's initial fill
//   litap_grid_write_raw  one result of the grid (ring removed, like remove_buffer(), R/LITAP_load.R:334-340) as a
//                         headerless little-endian binary in its own dtype (+ an ESRI-style .hdr that says which), or as
//                         float32 .flt; the device rows travel through the download pair of pinned buffers while the
//                         previous chunk is written to the file
#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>
#include "cellmath.h"
#include "litap_internal.h"

namespace {
#define RC(x) do { int _rc = (x); if (_rc) return _rc; } while (0)

std::string with_ext(const std::string& path, const char* ext) {
  const size_t dot = path.find_last_of('.'), slash = path.find_last_of('/');
  const bool has = dot != std::string::npos && (slash == std::string::npos || dot > slash);
  return (has ? path.substr(0, dot) : path) + ext;
}

// rows [row0, row0 + nrows) of the raster, float32 in `src` (dense), -> elev rows 1 + row0 .. (ring offset), fp64
__global__ void k_flt_rows(const float* __restrict__ src, double* __restrict__ elev, GridDims d, int row0, int nrows, int ncol,
                           float nodata, float missing, int swap) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= ncol || r >= nrows) return;
  uint32_t b = reinterpret_cast<const uint32_t*>(src)[(size_t)r * ncol + c];
  if (swap) b = __byte_perm(b, 0, 0x0123);
  const float v = __uint_as_float(b);
  double e = (double)v;
  if (v != v || v == nodata || v == missing) e = __longlong_as_double(0x7ff8000000000000LL);
  elev[(size_t)(1 + row0 + r) * d.pitch + 1 + c] = e;
}

// interior rows [row0, row0 + nrows) of a result -> dense buffer, optionally narrowed to float32
template <typename T, typename O>
__global__ void k_pack_rows(const T* __restrict__ a, O* __restrict__ dst, GridDims d, int row0, int nrows, int ncol) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= ncol || r >= nrows) return;
  dst[(size_t)r * ncol + c] = (O)a[(size_t)(1 + row0 + r) * d.pitch + 1 + c];
}

bool parse_hdr(FILE* f, litap_raster_info* o, std::string* err) {
  char key[64], val[128];
  bool have_nr = false, have_nc = false;
  o->xll = o->yll = 0;
  o->cellsize = 1;
  o->nodata = -9999;
  o->big_endian = 0;
  while (fscanf(f, "%63s %127s", key, val) == 2) {
    std::string k(key);
    std::transform(k.begin(), k.end(), k.begin(), [](unsigned char ch) { return (char)std::tolower(ch); });
    if (k == "ncols") { o->ncol = atoll(val); have_nc = true; }
    else if (k == "nrows") { o->nrow = atoll(val); have_nr = true; }
    else if (k == "xllcorner" || k == "xllcenter") o->xll = atof(val);
    else if (k == "yllcorner" || k == "yllcenter") o->yll = atof(val);
    else if (k == "cellsize") o->cellsize = atof(val);
    else if (k == "nodata_value" || k == "nodata") o->nodata = atof(val);
    else if (k == "byteorder") o->big_endian = (val[0] == 'M' || val[0] == 'm');
  }
  if (!have_nr || !have_nc || o->nrow < 1 || o->ncol < 1) { *err = "header lacks ncols / nrows"; return false; }
  return true;
}
}  // namespace

extern "C" int litap_flt_info(const char* path, litap_raster_info* out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!path || !out) return set_error(LITAP_EARG, "litap_flt_info: NULL");
  const std::string hdr = with_ext(path, ".hdr");
  FILE* f = fopen(hdr.c_str(), "r");
  if (!f) return set_error(LITAP_EARG, "litap_flt_info: cannot open " + hdr);
  std::string err;
  const bool ok = parse_hdr(f, out, &err);
  fclose(f);
  if (!ok) return set_error(LITAP_EARG, "litap_flt_info: " + hdr + ": " + err);
  return LITAP_OK;
}

extern "C" int litap_grid_read_flt(const char* path, double missing_value, litap_grid** out, litap_raster_info* info_out) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!path || !out) return set_error(LITAP_EARG, "litap_grid_read_flt: NULL");
  *out = nullptr;
  litap_raster_info info;
  RC(litap_flt_info(path, &info));
  if (info_out) *info_out = info;
  const std::string flt = with_ext(path, ".flt");
  const int fd = open(flt.c_str(), O_RDONLY);
  if (fd < 0) return set_error(LITAP_EARG, "litap_grid_read_flt: cannot open " + flt);
  struct Closer { int fd; ~Closer() { if (fd >= 0) close(fd); } } closer{fd};
  const int nthreads = std::max(1, std::min(8, (int)std::thread::hardware_concurrency()));
  litap_grid* g = nullptr;
  RC(litap_grid_create(info.nrow + 2, info.ncol + 2, &g));   // elev starts as NaN everywhere: the NA ring of add_buffer()
  struct Guard { litap_grid* g; ~Guard() { if (g) litap_grid_destroy(g); } } guard{g};
  Ctx& c = ctx();
  cudaStream_t st = c.stream;
  void* buf[2];
  cudaEvent_t ev[2];
  size_t cap = 0;
  RC(hostcopy_staging(0, buf, ev, &cap));
  const size_t rowb = (size_t)info.ncol * 4;
  if (rowb > cap) return set_error(LITAP_EARG, "litap_grid_read_flt: a raster row does not fit the staging buffer");
  const int per = (int)std::min<size_t>(32768, std::max<size_t>(1, cap / rowb));   // rows are blocks of the grid's y dimension
  PoolScratch S;
  float* dev[2] = {S.get<float>((size_t)per * info.ncol), S.get<float>((size_t)per * info.ncol)};
  if (!dev[0] || !dev[1]) return set_error(LITAP_ENOMEM, "device allocation failed (raster rows)");
  int b = 0;
  for (int64_t r = 0; r < info.nrow; r += per, b ^= 1) {
    const int nr = (int)std::min<int64_t>(per, info.nrow - r);
    CUDA_TRY(cudaEventSynchronize(ev[b]));   // the copy engine is done with this buffer (and the kernel with dev[b])
    {   // the chunk comes off the file through several pread()s at once (one reader manages ~5 GB/s from the page cache)
      const size_t bytes = rowb * nr, off0 = (size_t)r * rowb;
      const size_t slice = ((bytes + nthreads - 1) / nthreads + 4095) & ~(size_t)4095;
      std::vector<std::thread> th;
      std::vector<int> ok(nthreads, 1);
      for (int t = 0; t < nthreads; ++t) {
        const size_t lo = std::min(bytes, slice * t), hi = std::min(bytes, slice * (t + 1));
        if (hi <= lo) continue;
        th.emplace_back([&, t, lo, hi] {
          char* dst = (char*)buf[b] + lo;
          size_t done = 0;
          while (done < hi - lo) {
            const ssize_t got = pread(fd, dst + done, hi - lo - done, (off_t)(off0 + lo + done));
            if (got <= 0) { ok[t] = 0; return; }
            done += (size_t)got;
          }
        });
      }
      for (auto& x : th) x.join();
      for (int t = 0; t < nthreads; ++t)
        if (!ok[t]) return set_error(LITAP_EARG, "litap_grid_read_flt: " + flt + " is shorter than its header says");
    }
    CUDA_TRY(cudaMemcpyAsync(dev[b], buf[b], rowb * nr, cudaMemcpyHostToDevice, st));
    dim3 grid((unsigned)((info.ncol + 255) / 256), (unsigned)nr);
    KERNEL("k_flt_rows", k_flt_rows<<<grid, 256, 0, st>>>(dev[b], g->elev, g->d, (int)r, nr, (int)info.ncol, (float)info.nodata,
                                                         (float)missing_value, info.big_endian));
    CUDA_TRY(cudaEventRecord(ev[b], st));
  }
  CUDA_TRY(cudaStreamSynchronize(st));
  guard.g = nullptr;
  *out = g;
  return LITAP_OK;
}

// ---- writer ------------------------------------------------------------------------------------------------------
// One file = one worker with its own stream, pair of pinned buffers and events (nothing of the library's shared state
// is touched, so several workers run side by side).  The kernel instantiates the page-cache pages of ONE fresh file at
// ~4 GB/s whatever the number of threads storing into it (per-inode serialisation), so the way to write faster is to
// write several files at once: litap_grid_write_raw_many.
namespace {
struct WriteJob {
  const void* arr; int esize; const char* name; const char* path; int as_flt;
  int rc = LITAP_OK; std::string err;
};
struct WriteLane {            // per worker
  cudaStream_t st = nullptr;
  void* buf[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  char* dev[2] = {nullptr, nullptr};   // float32 packing buffers (as_flt), from the pool, allocated by the caller thread
  size_t cap = 0;
};
#define WJ_CUDA(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { J.rc = LITAP_ECUDA; J.err = std::string(#expr) + ": " + cudaGetErrorString(_e); return; } } while (0)

void write_one(const GridDims d, const litap_raster_info* georef, WriteLane& L, WriteJob& J) {
  const int nrow = d.NR - 2, ncol = d.NC - 2;
  const int osize = J.as_flt ? 4 : J.esize;
  const size_t rowb = (size_t)ncol * osize;
  if (rowb > L.cap) { J.rc = LITAP_EARG; J.err = "litap_grid_write_raw: a row does not fit the staging buffer"; return; }
  const int fd = open(J.path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
  if (fd < 0) { J.rc = LITAP_EARG; J.err = std::string("litap_grid_write_raw: cannot create ") + J.path; return; }
  struct Closer { int fd; ~Closer() { close(fd); } } closer{fd};
  const int per = (int)std::min<size_t>(32768, std::max<size_t>(1, L.cap / rowb));
  int pending_rows[2] = {0, 0};
  auto drain = [&](int b) -> bool {
    if (!pending_rows[b]) return true;
    if (cudaEventSynchronize(L.ev[b]) != cudaSuccess) { J.rc = LITAP_ECUDA; J.err = "litap_grid_write_raw: copy failed"; return false; }
    const char* src = (const char*)L.buf[b];
    size_t left = rowb * pending_rows[b];
    while (left) {
      const ssize_t w = write(fd, src, left);
      if (w <= 0) { J.rc = LITAP_EARG; J.err = std::string("litap_grid_write_raw: short write to ") + J.path; return false; }
      src += w; left -= (size_t)w;
    }
    pending_rows[b] = 0;
    return true;
  };
  int b = 0;
  for (int r = 0; r < nrow; r += per, b ^= 1) {
    const int nr = std::min(per, nrow - r);
    if (!drain(b)) return;   // the chunk this buffer held goes to the file while the other buffer's chunk crosses PCIe
    if (J.as_flt) {
      dim3 grid((unsigned)((ncol + 255) / 256), (unsigned)nr);
      if (J.esize == 8) k_pack_rows<<<grid, 256, 0, L.st>>>((const double*)J.arr, (float*)L.dev[b], d, r, nr, ncol);
      else if (J.esize == 4) k_pack_rows<<<grid, 256, 0, L.st>>>((const int32_t*)J.arr, (float*)L.dev[b], d, r, nr, ncol);
      else k_pack_rows<<<grid, 256, 0, L.st>>>((const uint8_t*)J.arr, (float*)L.dev[b], d, r, nr, ncol);
      WJ_CUDA(cudaGetLastError());
      WJ_CUDA(cudaMemcpyAsync(L.buf[b], L.dev[b], rowb * nr, cudaMemcpyDeviceToHost, L.st));
    } else {   // own dtype: the copy engine strips the ring and the pitch
      const char* src = (const char*)J.arr + ((size_t)(1 + r) * d.pitch + 1) * J.esize;
      WJ_CUDA(cudaMemcpy2DAsync(L.buf[b], rowb, src, (size_t)d.pitch * J.esize, rowb, nr, cudaMemcpyDeviceToHost, L.st));
    }
    WJ_CUDA(cudaEventRecord(L.ev[b], L.st));
    pending_rows[b] = nr;
  }
  if (!drain(b) || !drain(b ^ 1)) return;
  // header: what is in the file and where it lies
  const std::string hdr = with_ext(J.path, ".hdr");
  FILE* h = fopen(hdr.c_str(), "w");
  if (!h) { J.rc = LITAP_EARG; J.err = "litap_grid_write_raw: cannot create " + hdr; return; }
  const bool is_float = J.as_flt || J.esize == 8;
  fprintf(h, "ncols %d\nnrows %d\nxllcorner %.10g\nyllcorner %.10g\ncellsize %.10g\nNODATA_value %s\nbyteorder LSBFIRST\n", ncol, nrow,
          georef ? georef->xll : 0.0, georef ? georef->yll : 0.0, georef ? georef->cellsize : 1.0, is_float ? "nan" : "0");
  fprintf(h, "nbits %d\npixeltype %s\nlayer %s\n", osize * 8,
          is_float ? "FLOAT" : (J.esize == 4 && !strcmp(J.name, "shed") ? "SIGNEDINT" : "UNSIGNEDINT"), J.name);
  fclose(h);
}

// pinned buffers and streams of the writer lanes, kept between calls (freed by litap_shutdown through rasterio_forget)
constexpr int MAX_LANES = 8;
constexpr size_t LANE_BUF = 16u << 20;
WriteLane g_lanes[MAX_LANES];
int lanes_init(int n) {
  for (int i = 0; i < n; ++i) {
    WriteLane& L = g_lanes[i];
    if (!L.st) CUDA_TRY(cudaStreamCreateWithFlags(&L.st, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
      if (!L.buf[b]) CUDA_TRY(cudaHostAlloc(&L.buf[b], LANE_BUF, cudaHostAllocDefault));
      if (!L.ev[b]) CUDA_TRY(cudaEventCreateWithFlags(&L.ev[b], cudaEventDisableTiming));
    }
    L.cap = LANE_BUF;
  }
  return LITAP_OK;
}
}  // namespace

void rasterio_forget() {   // litap_shutdown
  for (WriteLane& L : g_lanes) {
    for (int b = 0; b < 2; ++b) {
      if (L.buf[b]) { cudaFreeHost(L.buf[b]); L.buf[b] = nullptr; }
      if (L.ev[b]) { cudaEventDestroy(L.ev[b]); L.ev[b] = nullptr; }
    }
    if (L.st) { cudaStreamDestroy(L.st); L.st = nullptr; }
  }
}

extern "C" int litap_grid_write_raw_many(litap_grid* g, int n, const char* const* names, const char* const* paths, int as_flt,
                                         const litap_raster_info* georef) {
  ApiLock api_lock_;   // one host thread at a time (include/litap.h)
  if (!g || n < 1 || !names || !paths) return set_error(LITAP_EARG, "litap_grid_write_raw_many: bad arguments");
  std::vector<WriteJob> jobs(n);
  for (int k = 0; k < n; ++k) {
    if (!names[k] || !paths[k]) return set_error(LITAP_EARG, "litap_grid_write_raw_many: NULL name / path");
    const ArrRef a = grid_find_array(g, names[k]);
    if (!a.p) return set_error(LITAP_EARG, std::string("litap_grid_write_raw: no such result on the grid: ") + names[k]);
    jobs[k].arr = a.p; jobs[k].esize = a.esize; jobs[k].name = names[k]; jobs[k].path = paths[k]; jobs[k].as_flt = as_flt;
  }
  const int nl = std::min(n, MAX_LANES);
  RC(lanes_init(nl));
  CUDA_TRY(cudaStreamSynchronize(ctx().stream));   // the results are complete before the lanes' streams read them
  PoolScratch S;
  const GridDims d = g->d;
  if (as_flt) {
    const size_t rowb = (size_t)(d.NC - 2) * 4;
    const size_t per = std::min<size_t>(32768, std::max<size_t>(1, LANE_BUF / std::max<size_t>(rowb, 1)));
    for (int i = 0; i < nl; ++i)
      for (int b = 0; b < 2; ++b) {
        g_lanes[i].dev[b] = S.get<char>(per * rowb);
        if (!g_lanes[i].dev[b]) return set_error(LITAP_ENOMEM, "device allocation failed (raster rows)");
      }
  }
  const int device = ctx().device;
  std::vector<std::thread> th;
  for (int i = 0; i < nl; ++i)
    th.emplace_back([&, i] {
      cudaSetDevice(device);
      for (int k = i; k < n; k += nl) write_one(d, georef, g_lanes[i], jobs[k]);
      cudaStreamSynchronize(g_lanes[i].st);
    });
  for (auto& x : th) x.join();
  for (int k = 0; k < n; ++k)
    if (jobs[k].rc) return set_error(jobs[k].rc, jobs[k].err);
  return LITAP_OK;
}

extern "C" int litap_grid_write_raw(litap_grid* g, const char* name, const char* path, int as_flt, const litap_raster_info* georef) {
  return litap_grid_write_raw_many(g, 1, &name, &path, as_flt, georef);
}
