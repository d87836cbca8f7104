// synth.cu — bench / test utilities that run on the device: the synthetic hash-fBm DEM generator
// (spec in DESIGN.md "workload"; bit-identical with the CPU generator of the test suite because only
// + - * / floor and uint32 hashing are used and the build has -fmad=false) and an order-independent
// checksum of a result array.
#include "litap_internal.h"

namespace {
__device__ __forceinline__ uint32_t hash32(uint32_t ix, uint32_t iy, uint32_t s) {
  uint32_t h = ix * 0x9E3779B1u ^ iy * 0x85EBCA77u ^ s * 0xC2B2AE3Du;
  h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ double vnoise(double x, double y, uint32_t s) {
  double fx0 = floor(x), fy0 = floor(y);
  uint32_t ix = (uint32_t)(long long)fx0, iy = (uint32_t)(long long)fy0;
  double fx = x - fx0, fy = y - fy0;
  double ux = fx * fx * (3.0 - 2.0 * fx), uy = fy * fy * (3.0 - 2.0 * fy);
  const double k = 1.0 / 4294967296.0;
  double a = hash32(ix, iy, s) * k, b = hash32(ix + 1, iy, s) * k;
  double c = hash32(ix, iy + 1, s) * k, d = hash32(ix + 1, iy + 1, s) * k;
  return ((a * (1.0 - ux) + b * ux) * (1.0 - uy)) + ((c * (1.0 - ux) + d * ux) * uy);
}

__global__ void k_synth(double* __restrict__ elev, GridDims d, long long row0, long long gnr, int O, uint32_t seed,
                        double quant, int na_permille) {
  const int C = blockIdx.x * blockDim.x + threadIdx.x + 1 - 32;
  const int R = blockIdx.y;
  if (C < 0 || C >= d.NC) return;
  const long long gr = row0 + R - 1, gc = C - 1;
  const bool ring = (C == 0 || C == d.NC - 1 || gr < 0 || gr >= gnr);
  double v;
  if (ring) {
    v = __longlong_as_double(0x7ff8000000000000LL);
  } else {
    double acc = 0.0, amp = 1.0;
    for (int o = 0; o < O; ++o) {
      double sc = (double)(1ll << (O - 1 - o));
      acc = acc + amp * vnoise((double)gr / sc, (double)gc / sc, seed + (uint32_t)o);
      amp = amp * 0.5;
    }
    v = 600.0 + 100.0 * acc;
    if (quant > 0) v = floor(v * quant + 0.5) / quant;
    if (na_permille > 0 && (int)(hash32((uint32_t)gr, (uint32_t)gc, seed + 99u) % 1000u) < na_permille)
      v = __longlong_as_double(0x7ff8000000000000LL);
  }
  elev[(size_t)R * d.pitch + C] = v;
}

// rows [row_lo, row_hi) of the array, hashed with the cell numbers of a DEM in which local row 0 is row `row_off`
template <typename T>
__global__ void k_checksum(const T* __restrict__ a, GridDims d, unsigned long long* out, int row_lo, int row_hi, long long row_off) {
  unsigned long long acc = 0;
  for (int r = row_lo + blockIdx.y; r < row_hi; r += gridDim.y)
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < d.NC; c += gridDim.x * blockDim.x) {
      unsigned long long v;
      if (sizeof(T) == 8) {
        v = (unsigned long long)a[(size_t)r * d.pitch + c];
        if ((v & 0x7ff0000000000000ull) == 0x7ff0000000000000ull && (v & 0x000fffffffffffffull)) v = 0x7ff8000000000000ull;
      } else {
        v = (unsigned long long)a[(size_t)r * d.pitch + c];
      }
      unsigned long long idx = (unsigned long long)(row_off + r) * (unsigned long long)d.NC + (unsigned long long)c + 1ull;
      unsigned long long h = (v + 0x9E3779B97F4A7C15ull) * (idx * 0xBF58476D1CE4E5B9ull | 1ull);
      h ^= h >> 29;
      acc += h;
    }
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}
}  // namespace

int run_synth(litap_grid* g, int64_t row0, int64_t gnr, uint32_t seed, double quant, int na_permille) {
  const GridDims d = g->d;
  int64_t m = gnr > (d.NC - 2) ? gnr : (d.NC - 2);
  int O = 0;
  while ((int64_t(1) << (O + 1)) <= m) ++O;
  O -= 2;
  if (O < 1) O = 1;
  dim3 grid((d.NC + 31 + 255) / 256, d.NR);
  KERNEL("k_synth", k_synth<<<grid, 256, 0, ctx().stream>>>(g->elev, d, row0, gnr, O, seed, quant, na_permille));
  return LITAP_OK;
}

int run_checksum(const void* arr, int esize, const GridDims& d, uint64_t* out) { return run_checksum_rows(arr, esize, d, 0, d.NR, 0, out); }

int run_checksum_rows(const void* arr, int esize, const GridDims& d, int row_lo, int row_hi, long long row_off, uint64_t* out) {
  cudaStream_t st = ctx().stream;
  unsigned long long* acc = (unsigned long long*)pool_alloc(sizeof(unsigned long long));
  if (!acc) return set_error(LITAP_ENOMEM, "device allocation failed (checksum)");
  CUDA_TRY(cudaMemsetAsync(acc, 0, sizeof(*acc), st));
  dim3 grid((d.NC + 255) / 256 < 32 ? (d.NC + 255) / 256 : 32, d.NR < 2048 ? d.NR : 2048);
  if (esize == 8) k_checksum<<<grid, 256, 0, st>>>((const unsigned long long*)arr, d, acc, row_lo, row_hi, row_off);
  else if (esize == 4) k_checksum<<<grid, 256, 0, st>>>((const uint32_t*)arr, d, acc, row_lo, row_hi, row_off);
  else k_checksum<<<grid, 256, 0, st>>>((const uint8_t*)arr, d, acc, row_lo, row_hi, row_off);
  ctx().launches++;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, acc, sizeof(*acc), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  pool_free(acc);
  if (e != cudaSuccess) return set_error(LITAP_ECUDA, std::string("checksum: ") + cudaGetErrorString(e));
  return LITAP_OK;
}
