// noalloc.cu — litap_checksum_noalloc: the checksum of litap_grid_checksum (synth.cu k_checksum) for callers that may not
// allocate device memory (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  It has no counterpart in
// the reference.
//
// The existing entry point takes its accumulator from the device pool (pool_alloc -> cudaMalloc on a cold pool) and reads
// the result back with a stream synchronise.  This variant allocates nothing and synchronises nothing:
//
//   k_rowsum_noalloc   one block per raster row: the hashes of the row's cells, summed, into g_row_partial[row]
//   k_fold_noalloc     one block: the sum of g_row_partial[0 .. NR) -> *out (a device pointer the caller owns)
//
// g_row_partial is a file-scope __device__ array of LITAP_NOALLOC_MAX_ROWS words (512 KB) that every GPU which loads the
// library holds until exit.  In host code `g_row_partial` is only a stand-in for the symbol; the device address comes from
// cudaGetSymbolAddress, looked up once per device with that device current.  Every slot the second kernel reads is written
// by the first in the same call, so nothing has to be cleared between calls -- but all calls share the array: two calls
// that overlap on the GPU (different streams, or graphs replayed concurrently) race on it.
#include <mutex>
#include <string>
#include "litap_internal.h"

namespace {

constexpr int MAX_DEVICES = 64;
constexpr int RT = 256;    // threads of a row block
constexpr int FT = 1024;   // threads of the folding block

static __device__ __align__(256) unsigned long long g_row_partial[LITAP_NOALLOC_MAX_ROWS];

std::mutex g_mu;
unsigned long long* g_partial_addr[MAX_DEVICES];   // device address of g_row_partial on each device that has asked

// the cell hash of k_checksum (synth.cu), restated: value bits (NaNs made canonical) mixed with the 1-based cell number
template <typename T>
__device__ __forceinline__ unsigned long long cell_hash(T x, unsigned long long idx) {
  unsigned long long v = (unsigned long long)x;
  if (sizeof(T) == 8 && (v & 0x7ff0000000000000ull) == 0x7ff0000000000000ull && (v & 0x000fffffffffffffull))
    v = 0x7ff8000000000000ull;
  unsigned long long h = (v + 0x9E3779B97F4A7C15ull) * (idx * 0xBF58476D1CE4E5B9ull | 1ull);
  return h ^ (h >> 29);
}

template <int T>
__device__ __forceinline__ unsigned long long block_sum(unsigned long long acc) {
  __shared__ unsigned long long warp_acc[T / 32];
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) warp_acc[threadIdx.x >> 5] = acc;
  __syncthreads();
  acc = threadIdx.x < T / 32 ? warp_acc[threadIdx.x] : 0ull;
  if (threadIdx.x < 32)
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  return acc;   // valid in thread 0
}

template <typename T>
__global__ void __launch_bounds__(RT) k_rowsum_noalloc(const T* __restrict__ a, long long NC, long long pitch,
                                                       unsigned long long* __restrict__ partial) {
  const long long r = blockIdx.x;
  const T* row = a + (size_t)r * (size_t)pitch;
  const unsigned long long base = (unsigned long long)r * (unsigned long long)NC + 1ull;
  unsigned long long acc = 0;
  for (long long c = threadIdx.x; c < NC; c += RT) acc += cell_hash(row[c], base + (unsigned long long)c);
  acc = block_sum<RT>(acc);
  if (threadIdx.x == 0) partial[r] = acc;
}

__global__ void __launch_bounds__(FT) k_fold_noalloc(const unsigned long long* __restrict__ partial, long long NR,
                                                     unsigned long long* __restrict__ out) {
  unsigned long long acc = 0;
  for (long long r = threadIdx.x; r < NR; r += FT) acc += partial[r];
  acc = block_sum<FT>(acc);
  if (threadIdx.x == 0) *out = acc;
}

// once per device, with that device current: where g_row_partial lives there
int partial_address(unsigned long long** out) {
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= MAX_DEVICES) return set_error(LITAP_EARG, "litap_checksum_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_partial_addr[dev]) {
    unsigned long long* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_row_partial);
    if (e != cudaSuccess || !p)
      return set_error(LITAP_ECUDA, std::string("litap_checksum_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    g_partial_addr[dev] = p;
  }
  *out = g_partial_addr[dev];
  return LITAP_OK;
}

}  // namespace

extern "C" int litap_checksum_noalloc(const void* dev_arr, int elem_size, int64_t NR, int64_t NC, int64_t pitch,
                                      uint64_t* dev_out, void* cuda_stream) {
  if (!dev_arr || !dev_out) return set_error(LITAP_EARG, "litap_checksum_noalloc: NULL");
  if (elem_size != 1 && elem_size != 4 && elem_size != 8)
    return set_error(LITAP_EARG, "litap_checksum_noalloc: elem_size must be 1, 4 or 8");
  if (NR < 1 || NC < 1 || pitch < NC) return set_error(LITAP_EARG, "litap_checksum_noalloc: needs NR >= 1, NC >= 1, pitch >= NC");
  if (NR > LITAP_NOALLOC_MAX_ROWS)
    return set_error(LITAP_EARG, "litap_checksum_noalloc: NR = " + std::to_string(NR) + " exceeds LITAP_NOALLOC_MAX_ROWS (" +
                                     std::to_string(LITAP_NOALLOC_MAX_ROWS) +
                                     "): the static scratch holds one partial sum per raster row");
  if ((uintptr_t)dev_arr % (size_t)elem_size || (uintptr_t)dev_out % 8)
    return set_error(LITAP_EARG, "litap_checksum_noalloc: misaligned pointer");
  unsigned long long* partial = nullptr;
  int rc = partial_address(&partial);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  const unsigned nblk = (unsigned)NR;
  if (elem_size == 8) k_rowsum_noalloc<<<nblk, RT, 0, st>>>((const unsigned long long*)dev_arr, NC, pitch, partial);
  else if (elem_size == 4) k_rowsum_noalloc<<<nblk, RT, 0, st>>>((const uint32_t*)dev_arr, NC, pitch, partial);
  else k_rowsum_noalloc<<<nblk, RT, 0, st>>>((const uint8_t*)dev_arr, NC, pitch, partial);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) {
    k_fold_noalloc<<<1, FT, 0, st>>>(partial, NR, (unsigned long long*)dev_out);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return set_error(LITAP_ECUDA, std::string("litap_checksum_noalloc: kernel launch: ") + cudaGetErrorString(e));
  return LITAP_OK;
}
