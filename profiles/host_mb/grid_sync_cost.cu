#include <cooperative_groups.h>
#include <cstdio>
namespace cg = cooperative_groups;
__global__ void k(int n, unsigned* out) {
  cg::grid_group g = cg::this_grid();
  unsigned acc = 0;
  for (int i = 0; i < n; ++i) { acc += i; g.sync(); }
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = acc;
}
int main() {
  unsigned* out; cudaMalloc(&out, 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int bps : {1, 2, 4, 5, 8}) for (int n : {0, 10, 100}) {
    int blocks = 148 * bps; void* args[] = {&n, &out};
    float best = 1e9;
    for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(a);
      cudaError_t e = cudaLaunchCooperativeKernel((const void*)k, dim3(blocks), dim3(256), args, 0, 0);
      cudaEventRecord(b); cudaEventSynchronize(b);
      if (e != cudaSuccess) { printf("launch failed %s\n", cudaGetErrorString(e)); break; }
      float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    printf("blocks/SM %d syncs %d: %.1f us\n", bps, n, best * 1e3);
  }
  return 0;
}
