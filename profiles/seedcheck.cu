// Measures the relative error of rcp.approx.ftz.f64 / rsqrt.approx.ftz.f64 (MUFU.RCP64H / RSQ64H):
// the quick path of calc_form (cellmath.h) assumes seeds with >= 20 good bits.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -o seedcheck seedcheck.cu
#include <cstdio>
#include <cmath>
__global__ void k(double* out) {
  double mr = 0, ms = 0;
  unsigned long long s = 88172645463325252ull + threadIdx.x + blockIdx.x * 1024ull;
  for (int i = 0; i < 20000; ++i) {
    s ^= s << 13; s ^= s >> 7; s ^= s << 17;
    double m = 1.0 + (double)(s >> 11) * (1.0 / 9007199254740992.0);
    int e = (int)((s >> 3) & 63) - 32;
    double x = ldexp(m, e);
    double r, q;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(q) : "d"(x));
    double er = fabs(r * x - 1.0), es = fabs(q * q * x - 1.0) * 0.5;
    if (er > mr) mr = er;
    if (es > ms) ms = es;
  }
  out[2 * (blockIdx.x * blockDim.x + threadIdx.x)] = mr;
  out[2 * (blockIdx.x * blockDim.x + threadIdx.x) + 1] = ms;
}
int main() {
  const int n = 64 * 256;
  double* d; cudaMalloc(&d, n * 16);
  k<<<64, 256>>>(d);
  double* h = new double[2 * n];
  cudaMemcpy(h, d, n * 16, cudaMemcpyDeviceToHost);
  double mr = 0, ms = 0;
  for (int i = 0; i < n; ++i) { if (h[2 * i] > mr) mr = h[2 * i]; if (h[2 * i + 1] > ms) ms = h[2 * i + 1]; }
  printf("rcp.approx.f64 max rel err %.3e (2^%.1f)  rsqrt.approx.f64 max rel err %.3e (2^%.1f)\n", mr, log2(mr), ms, log2(ms));
  return 0;
}
