// oracle_synth.cpp — synthetic stateless hash-fBm DEM (TEST/BENCH INFRASTRUCTURE; see oracle.h).
// Spec.  Only + - * / floor in fp64 and uint32 integer hashing, so the CUDA
// generator (litap_b200/csrc/synth.cu, compiled -fmad=false) produces bit-identical grids.
//
//   hash32(ix,iy,s): h = ix*0x9E3779B1 ^ iy*0x85EBCA77 ^ s*0xC2B2AE3D  (uint32 wrap)
//                    h ^= h>>16; h *= 0x7FEB352D; h ^= h>>15; h *= 0x846CA68B; h ^= h>>16
//   vnoise(x,y,s):   ix=floor(x) fx=x-ix ux=fx*fx*(3-2*fx)  (same for y)
//                    a=h(ix,iy) b=h(ix+1,iy) c=h(ix,iy+1) d=h(ix+1,iy+1), each /2^32
//                    ((a*(1-ux)+b*ux)*(1-uy)) + ((c*(1-ux)+d*ux)*uy)
//   O = max(1, floor(log2(max(gnr,nc))) - 2)
//   elev(r,c) = 600 + 100 * sum_{o=0}^{O-1} 2^-o * vnoise(r/2^(O-1-o), c/2^(O-1-o), seed+o)   (o ascending)
//   quantise: floor(elev*quant + 0.5)/quant   (quant = 1000 -> 0.001 m)
//   NA cells: hash32(r, c, seed+99) % 1000 < na_permille
// r,c are 0-based GLOBAL interior coordinates (x = r is the row coordinate).
#include "oracle.h"
#include <cmath>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {
inline uint32_t hash32(uint32_t ix, uint32_t iy, uint32_t s) {
  uint32_t h = ix * 0x9E3779B1u ^ iy * 0x85EBCA77u ^ s * 0xC2B2AE3Du;
  h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
  return h;
}
inline double vnoise(double x, double y, uint32_t s) {
  double fx0 = std::floor(x), fy0 = std::floor(y);
  uint32_t ix = (uint32_t)(int64_t)fx0, iy = (uint32_t)(int64_t)fy0;
  double fx = x - fx0, fy = y - fy0;
  double ux = fx * fx * (3.0 - 2.0 * fx), uy = fy * fy * (3.0 - 2.0 * fy);
  const double k = 1.0 / 4294967296.0;
  double a = hash32(ix, iy, s) * k, b = hash32(ix + 1, iy, s) * k;
  double c = hash32(ix, iy + 1, s) * k, d = hash32(ix + 1, iy + 1, s) * k;
  return ((a * (1.0 - ux) + b * ux) * (1.0 - uy)) + ((c * (1.0 - ux) + d * ux) * uy);
}
}  // namespace

extern "C" void oracle_synth_dem(double* elev, int64_t nr, int64_t nc, int64_t row0, int64_t gnr,
                                 uint32_t seed, double quant, int na_permille, int nthreads) {
  const int64_t NR = nr + 2, NC = nc + 2;
  int64_t m = gnr > nc ? gnr : nc;
  int O = 0;
  while ((int64_t(1) << (O + 1)) <= m) ++O;  // floor(log2(m))
  O -= 2;
  if (O < 1) O = 1;
  (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
#endif
  for (int64_t R = 0; R < NR; ++R)
    for (int64_t C = 0; C < NC; ++C) {
      int64_t gr = row0 + R - 1, gc = C - 1;  // global interior coords
      bool ring = (C == 0 || C == NC - 1 || gr < 0 || gr >= gnr);
      // rows above/below a band that are inside the global DEM are real halo rows
      double v;
      if (ring) {
        v = NAN;
      } else {
        double acc = 0.0, amp = 1.0;
        for (int o = 0; o < O; ++o) {
          double sc = (double)(int64_t(1) << (O - 1 - o));
          acc = acc + amp * vnoise((double)gr / sc, (double)gc / sc, seed + (uint32_t)o);
          amp = amp * 0.5;
        }
        v = 600.0 + 100.0 * acc;
        if (quant > 0) v = std::floor(v * quant + 0.5) / quant;
        if (na_permille > 0 && (int)(hash32((uint32_t)gr, (uint32_t)gc, seed + 99u) % 1000u) < na_permille) v = NAN;
      }
      elev[R * NC + C] = v;
    }
}
