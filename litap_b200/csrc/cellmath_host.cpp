// cellmath_host.cpp — host build of cellmath.h for CPU unit tests of the per-cell arithmetic the
// CUDA kernels use (tests/test_cellmath_host.py).  Not part of liblitap_cuda.so.
#include "cellmath.h"
#include <cmath>

extern "C" {
void host_fill_trig(double* sin_t, double* cos_t) {
  for (int A = 0; A <= 180; ++A) {
    double t = ((double)A * M_PI) / 180;
    sin_t[A] = std::sin(t);
    cos_t[A] = std::cos(t);
  }
}
// initial D8 stencil only (no plateau / patch handling)
void host_d8(const double* elev, long long NR, long long NC, unsigned char* ddir, int slow) {
  for (long long r = 0; r < NR; ++r)
    for (long long c = 0; c < NC; ++c) {
      double nb[10];
      for (int n = 1; n <= 9; ++n) {
        long long rr = r + nb_dr(n), cc = c + nb_dc(n);
        nb[n] = (rr < 0 || cc < 0 || rr >= NR || cc >= NC) ? NAN : elev[rr * NC + cc];
      }
      double e = elev[r * NC + c];
      ddir[r * NC + c] = (unsigned char)(slow ? (is_nan(e) ? 0 : d8_slow(e, nb)) : d8_dir(e, nb));
    }
}
int host_get_dir(long long row, long long col, long long row_f, long long col_f, unsigned opts) {
  return get_dir_int(row, col, row_f, col_f, opts);
}
double host_r_round3(double x) { return r_round3(x); }
// raw per-cell form values (before edge replication / NA->0)
void host_form_raw(const double* elev, long long NR, long long NC, double grid, double* out5) {
  double st[181], ct[181];
  host_fill_trig(st, ct);
  long long NB = NR * NC;
  for (long long r = 0; r < NR; ++r)
    for (long long c = 0; c < NC; ++c) {
      double e[10];
      for (int n = 1; n <= 9; ++n) {
        long long rr = r + nb_dr(n), cc = c + nb_dc(n);
        e[n] = (rr < 0 || cc < 0 || rr >= NR || cc >= NC) ? NAN : elev[rr * NC + cc];
      }
      FormOut o = form_cell(e, grid, st, ct);
      long long i = r * NC + c;
      out5[i] = o.slope_pct; out5[NB + i] = o.slope_deg; out5[2 * NB + i] = o.aspect;
      out5[3 * NB + i] = o.prof; out5[4 * NB + i] = o.plan;
    }
}
}
