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
// n windows of nine doubles (neighbour codes 1..9, code 5 = the cell): d8_dir and the reference loop side by side
void host_d8_windows(const double* win, long long n, unsigned char* fast, unsigned char* slow, long long* nrare) {
  long long rare = 0;
  for (long long i = 0; i < n; ++i) {
    double nb[10];
    for (int k = 1; k <= 9; ++k) nb[k] = win[i * 9 + k - 1];
    const double e = nb[5];
    rare += d8_fast(e, nb) < 0;
    fast[i] = (unsigned char)d8_dir(e, nb);
    slow[i] = (unsigned char)(is_nan(e) ? 0 : d8_slow(e, nb));
  }
  *nrare = rare;
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
// fast path vs reference tree on every window: out[0] = windows the fast path decided, out[1] = of those,
// windows whose outputs differ in any bit from form_cell()
void host_form_fast_check(const double* elev, long long NR, long long NC, double grid, long long* out) {
  double st[181], ct[181];
  host_fill_trig(st, ct);
  FormK K;
  K.grid = grid; K.inv2g = 1.0 / (2 * grid); K.invg2 = 1.0 / (grid * grid);
  long long nfast = 0, nbad = 0;
  for (long long r = 0; r < NR; ++r)
    for (long long c = 0; c < NC; ++c) {
      double e[10];
      for (int n = 1; n <= 9; ++n) {
        long long rr = r + nb_dr(n), cc = c + nb_dc(n);
        e[n] = (rr < 0 || cc < 0 || rr >= NR || cc >= NC) ? NAN : elev[rr * NC + cc];
      }
      FormOut a, b = form_cell(e, grid, st, ct);
      if (!form_cell_fast(e, K, st, ct, &a)) continue;
      ++nfast;
      const double av[5] = {a.slope_pct, a.slope_deg, a.aspect, a.prof, a.plan};
      const double bv[5] = {b.slope_pct, b.slope_deg, b.aspect, b.prof, b.plan};
      bool same = true;
      for (int k = 0; k < 5; ++k) same &= (av[k] == bv[k]) || (av[k] != av[k] && bv[k] != bv[k]);
      nbad += !same;
    }
  out[0] = nfast;
  out[1] = nbad;
}
// quick path (the one k_form runs first) vs reference tree: out[0] = windows it decided, out[1] = of
// those, windows that differ in any bit from form_cell(), out[2] = windows (with 9 valid cells and a
// slope) it left in doubt
void host_form_quick_check(const double* elev, long long NR, long long NC, double grid, long long* out) {
  double st[181], ct[181], qw[181 * 3], qat[QK_ATAN_N];
  host_fill_trig(st, ct);
  form_fill_quick_tables(grid, st, ct, qw, qat);
  FormQ Q;
  Q.inv2g = 1.0 / (2 * grid);
  Q.gz = 8.0 * 4.440892098500626e-16 * 57295.78 / (grid * grid);
  long long nq = 0, nbad = 0, ndoubt = 0;
  for (long long r = 0; r < NR; ++r)
    for (long long c = 0; c < NC; ++c) {
      double e[10], z[10];
      bool na = false;
      for (int n = 1; n <= 9; ++n) {
        long long rr = r + nb_dr(n), cc = c + nb_dc(n);
        e[n] = (rr < 0 || cc < 0 || rr >= NR || cc >= NC) ? NAN : elev[rr * NC + cc];
        z[n] = e[n] * 100;
        na |= is_nan(e[n]);
      }
      FormOut a, b = form_cell(e, grid, st, ct);
      if (!form_cell_quick(z, Q, qw, qat, &a)) { ndoubt += (!na && b.slope_pct > 0); continue; }
      ++nq;
      const double av[5] = {a.slope_pct, a.slope_deg, a.aspect, a.prof, a.plan};
      const double bv[5] = {b.slope_pct, b.slope_deg, b.aspect, b.prof, b.plan};
      bool same = true;
      for (int k = 0; k < 5; ++k) same &= __builtin_memcmp(&av[k], &bv[k], 8) == 0;
      nbad += !same;
    }
  out[0] = nq;
  out[1] = nbad;
  out[2] = ndoubt;
}
}

// emulated long-double accumulation vs the host's real long double (x86-64: x87 extended)
extern "C" double host_ld80_sum(const double* t, int n) {
  Ld80 s;
  s.m = 0; s.e = 0;
  for (int i = 0; i < n; ++i) s = ld80_add(s, ld80_from_double(t[i]));
  return ld80_to_double(s);
}
extern "C" double host_real_ld_sum(const double* t, int n) {
  long double s = 0.0L;
  for (int i = 0; i < n; ++i) s += t[i];
  return (double)s;
}

// div_rcp (Markstein division from a precomputed reciprocal) vs the IEEE division: number of operand pairs, out of
// n, whose results differ in any bit.  Operands: random mantissas over a range of exponents, mantissas with long
// runs of ones / zeros, quotients next to a power of two, and dividends that are (nearly) integer multiples of b.
extern "C" long long host_div_rcp_check(long long n, unsigned long long seed) {
  unsigned long long s = seed * 0x9E3779B97F4A7C15ull + 1;
  auto next = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
  auto mk = [&](unsigned long long mant, int ex) {
    unsigned long long b = ((unsigned long long)(ex + 1023) << 52) | (mant & 0x000FFFFFFFFFFFFFull);
    double x; __builtin_memcpy(&x, &b, 8); return x;
  };
  long long bad = 0;
  for (long long i = 0; i < n; ++i) {
    unsigned long long ma = next(), mb = next();
    const unsigned kind = (unsigned)(next() & 7);
    if (kind == 1) { ma |= ~0ull << (next() & 31); mb |= ~0ull << (next() & 31); }       // runs of ones
    if (kind == 2) { ma &= ~0ull << (next() & 47); mb &= ~0ull << (next() & 47); }       // few significant bits
    if (kind == 3) mb = 0x000FFFFFFFFFFFFFull - (next() & 15);                            // b just below a power of two
    const int ea = (int)(next() % 61) - 30, eb = (int)(next() % 41) - 20;
    double b = mk(mb, eb), a = mk(ma, ea);
    if (kind == 4) a = b * (double)(1 + (next() & 1023));                                 // exact or near-exact quotients
    if (kind == 5) a = b * (double)(1 + (next() & 1023)) + mk(0, eb - 52 + (int)(next() & 3));
    const double y = div_rcp_prepare(b);
    const double q1 = div_rcp(a, b, y), q2 = a / b;
    bad += __builtin_memcmp(&q1, &q2, 8) != 0;
  }
  return bad;
}

// the signed accumulator (calc_upslopes "uced") vs the host's long double
extern "C" double host_lds_sum(const double* t, int n) {
  Lds s;
  s.m = 0; s.e = 0; s.neg = 0;
  for (int i = 0; i < n; ++i) s = lds_add(s, lds_from_double(t[i]));
  return lds_to_double(s);
}
