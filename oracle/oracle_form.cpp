// oracle_form.cpp — CPU restatement of calc_form (TEST INFRASTRUCTURE; see oracle.h).
// Follows R/form_01_calc_form.R:4-115.  Expression trees are kept as written in R
// (left-to-right, no re-association; build with -ffp-contract=off).
#include "oracle.h"
#include <cmath>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

// base R round(x, digits), R >= 4.0.0 (src/nmath/fround.c, not vendored in the reference):
// of the two `digits`-decimal neighbours of x pick the nearer one in double arithmetic,
// ties -> even last digit.  digits == 0 -> nearbyint (ties to even).
extern "C" double oracle_r_round(double x, int dig) {
  if (std::isnan(x) || !std::isfinite(x) || x == 0.0) return x;
  if (dig == 0) return std::nearbyint(x);
  double sgn = 1.0;
  if (x < 0) { sgn = -1.0; x = -x; }
  // l10x + dig > DBL_DIG: already has fewer significant digits than can be rounded away
  double l10x = 0.301029995663981195 * (0.5 + std::logb(x));
  if (l10x + dig > 15) return sgn * x;
  double p10 = 1.0;
  for (int i = 0; i < dig; ++i) p10 *= 10.0;
  double x10 = p10 * x, i10 = std::floor(x10);
  double xd = i10 / p10, xu = std::ceil(x10) / p10;
  double du = xu - x, dd = x - xd;
  return sgn * ((dd < du || (dd == du && std::fmod(i10, 2.0) == 0.0)) ? xd : xu);
}

namespace {
inline double rad_deg(double x) { return (x * 180) / M_PI; }  // :114
inline double deg_rad(double x) { return (x * M_PI) / 180; }  // :115

// aspect() :69-81
inline double aspect_fn(double sx, double sy, double sp) {
  double la = rad_deg(std::acos(std::fabs(sx) / sp));
  if (sp <= 0) return 360;
  if (sx > 0 && sy > 0) return 270 + la;
  if (sx > 0 && sy < 0) return 270 - la;
  if (sx < 0 && sy > 0) return 90 - la;
  if (sx < 0 && sy < 0) return 90 + la;
  if (sx < 0 && sy == 0) return 90;
  if (sx > 0 && sy == 0) return 270;
  if (sx == 0 && sy < 0) return 180;
  if (sx == 0 && sy > 0) return 360;
  return NAN;
}

// prof_plan() :83-111 ; e[1..9] are elev_n*100
inline double prof_plan(double A, const double* e, double grid, double sp) {
  double x1 = 2 + std::sin(deg_rad(A));
  double y1 = 2 - std::cos(deg_rad(A));
  double x2 = 2 - std::sin(deg_rad(A));
  double y2 = 2 + std::cos(deg_rad(A));
  double z1, z2;
  if (A <= 90) {
    z1 = (2 - y1) * ((e[9] * (x1 - 2)) + (e[8] * (3 - x1)));
    z1 = z1 + ((y1 - 1) * ((e[6] * (x1 - 2)) + (e[5] * (3 - x1))));
    z2 = (3 - y2) * ((e[5] * (x2 - 1)) + (e[4] * (2 - x2)));
    z2 = z2 + ((y2 - 2) * ((e[2] * (x2 - 1)) + (e[1] * (2 - x2))));
  } else {
    z1 = (3 - y1) * ((e[6] * (x1 - 2)) + (e[5] * (3 - x1)));
    z1 = z1 + ((y1 - 2) * ((e[3] * (x1 - 2)) + (e[2] * (3 - x1))));
    z2 = (2 - y2) * ((e[8] * (x2 - 1)) + (e[7] * (2 - x2)));
    z2 = z2 + ((y2 - 1) * ((e[5] * (x2 - 1)) + (e[4] * (2 - x2))));
  }
  if (sp <= 0) return 0;
  return rad_deg(std::atan((((2 * e[5]) - z1 - z2) / (grid * grid))));
}
const int DR[10] = {0, +1, +1, +1, 0, 0, 0, -1, -1, -1};
const int DC[10] = {0, -1, 0, +1, -1, 0, +1, -1, 0, +1};
}  // namespace

extern "C" int oracle_form(const double* elev, int64_t NR, int64_t NC, double grid, double* slope_pct,
                           double* slope_deg, double* aspect, double* prof, double* plan, int nthreads) {
  const int64_t NB = NR * NC;
  (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
#endif
  for (int64_t i = 0; i < NB; ++i) {
    double e[10];
    int nna = 0;
    for (int n = 1; n <= 9; ++n) {
      int64_t j = i + DR[n] * NC + DC[n];
      double v = (j < 0 || j >= NB) ? NAN : elev[j];
      e[n] = v * 100;  // :10-11
      nna += std::isnan(v);
    }
    double sp = NAN, sd = NAN, as = NAN, pr = NAN, pl = NAN;
    if (nna > 0) {
      if (!std::isnan(elev[i])) as = 360;  // :23
    } else {
      double sx = (e[6] - e[4]) / (2 * grid);
      double sy = (e[2] - e[8]) / (2 * grid);
      sp = std::sqrt(sx * sx + sy * sy);
      sd = rad_deg(std::atan(sp / 100));
      as = aspect_fn(sx, sy, sp);
    }
    as = oracle_r_round(as, 0);  // :26
    if (nna == 0) {
      double pa = as > 180 ? as - 180 : as;
      double pla = (pa + 90) > 180 ? pa + 90 - 180 : pa + 90;
      pr = prof_plan(pa, e, grid, sp);
      pl = prof_plan(pla, e, grid, sp);
    }
    slope_pct[i] = oracle_r_round(sp, 3);  // :43-47
    slope_deg[i] = oracle_r_round(sd, 3);
    aspect[i] = oracle_r_round(as, 0);
    prof[i] = oracle_r_round(pr, 3);
    plan[i] = oracle_r_round(pl, 3);
  }
  // ---- edge replication (:54-57), buffered 1-based col 2<-3, NC-1<-NC-2, row 2<-3, NR-1<-NR-2
  double* outs[5] = {slope_pct, slope_deg, aspect, prof, plan};
  auto copy_col = [&](int64_t dst, int64_t src) {
    if (dst < 0 || src < 0 || dst >= NC || src >= NC) return;
    for (double* o : outs)
      for (int64_t r = 0; r < NR; ++r) o[r * NC + dst] = o[r * NC + src];
  };
  auto copy_row = [&](int64_t dst, int64_t src) {
    if (dst < 0 || src < 0 || dst >= NR || src >= NR) return;
    for (double* o : outs)
      for (int64_t c = 0; c < NC; ++c) o[dst * NC + c] = o[src * NC + c];
  };
  copy_col(1, 2);
  copy_col(NC - 2, NC - 3);
  copy_row(1, 2);
  copy_row(NR - 2, NR - 3);
  // ---- NA -> 0 where elev is not NA (:60-65); aspect is left alone
  for (int64_t i = 0; i < NB; ++i) {
    if (std::isnan(elev[i])) continue;
    if (std::isnan(prof[i])) prof[i] = 0;
    if (std::isnan(plan[i])) plan[i] = 0;
    if (std::isnan(slope_pct[i])) slope_pct[i] = 0;
    if (std::isnan(slope_deg[i])) slope_deg[i] = 0;
  }
  return 0;
}
