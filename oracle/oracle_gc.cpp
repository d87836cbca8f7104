// oracle_gc.cpp — CPU restatement of slope_gc() (R/flow_05_slope_gc.R:73-164): row / column slope
// gradients and curvatures and the hillslope run labels.  TEST INFRASTRUCTURE (see oracle.h).
//
// Written to follow the R code step by step, including what dplyr does with NA:
//   * nb_values (R/LITAP_functions.R:207-235) addresses neighbours by LINEAR seqno (+-1, +-max_cols)
//     with no wrap check; out of range -> NA.
//   * the zero-sign loops are synchronous (Jacobi) rounds with the "no progress -> +0.00001" exit.
//   * cumsum() turns NA for the rest of the group once one term is NA; factor() renumbers the
//     distinct non-NA values of a group 1..k; 1:n() counts inside (group, label) with NA as a label.
#include <cmath>
#include <cstdint>
#include <map>
#include <vector>
#include "oracle.h"

namespace {
const double NA = NAN;
inline bool na(double x) { return x != x; }
inline double sgn(double x) { return na(x) ? NA : (x > 0 ? 1.0 : (x < 0 ? -1.0 : 0.0)); }

// value of v at linear index i (0-based), NA outside [0, n)
inline double at(const std::vector<double>& v, int64_t i) { return (i < 0 || i >= (int64_t)v.size()) ? NA : v[(size_t)i]; }

// the zero-sign loop of :103-118 (prev = -1: n4 / next = +1: n6) and :122-135 (prev = +NC: n2 / next = -NC: n8)
void fix_zero_signs(std::vector<double>& g, int64_t prev, int64_t next) {
  const int64_t n = (int64_t)g.size();
  auto zeros = [&]() { int64_t k = 0; for (double x : g) k += (x == 0); return k; };
  int64_t nz = zeros();
  while (nz > 0) {
    std::vector<double> old(g);
    for (int64_t i = 0; i < n; ++i) {
      if (!(old[i] == 0)) continue;  // NA == 0 is NA: if_else keeps NA
      const double p = at(old, i + prev);
      const double s = na(p) ? sgn(at(old, i + next)) : sgn(p);
      g[i] = 0.00001 * s;  // may be NA or stay 0
    }
    const int64_t now = zeros();
    if (now == nz) {
      for (double& x : g) if (x == 0) x = 0.00001;
      nz = 0;
    } else nz = now;
  }
}

// one group (a row or a column) of the hillslope labelling, :145-160; idx = linear indices in data order
void label_group(const std::vector<int64_t>& idx, const double* elev, const std::vector<double>& grad,
                 int32_t* hill_n, int32_t* hill_cell) {
  const size_t m = idx.size();
  std::vector<double> pre(m);  // cumsum(...) then NA where elev is NA
  double lag = NA, cum = 0;
  bool poisoned = false;
  for (size_t k = 0; k < m; ++k) {
    const double s = sgn(grad[idx[k]]);
    // sign != lag_s | is.na(lag_s)
    double flag;
    if (na(lag)) flag = 1;                 // (NA | TRUE) = TRUE
    else if (na(s)) flag = NA;             // (NA | FALSE) = NA
    else flag = (s != lag) ? 1 : 0;
    if (poisoned || na(flag)) { poisoned = true; cum = NA; }
    else cum += flag;
    pre[k] = na(elev[idx[k]]) ? NA : cum;
    lag = s;
  }
  // as.numeric(factor(x)): dense rank of the distinct non-NA values
  std::map<double, int> lev;
  for (double v : pre) if (!na(v)) lev.emplace(v, 0);
  int r = 0;
  for (auto& kv : lev) kv.second = ++r;
  // 1:n() within (group, label); NA is one more label
  std::map<int, int> count;  // label (0 = NA) -> members so far
  for (size_t k = 0; k < m; ++k) {
    const int lb = na(pre[k]) ? 0 : lev[pre[k]];
    const int c = ++count[lb];
    hill_n[idx[k]] = lb;
    hill_cell[idx[k]] = na(elev[idx[k]]) ? 0 : c;
  }
}
}  // namespace

extern "C" int oracle_slope_gc(const double* elev, int64_t NR, int64_t NC, double grid,
                               double* sgre, double* sgr, double* sgcn, double* sgc, double* scr, double* scc,
                               uint8_t* hill_r_dir, uint8_t* hill_c_dir, int32_t* hill_r_n, int32_t* hill_r_cell,
                               int32_t* hill_c_n, int32_t* hill_c_cell) {
  const int64_t n = NR * NC;
  std::vector<double> e(elev, elev + n), ge((size_t)n), gn((size_t)n);
  for (int64_t i = 0; i < n; ++i) {
    const double e5 = e[i];
    double n2 = at(e, i + NC), n4 = at(e, i - 1), n6 = at(e, i + 1), n8 = at(e, i - NC);
    if (na(n2)) n2 = e5;
    if (na(n4)) n4 = e5;
    if (na(n6)) n6 = e5;
    if (na(n8)) n8 = e5;
    ge[i] = (n4 - n6) / (2 * grid);
    gn[i] = (n2 - n8) / (2 * grid);
    double r = std::fabs(ge[i]), c = std::fabs(gn[i]);
    scr[i] = (2 * e5 - n4 - n6) / (grid * grid);
    scc[i] = (2 * e5 - n2 - n8) / (grid * grid);
    if (r == 0) r = 0.00001;
    if (c == 0) c = 0.00001;
    sgr[i] = r;
    sgc[i] = c;
  }
  fix_zero_signs(ge, -1, +1);
  fix_zero_signs(gn, +NC, -NC);
  for (int64_t i = 0; i < n; ++i) {
    sgre[i] = ge[i];
    sgcn[i] = gn[i];
    hill_r_dir[i] = na(ge[i]) ? 0 : (ge[i] > 0 ? 2 : 4);
    hill_c_dir[i] = na(gn[i]) ? 0 : (gn[i] > 0 ? 1 : 3);
  }
  std::vector<int64_t> idx;
  for (int64_t r = 0; r < NR; ++r) {  // group_by(row)
    idx.clear();
    for (int64_t c = 0; c < NC; ++c) idx.push_back(r * NC + c);
    label_group(idx, elev, ge, hill_r_n, hill_r_cell);
  }
  for (int64_t c = 0; c < NC; ++c) {  // group_by(col)
    idx.clear();
    for (int64_t r = 0; r < NR; ++r) idx.push_back(r * NC + c);
    label_group(idx, elev, gn, hill_c_n, hill_c_cell);
  }
  return 0;
}
