// oracle_ddir.cpp — CPU restatement of calc_ddir2 (TEST INFRASTRUCTURE; see oracle.h).
// Follows R/flow_01_calc_ddir.R:2-118, R/LITAP_functions.R:207-273 and :576-617.
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace {
// neighbour code n -> (drow, dcol); R/LITAP_functions.R:212-220 with row growing southward
const int DR[10] = {0, +1, +1, +1, 0, 0, 0, -1, -1, -1};
const int DC[10] = {0, -1, 0, +1, -1, 0, +1, -1, 0, +1};
}  // namespace

extern "C" int oracle_get_dir(int64_t row, int64_t col, int64_t row_f, int64_t col_f, unsigned opts) {
  // R/LITAP_functions.R:576-617
  double h = (double)(col_f - col), v = (double)(row_f - row);
  double a = std::atan(std::fabs(h) / std::fabs(v)) * 180 / M_PI;
  int l = 0;
  if (h == 0 && v == 0) {
    l = 5;
  } else if (h > 0 && v <= 0) {
    if (a >= 67.5) l = 6;
    if (a >= 22.5 && a < 67.5) l = 9;
    if (a < 22.5) l = 8;
  } else if (h > 0 && v > 0) {
    if (a >= 67.5) l = 6;
    if (a >= 22.5 && a < 67.5) l = 3;
    if (a < 22.5) l = 2;
  } else if (h <= 0 && v <= 0) {
    if (a >= 67.5) l = 4;
    if (a >= 22.5 && a < 67.5) l = 7;
    if (a < 22.5) l = 8;
  } else {
    if (a >= 67.5) l = 4;
    if (a >= 22.5 && a < 67.5) l = 1;
    if (a < 22.5) l = 2;
  }
  if (l != 5 && !((opts >> l) & 1u)) {
    static const int closest[10][7] = {
        {0, 0, 0, 0, 0, 0, 0}, {4, 2, 7, 3, 8, 6, 9}, {1, 3, 4, 6, 7, 9, 8}, {6, 2, 9, 1, 8, 4, 7},
        {7, 1, 8, 2, 9, 3, 6}, {0, 0, 0, 0, 0, 0, 0}, {9, 3, 8, 2, 7, 1, 4}, {8, 4, 9, 1, 6, 2, 3},
        {7, 9, 4, 6, 1, 3, 2}, {8, 6, 7, 3, 4, 2, 1}};
    int best = 0;  // 0 = NA
    for (int k = 0; k < 7; ++k)
      if ((opts >> closest[l][k]) & 1u) { best = closest[l][k]; break; }
    l = best;
  }
  return l;
}

extern "C" void oracle_invert(const double* elev, int64_t n, double* out) {
  // R/flow_06_invert.R:1-5 : max(elev, na.rm=TRUE) - elev
  double m = -INFINITY;
  for (int64_t i = 0; i < n; ++i)
    if (!std::isnan(elev[i]) && elev[i] > m) m = elev[i];
  for (int64_t i = 0; i < n; ++i) out[i] = m - elev[i];
}

extern "C" int oracle_ddir(const double* elev, int64_t NR, int64_t NC, uint8_t* ddir, uint8_t* flatcell,
                           int64_t* drec, int* plateau_rounds, int64_t* patch_cells) {
  const int64_t NB = NR * NC;
  const double sqrt2 = std::sqrt(2.0);
  // nb_values has no wrap check and returns NA for seqno outside 1..NB (R/LITAP_functions.R:224);
  // with the NaN ring every in-range wrapped neighbour of a non-NaN cell is a real neighbour.
  auto nb = [&](int64_t i, int n) -> int64_t {
    int64_t j = i + DR[n] * NC + DC[n];
    return (j < 0 || j >= NB) ? -1 : j;
  };
  // ---- finddir2 (R/LITAP_functions.R:257-273)
  std::vector<int64_t> flats;
  for (int64_t i = 0; i < NB; ++i) {
    double e = elev[i];
    if (std::isnan(e)) { ddir[i] = 0; flatcell[i] = 2; continue; }  // ddir NA, flatcell NA
    double best = 0.0;
    int d = 5;
    for (int n = 1; n <= 9; ++n) {
      if (n == 5) continue;
      int64_t j = nb(i, n);
      if (j < 0) continue;
      double en = elev[j];
      if (std::isnan(en)) continue;  // replace_na(elev_diff, 0): never > 0
      double diff = e - en;
      if (n == 1 || n == 3 || n == 7 || n == 9) diff = diff / sqrt2;
      if (diff > best) { best = diff; d = n; }  // first (lowest n) among exact ties
    }
    ddir[i] = (uint8_t)d;
    flatcell[i] = (d == 5);
    if (d == 5) flats.push_back(i);
  }
  // ---- plateau rounds, synchronous (R/flow_01_calc_ddir.R:12-41)
  int rounds = 0;
  std::vector<std::pair<int64_t, uint8_t>> upd;
  while (true) {
    upd.clear();
    for (int64_t i : flats) {
      if (ddir[i] != 5) continue;
      double e = elev[i];
      for (int n = 1; n <= 9; ++n) {
        int64_t j = nb(i, n);
        if (j < 0) continue;
        double en = elev[j];
        if (std::isnan(en)) continue;
        if (!(e - en >= 0)) continue;
        if (ddir[j] == 5 || ddir[j] == 0) continue;
        upd.emplace_back(i, (uint8_t)n);
        break;
      }
    }
    if (upd.empty()) break;
    ++rounds;
    for (auto& u : upd) ddir[u.first] = u.second;
  }
  if (plateau_rounds) *plateau_rounds = rounds;
  // ---- flat-pit patches (R/flow_01_calc_ddir.R:48-107)
  std::vector<int64_t> P;  // ddir==5 cells with >=1 neighbour ddir==5, seqno order
  for (int64_t i : flats) {
    if (ddir[i] != 5) continue;
    for (int n = 1; n <= 9; ++n) {
      if (n == 5) continue;
      int64_t j = nb(i, n);
      if (j >= 0 && ddir[j] == 5) { P.push_back(i); break; }
    }
  }
  if (patch_cells) *patch_cells = (int64_t)P.size();
  if (!P.empty()) {
    std::vector<int64_t> comp(P.size(), -1);
    auto find = [&](int64_t cell) -> int64_t {
      auto it = std::lower_bound(P.begin(), P.end(), cell);
      return (it != P.end() && *it == cell) ? (it - P.begin()) : -1;
    };
    std::vector<uint8_t> newdir(P.size(), 5);
    std::vector<int64_t> stack, members;
    for (size_t s = 0; s < P.size(); ++s) {
      if (comp[s] >= 0) continue;
      members.clear();
      stack.assign(1, (int64_t)s);
      comp[s] = (int64_t)s;
      while (!stack.empty()) {
        int64_t k = stack.back();
        stack.pop_back();
        members.push_back(k);
        for (int n = 1; n <= 9; ++n) {
          if (n == 5) continue;
          int64_t j = nb(P[k], n);
          if (j < 0 || ddir[j] != 5) continue;
          int64_t q = find(j);
          if (q >= 0 && comp[q] < 0) { comp[q] = (int64_t)s; stack.push_back(q); }
        }
      }
      std::sort(members.begin(), members.end());  // seqno order
      // centre = (ceiling(median(col)), ceiling(median(row))), 1-based buffered coords (:78)
      std::vector<double> cols, rows;
      for (int64_t k : members) { cols.push_back((double)(P[k] % NC + 1)); rows.push_back((double)(P[k] / NC + 1)); }
      auto median = [](std::vector<double> v) {
        std::sort(v.begin(), v.end());
        size_t m = v.size();
        return (m % 2) ? v[m / 2] : (v[m / 2 - 1] + v[m / 2]) / 2.0;
      };
      double cc = std::ceil(median(cols)), cr = std::ceil(median(rows));
      double dmin = INFINITY;
      int64_t centre = -1;
      for (size_t t = 0; t < members.size(); ++t) {
        double dist = std::sqrt((cols[t] - cc) * (cols[t] - cc) + (rows[t] - cr) * (rows[t] - cr));
        if (dist < dmin) { dmin = dist; centre = members[t]; }  // first minimum = lowest seqno (:85)
      }
      int64_t row_f = P[centre] / NC + 1, col_f = P[centre] % NC + 1;
      for (int64_t k : members) {
        unsigned opts = 0;
        for (int n = 1; n <= 9; ++n) {
          if (n == 5) continue;
          int64_t j = nb(P[k], n);
          if (j >= 0 && ddir[j] == 5) opts |= 1u << n;
        }
        newdir[k] = (uint8_t)oracle_get_dir(P[k] / NC + 1, P[k] % NC + 1, row_f, col_f, opts);
      }
    }
    for (size_t s = 0; s < P.size(); ++s) ddir[P[s]] = newdir[s];
  }
  // ---- drec (R/LITAP_functions.R:237-254)
  if (drec)
    for (int64_t i = 0; i < NB; ++i)
      drec[i] = ddir[i] ? (i + DR[ddir[i]] * NC + DC[ddir[i]] + 1) : 0;
  return 0;
}
