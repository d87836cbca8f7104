// oracle_pitstat.cpp — CPU restatement of pit_stat1, the pour point and pit statistics of every watershed
// (TEST INFRASTRUCTURE; see oracle.h).  Follows R/flow_03_pit_stat_fast.R:1-162; neighbours as nb_values
// (R/LITAP_functions.R:207-235).  Called by flow_mapper (R/flow_mapper.R:181,222,378,416) and by every round of pit
// removal (R/flow_04_remove_pit_fast.R:19,44,90,136,240,270).
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <map>
#include <vector>

namespace {
const int DR[10] = {0, +1, +1, +1, 0, 0, 0, -1, -1, -1};
const int DC[10] = {0, -1, 0, +1, -1, 0, +1, -1, 0, +1};
struct Row {          // one row of `pp`: a ridge cell and one neighbour in another shed
  int64_t seq;        // 0-based index of the focal cell
  int n;              // neighbour code
  int32_t shed, shed_n;
  double elev, elev_n, pour;
  int64_t up;
};
}  // namespace

extern "C" int oracle_pit_stat(const double* elev, const uint8_t* ddir, const int32_t* shed, const int64_t* upslope,
                               const uint8_t* ridge, const uint8_t* edge_map, int64_t NR, int64_t NC, int method,
                               int64_t cap, double* const* out, int64_t* nrows) {
  const int64_t NB = NR * NC;
  std::map<int32_t, int> present;
  for (int64_t i = 0; i < NB; ++i)
    if (shed[i] > 0) present[shed[i]] = 1;
  auto put = [&](int64_t r, int col, double v) { if (out[col] && r < cap) out[col][r] = v; };
  if (present.size() <= 1) {            // :127-151, "If only one watershed"
    *nrows = present.empty() ? 0 : 1;
    if (present.empty()) return 0;
    for (int c = 0; c < ORACLE_PITSTAT_NCOL; ++c) put(0, c, NAN);
    put(0, 0, present.begin()->first);
    put(0, 1, 1);                        // shed_area = seq_along(pit_vol)
    put(0, 2, 1);                        // edge_pit = TRUE
    for (int64_t i = 0; i < NB; ++i)
      if (!std::isnan(elev[i]) && ddir[i] == 5) {
        put(0, 5, elev[i]); put(0, 6, (double)(i + 1)); put(0, 7, (double)(i / NC + 1)); put(0, 8, (double)(i % NC + 1));
        break;
      }
    return 0;
  }
  // ---- :24-33 ridge cells of the sheds with their neighbours in OTHER sheds
  std::vector<Row> pp;
  for (int64_t i = 0; i < NB; ++i) {
    if (shed[i] <= 0 || ridge[i] != 1) continue;
    const size_t first = pp.size();
    for (int n = 1; n <= 9; ++n) {       // row order of nb_values' long format: n ascending
      const int64_t j = i + DR[n] * NC + DC[n];
      if (j < 0 || j >= NB) continue;
      if (shed[j] <= 0 || shed[j] == shed[i]) continue;
      Row r;
      r.seq = i; r.n = n; r.shed = shed[i]; r.shed_n = shed[j]; r.elev = elev[i]; r.elev_n = elev[j];
      r.pour = std::max(elev[i], elev[j]);   // :41
      r.up = upslope[i];
      pp.push_back(r);
    }
    const size_t last = pp.size();
    if (last == first) continue;
    if (method == 0) {                   // "litap" :46-48: the lowest neighbour(s)
      double mn = INFINITY;
      for (size_t k = first; k < last; ++k) mn = std::min(mn, pp[k].elev_n);
      size_t w = first;
      for (size_t k = first; k < last; ++k)
        if (pp[k].elev_n == mn) pp[w++] = pp[k];
      pp.resize(w);
    } else {                             // "landmapr" :53-63
      static const int np[10] = {0, 7, 8, 9, 4, 5, 6, 1, 2, 3};
      std::vector<Row> rows(pp.begin() + first, pp.end());
      std::stable_sort(rows.begin(), rows.end(), [&](const Row& a, const Row& b) { return np[a.n] < np[b.n]; });
      double cmin = INFINITY;
      int best = -1;
      for (size_t k = 0; k < rows.size(); ++k) {
        const double thr = cmin;                       // lag(cummin(pour_elev), default = Inf)
        cmin = std::min(cmin, rows[k].pour);
        if (rows[k].elev_n < thr) best = (int)k;       // filter(...); lm_n == max(lm_n) -> the last one kept
      }
      pp.resize(first);
      if (best >= 0) pp.push_back(rows[best]);
    }
  }
  // ---- :65-67 lowest pour point of every shed, ties ordered by (elev, upslope, seqno)
  std::map<int32_t, double> minpour;
  for (const Row& r : pp) {
    auto it = minpour.find(r.shed);
    if (it == minpour.end() || r.pour < it->second) minpour[r.shed] = r.pour;
  }
  std::vector<Row> low;
  for (const Row& r : pp)
    if (r.pour == minpour[r.shed]) low.push_back(r);
  std::stable_sort(low.begin(), low.end(), [](const Row& a, const Row& b) {
    if (a.shed != b.shed) return a.shed < b.shed;
    if (a.elev != b.elev) return a.elev < b.elev;
    if (a.up != b.up) return a.up < b.up;
    return a.seq < b.seq;
  });
  struct Sel { Row r; int64_t omit_cell; double omit_elev; };
  std::map<int32_t, Sel> sel;
  for (size_t a = 0; a < low.size();) {
    size_t b = a;
    while (b < low.size() && low[b].shed == low[a].shed) ++b;
    Sel s;
    s.r = low[a];                                      // slice(1) :78
    std::vector<int64_t> seqs;
    long double oe = 0.0L;                             // :73-76 (landmapr)
    for (size_t k = a; k < b; ++k) {
      seqs.push_back(low[k].seq);
      if (k > a) oe += low[k].elev;
    }
    std::sort(seqs.begin(), seqs.end());
    s.omit_cell = (int64_t)(std::unique(seqs.begin(), seqs.end()) - seqs.begin()) - 1;
    s.omit_elev = (double)oe;
    sel[low[a].shed] = s;
    a = b;
  }
  // ---- :93-111 the statistics of every shed that has a pour point
  struct Acc { int64_t shed_area = 0, pit_area = 0; bool edge = false; long double vol = 0.0L; int64_t pit = -1; };
  std::map<int32_t, Acc> acc;
  for (int64_t i = 0; i < NB; ++i) {
    const int32_t s = shed[i];
    if (s <= 0) continue;
    auto it = sel.find(s);
    if (it == sel.end()) continue;
    Acc& a = acc[s];
    ++a.shed_area;
    const double pour = it->second.r.pour;
    if (!(elev[i] <= pour)) continue;
    ++a.pit_area;
    a.edge = a.edge || edge_map[i] == 1;
    const double term = pour - elev[i];
    a.vol += term;
    if (ddir[i] == 5 && a.pit < 0) a.pit = i;
  }
  int64_t row = 0;
  for (auto& kv : sel) {                                // groups come out in shedno order
    const Sel& s = kv.second;
    const Acc& a = acc[kv.first];
    const Row& r = s.r;
    const int64_t j = r.seq + DR[r.n] * NC + DC[r.n];
    double pit_area = (double)a.pit_area, pit_vol = (double)a.vol;
    double varatio = a.shed_area > 0 ? pit_vol / (double)a.shed_area * 1000 : 0;
    if (method == 1) {                                  // :114-123
      pit_area = pit_area - (double)s.omit_cell;
      pit_vol = pit_vol - (r.pour * (double)s.omit_cell - s.omit_elev);
      varatio = a.shed_area > 0 ? pit_vol / (double)a.shed_area * 1000 : 0;
    }
    put(row, 0, kv.first);
    put(row, 1, (double)a.shed_area);
    put(row, 2, a.edge ? 1 : 0);
    put(row, 3, pit_area);
    put(row, 4, pit_vol);
    put(row, 5, a.pit >= 0 ? elev[a.pit] : NAN);
    put(row, 6, a.pit >= 0 ? (double)(a.pit + 1) : NAN);
    put(row, 7, a.pit >= 0 ? (double)(a.pit / NC + 1) : NAN);
    put(row, 8, a.pit >= 0 ? (double)(a.pit % NC + 1) : NAN);
    put(row, 9, 0);
    put(row, 10, varatio);
    put(row, 11, (double)(r.seq + 1));
    put(row, 12, (double)(r.seq / NC + 1));
    put(row, 13, (double)(r.seq % NC + 1));
    put(row, 14, r.elev);
    put(row, 15, (double)(j + 1));
    put(row, 16, (double)(j / NC + 1));
    put(row, 17, (double)(j % NC + 1));
    put(row, 18, r.elev_n);
    put(row, 19, r.shed_n);
    put(row, 20, r.pour);
    ++row;
  }
  *nrows = row;
  return 0;
}
