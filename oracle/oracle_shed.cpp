// oracle_shed.cpp — CPU restatement of calc_shed4 / calc_upslopes("upslope")
// (TEST INFRASTRUCTURE; see oracle.h).  Follows R/flow_02_calc_shed.R:10-145.
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <numeric>
#include <vector>

namespace {
const int DR[10] = {0, +1, +1, +1, 0, 0, 0, -1, -1, -1};
const int DC[10] = {0, -1, 0, +1, -1, 0, +1, -1, 0, +1};
}  // namespace

extern "C" int oracle_shed(const double* elev, const uint8_t* ddir, int64_t NR, int64_t NC, int32_t* shed,
                           int64_t* upslope, uint8_t* ridge, int64_t* npits_out) {
  const int64_t NB = NR * NC;
  // drec (0-based, -1 = NA)
  std::vector<int64_t> drec(NB);
  for (int64_t i = 0; i < NB; ++i) {
    int d = ddir[i];
    int64_t j = d ? i + DR[d] * NC + DC[d] : -1;
    drec[i] = (j < 0 || j >= NB) ? -1 : j;
  }
  // ---- raw labels: pits numbered in seqno order (:14-20), propagated up the drec forest (:25-29).
  // shedno <- shedno[drec] until the non-NA count stops changing == label of the root of the chain;
  // chains that never reach a pit (cycles, NA drec) stay NA.  Done here with an explicit walk +
  // memo; state 0 = unknown, -1 = NA, -2 = on current walk.
  std::vector<int32_t> raw(NB, 0);
  int64_t npits = 0;
  for (int64_t i = 0; i < NB; ++i)
    if (ddir[i] == 5) raw[i] = (int32_t)(++npits);
  std::vector<int64_t> path;
  for (int64_t i = 0; i < NB; ++i) {
    if (raw[i] != 0) continue;
    path.clear();
    int64_t c = i;
    int32_t lab;
    while (true) {
      if (c < 0) { lab = -1; break; }
      if (raw[c] == -2) { lab = -1; break; }  // cycle
      if (raw[c] != 0) { lab = raw[c]; break; }
      raw[c] = -2;
      path.push_back(c);
      c = drec[c];
    }
    for (int64_t p : path) raw[p] = lab;
  }
  // ---- canonical numbering (:32-38): arrange(desc(elev), seqno), first appearance of each label.
  // == rank sheds by (max elev in shed desc, seqno of that cell asc).
  std::vector<double> kelev(npits + 1, -INFINITY);
  std::vector<int64_t> kseq(npits + 1, -1);
  for (int64_t i = 0; i < NB; ++i) {
    int32_t l = raw[i];
    if (l <= 0) continue;
    double e = elev[i];
    if (kseq[l] < 0 || e > kelev[l]) { kelev[l] = e; kseq[l] = i; }  // i ascending: ties keep lowest seqno
  }
  std::vector<int32_t> order(npits);
  std::iota(order.begin(), order.end(), 1);
  std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
    if (kelev[a] != kelev[b]) return kelev[a] > kelev[b];
    return kseq[a] < kseq[b];
  });
  std::vector<int32_t> rank(npits + 1, 0);
  for (int64_t k = 0; k < npits; ++k) rank[order[k]] = (int32_t)(k + 1);
  for (int64_t i = 0; i < NB; ++i) shed[i] = raw[i] > 0 ? rank[raw[i]] : 0;
  // ---- upslope (:60-116): number of cells whose drec path passes through the cell (incl. itself);
  // 0 for cells without a shed.  Topological accumulation by in-degree.
  std::vector<int32_t> indeg(NB, 0);
  for (int64_t i = 0; i < NB; ++i) {
    upslope[i] = shed[i] ? 1 : 0;
    if (shed[i] && drec[i] >= 0 && drec[i] != i) indeg[drec[i]]++;
  }
  for (int64_t i = 0; i < NB; ++i) {
    if (!shed[i] || indeg[i] != 0) continue;
    int64_t c = i;
    indeg[i] = -1;  // done
    while (true) {
      int64_t d = drec[c];
      if (d < 0 || d == c) break;
      upslope[d] += upslope[c];
      if (--indeg[d] != 0) break;
      indeg[d] = -1;  // complete: continue downstream with its final count
      c = d;
    }
  }
  // ---- ridge (:48-57): non-ring cells; >1 distinct non-NA shed in the 3x3 window
  for (int64_t r = 0; r < NR; ++r)
    for (int64_t c = 0; c < NC; ++c) {
      int64_t i = r * NC + c;
      if (r == 0 || c == 0 || r == NR - 1 || c == NC - 1) { ridge[i] = 2; continue; }
      int32_t first = 0;
      bool multi = false;
      for (int n = 1; n <= 9; ++n) {
        int32_t s = shed[i + DR[n] * NC + DC[n]];
        if (!s) continue;
        if (!first) first = s;
        else if (s != first) multi = true;
      }
      ridge[i] = multi;
    }
  if (npits_out) *npits_out = npits;
  return 0;
}
