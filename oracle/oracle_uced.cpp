// oracle_uced.cpp — CPU restatement of calc_upslopes(type = "uced"), the upslope cumulative elevation drop
// (TEST INFRASTRUCTURE; see oracle.h).  Follows R/flow_02_calc_shed.R:60-145 (calc_upslopes :60-79, get_upslope3
// :81-109, uced_values :118-130, get_all_flow :133-145); call sites R/flow_04_remove_pit_fast.R:73,345.
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <numeric>
#include <vector>

namespace {
const int DR[10] = {0, +1, +1, +1, 0, 0, 0, -1, -1, -1};
const int DC[10] = {0, -1, 0, +1, -1, 0, +1, -1, 0, +1};
}  // namespace

extern "C" int oracle_uced(const double* elev, const uint8_t* ddir, const int32_t* shed, int64_t NR, int64_t NC,
                           double* uced, int64_t* work_out) {
  const int64_t NB = NR * NC;
  // drec (R/LITAP_functions.R:237-254): the cell ddir points at; a pit (ddir 5) is its own drec
  std::vector<int64_t> drec(NB, -1);
  for (int64_t i = 0; i < NB; ++i) {
    const int d = ddir[i];
    if (!d) continue;
    const int64_t j = i + DR[d] * NC + DC[d];
    drec[i] = (j < 0 || j >= NB) ? -1 : j;
  }
  // :72 arrange(desc(elev), seqno), then one group per shedno (:73): the visiting order inside a shed.  Sheds do
  // not interact (a track never leaves its shed), so one global pass in that order is the same thing.  The NA
  // group (w = NA, :91) is skipped: its cells keep NA.
  std::vector<int64_t> order;
  order.reserve(NB);
  for (int64_t i = 0; i < NB; ++i) {
    uced[i] = NAN;                                   // :87
    if (shed[i] > 0) order.push_back(i);
  }
  std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return elev[a] > elev[b]; });
  std::vector<int64_t> track, jj;
  std::vector<double> cur, nw;
  int64_t work = 0;
  for (int64_t cell : order) {
    if (!std::isnan(uced[cell])) continue;           // :97
    // the cell's row of get_all_flow(): follow drec_shed until the pit (its own drec -> NA, :134) or a repeat (:140)
    track.clear();
    int64_t c = cell;
    while (true) {
      track.push_back(c);
      const int64_t n = drec[c];
      if (n < 0 || n == c || shed[n] != shed[cell]) break;
      if (std::find(track.begin(), track.end(), n) != track.end()) break;
      c = n;
    }
    const size_t L = track.size();
    if (L < 2) continue;                             // :98
    // uced_values(track, db) :118-130
    cur.resize(L);
    nw.assign(L, 0.0);
    jj.clear();
    for (size_t i = 0; i < L; ++i) {
      cur[i] = uced[track[i]];
      if (std::isnan(cur[i])) jj.push_back((int64_t)i);   // :123 which(is.na(current))
    }
    for (size_t i = 1; i < L; ++i) {                 // :124 (ii = 2:length(track))
      long double s = 0.0L;                          // sum(): long double accumulator, src/main/summary.c rsum
      const double ei = elev[track[i]];
      for (int64_t j : jj) {
        if (j >= (int64_t)i) break;
        const double term = elev[track[j]] - ei;     // the vector elev[jj[jj < i]] - elev[i] is double
        s += term;
        ++work;
      }
      nw[i] = nw[i] + (double)s;
    }
    for (size_t i = 0; i < L; ++i) uced[track[i]] = (std::isnan(cur[i]) ? 0.0 : cur[i]) + nw[i];   // :128-129
  }
  if (work_out) *work_out = work;
  return 0;
}
