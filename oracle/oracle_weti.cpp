// oracle_weti.cpp — CPU restatement of calc_weti2 (TEST INFRASTRUCTURE; see oracle.h).
// Follows R/form_02b_calc_weti.R:3-110 literally, including the variables qc1/qc2 that
// survive from one loop iteration to the next (:70-71,:80 vs :87-88).
#include "oracle.h"
#include <algorithm>
#include <cmath>
#include <vector>

namespace {
const int DR[10] = {0, +1, +1, +1, 0, 0, 0, -1, -1, -1};
const int DC[10] = {0, -1, 0, +1, -1, 0, +1, -1, 0, +1};
inline double trunc_dec2(double x) { return std::trunc(x * 100.0) / 100.0; }  // R/LITAP_utils.R:93-95
}  // namespace

extern "C" int oracle_weti(const double* elev, const uint8_t* ddir, const double* upslope, int64_t NR,
                           int64_t NC, double grid, double* qarea1, double* qarea2, double* qweti1,
                           double* qweti2) {
  const int64_t NB = NR * NC;
  const double l1 = grid * 0.5, l2 = grid * 0.354;  // :5-6
  const double orth = grid, diag = grid * std::sqrt(2.0);  // :8-9
  std::vector<double> qa1(NB, 1.0), qa2(NB, grid * grid), qw1(NB, NAN), qw2(NB, NAN);
  // arrange(desc(elev), upslope, seqno), NA last, ring dropped (:23-26); NaN-elev cells are
  // skipped in the loop (:61) so they are simply not listed.
  std::vector<int64_t> ord;
  ord.reserve(NB);
  for (int64_t r = 1; r < NR - 1; ++r)
    for (int64_t c = 1; c < NC - 1; ++c)
      if (!std::isnan(elev[r * NC + c])) ord.push_back(r * NC + c);
  std::sort(ord.begin(), ord.end(), [&](int64_t a, int64_t b) {
    if (elev[a] != elev[b]) return elev[a] > elev[b];
    double ua = upslope[a], ub = upslope[b];
    bool na = std::isnan(ua), nb = std::isnan(ub);
    if (na != nb) return nb;  // NA last
    if (!na && ua != ub) return ua < ub;
    return a < b;
  });
  // window visit order NW,N,NE,W,self,E,SW,S,SE (column-major walk of the [col,row] matrix, :62,:67)
  static const int W[9] = {7, 8, 9, 4, 5, 6, 1, 2, 3};
  double qc1 = NAN, qc2 = NAN;  // R would stop with "object 'qc1' not found" if read before set
  for (int64_t i : ord) {
    const double f = elev[i];
    double tn[9];
    long double s = 0.0L;  // R sum(): LDOUBLE accumulator (src/main/summary.c rsum)
    for (int k = 0; k < 9; ++k) {
      int n = W[k];
      double nv = elev[i + DR[n] * NC + DC[n]];
      bool dg = (n == 1 || n == 3 || n == 7 || n == 9);
      double t = ((f - nv) / (dg ? diag : orth)) * (dg ? l2 : l1);  // :64
      if (nv >= f) t = 0;                                            // :65 (NaN compares false)
      tn[k] = t;
      if (!std::isnan(t)) s += t;
    }
    const double S = (double)s;
    if (S != 0) {  // :69-77
      qc1 = qa1[i] / S;
      qc2 = qa2[i] / S;
      for (int k = 0; k < 9; ++k) {
        int64_t j = i + DR[W[k]] * NC + DC[W[k]];
        qa1[j] = qa1[j] + (qc1 * tn[k]);
        qa2[j] = qa2[j] + (qc2 * tn[k]);
      }
    } else if (ddir[i] != 5) {  // :79-84 ; qc2 is NOT assigned here
      qc1 = qa1[i] / 0.0001;
      int d = ddir[i];
      if (d) {
        int64_t j = i + DR[d] * NC + DC[d];
        qa1[j] = qa1[j] + qa1[i];
        qa2[j] = qa2[j] + qa2[i];
      }
    }
    qw1[i] = (qc1 > 1) ? std::log(qc1) : std::log(1 + qc1);  // :87-88
    qw2[i] = (qc2 > 1) ? std::log(qc2) : std::log(1 + qc2);
  }
  // :94-95 (second line indexes with the already-filled qw1, as written)
  for (int64_t i = 0; i < NB; ++i)
    if (std::isnan(qw1[i])) qw1[i] = qa1[i] / 0.0001;
  for (int64_t i = 0; i < NB; ++i)
    if (std::isnan(qw1[i])) qw2[i] = qa2[i] / 0.0001;
  for (int64_t i = 0; i < NB; ++i) {
    bool miss = std::isnan(elev[i]);  // :98-101
    qarea1[i] = miss ? NAN : trunc_dec2(qa1[i]);
    qarea2[i] = miss ? NAN : trunc_dec2(qa2[i]);
    qweti1[i] = miss ? NAN : trunc_dec2(qw1[i]);
    qweti2[i] = miss ? NAN : trunc_dec2(qw2[i]);
  }
  return 0;
}
