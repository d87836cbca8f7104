// oracle_relz.cpp — CPU restatement of calc_relz() + calc_length():
//   calc_stream3  R/form_03_calc_relz.R:194-276   first channel cell downstream (streams; ridges on the inverted DEM)
//   calc_pit3     R/form_03_calc_relz.R:424-479   pit of the cell's shed, steps to the end of the flow path
//   calc_relz     R/form_03_calc_relz.R:3-48      the four calls + joins
//   calc_relief   R/form_03_calc_relz.R:487-514   relief percentages
//   calc_length   R/form_04_calc_length.R:3-35    slope lengths
// TEST INFRASTRUCTURE (see oracle.h).  Follows the R code in its own order — the sorted start cells, the
// "first visited / first channel cell" cut of every track, the overwrites — so that it does not
// assume what the device path assumes (that the result only depends on the flow path of the cell).
// Conventions: grids are the buffered NR x NC grids of form_mapper (R/form_mapper.R:60-77); integer
// inputs 0 = NA; seqno / drec are 1-based; row / col outputs are buffered 1-based numbers like R's.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <limits>
#include <map>
#include <numeric>
#include <vector>
#include "oracle.h"

namespace {
const double NA = std::numeric_limits<double>::quiet_NaN();
inline bool na(double x) { return x != x; }

// trace_flow_all() + unique(): the flow path of a cell down to the cell that drains to itself
void track_of(int64_t start, const int32_t* drec, std::vector<int64_t>& t) {
  t.clear();
  int64_t c = start;
  for (;;) {
    t.push_back(c);
    const int64_t d = (int64_t)drec[c] - 1;
    if (d < 0 || d == c) break;
    c = d;
  }
}

struct Stream { std::vector<double> row, col, elev, z2, n2; };

// calc_stream3 :194-276
void stream3(const double* elev, const int32_t* drec, const int32_t* upslope, const int32_t* shedno, int64_t NR, int64_t NC,
             double str_val, Stream& o) {
  const int64_t n = NR * NC;
  o.row.assign(n, NA); o.col.assign(n, NA); o.elev.assign(n, NA); o.z2.assign(n, 0.0); o.n2.assign(n, NA);
  // filter(!is.na(elev)) %>% arrange(shedno, desc(elev), upslope): stable, NA last
  std::vector<int64_t> order;
  for (int64_t i = 0; i < n; ++i) if (!na(elev[i])) order.push_back(i);
  auto key_na_last = [](int32_t v) { return v == 0 ? std::numeric_limits<int64_t>::max() : (int64_t)v; };
  std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) {
    const int64_t sa = key_na_last(shedno[a]), sb = key_na_last(shedno[b]);
    if (sa != sb) return sa < sb;
    if (elev[a] != elev[b]) return elev[a] > elev[b];
    return key_na_last(upslope[a]) < key_na_last(upslope[b]);
  });
  std::vector<int64_t> t;
  for (int64_t i : order) {
    track_of(i, drec, t);
    const int64_t len = (int64_t)t.size();
    int64_t visited = -1, channel = -1;
    for (int64_t k = 0; k < len && visited < 0; ++k) if (o.z2[t[k]] > 0) visited = k;
    for (int64_t k = 0; k < len && channel < 0; ++k) if (upslope[t[k]] != 0 && (double)upslope[t[k]] >= str_val) channel = k;
    int64_t num_dn = len - 1;                           // index of `end` in the track
    if (visited >= 0) num_dn = std::min(num_dn, visited);
    if (channel >= 0) num_dn = std::min(num_dn, channel);
    const int64_t end = t[num_dn];
    // the cut track: is there a visited / channel cell in it
    const bool has_vis = visited >= 0 && visited <= num_dn, has_ch = channel >= 0 && channel <= num_dn;
    if ((!has_vis && !has_ch) || has_ch) {
      for (int64_t k = 0; k <= num_dn; ++k) {
        const int64_t c = t[k];
        o.row[c] = (double)(end / NC + 1); o.col[c] = (double)(end % NC + 1); o.elev[c] = elev[end];
        o.z2[c] = elev[c] - elev[end];
        o.n2[c] = (double)(num_dn - k);
      }
    } else {
      const double er = o.row[end], ec = o.col[end], ee = o.elev[end], en = o.n2[end];
      for (int64_t k = 0; k <= num_dn; ++k) {
        const int64_t c = t[k];
        o.row[c] = er; o.col[c] = ec; o.elev[c] = ee;
        o.z2[c] = elev[c] - ee;
        o.n2[c] = (double)(num_dn + 1) + en - 1 - (double)k;   // seq(num_dn + n2st[end] - 1, by = -1) with 1-based num_dn
      }
    }
  }
  for (double& z : o.z2) if (z < 0) z = 0;
}

struct Pit { std::vector<double> row, col, elev, z2, n2; };

// calc_pit3 :424-479.  pits: shedno -> (row, col, elev) from the pond table, or the ddir == 5 cells
void pit3(const double* elev, const int32_t* drec, const int32_t* upslope, const int32_t* shedno, int64_t NR, int64_t NC,
          const std::map<int32_t, std::array<double, 3>>& pits, Pit& o) {
  const int64_t n = NR * NC;
  o.row.assign(n, NA); o.col.assign(n, NA); o.elev.assign(n, NA); o.z2.assign(n, NA); o.n2.assign(n, NA);
  for (int64_t i = 0; i < n; ++i) {
    auto it = shedno[i] ? pits.find(shedno[i]) : pits.end();
    if (it == pits.end()) continue;
    o.row[i] = it->second[0]; o.col[i] = it->second[1]; o.elev[i] = it->second[2];
    double z = elev[i] - it->second[2];
    if (z < 0) z = 0;
    o.z2[i] = z;
  }
  // arrange(upslope, desc(elev)) %>% filter(!is.na(shedno), !is.na(elev))
  std::vector<int64_t> order;
  for (int64_t i = 0; i < n; ++i) if (shedno[i] != 0 && !na(elev[i])) order.push_back(i);
  std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) {
    const int64_t ua = upslope[a] ? upslope[a] : std::numeric_limits<int64_t>::max();
    const int64_t ub = upslope[b] ? upslope[b] : std::numeric_limits<int64_t>::max();
    if (ua != ub) return ua < ub;
    return elev[a] > elev[b];
  });
  std::vector<int64_t> t;
  for (int64_t i : order) {
    if (!na(o.n2[i])) continue;
    track_of(i, drec, t);
    const int64_t len = (int64_t)t.size();
    for (int64_t k = 0; k < len; ++k) o.n2[t[k]] = (double)(len - 1 - k);
  }
}

inline double sq(double x) { return x * x; }   // R: x^2 is x * x
inline double pct(double a, double b) { return std::trunc((a / b) * 100); }
}  // namespace

extern "C" int oracle_relz(int64_t NR, int64_t NC, const double* elev, const uint8_t* ddir, const int32_t* drec,
                           const int32_t* upslope, const int32_t* fill_shed, const int32_t* local_shed,
                           const double* ielev, const uint8_t* iddir, const int32_t* idrec, const int32_t* iupslope,
                           const int32_t* inv_local_shed, int64_t npond, const int32_t* pond_shedno,
                           const int32_t* pond_row, const int32_t* pond_col, const double* pond_elev, double str_val,
                           double ridge_val, double grid, double* const* out) {
  const int64_t n = NR * NC;
  Stream st, cr;
  stream3(elev, drec, upslope, fill_shed, NR, NC, str_val, st);              // calc_relz :7-10
  stream3(ielev, idrec, iupslope, inv_local_shed, NR, NC, ridge_val, cr);    // :18-24
  std::map<int32_t, std::array<double, 3>> pits, peaks;
  if (npond > 0) {
    for (int64_t k = 0; k < npond; ++k) pits[pond_shedno[k]] = {(double)pond_row[k], (double)pond_col[k], pond_elev[k]};
  } else {
    for (int64_t i = 0; i < n; ++i)
      if (ddir[i] == 5 && local_shed[i]) pits[local_shed[i]] = {(double)(i / NC + 1), (double)(i % NC + 1), elev[i]};
  }
  for (int64_t i = 0; i < n; ++i)
    if (iddir[i] == 5 && inv_local_shed[i]) peaks[inv_local_shed[i]] = {(double)(i / NC + 1), (double)(i % NC + 1), ielev[i]};
  Pit pt, pk;
  pit3(elev, drec, upslope, local_shed, NR, NC, pits, pt);                   // :12-15
  pit3(ielev, idrec, iupslope, inv_local_shed, NR, NC, peaks, pk);           // :28-34
  double mn = NA, mx = NA;
  for (int64_t i = 0; i < n; ++i) if (!na(elev[i])) { mn = na(mn) || elev[i] < mn ? elev[i] : mn; mx = na(mx) || elev[i] > mx ? elev[i] : mx; }
  std::map<int32_t, double> shed_max;   // max_na(elev) by fill_shed (0 = the NA group)
  for (int64_t i = 0; i < n; ++i) if (!na(elev[i])) {
    auto it = shed_max.find(fill_shed[i]);
    if (it == shed_max.end()) shed_max[fill_shed[i]] = elev[i];
    else if (elev[i] > it->second) it->second = elev[i];
  }
  const double range = mx - mn;
  for (int64_t i = 0; i < n; ++i) {
    double v[ORACLE_RELZ_NOUT];
    const double e = elev[i], row = (double)(i / NC + 1), col = (double)(i % NC + 1);
    const double z2st = na(e) ? NA : st.z2[i], z2cr = na(e) ? NA : cr.z2[i];          // :41
    const double cr_elev = mx - cr.elev[i], peak_elev = mx - pk.elev[i];             // :24, :34
    v[0] = st.row[i]; v[1] = st.col[i]; v[2] = st.elev[i]; v[3] = z2st; v[4] = st.n2[i];
    v[5] = cr.row[i]; v[6] = cr.col[i]; v[7] = cr_elev; v[8] = z2cr; v[9] = cr.n2[i];
    v[10] = pt.row[i]; v[11] = pt.col[i]; v[12] = pt.elev[i]; v[13] = pt.z2[i]; v[14] = pt.n2[i];
    v[15] = pk.row[i]; v[16] = pk.col[i]; v[17] = peak_elev; v[18] = pk.z2[i]; v[19] = pk.n2[i];
    // calc_relief :487-514
    auto sm = shed_max.find(fill_shed[i]);
    const double max_shed = sm == shed_max.end() ? NA : sm->second;
    const double z2pit = pt.z2[i], z2peak = pk.z2[i];
    const double zpit2peak = z2pit + z2peak, zcr2st = z2cr + z2st, zcr2pit = z2cr + z2pit;
    const double z2top = max_shed - e, ztop2pit = max_shed - pt.elev[i], ncr2st = cr.n2[i] + st.n2[i];
    double pctz2top = std::trunc(100 - (z2top / ztop2pit) * 100), pctz2st = pct(z2st, zcr2st);
    const double pctz2pit = pct(z2pit, zpit2peak);
    double pctn2st = pct(st.n2[i], ncr2st), pmin2max = pct(e - mn, range);
    if (ztop2pit == 0) pctz2top = 0;
    if (zcr2st == 0) pctz2st = 0;
    if (ncr2st == 0) pctn2st = 0;
    if (range == 0) pmin2max = 0;
    v[20] = z2top; v[21] = zpit2peak; v[22] = zcr2st; v[23] = zcr2pit; v[24] = ztop2pit; v[25] = ncr2st;
    v[26] = pmin2max; v[27] = pctz2st; v[28] = pctz2pit; v[29] = pctz2top; v[30] = pctn2st;
    // calc_length :9-33
    auto len3 = [&](double r2, double c2, double z) {
      const double l = std::sqrt(sq(c2 - col) + sq(r2 - row)) * grid;
      return std::sqrt(sq(l) + sq(z * grid));
    };
    const double l2pit = len3(pt.row[i], pt.col[i], z2pit), l2peak = len3(pk.row[i], pk.col[i], z2peak);
    const double l2str = len3(st.row[i], st.col[i], z2st), l2div = len3(cr.row[i], cr.col[i], z2cr);
    const double lpit2peak = l2pit + l2peak, lstr2div = l2str + l2div;
    double ppit2peakl = pct(l2pit, lpit2peak), pstr2divl = pct(l2str, lstr2div);
    const double r_lpit2peak = oracle_r_round(lpit2peak, 1), r_lstr2div = oracle_r_round(lstr2div, 1);
    if (r_lpit2peak == 0) ppit2peakl = 0;
    if (r_lstr2div == 0) pstr2divl = 0;
    v[31] = oracle_r_round(l2pit, 1); v[32] = oracle_r_round(l2peak, 1); v[33] = r_lpit2peak; v[34] = ppit2peakl;
    v[35] = oracle_r_round(l2str, 1); v[36] = oracle_r_round(l2div, 1); v[37] = r_lstr2div; v[38] = pstr2divl;
    for (int k = 0; k < ORACLE_RELZ_NOUT; ++k) if (out[k]) out[k][i] = v[k];
  }
  return 0;
}
