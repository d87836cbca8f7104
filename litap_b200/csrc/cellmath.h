// cellmath.h — per-cell arithmetic of the hot path, shared by the CUDA kernels and by the
// host-compiled unit test of the same functions (tests/test_cellmath_host.py builds
// csrc/cellmath_host.cpp).  Everything here must keep the reference's expression trees:
// compile with -fmad=false (nvcc) / -ffp-contract=off (g++).
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define LITAP_HD __host__ __device__ __forceinline__
#else
#define LITAP_HD inline
#endif

// neighbour code n (1..9) -> row / col step; R/LITAP_functions.R:212-220, rows grow southward
LITAP_HD int nb_dr(int n) { return n <= 3 ? 1 : (n <= 6 ? 0 : -1); }
LITAP_HD int nb_dc(int n) { return (int)((0x92490u >> (2 * n)) & 3u) - 1; }   // (n - 1) % 3 - 1 for n = 1..9, as a 2-bit table

LITAP_HD bool is_nan(double x) { return x != x; }

// ---------------------------------------------------------------------------------------------
// finddir2 (R/LITAP_functions.R:257-273): D8 steepest descent with diagonals divided by sqrt(2),
// exact ties resolved to the lowest direction code.  nb[n] = elevation of neighbour n (nb[5] unused).
// Returns 0 (NA) for a NaN centre, 5 for a flat cell / pit.
//
// d8_slow is the reference's loop verbatim (eight differences, four IEEE divisions); d8_dir gets the same
// answer from one exact division per cell.
LITAP_HD int d8_slow(double e, const double* nb) {
  const double sqrt2 = 1.4142135623730951;  // == sqrt(2.0) in binary64
  double best = 0.0;
  int d = 5;
#pragma unroll
  for (int n = 1; n <= 9; ++n) {
    if (n == 5) continue;
    double diff = e - nb[n];
    if (n == 1 || n == 3 || n == 7 || n == 9) diff = diff / sqrt2;
    if (diff > best) { best = diff; d = n; }  // NaN compares false: replace_na(.,0) never wins
  }
  return d;
}

// RN(a / RN(sqrt 2)) for 2^-900 < a < 2^900 (and a == 0) without the generic division: y = RN(1 / RN(sqrt 2)),
// two Newton corrections with exact residuals (Markstein 1990, as div_rcp below with the constants folded in).
#if defined(__CUDACC__)
__device__ __constant__ double c_sqrt2[2] = {1.4142135623730951, 0.7071067811865475};  // constant-bank operands
#endif
LITAP_HD double div_sqrt2(double a) {
#if defined(__CUDA_ARCH__)
  const double s = c_sqrt2[0], y = c_sqrt2[1];
#else
  const double s = 1.4142135623730951, y = 0.7071067811865475;  // 0x1.6a09e667f3bcdp+0, 0x1.6a09e667f3bccp-1
#endif
  double q = a * y;
  double r = fma(-q, s, a);
  q = fma(r, y, q);
  r = fma(-q, s, a);
  return fma(r, y, q);
}

// (m, i) = d > 0 ? (d, n) : (0, 5): the first step of a "first maximum" chain.  Spelled out in PTX on the device because
// the compiler turns the C form into an fmax with NaN quieting (six instructions instead of three).
LITAP_HD void d8_first(double d, int n, double& m, int& i) {
#if defined(__CUDA_ARCH__)
  asm("{\n .reg .pred p;\n setp.gt.f64 p, %2, 0d0000000000000000;\n selp.f64 %0, %2, 0d0000000000000000, p;\n"
      " selp.s32 %1, %3, 5, p;\n}"
      : "=d"(m), "=r"(i)
      : "d"(d), "r"(n));
#else
  const bool p = d > 0.0;
  m = p ? d : 0.0;
  i = p ? n : 5;
#endif
}

// Branch-free fast path: first maximum (strict >, lowest code on ties) of the four orthogonal and of the four diagonal
// raw differences; the division is monotone, so the largest diagonal quotient is RN(md / sqrt 2) - ONE exact division
// per cell.  A smaller diagonal difference can round to the same quotient only if it is the double just below md
// (the quotients of consecutive doubles are 0.707 or 1.414 ulp apart, so three in a row never share one); such a window
// - and one whose md is outside the range of exact residuals - takes the reference's loop.  Against the orthogonal
// maximum the exact quotient decides; on equality the lowest code of the two groups wins.
// d8_fast returns -1 for a window that needs d8_slow.
LITAP_HD int d8_fast(double e, const double* nb) {
  double mo, md;
  int io, id;
  const double d2 = e - nb[2], d4 = e - nb[4], d6 = e - nb[6], d8 = e - nb[8];
  const double d1 = e - nb[1], d3 = e - nb[3], d7 = e - nb[7], d9 = e - nb[9];
  d8_first(d2, 2, mo, io);
  if (d4 > mo) { mo = d4; io = 4; }
  if (d6 > mo) { mo = d6; io = 6; }
  if (d8 > mo) { mo = d8; io = 8; }
  d8_first(d1, 1, md, id);
  if (d3 > md) { md = d3; id = 3; }
  if (d7 > md) { md = d7; id = 7; }
  if (d9 > md) { md = d9; id = 9; }
  // the double below md (md == 0: a NaN, equal to nothing) and md's exponent field (md >= 0)
  unsigned long long mb;
#if defined(__CUDA_ARCH__)
  mb = (unsigned long long)__double_as_longlong(md);
  const double mdp = __longlong_as_double((long long)(mb - 1ull));
#else
  double mdp;
  {
    __builtin_memcpy(&mb, &md, 8);
    const unsigned long long u = mb - 1ull;
    __builtin_memcpy(&mdp, &u, 8);
  }
#endif
  const bool in_range = (unsigned)(mb >> 52) - 123u <= 1800u;   // 2^-900 <= md < 2^901; false for md == 0 (nothing is divided)
  const bool rare = (d1 == mdp) | (d3 == mdp) | (d7 == mdp) | (d9 == mdp) | ((id != 5) & !in_range);
  const double q = div_sqrt2(md);
  const int lowest = io < id ? io : id;
  int res = q > mo ? id : (q < mo ? io : lowest);
  if (rare) res = -1;
  return is_nan(e) ? 0 : res;   // a NaN centre: every difference is NaN, nothing above is taken (never rare)
}
LITAP_HD int d8_dir(double e, const double* nb) {
  const int d = d8_fast(e, nb);
  return d >= 0 ? d : d8_slow(e, nb);
}

// ---------------------------------------------------------------------------------------------
// get_dir (R/LITAP_functions.R:576-617) with integer arithmetic.  The reference compares
// a = atan(|h|/|v|)*180/pi against 22.5 and 67.5; tan(22.5 deg) = sqrt(2)-1 is irrational, so for
// integer h, v:  a >= 67.5 <=> |v| <= (sqrt(2)-1)|h| <=> (|v|+|h|)^2 <= 2 h^2 ;
//                a >= 22.5 <=> |h| >= (sqrt(2)-1)|v| <=> (|h|+|v|)^2 >= 2 v^2 (equalities unreachable).
LITAP_HD int get_dir_int(long long row, long long col, long long row_f, long long col_f, unsigned opts) {
  long long h = col_f - col, v = row_f - row;
  if (h == 0 && v == 0) return 5;
  unsigned long long ah = (unsigned long long)(h < 0 ? -h : h), av = (unsigned long long)(v < 0 ? -v : v);
  unsigned long long s = ah + av;
  bool ge675 = (av == 0) || (s * s <= 2 * ah * ah);
  bool ge225 = (av == 0) || (s * s >= 2 * av * av);
  int l;
  if (h > 0 && v <= 0) l = ge675 ? 6 : (ge225 ? 9 : 8);
  else if (h > 0 && v > 0) l = ge675 ? 6 : (ge225 ? 3 : 2);
  else if (h <= 0 && v <= 0) l = ge675 ? 4 : (ge225 ? 7 : 8);
  else l = ge675 ? 4 : (ge225 ? 1 : 2);
  if (!((opts >> l) & 1u)) {
    // closest[l], R/LITAP_functions.R:604-611, packed 7 nibbles per entry (first = lowest nibble)
    const unsigned closest[10] = {0, 0x9683724u, 0x8976431u, 0x7481926u, 0x6392817u, 0,
                                  0x4172839u, 0x3261948u, 0x2316497u, 0x1243768u};
    unsigned cl = closest[l];
    l = 0;
    for (int k = 0; k < 7; ++k) {
      int cand = (cl >> (4 * k)) & 15;
      if ((opts >> cand) & 1u) { l = cand; break; }
    }
  }
  return l;
}

// ---------------------------------------------------------------------------------------------
// base R round(x, 3) for R >= 4.0.0 (src/nmath/fround.c): nearer of the two 3-decimal neighbours
// evaluated in binary64, ties -> even.  NaN passes through.
LITAP_HD double r_round3(double x) {
  if (is_nan(x) || x == 0.0) return x;
  double ax = x < 0 ? -x : x;
  if (ax >= 1099511627776.0 /* 2^40: log10(2)*(0.5+logb(x)) + 3 > DBL_DIG */ || ax > 1.7976931348623157e308)
    return x;
  double x10 = 1000.0 * ax, i10 = floor(x10);
  double xd = i10 / 1000.0, xu = ceil(x10) / 1000.0;
  double du = xu - ax, dd = ax - xd;
  bool even = (((long long)i10) & 1LL) == 0;
  double r = (dd < du || (dd == du && even)) ? xd : xu;
  return x < 0 ? -r : r;
}

// base R round(x, 1), same algorithm (calc_length rounds its lengths to 0.1): the "enough digits already"
// cut-off log10(2)*(0.5+logb(x)) + 1 > DBL_DIG is logb(x) >= 47
LITAP_HD double r_round1(double x) {
  if (is_nan(x) || x == 0.0) return x;
  double ax = x < 0 ? -x : x;
  if (ax >= 140737488355328.0 /* 2^47 */) return x;
  double x10 = 10.0 * ax, i10 = floor(x10);
  double xd = i10 / 10.0, xu = ceil(x10) / 10.0;
  double du = xu - ax, dd = ax - xd;
  bool even = (((long long)i10) & 1LL) == 0;
  double r = (dd < du || (dd == du && even)) ? xd : xu;
  return x < 0 ? -r : r;
}

// ---------------------------------------------------------------------------------------------
// calc_form for one cell (R/form_01_calc_form.R:9-47).  e[n] = elev_n (NOT yet *100), n = 1..9,
// e[5] = centre.  sin_t / cos_t: tables of sin((A*pi)/180), cos((A*pi)/180) for integer A = 0..180,
// filled by the host libm (the GPU build uses the same table, so both follow glibc exactly).
// Outputs are the rounded values BEFORE edge replication and the final NA -> 0 step.
struct FormOut { double slope_pct, slope_deg, aspect, prof, plan; };

LITAP_HD double form_rad_deg(double x) { return (x * 180) / 3.14159265358979323846; }

// numerator (2*z5 - z1 - z2) of the curvature along aspect line A (R/form_01_calc_form.R:87-106)
LITAP_HD double form_curv_num(int A, const double* z, const double* sin_t, const double* cos_t) {
  double sn = sin_t[A], cs = cos_t[A];
  double x1 = 2 + sn, y1 = 2 - cs, x2 = 2 - sn, y2 = 2 + cs;
  // the two branches of the reference (A <= 90 / A > 90) are the same expression tree on different
  // operands: select the operands, evaluate the tree once (no divergence between lanes)
  const bool lo = A <= 90;
  const double a9 = lo ? z[9] : z[6], a8 = lo ? z[8] : z[5], a6 = lo ? z[6] : z[3], a5 = lo ? z[5] : z[2];
  const double w1 = (lo ? 2 : 3) - y1, w2 = y1 - (lo ? 1 : 2);
  const double b5 = lo ? z[5] : z[8], b4 = lo ? z[4] : z[7], b2 = lo ? z[2] : z[5], b1 = lo ? z[1] : z[4];
  const double v1 = (lo ? 3 : 2) - y2, v2 = y2 - (lo ? 2 : 1);
  double z1 = w1 * ((a9 * (x1 - 2)) + (a8 * (3 - x1)));
  z1 = z1 + (w2 * ((a6 * (x1 - 2)) + (a5 * (3 - x1))));
  double z2 = v1 * ((b5 * (x2 - 1)) + (b4 * (2 - x2)));
  z2 = z2 + (v2 * ((b2 * (x2 - 1)) + (b1 * (2 - x2))));
  return (2 * z[5]) - z1 - z2;
}

LITAP_HD double form_curv(int A, const double* z, double grid, double sp, const double* sin_t, const double* cos_t) {
  double sn = sin_t[A], cs = cos_t[A];
  double x1 = 2 + sn, y1 = 2 - cs, x2 = 2 - sn, y2 = 2 + cs;
  double z1, z2;
  if (A <= 90) {
    z1 = (2 - y1) * ((z[9] * (x1 - 2)) + (z[8] * (3 - x1)));
    z1 = z1 + ((y1 - 1) * ((z[6] * (x1 - 2)) + (z[5] * (3 - x1))));
    z2 = (3 - y2) * ((z[5] * (x2 - 1)) + (z[4] * (2 - x2)));
    z2 = z2 + ((y2 - 2) * ((z[2] * (x2 - 1)) + (z[1] * (2 - x2))));
  } else {
    z1 = (3 - y1) * ((z[6] * (x1 - 2)) + (z[5] * (3 - x1)));
    z1 = z1 + ((y1 - 2) * ((z[3] * (x1 - 2)) + (z[2] * (3 - x1))));
    z2 = (2 - y2) * ((z[8] * (x2 - 1)) + (z[7] * (2 - x2)));
    z2 = z2 + ((y2 - 1) * ((z[5] * (x2 - 1)) + (z[4] * (2 - x2))));
  }
  if (sp <= 0) return 0;
  return form_rad_deg(atan((((2 * z[5]) - z1 - z2) / (grid * grid))));
}

LITAP_HD FormOut form_cell(const double* e, double grid, const double* sin_t, const double* cos_t) {
  const double NA = NAN;
  FormOut o;
  bool any_na = false;
  double z[10];
#pragma unroll
  for (int n = 1; n <= 9; ++n) {
    any_na |= is_nan(e[n]);
    z[n] = e[n] * 100;
  }
  if (any_na) {
    o.slope_pct = NA; o.slope_deg = NA; o.prof = NA; o.plan = NA;
    o.aspect = is_nan(e[5]) ? NA : 360.0;
    return o;
  }
  double sx = (z[6] - z[4]) / (2 * grid);
  double sy = (z[2] - z[8]) / (2 * grid);
  double sp = sqrt(sx * sx + sy * sy);
  double sd = form_rad_deg(atan(sp / 100));
  double as;
  if (sp <= 0) as = 360;
  else {
    double la = form_rad_deg(acos(fabs(sx) / sp));
    if (sx > 0 && sy > 0) as = 270 + la;
    else if (sx > 0 && sy < 0) as = 270 - la;
    else if (sx < 0 && sy > 0) as = 90 - la;
    else if (sx < 0 && sy < 0) as = 90 + la;
    else if (sx < 0) as = 90;
    else if (sx > 0) as = 270;
    else if (sy < 0) as = 180;
    else as = 360;
  }
  as = nearbyint(as);  // round(aspect): ties to even
  double pa = as > 180 ? as - 180 : as;
  double pl = (pa + 90) > 180 ? pa + 90 - 180 : pa + 90;
  o.prof = r_round3(form_curv((int)pa, z, grid, sp, sin_t, cos_t));
  o.plan = r_round3(form_curv((int)pl, z, grid, sp, sin_t, cos_t));
  o.slope_pct = r_round3(sp);
  o.slope_deg = r_round3(sd);
  o.aspect = as;
  return o;
}

// ---------------------------------------------------------------------------------------------
// Fast path of calc_form.  Every output of calc_form is a rounded number (3 decimals, or an integer
// for aspect), so the reference's value is determined by the INTEGER k = round(1000 v) (or
// round(aspect)).  The fast path evaluates v with cheap fp64 arithmetic (reciprocal multiplies, a
// degree-9 arctangent polynomial; relative error < 1e-13, far below the margins used here), decides
// k, and returns false ("in doubt") whenever v is so close to a rounding boundary that the reference
// expression tree could decide differently; the caller then runs form_cell() (the reference's
// tree) for that cell.  When it returns true the outputs are bit-identical with form_cell():
// k/1000 is produced with the IEEE-correct quotient (div1000, verified exhaustively for k < 2^32).
#if defined(__CUDA_ARCH__)
#define LITAP_WARP_ANY(c) __any_sync(__activemask(), (c))
#else
#define LITAP_WARP_ANY(c) (c)
#endif
struct FormK {
  double grid, inv2g, invg2;  // grid, 1/(2 grid), 1/(grid*grid)
};

// RN(k / 1000) for an integer-valued double 0 <= k < 2^32, without a division
LITAP_HD double div1000(double k) {
  double q = k * 0.001;
  double r = fma(-q, 1000.0, k);
  return fma(r, 0.001, q);
}
// reciprocal with relative error ~1e-16 for 1e-30 <= x <= 1e30
LITAP_HD double fast_rcp(double x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"((float)x));
#else
  float r = 1.0f / (float)x;
#endif
  double y = (double)r;
  double e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  return y;
}
// atan(y) for |y| <= tan(pi/8) * 1.01: y * P(y^2), P of degree 9 (max relative error 1.3e-15)
LITAP_HD double fast_atan_core(double y) {
  double t = y * y;
  double p = -0.02496927912537512;
  p = fma(p, t, 0.04999137488007895);
  p = fma(p, t, -0.0649812787648163);
  p = fma(p, t, 0.0767224153108192);
  p = fma(p, t, -0.09089406239578209);
  p = fma(p, t, 0.11111041936867709);
  p = fma(p, t, -0.14285712445599388);
  p = fma(p, t, 0.1999999997503182);
  p = fma(p, t, -0.33333333333201454);
  p = fma(p, t, 0.9999999999999989);
  return y * p;
}
// atan(a) for 0 <= a <= 1e30
LITAP_HD double fast_atan_pos(double a) {
  // slope and curvature arguments are almost always below tan(pi/8): the two reductions sit behind
  // warp-uniform branches so that they cost nothing when no lane needs them
  bool inv = false;
  double off = 0.0;
  if (LITAP_WARP_ANY(a > 0.41421356237309503)) {
    if (LITAP_WARP_ANY(a > 1.0)) {
      inv = a > 1.0;
      if (inv) a = fast_rcp(a);
    }
    if (a > 0.41421356237309503) {
      a = (a - 1.0) * fast_rcp(a + 1.0);
      off = 0.78539816339744831;
    }
  }
  double r = off + fast_atan_core(a);
  return inv ? 1.5707963267948966 - r : r;
}
// round(v, 3) when v is safely away from a rounding boundary; false = in doubt
LITAP_HD bool fast_round3(double v, double* out) {
  const double a = fabs(v);
  const double t = a * 1000.0;
  const double k = rint(t);
  const double dist = fabs(t - k);
  // in doubt: near a half, beyond the verified range of div1000, or NaN (every comparison false)
  if (!(dist <= 0.5 - (1e-9 + t * 2e-11) && t < 4.0e9)) return false;
  *out = copysign(div1000(k), v);  // v == 0 keeps its sign like the reference (returns x itself)
  return true;
}

LITAP_HD bool form_cell_fast(const double* e, const FormK& K, const double* sin_t, const double* cos_t, FormOut* o) {
  const double NA = NAN;
  bool any_na = false;
  double z[10];
#pragma unroll
  for (int n = 1; n <= 9; ++n) {
    any_na |= is_nan(e[n]);
    z[n] = e[n] * 100;
  }
  if (any_na) {
    o->slope_pct = NA; o->slope_deg = NA; o->prof = NA; o->plan = NA;
    o->aspect = is_nan(e[5]) ? NA : 360.0;
    return true;
  }
  const double sx = (z[6] - z[4]) * K.inv2g, sy = (z[2] - z[8]) * K.inv2g;
  const double ax = sx < 0 ? -sx : sx, ay = sy < 0 ? -sy : sy;
  const double mn = ax < ay ? ax : ay, mx = ax < ay ? ay : ax;
  if (!(mx > 1e-30 && mx < 1e30)) return false;  // flat window (slope 0) or absurd values: reference tree
  const double sp = sqrt(sx * sx + sy * sy);
  const double DEG = 57.295779513082323;
  if (!fast_round3(sp, &o->slope_pct)) return false;
  if (!fast_round3(fast_atan_pos(sp * 0.01) * DEG, &o->slope_deg)) return false;
  // aspect (R/form_01_calc_form.R:71-80): la = acos(|sx|/sp) = atan(|sy|/|sx|); with sx or sy exactly 0
  // (common on quantised DEMs) the table gives 90 / 270 / 180 / 360 = the same formula with la = 0 or 90
  double la = 0.0;
  if (mn > 0) la = fast_atan_pos(mn * fast_rcp(mx));
  if (ay > ax) la = 1.5707963267948966 - la;
  la *= DEG;
  double as_hat = sx > 0 ? (sy > 0 ? 270 + la : 270 - la) : (sy > 0 ? 90 - la : 90 + la);
  if (sx == 0) as_hat = sy < 0 ? 180 : 360;
  if (sy == 0) as_hat = sx < 0 ? 90 : 270;
  const double as = rint(as_hat);
  double da = as_hat - as;
  if (da < 0) da = -da;
  if (da > 0.5 - 1e-6) return false;
  const double pa = as > 180 ? as - 180 : as;
  const double pl = (pa + 90) > 180 ? pa + 90 - 180 : pa + 90;
  {
    double c = form_curv_num((int)pa, z, sin_t, cos_t) * K.invg2;
    double a = c < 0 ? -c : c;
    if (!(a < 1e30)) return false;
    double v = fast_atan_pos(a) * DEG;
    if (!fast_round3(c < 0 ? -v : v, &o->prof)) return false;
  }
  {
    double c = form_curv_num((int)pl, z, sin_t, cos_t) * K.invg2;
    double a = c < 0 ? -c : c;
    if (!(a < 1e30)) return false;
    double v = fast_atan_pos(a) * DEG;
    if (!fast_round3(c < 0 ? -v : v, &o->plan)) return false;
  }
  o->aspect = as;
  return true;
}

// ---------------------------------------------------------------------------------------------
// Quick path of calc_form (the one k_form runs on every cell; form_cell_fast / form_cell take the
// cells it leaves in doubt).  Same contract as form_cell_fast: when it returns true the five outputs
// are bit-identical with form_cell(); it differs in HOW the integer k = round(1000 v) is found:
//   * slope_pct: fp64 sqrt from an f32 rsqrt seed and one Newton step (relative error < 2e-13);
//   * the three arctangents (slope_deg, prof, plan) come out directly in units of 0.001 degree as
//     T[i] + corr (after a -> 1/a for a > 1): T[i] = 1000*deg(atan(i/512)) from a 513-entry fp64
//     table, corr = the f32 evaluation of atan((x - i/512) / (1 + x*i/512)), |argument| <= 1/1024,
//     so that the ABSOLUTE error of an f32 chain (9 roundings, ~5.4e-7 relative) is < 4e-5 units;
//   * curvature numerators use the identity (R/form_01_calc_form.R:87-106, both branches)
//       2*z5 - z1 - z2 = -( |c| s D_diag + |c| (1-s) D_82 + (1-|c|) s D_64 ),  s = sin A, c = cos A,
//       D_ab = z_a + z_b - 2 z5,  D_diag = D_91 for A <= 90 and D_37 for A > 90,
//     with the three weights (times -1/grid^2) tabulated per integer aspect A = 0..180;
//   * aspect is rounded to a whole degree, so its arctangent is a plain f32 polynomial.
// Guards (in units of 0.001 of the output): QK_GUARD_T = 2e-4 for the table arctangents (5x the
// bound above), plus |z5| * gz for the rounding noise of z = elev*100 in second differences
// (both this path's and the reference's: <= 4.5 * 2^-51 |z| / grid^2 radians); 3e-4 degree for
// aspect (error budget 5e-5).  Values that round to k = 0 must also have a certain sign (-0.0).
struct FormQ {
  double inv2g;   // 1 / (2 grid)
  double gz;      // guard per unit |z5|: 8 * 2^-51 * 57295.78 / grid^2
};
#define QK_ATAN_N 513          // x in [0, 1]: i = rint(512 x) <= 512
#define QK_GUARD_T 2.0e-5

// tables for form_cell_quick, filled by the host (libm in long double):
//   w[3*A + {0,1,2}] = -{ |c| s, |c| (1-s), (1-|c|) s } / grid^2 ; at[i] = 1000 * deg(atan(i/512))
inline void form_fill_quick_tables(double grid, const double* sin_t, const double* cos_t, double* w, double* at) {
  const double invg2 = 1.0 / (grid * grid);
  for (int A = 0; A <= 180; ++A) {
    double s = sin_t[A], c = fabs(cos_t[A]);
    w[3 * A + 0] = -(c * s) * invg2;
    w[3 * A + 1] = -(c * (1 - s)) * invg2;
    w[3 * A + 2] = -((1 - c) * s) * invg2;
  }
  for (int i = 0; i < QK_ATAN_N; ++i)
    at[i] = (double)(atanl((long double)i / 512.0L) * (180000.0L / 3.14159265358979323846264338327950288L));
}

LITAP_HD float qk_rcp(float x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / x;
#endif
}
// Seeds with >= 20 good bits (MUFU.RCP64H / MUFU.RSQ64H work on the high word of the double: one
// conversion-pipe instruction instead of f64 -> f32 -> MUFU -> f64).  The host build keeps 20 bits
// of the exact value so that the CPU test exercises the same refinement steps.
LITAP_HD double qk_seed20(double v) {
#if defined(__CUDA_ARCH__)
  return v;
#else
  unsigned long long b;
  __builtin_memcpy(&b, &v, 8);
  b &= 0xFFFFFFFF00000000ull;
  __builtin_memcpy(&v, &b, 8);
  return v;
#endif
}
LITAP_HD double qk_rcp64(double x) {
#if defined(__CUDA_ARCH__)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  return r;
#else
  return qk_seed20(1.0 / x);
#endif
}
LITAP_HD double qk_rsqrt64(double x) {
#if defined(__CUDA_ARCH__)
  double r;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  return r;
#else
  return qk_seed20(1.0 / sqrt(x));
#endif
}
LITAP_HD int qk_lo32(double u) {
#if defined(__CUDA_ARCH__)
  return __double2loint(u);
#else
  unsigned long long b;
  __builtin_memcpy(&b, &u, 8);
  return (int)(unsigned)(b & 0xFFFFFFFFull);
#endif
}
#define QK_MAGIC 6755399441055744.0   // 2^52 + 2^51: (v + M) - M = rint(v) for |v| < 2^51, low word of v + M = the integer
// t = 1000 * deg(atan(a)) for 0 <= a < 1e30, absolute error < 1e-7; *ok cleared outside the range.
// a > 1 goes through x = 1/a and 90 deg - atan(x).  With i = rint(512 x), c0 = i/512:
// atan(x) = atan(c0) + atan(dl), dl = (x - c0) / (1 + x c0), |dl| <= 1/1024, atan(dl) = dl (1 - dl^2/3)
// up to 2e-16.  Both reciprocals: 20-bit seed + one Newton step (relative error < 1e-12).
LITAP_HD double qk_atan_t(double a, const double* at, bool* ok) {
  if (!(a < 1e30)) { *ok = false; a = 0.0; }   // also NaN
  const bool big = a > 1.0;
  double r = qk_rcp64(a);
  r = fma(fma(-a, r, 1.0), r, r);
  const double x = big ? r : a;
  const double u = fma(x, 512.0, QK_MAGIC);
  const int i = qk_lo32(u);
  const double c0 = (u - QK_MAGIC) * 0.001953125;
  const double dd = x - c0;
  const double den = fma(x, c0, 1.0);
  double r2 = qk_rcp64(den);
  r2 = fma(fma(-den, r2, 1.0), r2, r2);
  const double dl = dd * r2;
  const double corr = dl * fma(dl * dl, -0.33333333333333331, 1.0);
  const double t = fma(corr, 57295.779513082323, at[i]);
  return big ? 90000.0 - t : t;
}
// decide round(v, 3) from t = 1000 |v| with an absolute guard; sign of the result from sgn
LITAP_HD bool qk_round_t(double t, double guard, double sgn, double* out) {
  const double k = (t + QK_MAGIC) - QK_MAGIC;   // rint(t), t < 2^51
  const double dist = fabs(t - k);
  *out = copysign(div1000(k), sgn);
  return dist <= 0.5 - guard && t > guard;  // away from a half, and a certain sign when k == 0
}

// z[n] = elev_n * 100 (n = 1..9, the reference's z values); qw / qat: the tables above
LITAP_HD bool form_cell_quick(const double* z, const FormQ& Q, const double* qw, const double* qat, FormOut* o) {
  bool ok = true;
  const double dx = z[6] - z[4], dy = z[2] - z[8];
  const double sx = dx * Q.inv2g, sy = dy * Q.inv2g;
  const double ax = fabs(sx), ay = fabs(sy);
  const double mx = ax < ay ? ay : ax;
  ok = mx > 1e-15 && mx < 1e15;            // flat window, absurd values or NaN: not decided here
  const double s2 = fma(sx, sx, sy * sy);
  double sp;
  {
    const double y = qk_rsqrt64(s2);
    const double g = s2 * y;
    sp = fma(fma(-g, g, s2), 0.5 * y, g);   // one Newton step: relative error 1.5 seed^2 < 2e-12
  }
  {
    const double t = sp * 1000.0;
    ok &= qk_round_t(t, fma(t, 4e-11, 1e-9), 1.0, &o->slope_pct) && t < 4.0e9;
  }
  // second differences through the centre; the sum is NaN iff a corner or the centre is
  const double tz5 = z[5] + z[5];
  const double d91 = (z[9] + z[1]) - tz5, d37 = (z[3] + z[7]) - tz5;
  const double d82 = (z[8] + z[2]) - tz5, d64 = (z[6] + z[4]) - tz5;
  const double dsum = d91 + d37;
  ok &= dsum == dsum;
  const double guard = fma(fabs(z[5]), Q.gz, QK_GUARD_T);
  {
    const double t = qk_atan_t(sp * 0.01, qat, &ok);
    ok &= qk_round_t(t, guard, 1.0, &o->slope_deg);
  }
  // aspect (R/form_01_calc_form.R:71-80): la = acos(|sx|/sp) = atan(|sy|/|sx|) in f32
  int A;
  {
    const float fx = (float)ax, fy = (float)ay;
    const float q = fminf(fx, fy) * qk_rcp(fmaxf(fx, fy));
    const float t = q * q;
    float p = 0.00680924067273736f;
    p = fmaf(p, t, -0.03359666094183922f);
    p = fmaf(p, t, 0.07961514592170715f);
    p = fmaf(p, t, -0.1323288530111313f);
    p = fmaf(p, t, 0.198076993227005f);
    p = fmaf(p, t, -0.333173543214798f);
    p = fmaf(p, t, 0.9999961256980896f);
    float la = q * p;                           // |error| < 4e-7 rad
    if (fy > fx) la = 1.5707964f - la;
    la *= 57.29578f;
    float ah = dx > 0 ? (dy > 0 ? 270.0f + la : 270.0f - la) : (dy > 0 ? 90.0f - la : 90.0f + la);
    if (dx == 0) ah = dy < 0 ? 180.0f : 360.0f;
    if (dy == 0) ah = dx < 0 ? 90.0f : 270.0f;
    const float am = ah + 12582912.0f;          // 1.5 * 2^23: rounds to a whole number, ties to even
    const float as = am - 12582912.0f;
    ok &= fabsf(ah - as) < 0.5f - 3e-4f;
#if defined(__CUDA_ARCH__)
    A = __float_as_int(am) - 0x4B400000;
    if (!(A >= 0 && A <= 360)) A = 0;           // NaN window: keeps the table index legal
    o->aspect = __hiloint2double(0x43300000, A) - 4503599627370496.0;  // (double)A without a conversion
#else
    A = (as >= 0 && as <= 360) ? (int)as : 0;
    o->aspect = (double)A;
#endif
  }
  const int pa = A > 180 ? A - 180 : A;
  const int pl = pa + 90 > 180 ? pa - 90 : pa + 90;
  // weights of pl = pa +- 90: sin and |cos| swap, so {w0, w1, w2}(pl) = {w0, w2, w1}(pa)
  const double* w = qw + 3 * pa;
  const double w0 = w[0], w1 = w[1], w2 = w[2];
  {
    const double c = fma(w0, pa <= 90 ? d91 : d37, fma(w1, d82, w2 * d64));
    const double t = qk_atan_t(fabs(c), qat, &ok);
    ok &= qk_round_t(t, guard, c, &o->prof);
  }
  {
    const double c = fma(w0, pl <= 90 ? d91 : d37, fma(w2, d82, w1 * d64));
    const double t = qk_atan_t(fabs(c), qat, &ok);
    ok &= qk_round_t(t, guard, c, &o->plan);
  }
  return ok;
}

// ---------------------------------------------------------------------------------------------
// R's sum() accumulates in long double (x87 80-bit extended, 64-bit significand) and rounds the
// result to double once (src/main/summary.c rsum; call site R/form_02b_calc_weti.R:67).  The terms
// summed on this path are all >= 0, so the emulation only needs the addition of non-negative
// values: round-to-nearest-even at 64 bits after every add, then 64 -> 53 bits at the end.
struct Ld80 {          // value = m * 2^(e - 63), m normalised (bit 63 set) or 0
  unsigned long long m;
  int e;
};
LITAP_HD int ld80_clz64(unsigned long long x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return __builtin_clzll(x);
#endif
}
LITAP_HD unsigned long long ld80_bits(double x) {
#if defined(__CUDA_ARCH__)
  return (unsigned long long)__double_as_longlong(x);
#else
  unsigned long long b;
  __builtin_memcpy(&b, &x, 8);
  return b;
#endif
}
LITAP_HD double ld80_from_bits(unsigned long long b) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double((long long)b);
#else
  double x;
  __builtin_memcpy(&x, &b, 8);
  return x;
#endif
}
// x must be finite and >= 0
LITAP_HD Ld80 ld80_from_double(double x) {
  Ld80 r;
  unsigned long long b = ld80_bits(x);
  int ex = (int)((b >> 52) & 0x7FF);
  unsigned long long fr = b & 0x000FFFFFFFFFFFFFull;
  if (ex == 0) {
    if (fr == 0) { r.m = 0; r.e = 0; return r; }
    int s = ld80_clz64(fr);  // subnormal: value = fr * 2^-1074
    r.m = fr << s;
    r.e = 63 - s - 1074;
    return r;
  }
  r.m = 0x8000000000000000ull | (fr << 11);
  r.e = ex - 1023;
  return r;
}
LITAP_HD Ld80 ld80_add(Ld80 a, Ld80 b) {
  if (b.m == 0) return a;
  if (a.m == 0) return b;
  if (a.e < b.e) { Ld80 t = a; a = b; b = t; }
  const int sh = a.e - b.e;
  unsigned long long hi, lo;
  bool sticky = false;
  if (sh == 0) { hi = b.m; lo = 0; }
  else if (sh < 64) { hi = b.m >> sh; lo = b.m << (64 - sh); }
  else if (sh == 64) { hi = 0; lo = b.m; }
  else if (sh < 128) { hi = 0; lo = b.m >> (sh - 64); sticky = (b.m << (128 - sh)) != 0; }
  else { hi = 0; lo = 0; sticky = true; }
  unsigned long long sh_ = a.m + hi;
  int e = a.e;
  if (sh_ < a.m) {  // carry out of bit 63
    sticky = sticky || (lo & 1ull);
    lo = (lo >> 1) | (sh_ << 63);
    sh_ = (sh_ >> 1) | 0x8000000000000000ull;
    ++e;
  }
  const unsigned long long half = 0x8000000000000000ull;
  if (lo > half || (lo == half && (sticky || (sh_ & 1ull)))) {
    ++sh_;
    if (sh_ == 0) { sh_ = half; ++e; }
  }
  Ld80 r;
  r.m = sh_;
  r.e = e;
  return r;
}
// (double) of the extended value, round to nearest even; normal range only (the sums here are
// slopes times cell sizes)
LITAP_HD double ld80_to_double(Ld80 a) {
  if (a.m == 0) return 0.0;
  unsigned long long m53 = a.m >> 11, rem = a.m & 0x7FFull;
  int e = a.e;
  if (rem > 0x400ull || (rem == 0x400ull && (m53 & 1ull))) {
    ++m53;
    if (m53 == (1ull << 53)) { m53 >>= 1; ++e; }
  }
  if (e > 1023) return ld80_from_bits(0x7FF0000000000000ull);
  if (e < -1022) return 0.0;  // not reachable for this path's magnitudes
  return ld80_from_bits(((unsigned long long)(e + 1023) << 52) | (m53 & 0x000FFFFFFFFFFFFFull));
}

// signed variant (calc_upslopes(type = "uced") sums elevation differences of either sign, R/flow_02_calc_shed.R:126):
// value = (-1)^neg * m * 2^(e - 63)
struct Lds {
  unsigned long long m;
  int e;
  int neg;
};
LITAP_HD Lds lds_from_double(double x) {   // x finite
  const Ld80 u = ld80_from_double(x < 0 ? -x : x);
  Lds r;
  r.m = u.m; r.e = u.e; r.neg = x < 0;
  return r;
}
LITAP_HD Lds lds_add(Lds a, Lds b) {
  if (b.m == 0) return a;
  if (a.m == 0) return b;
  if (a.neg == b.neg) {
    Ld80 x, y;
    x.m = a.m; x.e = a.e; y.m = b.m; y.e = b.e;
    const Ld80 r = ld80_add(x, y);
    Lds o;
    o.m = r.m; o.e = r.e; o.neg = a.neg;
    return o;
  }
  // |a| - |b| with |a| >= |b|, exact in 128 bits (+ a sticky fraction when b lies more than 64 bits below a)
  if (a.e < b.e || (a.e == b.e && a.m < b.m)) { const Lds t = a; a = b; b = t; }
  const int sh = a.e - b.e;
  unsigned long long yh, yl;
  bool sticky = false;
  if (sh == 0) { yh = b.m; yl = 0; }
  else if (sh < 64) { yh = b.m >> sh; yl = b.m << (64 - sh); }
  else if (sh == 64) { yh = 0; yl = b.m; }
  else if (sh < 128) { yh = 0; yl = b.m >> (sh - 64); sticky = (b.m << (128 - sh)) != 0; }
  else { yh = 0; yl = 0; sticky = true; }
  unsigned long long dl = 0ull - yl;
  unsigned long long dh = a.m - yh - (yl != 0 ? 1ull : 0ull);
  if (sticky) {   // true difference = (dh:dl - 1) + a fraction in (0, 1)
    if (dl == 0) --dh;
    --dl;
  }
  Lds o;
  o.neg = a.neg;
  if (dh == 0 && dl == 0) { o.m = 0; o.e = 0; o.neg = 0; return o; }   // x + (-x) = +0
  const int s = dh ? ld80_clz64(dh) : 64 + ld80_clz64(dl);   // sticky implies s <= 1: the fraction stays below the rounding bit
  if (s >= 64) { dh = dl << (s - 64); dl = 0; }
  else if (s) { dh = (dh << s) | (dl >> (64 - s)); dl <<= s; }
  int e = a.e - s;
  const unsigned long long half = 0x8000000000000000ull;
  if (dl > half || (dl == half && (sticky || (dh & 1ull)))) {
    ++dh;
    if (dh == 0) { dh = half; ++e; }
  }
  o.m = dh;
  o.e = e;
  return o;
}
LITAP_HD double lds_to_double(Lds a) {
  Ld80 u;
  u.m = a.m; u.e = a.e;
  const double v = ld80_to_double(u);
  return a.neg ? -v : v;
}

// ---------------------------------------------------------------------------------------------
// calc_weti2 per-cell pieces (R/form_02b_calc_weti.R:47-67).  Window visit order of the reference:
// NW,N,NE,W,self,E,SW,S,SE = neighbour codes 7,8,9,4,5,6,1,2,3.
struct WetiK {
  double l1, l2, orth, diag, g2;   // grid*0.5, grid*0.354, grid, grid*sqrt(2), grid^2
  double rorth, rdiag, r1e4;       // RN(1/orth), RN(1/diag), RN(1/0.0001) for div_rcp (0 = use the IEEE division)
};
// a / b, correctly rounded, from y = RN(1/b) computed once: two Newton corrections with exact residuals
// (Markstein 1990: if y is the correctly rounded reciprocal and q is within 1 ulp of a/b, then
// RN(q + (a - b q) y) = RN(a/b); the first correction makes q faithful, the second one applies the theorem).
// Five dependent operations instead of the ~12 + range check of the generic division; bit-identical results
// (checked against the division on 4e9 random and structured operands by tests/test_host_cpu.py).  Operands
// outside the range in which the residuals are exact (|a| outside 2^-900 .. 2^900, zero, inf, NaN) and
// callers without a usable reciprocal (y == 0) take the IEEE division.
LITAP_HD double div_rcp(double a, double b, double y) {
  const unsigned ex = (unsigned)((ld80_bits(a) >> 52) & 0x7FFull);
  if (ex - 123u > 1800u || y == 0.0) return a / b;
  double q = a * y;
  double r = fma(-q, b, a);
  q = fma(r, y, q);
  r = fma(-q, b, a);
  return fma(r, y, q);
}
// reciprocal usable by div_rcp, else 0: b must be far from the ends of the exponent range
LITAP_HD double div_rcp_prepare(double b) {
  const double ab = b < 0 ? -b : b;
  return (ab > 1e-30 && ab < 1e30) ? 1.0 / b : 0.0;
}
LITAP_HD bool weti_is_diag(int n) { return (0x28Au >> n) & 1u; }   // n in {1, 3, 7, 9}
// tan_k of a cell at elevation f towards a neighbour at elevation nv in direction n (:64-65);
// NaN neighbour -> NaN, neighbour not lower -> 0
LITAP_HD double weti_tan(double f, double nv, int n, const WetiK& K) {
  if (!(nv < f)) return is_nan(nv) ? nv : 0.0;
  const bool dg = weti_is_diag(n);
  return div_rcp(f - nv, dg ? K.diag : K.orth, dg ? K.rdiag : K.rorth) * (dg ? K.l2 : K.l1);
}
// S = sum(tan, na.rm = TRUE) with R's long double accumulator; e[n] = elevation of neighbour n
// (e[5] = the cell itself).  *lower gets bit n set for every strictly lower neighbour.
LITAP_HD double weti_S(const double* e, const WetiK& K, unsigned* lower) {
  Ld80 s;
  s.m = 0; s.e = 0;
  unsigned lo = 0;
  // window order 7 8 9 4 5 6 1 2 3.  On the device the loop stays a loop: unrolled (eight emulated long double additions
  // and eight divisions in a row) it was most of a 65 KB kernel that waited for instructions
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int k = 0; k < 9; ++k) {
    const int n = (int)((0x321654987ull >> (4 * k)) & 15ull);
    if (n == 5) continue;  // self: tan = 0
    double t = weti_tan(e[5], e[n], n, K);
    if (is_nan(t)) continue;
    if (t > 0) { lo |= 1u << n; s = ld80_add(s, ld80_from_double(t)); }
  }
  *lower = lo;
  return ld80_to_double(s);
}
