/*
 * litap_call.c — the R side of the boundary: `.Call` entry points that hand R-owned vectors to
 * liblitap_cuda.so (include/litap.h).  This is the ONLY file that includes Rinternals.h; it holds
 * no algorithm.  It is what a LITAP maintainer adds as src/litap_call.c (plus
 * `useDynLib(LITAP, .registration = TRUE)` in NAMESPACE and `PKG_LIBS = -llitap_cuda` in
 * src/Makevars); see INTEGRATION.md.  R is not installed in the build image, so this file is
 * compiled only where R exists (R CMD SHLIB litap_call.c -llitap_cuda).
 *
 * Conventions
 *   - every column arrives as the full buffered grid in seqno order (length NR*NC), exactly the
 *     tibble columns the reference's internals receive (R/LITAP_load.R:273-332);
 *   - NA_real_ is a NaN, so REAL() pointers go to the library unchanged;
 *   - ddir travels as uint8 (0 = NA); R's numeric ddir column is converted here;
 *   - results are returned as a named list of R vectors with the reference's column types
 *     (ddir/upslope/initial_shed numeric, drec integer (numeric above 2^31), ridge/flatcell logical);
 *   - errors: every non-zero status becomes Rf_error() AFTER all temporary buffers are released
 *     (R_alloc memory is reclaimed by R itself, so nothing leaks across the longjmp).
 */
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include <stdint.h>
#include <string.h>
#include "litap.h"

static void fail_if(int rc, const char* what) {
  if (rc != LITAP_OK) Rf_error("%s: %s (liblitap_cuda status %d)", what, litap_last_error(), rc);
}

static SEXP named_list(int n, const char** names) {
  SEXP out = PROTECT(Rf_allocVector(VECSXP, n));
  SEXP nm = PROTECT(Rf_allocVector(STRSXP, n));
  for (int i = 0; i < n; ++i) SET_STRING_ELT(nm, i, Rf_mkChar(names[i]));
  Rf_setAttrib(out, R_NamesSymbol, nm);
  UNPROTECT(2);
  return out;
}

static uint8_t* ddir_to_u8(SEXP ddir, R_xlen_t n) {
  uint8_t* d = (uint8_t*)R_alloc((size_t)n, 1);
  if (TYPEOF(ddir) == REALSXP) {
    const double* p = REAL(ddir);
    for (R_xlen_t i = 0; i < n; ++i) d[i] = (ISNAN(p[i]) || p[i] < 1 || p[i] > 9) ? 0 : (uint8_t)p[i];
  } else {
    const int* p = INTEGER(ddir);
    for (R_xlen_t i = 0; i < n; ++i) d[i] = (p[i] == NA_INTEGER || p[i] < 1 || p[i] > 9) ? 0 : (uint8_t)p[i];
  }
  return d;
}

/* offsets of direction codes 1..9 on a grid with NC columns (R/LITAP_functions.R:237-254 calc_seq) */
static double drec_of(R_xlen_t seqno, int d, R_xlen_t NC) {
  static const int dr[10] = {0, 1, 1, 1, 0, 0, 0, -1, -1, -1};
  static const int dc[10] = {0, -1, 0, 1, -1, 0, 1, -1, 0, 1};
  return (double)(seqno + (R_xlen_t)dr[d] * NC + dc[d]);
}

/* calc_ddir2(db): .Call("litap_R_ddir", db$elev, nrow, ncol, invert) -> list(ddir, flatcell, drec) */
SEXP litap_R_ddir(SEXP elev, SEXP nrow, SEXP ncol, SEXP invert) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  if (TYPEOF(elev) != REALSXP || XLENGTH(elev) != n) Rf_error("litap_R_ddir: elev must be a double vector of nrow*ncol");
  uint8_t* d = (uint8_t*)R_alloc((size_t)n, 1);
  uint8_t* f = (uint8_t*)R_alloc((size_t)n, 1);
  fail_if(litap_ddir(REAL(elev), NR, NC, Rf_asLogical(invert) == TRUE, d, f), "calc_ddir2");
  const char* nm[3] = {"ddir", "flatcell", "drec"};
  SEXP out = PROTECT(named_list(3, nm));
  SEXP s_d = PROTECT(Rf_allocVector(REALSXP, n)), s_f = PROTECT(Rf_allocVector(LGLSXP, n));
  const int big = n > 2147483647;  /* seqno does not fit INTSXP: emit doubles (SURVEY.md 8b Limits) */
  SEXP s_r = PROTECT(Rf_allocVector(big ? REALSXP : INTSXP, n));
  for (R_xlen_t i = 0; i < n; ++i) {
    REAL(s_d)[i] = d[i] ? (double)d[i] : NA_REAL;
    LOGICAL(s_f)[i] = f[i] == 2 ? NA_LOGICAL : (int)f[i];
    if (big) REAL(s_r)[i] = d[i] ? drec_of(i + 1, d[i], NC) : NA_REAL;
    else INTEGER(s_r)[i] = d[i] ? (int)drec_of(i + 1, d[i], NC) : NA_INTEGER;
  }
  SET_VECTOR_ELT(out, 0, s_d);
  SET_VECTOR_ELT(out, 1, s_f);
  SET_VECTOR_ELT(out, 2, s_r);
  UNPROTECT(4);
  return out;
}

/* calc_shed4(db): .Call("litap_R_shed", db$elev, db$ddir, nrow, ncol) -> list(initial_shed, upslope, ridge) */
SEXP litap_R_shed(SEXP elev, SEXP ddir, SEXP nrow, SEXP ncol) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  if (TYPEOF(elev) != REALSXP || XLENGTH(elev) != n || XLENGTH(ddir) != n)
    Rf_error("litap_R_shed: elev/ddir must have nrow*ncol entries");
  uint8_t* d = ddir_to_u8(ddir, n);
  int32_t* sh = (int32_t*)R_alloc((size_t)n, sizeof(int32_t));
  uint32_t* up = (uint32_t*)R_alloc((size_t)n, sizeof(uint32_t));
  uint8_t* rg = (uint8_t*)R_alloc((size_t)n, 1);
  int64_t npits = 0;
  fail_if(litap_shed(REAL(elev), d, NR, NC, sh, up, rg, &npits), "calc_shed4");
  const char* nm[3] = {"initial_shed", "upslope", "ridge"};
  SEXP out = PROTECT(named_list(3, nm));
  SEXP s_s = PROTECT(Rf_allocVector(REALSXP, n)), s_u = PROTECT(Rf_allocVector(REALSXP, n));
  SEXP s_r = PROTECT(Rf_allocVector(LGLSXP, n));
  for (R_xlen_t i = 0; i < n; ++i) {
    REAL(s_s)[i] = sh[i] ? (double)sh[i] : NA_REAL;
    REAL(s_u)[i] = (double)up[i];
    LOGICAL(s_r)[i] = rg[i] == 2 ? NA_LOGICAL : (int)rg[i];
  }
  SET_VECTOR_ELT(out, 0, s_s);
  SET_VECTOR_ELT(out, 1, s_u);
  SET_VECTOR_ELT(out, 2, s_r);
  UNPROTECT(4);
  return out;
}

/* calc_form(db, grid): results are written straight into freshly allocated REALSXPs */
SEXP litap_R_form(SEXP elev, SEXP nrow, SEXP ncol, SEXP grid) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  if (TYPEOF(elev) != REALSXP || XLENGTH(elev) != n) Rf_error("litap_R_form: elev must be a double vector of nrow*ncol");
  const char* nm[5] = {"slope_pct", "slope_deg", "aspect", "prof", "plan"};
  SEXP out = PROTECT(named_list(5, nm));
  double* p[5];
  for (int k = 0; k < 5; ++k) {
    SEXP v = Rf_allocVector(REALSXP, n);
    SET_VECTOR_ELT(out, k, v);  /* protected through `out` */
    p[k] = REAL(v);
  }
  fail_if(litap_form(REAL(elev), NR, NC, Rf_asReal(grid), p[0], p[1], p[2], p[3], p[4]), "calc_form");
  UNPROTECT(1);
  return out;
}

/* calc_weti2(db, grid) */
SEXP litap_R_weti(SEXP elev, SEXP ddir, SEXP upslope, SEXP nrow, SEXP ncol, SEXP grid) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  if (TYPEOF(elev) != REALSXP || XLENGTH(elev) != n || XLENGTH(ddir) != n || XLENGTH(upslope) != n)
    Rf_error("litap_R_weti: elev/ddir/upslope must have nrow*ncol entries");
  uint8_t* d = ddir_to_u8(ddir, n);
  uint32_t* up = (uint32_t*)R_alloc((size_t)n, sizeof(uint32_t));
  SEXP upr = PROTECT(Rf_coerceVector(upslope, REALSXP));
  for (R_xlen_t i = 0; i < n; ++i) up[i] = ISNAN(REAL(upr)[i]) ? 0u : (uint32_t)REAL(upr)[i];
  const char* nm[4] = {"qarea1", "qarea2", "qweti1", "qweti2"};
  SEXP out = PROTECT(named_list(4, nm));
  double* p[4];
  for (int k = 0; k < 4; ++k) {
    SEXP v = Rf_allocVector(REALSXP, n);
    SET_VECTOR_ELT(out, k, v);
    p[k] = REAL(v);
  }
  fail_if(litap_weti(REAL(elev), d, up, NR, NC, Rf_asReal(grid), p[0], p[1], p[2], p[3]), "calc_weti2");
  UNPROTECT(2);
  return out;
}

/* integer-like column (numeric or integer, NA allowed) -> int32 with 0 = NA */
static int32_t* col_to_i32(SEXP v, R_xlen_t n) {
  int32_t* o = (int32_t*)R_alloc((size_t)n, sizeof(int32_t));
  if (TYPEOF(v) == REALSXP) {
    const double* p = REAL(v);
    for (R_xlen_t i = 0; i < n; ++i) o[i] = ISNAN(p[i]) ? 0 : (int32_t)p[i];
  } else {
    const int* p = INTEGER(v);   /* INTSXP and LGLSXP share the layout */
    for (R_xlen_t i = 0; i < n; ++i) o[i] = p[i] == NA_INTEGER ? 0 : (int32_t)p[i];
  }
  return o;
}
static uint8_t* col_to_u8(SEXP v, R_xlen_t n, uint8_t na) {
  uint8_t* o = (uint8_t*)R_alloc((size_t)n, 1);
  if (TYPEOF(v) == REALSXP) {
    const double* p = REAL(v);
    for (R_xlen_t i = 0; i < n; ++i) o[i] = ISNAN(p[i]) ? na : (uint8_t)p[i];
  } else {
    const int* p = INTEGER(v);
    for (R_xlen_t i = 0; i < n; ++i) o[i] = p[i] == NA_INTEGER ? na : (uint8_t)p[i];
  }
  return o;
}

/* calc_upslopes(db, type = "uced"): .Call("litap_R_uced", db$elev, db$ddir, db$shedno, nrow, ncol) -> uced */
SEXP litap_R_uced(SEXP elev, SEXP ddir, SEXP shedno, SEXP nrow, SEXP ncol) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  if (TYPEOF(elev) != REALSXP || XLENGTH(elev) != n || XLENGTH(ddir) != n || XLENGTH(shedno) != n)
    Rf_error("litap_R_uced: elev/ddir/shedno must have nrow*ncol entries");
  uint8_t* d = ddir_to_u8(ddir, n);
  int32_t* s = col_to_i32(shedno, n);
  SEXP out = PROTECT(Rf_allocVector(REALSXP, n));
  fail_if(litap_uced(REAL(elev), d, s, NR, NC, REAL(out)), "calc_upslopes(type = \"uced\")");
  UNPROTECT(1);
  return out;
}

/* pit_stat1(db, shed = , method = ): .Call("litap_R_pit_stat", elev, ddir, shed column, upslope, ridge, edge_map, nrow,
 * ncol, method (0 litap / 1 landmapr)) -> named list of the 21 columns (doubles; seqno / row / col buffered) */
SEXP litap_R_pit_stat(SEXP elev, SEXP ddir, SEXP shed, SEXP upslope, SEXP ridge, SEXP edge_map, SEXP nrow, SEXP ncol,
                      SEXP method) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  if (TYPEOF(elev) != REALSXP || XLENGTH(elev) != n || XLENGTH(ddir) != n || XLENGTH(shed) != n || XLENGTH(upslope) != n ||
      XLENGTH(ridge) != n || XLENGTH(edge_map) != n)
    Rf_error("litap_R_pit_stat: every column must have nrow*ncol entries");
  uint8_t* d = ddir_to_u8(ddir, n);
  int32_t* s = col_to_i32(shed, n);
  uint32_t* up = (uint32_t*)col_to_i32(upslope, n);
  uint8_t* rg = col_to_u8(ridge, n, 2);
  uint8_t* em = col_to_u8(edge_map, n, 0);
  int32_t mx = 0;
  for (R_xlen_t i = 0; i < n; ++i) if (s[i] > mx) mx = s[i];
  const int64_t cap = mx > 0 ? mx : 1;            /* at most one row per shed number */
  int ncolumns = 0;
  const char* const* names = litap_pit_stat_names(&ncolumns);
  double** col = (double**)R_alloc((size_t)ncolumns, sizeof(double*));
  for (int k = 0; k < ncolumns; ++k) col[k] = (double*)R_alloc((size_t)cap, sizeof(double));
  int64_t rows = 0;
  fail_if(litap_pit_stat(REAL(elev), d, s, up, rg, em, NR, NC, Rf_asInteger(method), cap, col, &rows), "pit_stat1");
  SEXP out = PROTECT(named_list(ncolumns, (const char**)names));
  for (int k = 0; k < ncolumns; ++k) {
    SEXP v = Rf_allocVector(REALSXP, (R_xlen_t)rows);
    SET_VECTOR_ELT(out, k, v);
    memcpy(REAL(v), col[k], (size_t)rows * sizeof(double));
  }
  UNPROTECT(1);
  return out;
}

/* slope_gc(db, grid): the six gradients / curvatures as doubles, hill_*_dir as doubles (2/4, 1/3),
 * hill_*_n / hill_*_cell as integers; NA where the library reports 0 / NaN */
SEXP litap_R_slope_gc(SEXP elev, SEXP nrow, SEXP ncol, SEXP grid) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  if (TYPEOF(elev) != REALSXP || XLENGTH(elev) != n) Rf_error("litap_R_slope_gc: elev must be double with nrow*ncol entries");
  const char* nm[12] = {"sgre", "sgr", "sgcn", "sgc", "scr", "scc", "hill_r_dir", "hill_c_dir",
                        "hill_r_n", "hill_r_cell", "hill_c_n", "hill_c_cell"};
  SEXP out = PROTECT(named_list(12, nm));
  double* f[8];
  int* iv[4];
  for (int k = 0; k < 8; ++k) {
    SEXP v = Rf_allocVector(REALSXP, n);
    SET_VECTOR_ELT(out, k, v);
    f[k] = REAL(v);
  }
  for (int k = 0; k < 4; ++k) {
    SEXP v = Rf_allocVector(INTSXP, n);
    SET_VECTOR_ELT(out, 8 + k, v);
    iv[k] = INTEGER(v);
  }
  uint8_t* rd = (uint8_t*)R_alloc((size_t)n, 1);
  uint8_t* cd = (uint8_t*)R_alloc((size_t)n, 1);
  fail_if(litap_slope_gc(REAL(elev), NR, NC, Rf_asReal(grid), f[0], f[1], f[2], f[3], f[4], f[5], rd, cd,
                         (int32_t*)iv[0], (int32_t*)iv[1], (int32_t*)iv[2], (int32_t*)iv[3]), "slope_gc");
  for (R_xlen_t i = 0; i < n; ++i) {
    f[6][i] = rd[i] ? (double)rd[i] : NA_REAL;
    f[7][i] = cd[i] ? (double)cd[i] : NA_REAL;
    for (int k = 0; k < 4; ++k) if (iv[k][i] == 0) iv[k][i] = NA_INTEGER;
  }
  UNPROTECT(1);
  return out;
}

/* calc_relz + calc_length: db / idb columns as vectors over the buffered grid (integers: NA -> 0 here),
 * the pond table (may have 0 rows); returns the 39 columns of litap_relz_names() as doubles (NaN -> NA) */
static int* na0(SEXP v, R_xlen_t n, const char* what) {
  if (XLENGTH(v) != n) Rf_error("litap_R_relz: %s has the wrong length", what);
  SEXP iv = PROTECT(Rf_coerceVector(v, INTSXP));
  int* o = (int*)R_alloc((size_t)(n ? n : 1), sizeof(int));
  for (R_xlen_t i = 0; i < n; ++i) o[i] = INTEGER(iv)[i] == NA_INTEGER ? 0 : INTEGER(iv)[i];
  UNPROTECT(1);
  return o;
}
static uint8_t* na0_u8(SEXP v, R_xlen_t n, const char* what) {
  int* t = na0(v, n, what);
  uint8_t* o = (uint8_t*)R_alloc((size_t)(n ? n : 1), 1);
  for (R_xlen_t i = 0; i < n; ++i) o[i] = (uint8_t)t[i];
  return o;
}
SEXP litap_R_relz(SEXP nrow, SEXP ncol, SEXP db, SEXP idb, SEXP pond, SEXP str_val, SEXP ridge_val, SEXP grid) {
  const R_xlen_t NR = (R_xlen_t)Rf_asReal(nrow), NC = (R_xlen_t)Rf_asReal(ncol), n = NR * NC;
  /* db: elev, ddir, drec, upslope, fill_shed, local_shed; idb: elev, ddir, drec, upslope, inv_local_shed;
   * pond: shedno, pit_row, pit_col, pit_elev (lists in this order, see calc_relz() in litap_cuda.R) */
  SEXP e = VECTOR_ELT(db, 0), ie = VECTOR_ELT(idb, 0);
  if (TYPEOF(e) != REALSXP || XLENGTH(e) != n || TYPEOF(ie) != REALSXP || XLENGTH(ie) != n)
    Rf_error("litap_R_relz: elev must be double with nrow*ncol entries");
  const R_xlen_t np = XLENGTH(VECTOR_ELT(pond, 0));
  SEXP pe = PROTECT(Rf_coerceVector(VECTOR_ELT(pond, 3), REALSXP));
  const char* names = litap_relz_names();
  SEXP out = PROTECT(Rf_allocVector(VECSXP, LITAP_RELZ_NOUT));
  SEXP nms = PROTECT(Rf_allocVector(STRSXP, LITAP_RELZ_NOUT));
  double* o[LITAP_RELZ_NOUT];
  for (int k = 0; k < LITAP_RELZ_NOUT; ++k) {
    const char* end = strchr(names, ',');
    const size_t len = end ? (size_t)(end - names) : strlen(names);
    SET_STRING_ELT(nms, k, Rf_mkCharLen(names, (int)len));
    names += len + 1;
    SEXP v = Rf_allocVector(REALSXP, n);
    SET_VECTOR_ELT(out, k, v);
    o[k] = REAL(v);
  }
  Rf_setAttrib(out, R_NamesSymbol, nms);
  fail_if(litap_relz(NR, NC, REAL(e), na0_u8(VECTOR_ELT(db, 1), n, "db$ddir"), na0(VECTOR_ELT(db, 2), n, "db$drec"),
                     na0(VECTOR_ELT(db, 3), n, "db$upslope"), na0(VECTOR_ELT(db, 4), n, "db$fill_shed"),
                     na0(VECTOR_ELT(db, 5), n, "db$local_shed"), REAL(ie), na0_u8(VECTOR_ELT(idb, 1), n, "idb$ddir"),
                     na0(VECTOR_ELT(idb, 2), n, "idb$drec"), na0(VECTOR_ELT(idb, 3), n, "idb$upslope"),
                     na0(VECTOR_ELT(idb, 4), n, "idb$inv_local_shed"), np, na0(VECTOR_ELT(pond, 0), np, "pond$shedno"),
                     na0(VECTOR_ELT(pond, 1), np, "pond$pit_row"), na0(VECTOR_ELT(pond, 2), np, "pond$pit_col"), REAL(pe),
                     Rf_asReal(str_val), Rf_asReal(ridge_val), Rf_asReal(grid), o), "relz");
  for (int k = 0; k < LITAP_RELZ_NOUT; ++k)
    for (R_xlen_t i = 0; i < n; ++i) if (o[k][i] != o[k][i]) o[k][i] = NA_REAL;
  UNPROTECT(3);
  return out;
}

SEXP litap_R_init(SEXP device) {
  fail_if(litap_init(Rf_asInteger(device)), "litap_init");
  return Rf_mkString(litap_version());
}
SEXP litap_R_shutdown(void) {
  litap_shutdown();
  return R_NilValue;
}
SEXP litap_R_timings(void) {
  litap_timings t;
  fail_if(litap_get_timings(&t), "litap_get_timings");
  const char* nm[8] = {"h2d_ms", "d2h_ms", "ddir_ms", "shed_ms", "form_ms", "weti_ms", "total_ms", "kernel_launches"};
  const double v[8] = {t.h2d_ms, t.d2h_ms, t.ddir_ms, t.shed_ms, t.form_ms, t.weti_ms, t.total_ms, (double)t.kernel_launches};
  SEXP out = PROTECT(Rf_allocVector(REALSXP, 8));
  SEXP names = PROTECT(Rf_allocVector(STRSXP, 8));
  for (int i = 0; i < 8; ++i) { REAL(out)[i] = v[i]; SET_STRING_ELT(names, i, Rf_mkChar(nm[i])); }
  Rf_setAttrib(out, R_NamesSymbol, names);
  UNPROTECT(2);
  return out;
}

static const R_CallMethodDef call_methods[] = {
    {"litap_R_ddir", (DL_FUNC)&litap_R_ddir, 4},     {"litap_R_shed", (DL_FUNC)&litap_R_shed, 4},
    {"litap_R_form", (DL_FUNC)&litap_R_form, 4},     {"litap_R_weti", (DL_FUNC)&litap_R_weti, 6},
    {"litap_R_slope_gc", (DL_FUNC)&litap_R_slope_gc, 4}, {"litap_R_relz", (DL_FUNC)&litap_R_relz, 8},
    {"litap_R_uced", (DL_FUNC)&litap_R_uced, 5},     {"litap_R_pit_stat", (DL_FUNC)&litap_R_pit_stat, 9},
    {"litap_R_init", (DL_FUNC)&litap_R_init, 1},     {"litap_R_shutdown", (DL_FUNC)&litap_R_shutdown, 0},
    {"litap_R_timings", (DL_FUNC)&litap_R_timings, 0}, {NULL, NULL, 0}};

void R_init_LITAP(DllInfo* dll) {
  R_registerRoutines(dll, NULL, call_methods, NULL, NULL);
  R_useDynamicSymbols(dll, FALSE);
}
