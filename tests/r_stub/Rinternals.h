/* Stand-in for <Rinternals.h>: declarations only (see R.h in this directory). */
#ifndef R_STUB_RINTERNALS_H
#define R_STUB_RINTERNALS_H
#include <stddef.h>
typedef struct SEXPREC* SEXP;
typedef ptrdiff_t R_xlen_t;
#define REALSXP 14
#define INTSXP 13
#define LGLSXP 10
#define STRSXP 16
#define VECSXP 19
#define TRUE 1
#define FALSE 0
extern SEXP R_NilValue, R_NamesSymbol;
extern double R_NaReal;
extern int R_NaInt;
#define NA_REAL R_NaReal
#define NA_INTEGER R_NaInt
#define NA_LOGICAL R_NaInt
int R_IsNaN(double);
#define ISNAN(x) ((x) != (x))
SEXP Rf_allocVector(unsigned int, R_xlen_t);
SEXP Rf_protect(SEXP);
void Rf_unprotect(int);
#define PROTECT(s) Rf_protect(s)
#define UNPROTECT(n) Rf_unprotect(n)
SEXP Rf_mkChar(const char*);
SEXP Rf_mkString(const char*);
SEXP Rf_setAttrib(SEXP, SEXP, SEXP);
SEXP Rf_coerceVector(SEXP, unsigned int);
void SET_STRING_ELT(SEXP, R_xlen_t, SEXP);
SEXP SET_VECTOR_ELT(SEXP, R_xlen_t, SEXP);
SEXP VECTOR_ELT(SEXP, R_xlen_t);
SEXP Rf_mkCharLen(const char*, int);
int TYPEOF(SEXP);
R_xlen_t XLENGTH(SEXP);
double* REAL(SEXP);
int* INTEGER(SEXP);
int* LOGICAL(SEXP);
double Rf_asReal(SEXP);
int Rf_asInteger(SEXP);
int Rf_asLogical(SEXP);
#endif
