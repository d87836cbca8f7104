/* Stand-in for R's <R.h>: DECLARATIONS ONLY, just enough to type-check
 * litap_b200/r/src/litap_call.c where R is not installed (tests/test_host_cpu.py). */
#ifndef R_STUB_R_H
#define R_STUB_R_H
#include <stddef.h>
char* R_alloc(size_t n, int size);
void Rf_error(const char* fmt, ...);
#endif
