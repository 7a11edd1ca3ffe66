/* Minimal stand-in for the declarations of R's C API that call_stsparse_b200.c uses.  ONLY for the syntax check in
 * tests/test_host_api.py (there is no R in this image); never linked, never shipped to an R build. */
#ifndef MOCK_R_H
#define MOCK_R_H
#include <stddef.h>
typedef struct SEXPREC *SEXP;
typedef void *(*DL_FUNC)(void);
#define REALSXP 14
#define INTSXP 13
double *REAL(SEXP);
int *INTEGER(SEXP);
int LENGTH(SEXP);
int isNull(SEXP);
SEXP allocVector(unsigned int, long);
SEXP PROTECT(SEXP);
void UNPROTECT(int);
SEXP install(const char *);
SEXP setAttrib(SEXP, SEXP, SEXP);
char *R_alloc(size_t, int);
DL_FUNC R_ExternalPtrAddrFn(SEXP);
void error(const char *, ...);
#endif
