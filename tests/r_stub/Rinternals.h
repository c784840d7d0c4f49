/* TEST STUB -- see R.h in this directory. */
#ifndef ATE_TEST_RINTERNALS_H
#define ATE_TEST_RINTERNALS_H
#include <stddef.h>
typedef struct SEXPREC *SEXP;
typedef ptrdiff_t R_xlen_t;
typedef unsigned int SEXPTYPE;
typedef enum { FALSE = 0, TRUE } Rboolean;
#define NILSXP 0
#define INTSXP 13
#define REALSXP 14
#define VECSXP 19
#define RAWSXP 24
extern SEXP R_NilValue;
extern double R_NaReal;
#define NA_REAL R_NaReal
#define ISNAN(x) ((x) != (x))
SEXP Rf_protect(SEXP);
void Rf_unprotect(int);
#define PROTECT(s) Rf_protect(s)
#define UNPROTECT(n) Rf_unprotect(n)
double *REAL(SEXP);
int *INTEGER(SEXP);
unsigned char *RAW(SEXP);
R_xlen_t XLENGTH(SEXP);
SEXP Rf_allocVector(SEXPTYPE, R_xlen_t);
SEXP Rf_allocMatrix(SEXPTYPE, int, int);
SEXP Rf_ScalarInteger(int);
SEXP Rf_ScalarReal(double);
int Rf_asInteger(SEXP);
int Rf_asLogical(SEXP);
double Rf_asReal(SEXP);
SEXP Rf_duplicate(SEXP);
SEXP Rf_install(const char *);
Rboolean Rf_isNull(SEXP);
SEXP Rf_mkNamed(SEXPTYPE, const char **);
int Rf_ncols(SEXP);
int Rf_nrows(SEXP);
SEXP Rf_setAttrib(SEXP, SEXP, SEXP);
SEXP SET_VECTOR_ELT(SEXP, R_xlen_t, SEXP);
void Rf_error(const char *, ...) __attribute__((noreturn));
void Rf_warning(const char *, ...);
void *R_ExternalPtrAddr(SEXP);
void R_ClearExternalPtr(SEXP);
SEXP R_MakeExternalPtr(void *, SEXP, SEXP);
typedef void (*R_CFinalizer_t)(SEXP);
void R_RegisterCFinalizerEx(SEXP, R_CFinalizer_t, Rboolean);
#endif
