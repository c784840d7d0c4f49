/* TEST STUB -- see ../R.h. */
#ifndef ATE_TEST_RDYNLOAD_H
#define ATE_TEST_RDYNLOAD_H
#include <Rinternals.h>
typedef void *(*DL_FUNC)(void);
typedef struct { const char *name; DL_FUNC fun; int numArgs; } R_CallMethodDef;
typedef R_CallMethodDef R_CMethodDef;
typedef R_CallMethodDef R_FortranMethodDef;
typedef R_CallMethodDef R_ExternalMethodDef;
typedef struct _DllInfo DllInfo;
int R_registerRoutines(DllInfo *, const R_CMethodDef *, const R_CallMethodDef *, const R_FortranMethodDef *,
                       const R_ExternalMethodDef *);
Rboolean R_useDynamicSymbols(DllInfo *, Rboolean);
#endif
