/* TEST STUB -- not R.  Just enough declarations of R's C API for `gcc -fsyntax-only` to type-check
 * r/ate_b200_shim.c in an image without R (tests/test_r_shim_syntax.py).  Signatures follow
 * Rinternals.h / Rdynload.h of R 3.x-4.x. */
#ifndef ATE_TEST_R_H
#define ATE_TEST_R_H
#include <stddef.h>
#endif
