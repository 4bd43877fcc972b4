/* Stub for the WFA submodule header, which is absent from the reference checkout.
 * src/utils.h includes it unconditionally; only uvaialign (out of scope) uses WFA.
 * These three declarations are just enough for src/utils.c to compile. */
#ifndef ORACLE_WFA_STUB_H
#define ORACLE_WFA_STUB_H
#include <stdlib.h>
typedef struct { char *operations; int max_operations, begin_offset, end_offset, score; } edit_cigar_t;
typedef struct { int dummy; } mm_allocator_t;
#define mm_allocator_calloc(a,n,type,zero) ((type*)calloc((n),sizeof(type)))
#endif
