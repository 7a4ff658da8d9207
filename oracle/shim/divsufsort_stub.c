/* TEST INFRASTRUCTURE ONLY — see divsufsort.h in this directory. */
#include <stdio.h>
#include <stdlib.h>
#include "divsufsort.h"
#include "divsufsort64.h"
int32_t divsufsort(const uint8_t* T, int32_t* SA, int32_t n)
{ (void)T; (void)SA; (void)n; fprintf(stderr, "oracle shim: divsufsort must not be reached on the VLG path\n"); abort(); }
int64_t divsufsort64(const uint8_t* T, int64_t* SA, int64_t n)
{ (void)T; (void)SA; (void)n; fprintf(stderr, "oracle shim: divsufsort64 must not be reached on the VLG path\n"); abort(); }
