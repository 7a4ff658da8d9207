/* TEST INFRASTRUCTURE ONLY — see divsufsort.h in this directory. */
#ifndef VLGB_ORACLE_DIVSUFSORT64_SHIM_H
#define VLGB_ORACLE_DIVSUFSORT64_SHIM_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
int64_t divsufsort64(const uint8_t* T, int64_t* SA, int64_t n);
#ifdef __cplusplus
}
#endif
#endif
