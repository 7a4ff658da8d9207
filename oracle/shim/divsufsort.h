/* TEST INFRASTRUCTURE ONLY (oracle build shim, not product code).
 * The reference's construct_sa.hpp includes "divsufsort.h"/"divsufsort64.h", which its CMake
 * build generates from external/libdivsufsort/include/divsufsort.h.cmake.  The VLG path never
 * calls them (vlg_index.hpp:383 builds csa_wt<wt_int<>>, an int-alphabet CSA, which goes through
 * qsufsort: construct_sa.hpp:162-166), so this shim only declares the two prototypes the header
 * needs to compile; divsufsort_stub.c aborts if either is ever reached. */
#ifndef VLGB_ORACLE_DIVSUFSORT_SHIM_H
#define VLGB_ORACLE_DIVSUFSORT_SHIM_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
int32_t divsufsort(const uint8_t* T, int32_t* SA, int32_t n);
#ifdef __cplusplus
}
#endif
#endif
