/* Test-infrastructure shim (NOT product code).
 * Stands in for the Mako-generated <volk_gnsssdr/volk_gnsssdr.h> so that the
 * reference's header-only protokernels under
 * /root/reference/src/algorithms/libs/volk_gnsssdr_module/volk_gnsssdr/kernels/volk_gnsssdr/
 * compile as plain C from where they lie (see oracle/Makefile). */
#ifndef ORACLE_SHIM_VOLK_GNSSSDR_C_H
#define ORACLE_SHIM_VOLK_GNSSSDR_C_H
#include <immintrin.h>
#include <string.h>
#include <volk_gnsssdr/volk_gnsssdr_common.h>
#include <volk_gnsssdr/volk_gnsssdr_complex.h>
#include <volk_gnsssdr/volk_gnsssdr_malloc.h>
size_t volk_gnsssdr_get_alignment(void);
#endif
