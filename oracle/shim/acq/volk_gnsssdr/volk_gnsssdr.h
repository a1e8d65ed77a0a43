// Test-infrastructure shim (NOT product code): the volk_gnsssdr dispatchers pcps_acquisition.cc calls, bound to the reference's own
// GENERIC protokernels (compiled as C in oracle/ref_kernels.c: ref_acq_* below) — the generic variant is the normative arithmetic.
#ifndef ORACLE_SHIM_ACQ_VOLK_GNSSSDR_H
#define ORACLE_SHIM_ACQ_VOLK_GNSSSDR_H
#include <volk/volk_complex.h>
#include <cstddef>
#include <cstdint>
extern "C" {
void* volk_gnsssdr_malloc(size_t size, size_t alignment);
void volk_gnsssdr_free(void* p);
size_t volk_gnsssdr_get_alignment(void);
void ref_sincos_32fc(int arch, float* out_ri, float phase_inc, float* phase, unsigned int n);
void ref_index_max_32u(int arch, uint32_t* target, const float* src, uint32_t n);
void ref_16ic_convert_32fc(float* out_ri, const int16_t* in_ri, unsigned int n);
}
inline void volk_gnsssdr_s32f_sincos_32fc(lv_32fc_t* out, float phase_inc, float* phase, unsigned int n) { ref_sincos_32fc(0, reinterpret_cast<float*>(out), phase_inc, phase, n); }
inline void volk_gnsssdr_32f_index_max_32u(uint32_t* target, const float* src, uint32_t n) { ref_index_max_32u(0, target, src, n); }
inline void volk_gnsssdr_16ic_convert_32fc(lv_32fc_t* out, const lv_16sc_t* in, unsigned int n) { ref_16ic_convert_32fc(reinterpret_cast<float*>(out), reinterpret_cast<const int16_t*>(in), n); }
#endif
