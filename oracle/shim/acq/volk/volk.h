// Test-infrastructure stub (NOT product code): the five VOLK kernels pcps_acquisition.cc calls, as the generic loops of VOLK's own
// *_generic protokernels — all element-wise (no reductions), so SIMD variants give the same bits; the products are written out as the
// separately rounded real operations of the generic C (the library is compiled with -ffp-contract=off).
#ifndef ORACLE_SHIM_ACQ_VOLK_H
#define ORACLE_SHIM_ACQ_VOLK_H
#include <volk/volk_complex.h>
#include <cstring>  // (the real headers pull these in transitively; pcps_acquisition.cc relies on it)
#include <numeric>
inline void volk_32fc_x2_multiply_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t* b, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++)
        {
            const float ar = a[i].real(), ai = a[i].imag(), br = b[i].real(), bi = b[i].imag();
            c[i] = lv_32fc_t(ar * br - ai * bi, ar * bi + ai * br);
        }
}
inline void volk_32fc_conjugate_32fc(lv_32fc_t* c, const lv_32fc_t* a, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) c[i] = std::conj(a[i]);
}
inline void volk_32fc_magnitude_squared_32f(float* m, const lv_32fc_t* a, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++)
        {
            const float r = a[i].real(), q = a[i].imag();
            m[i] = (r * r) + (q * q);
        }
}
inline void volk_32f_x2_add_32f(float* c, const float* a, const float* b, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) c[i] = a[i] + b[i];
}
// (galileo_e5a_noncoherentIQ_acquisition_caf_cc::set_local_code: the sign flip of the B replicas; VOLK's generic loop, a full complex product)
inline void volk_32fc_s32fc_multiply_32fc(lv_32fc_t* c, const lv_32fc_t* a, const lv_32fc_t s, unsigned int n)
{
    for (unsigned int i = 0; i < n; i++) c[i] = a[i] * s;
}
// (pcps_acquisition_fine_doppler_cc::estimate_input_power only, which general_work never calls: VOLK's generic running sum)
inline void volk_32f_accumulator_s32f(float* result, const float* a, unsigned int n)
{
    float acc = 0.0F;
    for (unsigned int i = 0; i < n; i++) acc += a[i];
    *result = acc;
}
#endif
