/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 * Compiles the reference's own volk_gnsssdr protokernels, as C, from the
 * headers where they lie under /root/reference (never copied into this repo)
 * and exposes them (a) under ref_* names with an explicit protokernel choice
 * and (b) under the public dispatcher names so that the reference's
 * Cpu_Multicorrelator*.cc link unmodified (oracle/ref_cxx.cc).
 * Built only by oracle/Makefile into oracle/_ref/libgnssref.so.
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load it. */
#include <volk_gnsssdr/volk_gnsssdr.h>
#include <stdlib.h>
#include <stdint.h>

#include <volk_gnsssdr/volk_gnsssdr_32f_xn_resampler_32f_xn.h>
#include <volk_gnsssdr/volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h>
#include <volk_gnsssdr/volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h>
#include <volk_gnsssdr/volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h>
#include <volk_gnsssdr/volk_gnsssdr_32fc_xn_resampler_32fc_xn.h>
#include <volk_gnsssdr/volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn.h>
#include <volk_gnsssdr/volk_gnsssdr_s32f_sincos_32fc.h>
#include <volk_gnsssdr/volk_gnsssdr_32f_index_max_32u.h>
#include <volk_gnsssdr/volk_gnsssdr_16ic_convert_32fc.h>
#include <volk_gnsssdr/volk_gnsssdr_16ic_xn_resampler_16ic_xn.h>
#include <volk_gnsssdr/volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn.h>

/* 0 = generic protokernels (normative for parity), 1 = AVX protokernels (what an x86 host runs) */
static int g_arch = 0;
void ref_set_arch(int arch) { g_arch = arch; }
int ref_get_arch(void) { return g_arch; }

void* volk_gnsssdr_malloc(size_t size, size_t alignment)
{
    void* p = NULL;
    if (alignment < sizeof(void*)) alignment = sizeof(void*);
    if (posix_memalign(&p, alignment, size ? size : alignment) != 0) return NULL;
    return p;
}
void volk_gnsssdr_free(void* p) { free(p); }
size_t volk_gnsssdr_get_alignment(void) { return 32; }

void volk_gnsssdr_32f_xn_resampler_32f_xn(float** result, const float* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    if (g_arch == 1)
        volk_gnsssdr_32f_xn_resampler_32f_xn_a_avx(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
    else
        volk_gnsssdr_32f_xn_resampler_32f_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn(float** result, const float* local_code, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    if (g_arch == 1)
        volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_a_avx(result, local_code, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
    else
        volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const float** in_a, int num_a_vectors, unsigned int num_points)
{
    if (g_arch == 1)
        volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_u_avx(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
    else
        volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn_generic(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
}

void volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, const lv_32fc_t phase_inc_rate, lv_32fc_t* phase, const float** in_a, int num_a_vectors, unsigned int num_points)
{
    volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn_generic(result, in_common, phase_inc, phase_inc_rate, phase, in_a, num_a_vectors, num_points);
}

void volk_gnsssdr_32fc_xn_resampler_32fc_xn(lv_32fc_t** result, const lv_32fc_t* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    if (g_arch == 1)
        volk_gnsssdr_32fc_xn_resampler_32fc_xn_a_avx(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
    else
        volk_gnsssdr_32fc_xn_resampler_32fc_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn(lv_32fc_t* result, const lv_32fc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const lv_32fc_t** in_a, int num_a_vectors, unsigned int num_points)
{
    if (g_arch == 1)
        volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn_u_avx(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
    else
        volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn_generic(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
}

/* the int16 path of Cpu_Multicorrelator_16sc (cpu_multicorrelator_16sc.cc:64-91): generic protokernels (the saturating int16 law) */
void volk_gnsssdr_16ic_xn_resampler_16ic_xn(lv_16sc_t** result, const lv_16sc_t* local_code, float rem_code_phase_chips, float code_phase_step_chips, float* shifts_chips, unsigned int code_length_chips, int num_out_vectors, unsigned int num_points)
{
    volk_gnsssdr_16ic_xn_resampler_16ic_xn_generic(result, local_code, rem_code_phase_chips, code_phase_step_chips, shifts_chips, code_length_chips, num_out_vectors, num_points);
}

void volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn(lv_16sc_t* result, const lv_16sc_t* in_common, const lv_32fc_t phase_inc, lv_32fc_t* phase, const lv_16sc_t** in_a, int num_a_vectors, unsigned int num_points)
{
    volk_gnsssdr_16ic_x2_rotator_dot_prod_16ic_xn_generic(result, in_common, phase_inc, phase, in_a, num_a_vectors, num_points);
}

/* ---- explicit-protokernel entry points used by the tests ---- */
void ref_resampler_32f(int arch, float** result, const float* code, float rem, float step, float* shifts, unsigned int L, int ntaps, unsigned int n)
{
    int keep = g_arch; g_arch = arch;
    volk_gnsssdr_32f_xn_resampler_32f_xn(result, code, rem, step, shifts, L, ntaps, n);
    g_arch = keep;
}

/* phase_inc / phase passed as (re,im) float pairs to stay ABI-neutral for ctypes */
void ref_rotator_dot_prod_32fc_32f(int arch, float* result_ri, const float* in_ri, float inc_re, float inc_im, float* phase_ri, const float** in_a, int ntaps, unsigned int n)
{
    lv_32fc_t inc = lv_cmake(inc_re, inc_im);
    lv_32fc_t ph = lv_cmake(phase_ri[0], phase_ri[1]);
    int keep = g_arch; g_arch = arch;
    volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn((lv_32fc_t*)result_ri, (const lv_32fc_t*)in_ri, inc, &ph, in_a, ntaps, n);
    g_arch = keep;
    phase_ri[0] = lv_creal(ph); phase_ri[1] = lv_cimag(ph);
}

/* arch: 0 generic (libm), 1 a_sse2 (cephes polynomial, 4-lane phase accumulation), 2 u_avx2 */
void ref_sincos_32fc(int arch, float* out_ri, float phase_inc, float* phase, unsigned int n)
{
    if (arch == 1) volk_gnsssdr_s32f_sincos_32fc_u_sse2((lv_32fc_t*)out_ri, phase_inc, phase, n);
    else if (arch == 2) volk_gnsssdr_s32f_sincos_32fc_u_avx2((lv_32fc_t*)out_ri, phase_inc, phase, n);
    else volk_gnsssdr_s32f_sincos_32fc_generic((lv_32fc_t*)out_ri, phase_inc, phase, n);
}

void ref_index_max_32u(int arch, uint32_t* target, const float* src, uint32_t n)
{
    if (arch == 1) volk_gnsssdr_32f_index_max_32u_u_avx(target, src, n);
    else volk_gnsssdr_32f_index_max_32u_generic(target, src, n);
}

void ref_16ic_convert_32fc(float* out_ri, const int16_t* in_ri, unsigned int n)
{
    volk_gnsssdr_16ic_convert_32fc_generic((lv_32fc_t*)out_ri, (const lv_16sc_t*)in_ri, n);
}
