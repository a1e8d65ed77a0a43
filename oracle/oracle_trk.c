/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See oracle.h.
 * Scalar C restatement of the tracking multicorrelator
 * (Cpu_Multicorrelator_Real_Codes / Cpu_Multicorrelator and the generic
 * volk_gnsssdr protokernels they call).  Compiled with -ffp-contract=off so
 * every fp32 operation rounds exactly as the reference's generic C does. */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* index law: kernels/volk_gnsssdr/volk_gnsssdr_32f_xn_resampler_32f_xn.h:73-76 */
static inline int chip_index(float step, unsigned n, float shift, float rem, unsigned L)
{
    int idx = (int)floor(step * (float)n + shift - rem);
    if (idx < 0) idx += (int)L * (abs(idx) / L + 1);
    return idx % L;
}

void orc_resampler_index(int32_t* idx, float rem, float step, const float* shifts, unsigned L, int ntaps, unsigned n)
{
    for (int t = 0; t < ntaps; t++)
        for (unsigned i = 0; i < n; i++) idx[(size_t)t * n + i] = chip_index(step, i, shifts[t], rem, L);
}

void orc_resampler_32f(float** result, const float* code, float rem, float step, const float* shifts, unsigned L, int ntaps, unsigned n)
{
    for (int t = 0; t < ntaps; t++)
        for (unsigned i = 0; i < n; i++) result[t][i] = code[chip_index(step, i, shifts[t], rem, L)];
}

/* kernels/volk_gnsssdr/volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn.h:67-91 */
void orc_resampler_high_dyn_32f(float** result, const float* code, float rem, float step, float rate, const float* shifts, unsigned L, int ntaps, unsigned n)
{
    for (unsigned i = 0; i < n; i++)
        {
            /* (n * n) is an unsigned int product in the reference and wraps above 65535 */
            int idx = (int)floor(step * (float)i + rate * (float)(i * i) + shifts[0] - rem);
            if (idx < 0) idx += (int)L * (abs(idx) / L + 1);
            result[0][i] = code[idx % L];
        }
    unsigned shift_samples = 0;
    for (int t = 1; t < ntaps; t++)
        {
            shift_samples += (int)round((shifts[t] - shifts[t - 1]) / step);
            memcpy(&result[t][0], &result[0][shift_samples], (n - shift_samples) * sizeof(float));
            memcpy(&result[t][n - shift_samples], &result[0][0], shift_samples * sizeof(float));
        }
}

/* complex helpers written out so that no libgcc __mulsc3 special-casing is involved */
typedef struct { float re, im; } cf32;
static inline cf32 cmul(cf32 a, cf32 b) { cf32 r = {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; return r; }

/* kernels/volk_gnsssdr/volk_gnsssdr_32fc_32f_rotator_dot_prod_32fc_xn.h:66-98 */
void orc_rotator_dot_prod_32fc_32f(float* result_ri, const float* in_ri, float inc_re, float inc_im, float* phase_ri, const float* const* in_a, int ntaps, unsigned n)
{
    cf32 ph = {phase_ri[0], phase_ri[1]};
    const cf32 inc = {inc_re, inc_im};
    for (int t = 0; t < ntaps; t++) result_ri[2 * t] = result_ri[2 * t + 1] = 0.0f;
    for (unsigned i = 0; i < n; i++)
        {
            cf32 x = {in_ri[2 * i], in_ri[2 * i + 1]};
            cf32 w = cmul(x, ph);
            if (i % 256 == 0)
                {
                    float h = hypotf(ph.re, ph.im);
                    ph.re /= h;
                    ph.im /= h;
                }
            ph = cmul(ph, inc);
            for (int t = 0; t < ntaps; t++)
                {
                    float c = in_a[t][i];
                    result_ri[2 * t] += w.re * c;
                    result_ri[2 * t + 1] += w.im * c;
                }
        }
    phase_ri[0] = ph.re;
    phase_ri[1] = ph.im;
}

/* kernels/volk_gnsssdr/volk_gnsssdr_32fc_x2_rotator_dot_prod_32fc_xn.h:67-101 */
void orc_rotator_dot_prod_32fc_x2(float* result_ri, const float* in_ri, float inc_re, float inc_im, float* phase_ri, const float* const* in_a_ri, int ntaps, unsigned n)
{
    cf32 ph = {phase_ri[0], phase_ri[1]};
    const cf32 inc = {inc_re, inc_im};
    for (int t = 0; t < ntaps; t++) result_ri[2 * t] = result_ri[2 * t + 1] = 0.0f;
    for (unsigned i = 0; i < n; i++)
        {
            cf32 x = {in_ri[2 * i], in_ri[2 * i + 1]};
            cf32 w = cmul(x, ph);
            if (i % 256 == 0)
                {
                    float h = hypotf(ph.re, ph.im);
                    ph.re /= h;
                    ph.im /= h;
                }
            ph = cmul(ph, inc);
            for (int t = 0; t < ntaps; t++)
                {
                    cf32 c = {in_a_ri[t][2 * i], in_a_ri[t][2 * i + 1]};
                    cf32 p = cmul(w, c);
                    result_ri[2 * t] += p.re;
                    result_ri[2 * t + 1] += p.im;
                }
        }
    phase_ri[0] = ph.re;
    phase_ri[1] = ph.im;
}

/* kernels/volk_gnsssdr/volk_gnsssdr_32fc_32f_high_dynamic_rotator_dot_prod_32fc_xn.h:68-109.
 * The reference evaluates cpowf(phase_inc_rate, n*n) per sample; cpowf(z, k) for |z|=1 is
 * exp(j*k*arg z) up to libm rounding, which is what the reference's own _WIN32 branch (:93-95)
 * and _generic_arg protokernel (:114-160) compute.  We follow that closed form. */
static void rotator_dot_prod_high_dyn(float* result_ri, const float* in_ri, cf32 inc, cf32 inc_rate, cf32* phase, float* const* in_a, int ntaps, unsigned n)
{
    cf32 ph = *phase;
    cf32 ph_doppler = ph;
    const float arga = atan2f(inc_rate.im, inc_rate.re);
    for (int t = 0; t < ntaps; t++) result_ri[2 * t] = result_ri[2 * t + 1] = 0.0f;
    for (unsigned i = 0; i < n; i++)
        {
            if (i % 256 == 0)
                {
                    float h = hypotf(ph.re, ph.im);
                    ph.re /= h;
                    ph.im /= h;
                }
            cf32 x = {in_ri[2 * i], in_ri[2 * i + 1]};
            cf32 w = cmul(x, ph);
            ph_doppler = cmul(ph_doppler, inc);
            const float theta = (float)(i * i) * arga;
            cf32 pr = {cosf(theta), sinf(theta)};
            ph = cmul(ph_doppler, pr);
            for (int t = 0; t < ntaps; t++)
                {
                    float c = in_a[t][i];
                    result_ri[2 * t] += w.re * c;
                    result_ri[2 * t + 1] += w.im * c;
                }
        }
    *phase = ph;
}

/* std::exp(lv_32fc_t(0, -step)) as libstdc++ evaluates it for complex<float>: polar(exp(0), -step) */
static inline cf32 phase_inc_of(float step)
{
    cf32 r = {cosf(-step), sinf(-step)};
    return r;
}

/* src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:75-126 */
int orc_mc_real_codes(int high_dyn, const float* sig_ri, int n, const float* code, int L, const float* shifts, int ntaps,
    float rem_carr, float phase_step, float phase_rate, float rem_code, float code_step, float code_rate, float* out_ri)
{
    float** rep = (float**)malloc(sizeof(float*) * ntaps);
    if (!rep) return -1;
    for (int t = 0; t < ntaps; t++) rep[t] = (float*)malloc(sizeof(float) * (size_t)n);
    float ph[2] = {cosf(rem_carr), -sinf(rem_carr)};
    cf32 inc = phase_inc_of(phase_step);
    if (high_dyn)
        {
            orc_resampler_high_dyn_32f(rep, code, rem_code, code_step, code_rate, shifts, L, ntaps, n);
            cf32 p = {ph[0], ph[1]};
            rotator_dot_prod_high_dyn(out_ri, sig_ri, inc, phase_inc_of(phase_rate), &p, rep, ntaps, n);
        }
    else
        {
            orc_resampler_32f(rep, code, rem_code, code_step, shifts, L, ntaps, n);
            orc_rotator_dot_prod_32fc_32f(out_ri, sig_ri, inc.re, inc.im, ph, (const float* const*)rep, ntaps, n);
        }
    for (int t = 0; t < ntaps; t++) free(rep[t]);
    free(rep);
    return 0;
}

/* src/algorithms/tracking/libs/cpu_multicorrelator.cc:73-100; resampler law
 * kernels/volk_gnsssdr/volk_gnsssdr_32fc_xn_resampler_32fc_xn.h:63-80 (same index law, complex elements) */
int orc_mc_complex_codes(const float* sig_ri, int n, const float* code_ri, int L, const float* shifts, int ntaps,
    float rem_carr, float phase_step, float rem_code, float code_step, float* out_ri)
{
    float** rep = (float**)malloc(sizeof(float*) * ntaps);
    if (!rep) return -1;
    for (int t = 0; t < ntaps; t++)
        {
            rep[t] = (float*)malloc(sizeof(float) * 2 * (size_t)n);
            for (int i = 0; i < n; i++)
                {
                    int k = chip_index(code_step, i, shifts[t], rem_code, L);
                    rep[t][2 * i] = code_ri[2 * k];
                    rep[t][2 * i + 1] = code_ri[2 * k + 1];
                }
        }
    float ph[2] = {cosf(rem_carr), -sinf(rem_carr)};
    cf32 inc = phase_inc_of(phase_step);
    orc_rotator_dot_prod_32fc_x2(out_ri, sig_ri, inc.re, inc.im, ph, (const float* const*)rep, ntaps, n);
    for (int t = 0; t < ntaps; t++) free(rep[t]);
    free(rep);
    return 0;
}

/* float64 truth of the same call (identical fp32 chip-index law) */
int orc_mc_real_codes_f64(const float* sig_ri, int n, const float* code, int L, const float* shifts, int ntaps,
    float rem_carr, float phase_step, float rem_code, float code_step, int mimic_inc, double* out_ri)
{
    double step = (double)phase_step;
    double phi0 = (double)rem_carr;
    if (mimic_inc)
        {
            cf32 inc = phase_inc_of(phase_step);
            step = -atan2((double)inc.im, (double)inc.re);
            phi0 = -atan2((double)(-sinf(rem_carr)), (double)cosf(rem_carr));
        }
    for (int t = 0; t < ntaps; t++) out_ri[2 * t] = out_ri[2 * t + 1] = 0.0;
    for (int i = 0; i < n; i++)
        {
            double th = phi0 + step * (double)i;
            double c = cos(th), s = -sin(th);
            double xr = sig_ri[2 * i], xi = sig_ri[2 * i + 1];
            double wr = xr * c - xi * s, wi = xr * s + xi * c;
            for (int t = 0; t < ntaps; t++)
                {
                    double cc = code[chip_index(code_step, i, shifts[t], rem_code, L)];
                    out_ri[2 * t] += wr * cc;
                    out_ri[2 * t + 1] += wi * cc;
                }
        }
    return 0;
}

/* Scalar CPU baseline ("port"): every channel correlates its own window of one shared IQ ring,
 * exactly one Cpu_Multicorrelator_Real_Codes call per channel (cpu_multicorrelator_real_codes.cc:103-126),
 * one OpenMP thread per group of channels like the reference's thread-per-channel scheduling. */
int orc_mc_batch(const float* ring_ri, int64_t ring_len, int nch, int ntaps, const int64_t* offsets, const int32_t* lens,
    const float* codes, int L, const float* shifts, const float* rem_carr, const float* phase_step, const float* rem_code,
    const float* code_step, float* out_ri, int nthreads)
{
    int rc = 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic, 1)
    for (int ch = 0; ch < nch; ch++)
        {
            if (offsets[ch] < 0 || offsets[ch] + lens[ch] > ring_len)
                {
                    rc = -2;
                    continue;
                }
            orc_mc_real_codes(0, ring_ri + 2 * offsets[ch], lens[ch], codes + (size_t)ch * L, L, shifts + (size_t)ch * ntaps, ntaps,
                rem_carr[ch], phase_step[ch], 0.0f, rem_code[ch], code_step[ch], 0.0f, out_ri + (size_t)ch * ntaps * 2);
        }
    return rc;
}
