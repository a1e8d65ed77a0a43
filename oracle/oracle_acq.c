/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See oracle.h.
 * CPU restatement of pcps_acquisition (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc).
 * Element-wise steps that the reference delegates to upstream VOLK (not in the tree) are plain fp32 loops:
 *   volk_32fc_x2_multiply_32fc      -> (ar*br - ai*bi, ar*bi + ai*br)
 *   volk_32fc_conjugate_32fc        -> (re, -im)
 *   volk_32fc_magnitude_squared_32f -> re*re + im*im
 *   volk_32f_x2_add_32f             -> a + b
 * compiled with -ffp-contract=off (exactly rounded; VOLK's SIMD protokernels give the same bits except where a
 * build fuses multiply-add).  Pinned since round 2 against the reference's own pcps_acquisition.cc compiled unmodified into oracle/_ref
 * (oracle/ref_acq.cc; only the DFT library behind gr::fft is a stand-in): tests/test_acq_refblock_cpu.py — frozen vectors reproduced by
 * the reference block, search grids equal bit for bit.  See oracle.h / DESIGN.md section 2. */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int orc_acq_neff(const orc_acq_conf* c) { return c->bit_transition ? c->fft_size / 2 : c->fft_size; }

static int do_fft(const orc_acq_conf* c, const float* in, float* out, int inverse)
{
    const int n = c->fft_size;
    if (!c->fft_f64) return orc_fft_c32(in, out, n, inverse);
    /* cross-check mode: same data promoted to float64, transformed, rounded back */
    double* a = (double*)malloc(sizeof(double) * 2 * (size_t)n);
    double* b = (double*)malloc(sizeof(double) * 2 * (size_t)n);
    if (!a || !b) { free(a); free(b); return -2; }
    for (int i = 0; i < 2 * n; i++) a[i] = in[i];
    int rc = orc_fft_c64(a, b, n, inverse);
    for (int i = 0; i < 2 * n; i++) out[i] = (float)b[i];
    free(a);
    free(b);
    return rc;
}

/* pcps_acquisition.cc:175-208 */
int orc_acq_set_local_code(const orc_acq_conf* c, const float* code_ri, float* fft_codes_ri)
{
    const int n = c->fft_size;
    float* buf = (float*)calloc(2 * (size_t)n, sizeof(float));
    if (!buf) return -2;
    if (c->bit_transition)
        {
            const int offset = n / 2; /* :188-190: zeros first, first fft_size/2 code samples in the second half */
            memcpy(buf + 2 * offset, code_ri, sizeof(float) * 2 * (size_t)offset);
        }
    else if (n == c->consumed)
        {
            memcpy(buf, code_ri, sizeof(float) * 2 * (size_t)n); /* :196 */
        }
    else
        {
            /* :200-201: fft_size - consumed zeros, then `consumed` code samples placed at offset `consumed` */
            memcpy(buf + 2 * (size_t)c->consumed, code_ri, sizeof(float) * 2 * (size_t)c->consumed);
        }
    int rc = do_fft(c, buf, fft_codes_ri, 0);
    for (int i = 0; i < n; i++) fft_codes_ri[2 * i + 1] = -fft_codes_ri[2 * i + 1]; /* :207 conjugate */
    free(buf);
    return rc;
}

/* pcps_acquisition.cc:232-245 -> volk_gnsssdr_s32f_sincos_32fc_generic
 * (kernels/volk_gnsssdr/volk_gnsssdr_s32f_sincos_32fc.h:390-401): fp32 phase accumulated by repeated addition from 0 */
void orc_acq_wipeoff(float* out_ri, int n, float freq, double fs)
{
    const float two_pi = (float)6.283185307179586; /* static_cast<float>(TWO_PI) */
    const float phase_step_rad = two_pi * freq / (float)fs;
    const float inc = -phase_step_rad;
    float ph = 0.0f;
    for (int i = 0; i < n; i++)
        {
            out_ri[2 * i] = cosf(ph);
            out_ri[2 * i + 1] = sinf(ph);
            ph += inc;
        }
}

/* pcps_acquisition.cc:600-672, one call of acquisition_core's Doppler loop (not step two) */
int orc_acq_dwell(const orc_acq_conf* c, const float* in_ri, const float* fft_codes_ri, float* grid, int dwell_index, int nthreads)
{
    const int n = c->fft_size, neff = orc_acq_neff(c), B = c->num_doppler_bins;
    const size_t offset = c->bit_transition ? (size_t)neff : 0; /* :656 */
    int rc_all = 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        float* sig = (float*)calloc(2 * (size_t)n, sizeof(float));
        float* wip = (float*)malloc(sizeof(float) * 2 * (size_t)n);
        float* a = (float*)malloc(sizeof(float) * 2 * (size_t)n);
        float* b = (float*)malloc(sizeof(float) * 2 * (size_t)n);
        /* :612-619 copy consumed samples, zero-pad to fft_size */
        memcpy(sig, in_ri, sizeof(float) * 2 * (size_t)c->consumed);
#pragma omp for schedule(dynamic, 1)
        for (int k = 0; k < B; k++)
            {
                /* :299-300 doppler of the bin; GLONASS bias is 0 for every signal in scope */
                const int doppler = -c->doppler_max + c->doppler_center + c->doppler_step * k;
                orc_acq_wipeoff(wip, n, (float)doppler, c->fs_in);
                for (int i = 0; i < n; i++) /* :643 */
                    {
                        const float xr = sig[2 * i], xi = sig[2 * i + 1], wr = wip[2 * i], wi = wip[2 * i + 1];
                        a[2 * i] = xr * wr - xi * wi;
                        a[2 * i + 1] = xr * wi + xi * wr;
                    }
                int rc = do_fft(c, a, b, 0); /* :647 */
                for (int i = 0; i < n; i++) /* :650 */
                    {
                        const float xr = b[2 * i], xi = b[2 * i + 1], wr = fft_codes_ri[2 * i], wi = fft_codes_ri[2 * i + 1];
                        a[2 * i] = xr * wr - xi * wi;
                        a[2 * i + 1] = xr * wi + xi * wr;
                    }
                rc |= do_fft(c, a, b, 1); /* :653 */
                float* g = grid + (size_t)k * neff;
                const float* o = b + 2 * offset;
                if (dwell_index == 0) /* :657-660 */
                    for (int i = 0; i < neff; i++) g[i] = o[2 * i] * o[2 * i] + o[2 * i + 1] * o[2 * i + 1];
                else /* :661-665 */
                    for (int i = 0; i < neff; i++)
                        {
                            const float m = o[2 * i] * o[2 * i] + o[2 * i + 1] * o[2 * i + 1];
                            g[i] = g[i] + m;
                        }
                if (rc)
                    {
#pragma omp atomic write
                        rc_all = rc;
                    }
            }
        free(sig);
        free(wip);
        free(a);
        free(b);
    }
    return rc_all;
}

/* kernels/volk_gnsssdr/volk_gnsssdr_32f_index_max_32u.h:446-465 (first index wins ties) */
static uint32_t index_max(const float* s, uint32_t n)
{
    float mx = s[0];
    uint32_t idx = 0;
    for (uint32_t i = 1; i < n; i++)
        if (s[i] > mx)
            {
                idx = i;
                mx = s[i];
            }
    return idx;
}

/* pcps_acquisition.cc:496-528 (CFAR) / :531-597 (first vs second peak) and :691-696 */
int orc_acq_statistic(const orc_acq_conf* c, const float* grid, int num_dwells, orc_acq_result* r)
{
    const int neff = orc_acq_neff(c), B = c->num_doppler_bins;
    /* first_vs_second_peak_statistic searches d_fft_size entries per bin (:543): the tail beyond neff is all zeros
     * (never written), so with strict '>' the result equals a search over neff. */
    float grid_max = 0.0f;
    uint32_t index_doppler = 0, index_time = 0;
    for (int i = 0; i < B; i++)
        {
            const float* g = grid + (size_t)i * neff;
            uint32_t t = index_max(g, neff);
            if (g[t] > grid_max)
                {
                    grid_max = g[t];
                    index_doppler = i;
                    index_time = t;
                }
        }
    r->indext = index_time;
    r->doppler_index = (int32_t)index_doppler;
    r->doppler_hz = -c->doppler_max + c->doppler_center + c->doppler_step * (int32_t)index_doppler;
    r->peak = grid_max;
    if (c->use_cfar)
        {
            const uint32_t opp = (index_doppler + B / 2) % B; /* :518 */
            const float* g = grid + (size_t)opp * neff;
            float acc = 0.0f; /* std::accumulate with a float init: sequential fp32 sum (:519) */
            for (int i = 0; i < neff; i++) acc += g[i];
            r->input_power = (float)(acc / neff / 2.0 / num_dwells);
            r->test_statistics = grid_max / r->input_power;
        }
    else
        {
            /* :543,:565-593 work on d_fft_size entries per bin; with bit_transition only the first neff were ever
             * written (rest stay 0 from init, :281-284), so the exclusion wrap lands in that zero tail. */
            const int32_t nfull = c->fft_size;
            int32_t ex1 = (int32_t)index_time - c->samples_per_chip, ex2 = (int32_t)index_time + c->samples_per_chip;
            if (ex1 < 0) ex1 = nfull + ex1;
            else if (ex2 >= nfull) ex2 = ex2 - nfull;
            float* tmp = (float*)calloc((size_t)nfull, sizeof(float));
            if (!tmp) return -2;
            memcpy(tmp, grid + (size_t)index_doppler * neff, sizeof(float) * (size_t)neff);
            int32_t idx = ex1;
            do
                {
                    tmp[idx] = 0.0f;
                    idx++;
                    if (idx == nfull) idx = 0;
                }
            while (idx != ex2);
            const uint32_t t2 = index_max(tmp, nfull);
            r->input_power = tmp[t2];
            r->test_statistics = grid_max / tmp[t2];
            free(tmp);
        }
    r->acq_delay_samples = fmodf((float)index_time, c->samples_per_code); /* :693 */
    r->acq_doppler_hz = (float)r->doppler_hz;                             /* :694 */
    return 0;
}

/* regularised upper incomplete gamma Q(a, x) for integer a >= 1: exp(-x) * sum_{k<a} x^k / k! */
static double gamma_q_int(int a, double x)
{
    double term = 1.0, sum = 1.0;
    for (int k = 1; k < a; k++)
        {
            term *= x / k;
            sum += term;
        }
    return exp(-x) * sum;
}

/* pcps_acquisition.cc:884-899: threshold = 2 * gamma_p_inv(2*dwells, (1-pfa)^(1/num_bins)).
 * Boost.Math is not in the tree; P(a,x) = p is solved through Q(a,x) = 1 - p by bisection on log Q (a integer). */
float orc_acq_threshold(const orc_acq_conf* c, float pfa)
{
    if (pfa <= 0.0f) return 0.0f;
    const int num_bins = orc_acq_neff(c) * c->num_doppler_bins;
    const int a = 2 * (c->bit_transition ? 1 : c->max_dwells);
    const double p = pow(1.0 - pfa, 1.0 / (float)num_bins); /* double pow of (double - float), float reciprocal: as written at :898 */
    const double q = 1.0 - p;
    double lo = 0.0, hi = 1.0;
    while (gamma_q_int(a, hi) > q) hi *= 2.0;
    for (int it = 0; it < 200; it++)
        {
            const double mid = 0.5 * (lo + hi);
            if (gamma_q_int(a, mid) > q) lo = mid;
            else hi = mid;
        }
    return (float)(2.0 * 0.5 * (lo + hi));
}
