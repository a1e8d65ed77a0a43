/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See oracle.h.
 * Unnormalised complex DFT, any length whose prime factors are <= 61
 * (the d_fft_size of the reference's configurations: 4000, 16000, 20000,
 * 25000, 32000, 100000 ... and 6625 = 5^3 * 53, 5456 = 2^4 * 11 * 31 ...).  Stand-in for FFTW3f reached through
 * gr::fft::fft_complex_fwd/rev (src/algorithms/libs/gnss_sdr_fft.h:27-61;
 * call sites pcps_acquisition.cc:113-114,206,647,653) which is not part of
 * the reference tree.  Stockham autosort, twiddles from a float64 table. */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define MAXF 40

static int factorize(int n, int* radix)
{
    int nf = 0;
    while (n % 4 == 0) { radix[nf++] = 4; n /= 4; }
    while (n % 2 == 0) { radix[nf++] = 2; n /= 2; }
    while (n % 3 == 0) { radix[nf++] = 3; n /= 3; }
    while (n % 5 == 0) { radix[nf++] = 5; n /= 5; }
    while (n % 7 == 0) { radix[nf++] = 7; n /= 7; }
    for (int p = 11; p <= 61; p += 2)
        while (n % p == 0) { radix[nf++] = p; n /= p; }
    return n == 1 ? nf : -1;
}

/* twiddle tables exp(+-2*pi*i*k/n), computed in float64 and rounded once; cached per (n, direction, type) */
#define NCACHE 64
static struct { int n, inverse, is_double; void* w; } g_cache[NCACHE];
static int g_ncache = 0;

static const void* twiddles(int n, int inverse, int is_double)
{
    const void* found = NULL;
#pragma omp critical(orc_fft_cache)
    {
        for (int i = 0; i < g_ncache; i++)
            if (g_cache[i].n == n && g_cache[i].inverse == inverse && g_cache[i].is_double == is_double) found = g_cache[i].w;
        if (!found)
            {
                void* w = malloc((is_double ? sizeof(double) : sizeof(float)) * 2 * (size_t)n);
                const double sgn = inverse ? 1.0 : -1.0;
                for (int k = 0; w && k < n; k++)
                    {
                        const double ang = sgn * 2.0 * M_PI * (double)k / (double)n;
                        if (is_double) { ((double*)w)[2 * k] = cos(ang); ((double*)w)[2 * k + 1] = sin(ang); }
                        else { ((float*)w)[2 * k] = (float)cos(ang); ((float*)w)[2 * k + 1] = (float)sin(ang); }
                    }
                if (g_ncache < NCACHE) /* never evicts (readers hold raw pointers); a full cache just stops caching */
                    {
                        const int slot = g_ncache++;
                        g_cache[slot].n = n; g_cache[slot].inverse = inverse; g_cache[slot].is_double = is_double; g_cache[slot].w = w;
                    }
                found = w;
            }
    }
    return found;
}

#define DEFINE_FFT(NAME, T)                                                                        \
    static int NAME(const T* in, T* out, int n, int inverse)                                       \
    {                                                                                              \
        int radix[MAXF];                                                                           \
        const int nf = factorize(n, radix);                                                        \
        if (nf < 0) return -1;                                                                     \
        const T* w = (const T*)twiddles(n, inverse, sizeof(T) == sizeof(double));                  \
        T* a = (T*)malloc(sizeof(T) * 2 * (size_t)n);                                              \
        T* b = (T*)malloc(sizeof(T) * 2 * (size_t)n);                                              \
        if (!w || !a || !b) { free(a); free(b); return -2; }                                      \
        memcpy(a, in, sizeof(T) * 2 * (size_t)n);                                                  \
        int ns = 1;                                                                                \
        for (int s = 0; s < nf; s++)                                                               \
            {                                                                                      \
                const int R = radix[s];                                                            \
                const int m = n / R;                                                               \
                const int tw_stride = n / (ns * R);                                                \
                for (int j = 0; j < m; j++)                                                        \
                    {                                                                              \
                        const int k = j % ns;                                                      \
                        T vr[64], vi[64];                                                            \
                        for (int r = 0; r < R; r++)                                                \
                            {                                                                      \
                                const T xr = a[2 * (j + r * m)], xi = a[2 * (j + r * m) + 1];      \
                                const size_t ti = (size_t)k * r * tw_stride;                       \
                                const T c = w[2 * ti], sn = w[2 * ti + 1];                         \
                                vr[r] = xr * c - xi * sn;                                          \
                                vi[r] = xr * sn + xi * c;                                          \
                            }                                                                      \
                        const int j0 = (j / ns) * ns * R + k;                                      \
                        const int dft_stride = n / R;                                              \
                        for (int q = 0; q < R; q++)                                                \
                            {                                                                      \
                                T sr = vr[0], si = vi[0];                                          \
                                for (int r = 1; r < R; r++)                                        \
                                    {                                                              \
                                        const size_t ti = (size_t)((q * r) % R) * dft_stride;      \
                                        const T c = w[2 * ti], sn = w[2 * ti + 1];                 \
                                        sr += vr[r] * c - vi[r] * sn;                              \
                                        si += vr[r] * sn + vi[r] * c;                              \
                                    }                                                              \
                                b[2 * (j0 + q * ns)] = sr;                                         \
                                b[2 * (j0 + q * ns) + 1] = si;                                     \
                            }                                                                      \
                    }                                                                              \
                ns *= R;                                                                           \
                T* t = a; a = b; b = t;                                                            \
            }                                                                                      \
        memcpy(out, a, sizeof(T) * 2 * (size_t)n);                                                 \
        free(a); free(b);                                                                          \
        return 0;                                                                                  \
    }

DEFINE_FFT(fft_f32, float)
DEFINE_FFT(fft_f64, double)

int orc_fft_c32(const float* in_ri, float* out_ri, int n, int inverse) { return fft_f32(in_ri, out_ri, n, inverse); }
int orc_fft_c64(const double* in_ri, double* out_ri, int n, int inverse) { return fft_f64(in_ri, out_ri, n, inverse); }
