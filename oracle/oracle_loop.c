/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See oracle.h.
 * Synthetic IF generator for the closed-loop tests (SURVEY 8d signal model, the one tools/gen_if.py states in numpy):
 *   s[i] = sum_k A_k * d_k(bit) * c_k(floor(chips_k(i)) mod L) * exp(j(2 pi fd_k t_i + phi_k)) + CN(0, 1),   t_i = i / fs
 *   chips_k(i) = t_i * f_chip * (1 + fd_k / f_carrier) - delay_chips_k,   bit = floor(chips_k / (L * periods_per_bit))
 * Stateless in the absolute sample index (counter-based noise), so a stream can be produced chunk by chunk.
 * The loop-closure pieces themselves are checked against the reference's own objects in oracle/_ref (ref_loop.cc). */
#include "oracle.h"
#include <math.h>
#include <stdint.h>

int orc_version(void) { return 2; }

static inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

/* codes: nsat tables of L floats; nav_bits: nsat rows of nbits entries (+-1), wrapped; amp from CN0: sqrt(10^(cn0/10) / fs) */
void orc_gen_if(float* out_ri, int64_t n, int64_t first_index, double fs, int nsat, const float* codes, int L, double f_chip, double f_carrier,
    const double* doppler_hz, const double* delay_chips, const double* phase0, const double* cn0_dbhz, const signed char* nav_bits, int nbits,
    int periods_per_bit, double noise_sigma, uint64_t seed)
{
    const double two_pi = 6.283185307179586476925286766559;
    double amp_k[64];
    if (nsat > 64) nsat = 64;
    for (int k = 0; k < nsat; k++) amp_k[k] = sqrt(pow(10.0, cn0_dbhz[k] / 10.0) / fs);
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < n; j++)
        {
            const int64_t i = first_index + j;
            const double t = (double)i / fs;
            double re = 0.0, im = 0.0;
            for (int k = 0; k < nsat; k++)
                {
                    const double amp = amp_k[k];
                    const double chips = t * f_chip * (1.0 + doppler_hz[k] / f_carrier) - delay_chips[k];
                    const double fl = floor(chips);
                    int64_t ci = (int64_t)fl % L;
                    if (ci < 0) ci += L;
                    double v = amp * codes[(int64_t)k * L + ci];
                    if (nav_bits && nbits > 0)
                        {
                            int64_t b = (int64_t)floor(chips / ((double)L * periods_per_bit)) % nbits;
                            if (b < 0) b += nbits;
                            v *= nav_bits[(int64_t)k * nbits + b];
                        }
                    /* phase reduced in turns before the sincos to keep precision over minutes of signal */
                    double turns = doppler_hz[k] * t;
                    turns -= floor(turns);
                    const double ph = two_pi * turns + phase0[k];
                    re += v * cos(ph);
                    im += v * sin(ph);
                }
            if (noise_sigma > 0.0)
                {
                    const uint64_t r1 = splitmix64(seed ^ (uint64_t)i * 0xD6E8FEB86659FD93ULL);
                    const uint64_t r2 = splitmix64(r1);
                    const double u1 = ((double)(r1 >> 11) + 1.0) / 9007199254740993.0; /* (0,1] */
                    const double u2 = (double)(r2 >> 11) / 9007199254740992.0;
                    const double r = noise_sigma * sqrt(-log(u1)); /* complex variance sigma^2: each part sigma^2 / 2 */
                    re += r * cos(two_pi * u2);
                    im += r * sin(two_pi * u2);
                }
            out_ri[2 * j] = (float)re;
            out_ri[2 * j + 1] = (float)im;
        }
}
