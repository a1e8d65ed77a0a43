/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See oracle.h.
 * PRN / sampled-replica generators restated from the reference; pinned
 * bit-for-bit against oracle/_ref (tests/test_oracle_pin.py). */
#include "oracle.h"
#include <math.h>
#include <string.h>

/* G2 delay per PRN 1..210 (IS-GPS-200 Table 3-Ia/3-Ib; the reference carries the same ICD table at
 * src/algorithms/libs/gps_sdr_signal_replica.cc:41-53 and indexes it with prn-1 for every PRN, :64) */
static const int16_t k_g2_delay[210] = {
    5, 6, 7, 8, 17, 18, 139, 140, 141, 251, 252, 254, 255, 256, 257, 258,
    469, 470, 471, 472, 473, 474, 509, 512, 513, 514, 515, 516, 859, 860, 861, 862,
    863, 950, 947, 948, 950, 67, 103, 91, 19, 679, 225, 625, 946, 638, 161, 1001,
    554, 280, 710, 709, 775, 864, 558, 220, 397, 55, 898, 759, 367, 299, 1018, 729,
    695, 780, 801, 788, 732, 34, 320, 327, 389, 407, 525, 405, 221, 761, 260, 326,
    955, 653, 699, 422, 188, 438, 959, 539, 879, 677, 586, 153, 792, 814, 446, 264,
    1015, 278, 536, 819, 156, 957, 159, 712, 885, 461, 248, 713, 126, 807, 279, 122,
    197, 693, 632, 771, 467, 647, 203, 145, 175, 52, 21, 237, 235, 886, 657, 634,
    762, 355, 1012, 176, 603, 130, 359, 595, 68, 386, 797, 456, 499, 883, 307, 127,
    211, 121, 118, 163, 628, 853, 484, 289, 811, 202, 1021, 463, 568, 904, 670, 230,
    911, 684, 309, 644, 932, 12, 314, 891, 212, 185, 675, 503, 150, 395, 345, 846,
    798, 992, 357, 995, 877, 112, 144, 476, 193, 109, 445, 291, 87, 399, 292, 901,
    339, 208, 711, 189, 263, 537, 663, 942, 173, 900, 30, 500, 935, 556, 373, 85,
    652, 310};

/* src/algorithms/libs/gps_sdr_signal_replica.cc:25-110.  Registers are kept as bit arrays r[0..9] with
 * output r[0], G1 feedback r[7]^r[0], G2 feedback r[8]^r[7]^r[4]^r[2]^r[1]^r[0], shifted towards index 0. */
void orc_gps_l1_ca_code_gen_int(int32_t* dest, int prn, unsigned chip_shift)
{
    enum { L = 1023 };
    uint8_t g1[L], g2[L], r1[10], r2[10];
    const int prn_idx = prn - 1;
    if (prn_idx < 0 || prn_idx > 209) return; /* reference allows 210 and reads past the table; we stop at the table end */
    memset(r1, 1, sizeof r1);
    memset(r2, 1, sizeof r2);
    for (int i = 0; i < L; i++)
        {
            g1[i] = r1[0];
            g2[i] = r2[0];
            uint8_t f1 = r1[7] ^ r1[0];
            uint8_t f2 = r2[8] ^ r2[7] ^ r2[4] ^ r2[2] ^ r2[1] ^ r2[0];
            memmove(r1, r1 + 1, 9);
            memmove(r2, r2 + 1, 9);
            r1[9] = f1;
            r2[9] = f2;
        }
    unsigned delay = (L - k_g2_delay[prn_idx] + chip_shift) % L;
    for (unsigned i = 0; i < L; i++)
        {
            dest[i] = (g1[(i + chip_shift) % L] ^ g2[delay]) ? 1 : -1;
            delay = (delay + 1) % L;
        }
}

/* gps_sdr_signal_replica.cc:113-124 */
void orc_gps_l1_ca_code_gen_float(float* dest, int prn, unsigned chip_shift)
{
    int32_t c[1023];
    memset(c, 0, sizeof c);
    orc_gps_l1_ca_code_gen_int(c, prn, chip_shift);
    for (int i = 0; i < 1023; i++) dest[i] = (float)c[i];
}

/* gps_sdr_signal_replica.cc:127-187.  NOTE (fork-specific): gps_l1_ca_code_gen_complex (:127-139) puts the
 * chips in the IMAGINARY part, complex(0, +-1); the sampled version inherits that.  Index law (:170-171):
 * aux = (ts*(i+1))/tc in fp32; index = (int32)(int64)(aux + 1) - 1 (AUX_CEIL, :23); last sample forced to chip 1022. */
void orc_gps_l1_ca_code_gen_complex_sampled(float* dest_ri, int n_dest, unsigned prn, int fs, unsigned chip_shift)
{
    const int code_freq = 1023000, L = 1023;
    const float tc = 1.0f / (float)code_freq;
    const int samples_per_code = (int)((double)fs / ((double)code_freq / (double)L));
    const float ts = 1.0f / (float)fs;
    int32_t c[1023];
    memset(c, 0, sizeof c);
    orc_gps_l1_ca_code_gen_int(c, (int)prn, chip_shift);
    for (int i = 0; i < samples_per_code && i < n_dest; i++)
        {
            float aux = (ts * ((float)i + 1)) / tc;
            int k = (int32_t)((int64_t)(aux + 1)) - 1;
            if (i == samples_per_code - 1) k = L - 1;
            dest_ri[2 * i] = 0.0f;
            dest_ri[2 * i + 1] = (float)c[k];
        }
}

/* src/algorithms/libs/gnss_signal_replica.cc:257-273 (AUX_CEIL2(x) = (uint32)(int64)(x + 1), :27) */
void orc_code_resampler_32f(const float* from, int from_len, float* dest, int dest_len, float fs_in, float fs_out)
{
    const float t_out = 1.0f / fs_out;
    for (int i = 0; i < dest_len - 1; i++)
        {
            float aux = (t_out * ((float)i + 1.0f)) * fs_in;
            uint32_t k = (uint32_t)((int64_t)(aux + 1)) - 1;
            dest[i] = from[k];
        }
    dest[dest_len - 1] = from[from_len - 1];
}
