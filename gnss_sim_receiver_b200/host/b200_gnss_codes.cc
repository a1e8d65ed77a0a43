// GPS L1 C/A code tables for the standalone host layer (inside GNSS-SDR the adapters call the reference's own generators,
// see INTEGRATION.md).  Behaviour follows src/algorithms/libs/gps_sdr_signal_replica.cc:25-187 of the reference:
//   gps_l1_ca_code_gen_float           one +-1 entry per chip, G1 xor delayed G2 (IS-GPS-200 table 3-Ia delays)
//   gps_l1_ca_code_gen_complex_sampled the acquisition replica: chips in the IMAGINARY part (0, +-1), sampled with the
//                                      Borre index law  idx = int(fl(ts*(i+1))/tc + 1) - 1  in fp32, last sample forced to
//                                      the last chip (:170-186)
// tests/test_host_cpu.py checks both bit-exactly against the reference's generators (built by the test suite).
#include "b200_gnss_codes.h"
#include <array>
#include <stdexcept>

namespace gb200h
{
namespace
{
constexpr int L1CA_LEN = 1023;
// G2 delay in chips for PRN 1..37 (IS-GPS-200, table 3-Ia); SBAS PRNs 120..158 are not tabulated here
constexpr int G2_DELAY[37] = {5, 6, 7, 8, 17, 18, 139, 140, 141, 251, 252, 254, 255, 256, 257, 258, 469, 470, 471, 472, 473, 474, 509, 512, 513, 514, 515, 516,
    859, 860, 861, 862, 863, 950, 947, 948, 950};

void l1ca_chips(std::array<int, L1CA_LEN>& chips, uint32_t prn, uint32_t chip_shift)
{
    if (prn < 1 || prn > 37) throw std::invalid_argument("GPS L1 C/A PRN outside 1..37");
    std::array<int, L1CA_LEN> g1{}, g2{};
    // 10-stage LFSRs, all ones at start; G1 = 1 + x^3 + x^10, G2 = 1 + x^2 + x^3 + x^6 + x^8 + x^9 + x^10
    int r1[10], r2[10];
    for (int i = 0; i < 10; i++) r1[i] = r2[i] = 1;
    for (int n = 0; n < L1CA_LEN; n++)
        {
            g1[n] = r1[9];
            g2[n] = r2[9];
            const int f1 = r1[2] ^ r1[9];
            const int f2 = r2[1] ^ r2[2] ^ r2[5] ^ r2[7] ^ r2[8] ^ r2[9];
            for (int i = 9; i > 0; i--)
                {
                    r1[i] = r1[i - 1];
                    r2[i] = r2[i - 1];
                }
            r1[0] = f1;
            r2[0] = f2;
        }
    const int delay = (L1CA_LEN - G2_DELAY[prn - 1] + static_cast<int>(chip_shift % L1CA_LEN)) % L1CA_LEN;
    for (int n = 0; n < L1CA_LEN; n++)
        {
            const int a = g1[(n + chip_shift) % L1CA_LEN];
            const int b = g2[(n + delay) % L1CA_LEN];
            chips[n] = (a ^ b) ? 1 : -1;  // logic 1 -> +1, as the reference maps (G1 == G2 -> -1)
        }
}
}  // namespace

void gps_l1_ca_code_gen_float(float* dest, uint32_t prn, uint32_t chip_shift)
{
    std::array<int, L1CA_LEN> c{};
    l1ca_chips(c, prn, chip_shift);
    for (int i = 0; i < L1CA_LEN; i++) dest[i] = static_cast<float>(c[i]);
}

void gps_l1_ca_code_gen_complex_sampled(std::complex<float>* dest, int n_dest, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift)
{
    std::array<int, L1CA_LEN> c{};
    l1ca_chips(c, prn, chip_shift);
    constexpr float tc = 1.0F / 1023000.0F;
    const auto samples_per_code = static_cast<int32_t>(static_cast<double>(sampling_freq) / (1023000.0 / 1023.0));
    const float ts = 1.0F / static_cast<float>(sampling_freq);
    for (int32_t i = 0; i < samples_per_code && i < n_dest; i++)
        {
            const float aux = (ts * (static_cast<float>(i) + 1)) / tc;
            int idx = static_cast<int32_t>(static_cast<int64_t>(aux + 1)) - 1;
            if (i == samples_per_code - 1) idx = L1CA_LEN - 1;
            dest[i] = std::complex<float>(0.0F, static_cast<float>(c[idx]));
        }
}
}  // namespace gb200h
