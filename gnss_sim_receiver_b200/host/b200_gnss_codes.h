// GPS L1 C/A code tables (see b200_gnss_codes.cc); same names and arguments as
// src/algorithms/libs/gps_sdr_signal_replica.h of the reference, raw pointers instead of own::span.
#ifndef GB200_GNSS_CODES_H
#define GB200_GNSS_CODES_H
#include <complex>
#include <cstdint>
namespace gb200h
{
// 1023 entries of +-1; throws std::invalid_argument for a PRN outside 1..37
void gps_l1_ca_code_gen_float(float* dest, uint32_t prn, uint32_t chip_shift);
// floor(sampling_freq / 1000) samples of complex(0, +-1) (at most n_dest are written)
void gps_l1_ca_code_gen_complex_sampled(std::complex<float>* dest, int n_dest, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift);
}  // namespace gb200h
#endif
