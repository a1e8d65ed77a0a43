// GPS L1 C/A code tables (see b200_gnss_codes.cc); same names and arguments as
// src/algorithms/libs/gps_sdr_signal_replica.h of the reference, raw pointers instead of own::span.
#ifndef GB200_GNSS_CODES_H
#define GB200_GNSS_CODES_H
#include <complex>
#include <cstddef>
#include <cstdint>
#include <string>
namespace gb200h
{
// 1023 entries of +-1; throws std::invalid_argument for a PRN outside 1..37
void gps_l1_ca_code_gen_float(float* dest, uint32_t prn, uint32_t chip_shift);
// floor(sampling_freq / 1000) samples of complex(0, +-1) (at most n_dest are written)
void gps_l1_ca_code_gen_complex_sampled(std::complex<float>* dest, int n_dest, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift);

// ---- Galileo E1 / GPS L5 / Galileo E5a (b200_gnss_codes2.cc); names and arguments as galileo_e1_signal_replica.h,
// gps_l5_signal_replica.h, galileo_e5_signal_replica.h, gnss_signal_replica.h of the reference.  PRNs outside the tabulated
// range leave `dest` untouched (E1 / E5a, like the reference's early return) or produce the all-zero-register code (L5).
// galileo_e1_signal_replica.cc:98-110: 8184 entries, chip c -> (c, -c); signal_id "1B" or "1C"
void galileo_e1_code_gen_sinboc11_float(float* dest, const char* signal_id, uint32_t prn);
// :141-196: sinBOC(1,1) at 2 samples per chip or CBOC at 12, resampled to sampling_freq, optionally x the 25-chip E1-C
// secondary code (samples per code x 25), rotated by chip_shift; returns the number of samples written
int galileo_e1_code_gen_float_sampled(float* dest, const char* signal_id, bool cboc, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift, bool secondary_flag);
int galileo_e1_code_gen_complex_sampled(std::complex<float>* dest, const char* signal_id, bool cboc, uint32_t prn, int32_t sampling_freq, uint32_t chip_shift, bool secondary_flag);
// gps_l5_signal_replica.cc:175-262: 10230 entries of +-1 (XA xor advanced XB, IS-GPS-705 3.2.1.1), PRN 1..50
void gps_l5i_code_gen_float(float* dest, uint32_t prn);
void gps_l5q_code_gen_float(float* dest, uint32_t prn);
// floor(sampling_freq / 1000) samples: L5I complex(+-1, 0), L5Q complex(0, +-1)
int gps_l5i_code_gen_complex_sampled(std::complex<float>* dest, uint32_t prn, int32_t sampling_freq);
int gps_l5q_code_gen_complex_sampled(std::complex<float>* dest, uint32_t prn, int32_t sampling_freq);
// galileo_e5_signal_replica.cc:31-93: 10230 entries; "5I" -> (I, 0), "5Q" -> (Q, 0), "5X" -> (I, Q)
void galileo_e5_a_code_gen_complex_primary(std::complex<float>* dest, int32_t prn, const char* signal_id);
// :96-121: resampled to sampling_freq and rotated by chip_shift; returns the number of samples written
int galileo_e5_a_code_gen_complex_sampled(std::complex<float>* dest, uint32_t prn, const char* signal_id, int32_t sampling_freq, uint32_t chip_shift);
// GALILEO_E5A_Q_SECONDARY_CODE[prn - 1] (Galileo_E5a.h:3585-3633): the 100-chip E5a-Q secondary code as a '0'/'1' string; the
// reference's table initialises PRN 1..47 only (rows 48..50 are empty strings there): all '0' is returned for those
std::string galileo_e5a_q_secondary_code(uint32_t prn);
// gnss_signal_replica.cc:257-291: dest[i] = from[int(fl(t_out (i + 1)) fs_in + 1) - 1] in fp32, last entry forced to the last one
void resampler(const float* from, size_t n_from, float* dest, size_t n_dest, float fs_in, float fs_out);
void resampler(const std::complex<float>* from, size_t n_from, std::complex<float>* dest, size_t n_dest, float fs_in, float fs_out);
}  // namespace gb200h
#endif
