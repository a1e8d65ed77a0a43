// Test-infrastructure stub (NOT product code): gflags is not installed.  The command-line flags dll_pll_conf.cc reads, as plain
// variables holding the defaults of the reference's gnss_sdr_flags.cc (:48-57): no flag is ever given in the oracle runs.
#ifndef ORACLE_SHIM_GNSS_SDR_FLAGS_H
#define ORACLE_SHIM_GNSS_SDR_FLAGS_H
#include <cstdint>
inline int32_t FLAGS_cn0_samples = 20;
inline int32_t FLAGS_cn0_min = 25;
inline int32_t FLAGS_max_lock_fail = 50;
inline int32_t FLAGS_max_carrier_lock_fail = 5000;
inline double FLAGS_carrier_lock_th = 0.7;
inline double FLAGS_dll_bw_hz = 0.0;
inline double FLAGS_pll_bw_hz = 0.0;
inline int32_t FLAGS_doppler_max = 0;
inline int32_t FLAGS_doppler_step = 0;
#endif
