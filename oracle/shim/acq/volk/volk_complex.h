// Test-infrastructure stub (NOT product code)
#ifndef ORACLE_SHIM_ACQ_VOLK_COMPLEX_H
#define ORACLE_SHIM_ACQ_VOLK_COMPLEX_H
#include <complex>
#include <cstdint>
typedef std::complex<float> lv_32fc_t;
typedef std::complex<int16_t> lv_16sc_t;
typedef std::complex<int8_t> lv_8sc_t;
template <typename T>
inline std::complex<T> lv_cmake(const T& r, const T& i) { return std::complex<T>(r, i); }
#endif
