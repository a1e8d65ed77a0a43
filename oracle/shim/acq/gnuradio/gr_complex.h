// Test-infrastructure stub (NOT product code)
#ifndef ORACLE_SHIM_ACQ_GR_COMPLEX_H
#define ORACLE_SHIM_ACQ_GR_COMPLEX_H
#include <complex>
typedef std::complex<float> gr_complex;
#endif
