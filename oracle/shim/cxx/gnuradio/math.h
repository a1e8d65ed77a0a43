// Test-infrastructure stub (NOT product code): gr::fast_atan2f is GNU Radio's
// table-driven atan2 approximation (absent here); the stub uses libm atan2f.
#ifndef ORACLE_SHIM_GR_MATH_H
#define ORACLE_SHIM_GR_MATH_H
#include <cmath>
namespace gr { inline float fast_atan2f(float y, float x) { return std::atan2(y, x); } }
#endif
