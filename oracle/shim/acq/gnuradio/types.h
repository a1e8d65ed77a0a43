// Test-infrastructure stub (NOT product code)
#ifndef ORACLE_SHIM_ACQ_GR_TYPES_H
#define ORACLE_SHIM_ACQ_GR_TYPES_H
#include <gnuradio/gr_complex.h>
#include <vector>
typedef std::vector<int> gr_vector_int;
typedef std::vector<const void*> gr_vector_const_void_star;
typedef std::vector<void*> gr_vector_void_star;
#endif
