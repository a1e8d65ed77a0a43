// Test-infrastructure stub (NOT product code): the reference's gnss_block_interface.h pulls in GNU Radio's basic_block; the one thing
// unpack_2bit_samples.h needs from it is the gnss_shared_ptr alias (gnss_block_interface.h: std::shared_ptr with GNU Radio >= 3.9).
#ifndef ORACLE_SHIM_GNSS_BLOCK_INTERFACE_H
#define ORACLE_SHIM_GNSS_BLOCK_INTERFACE_H
#include <memory>
template <typename T>
using gnss_shared_ptr = std::shared_ptr<T>;
#endif
