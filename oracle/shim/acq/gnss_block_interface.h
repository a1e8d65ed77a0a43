// Test-infrastructure stub (NOT product code): only the gnss_shared_ptr alias of the reference's gnss_block_interface.h
#ifndef ORACLE_SHIM_ACQ_GNSS_BLOCK_INTERFACE_H
#define ORACLE_SHIM_ACQ_GNSS_BLOCK_INTERFACE_H
#include <memory>
template <typename T>
using gnss_shared_ptr = std::shared_ptr<T>;
#endif
