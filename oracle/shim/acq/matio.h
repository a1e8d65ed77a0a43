// Test-infrastructure stub (NOT product code): the .mat dump is never taken in the oracle runs (Acq_Conf::dump = false); the
// declarations only let dump_results() compile.
#ifndef ORACLE_SHIM_ACQ_MATIO_H
#define ORACLE_SHIM_ACQ_MATIO_H
#include <cstddef>
struct mat_t;
struct matvar_t;
enum mat_ft { MAT_FT_MAT73 = 0x0200 };
enum matio_classes { MAT_C_DOUBLE = 6, MAT_C_SINGLE = 7, MAT_C_INT32 = 12, MAT_C_UINT32 = 13, MAT_C_UINT64 = 15 };
enum matio_types { MAT_T_SINGLE = 7, MAT_T_DOUBLE = 9, MAT_T_INT32 = 5, MAT_T_UINT32 = 6, MAT_T_UINT64 = 13 };
enum matio_compression { MAT_COMPRESSION_NONE = 0, MAT_COMPRESSION_ZLIB = 1 };
inline mat_t* Mat_CreateVer(const char*, const char*, mat_ft) { return nullptr; }
inline matvar_t* Mat_VarCreate(const char*, matio_classes, matio_types, int, size_t*, void*, int) { return nullptr; }
inline int Mat_VarWrite(mat_t*, matvar_t*, matio_compression) { return 0; }
inline void Mat_VarFree(matvar_t*) {}
inline int Mat_Close(mat_t*) { return 0; }
#endif
