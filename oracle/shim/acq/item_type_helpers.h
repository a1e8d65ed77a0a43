// Test-infrastructure stub (NOT product code): acq_conf.cc only asks whether an item type name is valid and how large an item is
// (item_type_helpers.cc also builds VOLK-based converters, which nothing here uses).
#ifndef ORACLE_SHIM_ACQ_ITEM_TYPE_HELPERS_H
#define ORACLE_SHIM_ACQ_ITEM_TYPE_HELPERS_H
#include <complex>
#include <cstddef>
#include <cstdint>
#include <string>
inline size_t item_type_size(const std::string& t)
{
    if (t == "byte") return sizeof(int8_t);
    if (t == "ibyte") return sizeof(int8_t);
    if (t == "cbyte") return 2 * sizeof(int8_t);
    if (t == "short") return sizeof(int16_t);
    if (t == "ishort") return sizeof(int16_t);
    if (t == "cshort") return 2 * sizeof(int16_t);
    if (t == "float") return sizeof(float);
    if (t == "gr_complex") return sizeof(std::complex<float>);
    return 0;
}
inline bool item_type_valid(const std::string& t) { return item_type_size(t) != 0; }
inline bool item_type_is_complex(const std::string& t) { return t == "ibyte" || t == "cbyte" || t == "ishort" || t == "cshort" || t == "gr_complex"; }
#endif
