#include "b200_mat_writer.h"
#include <cstring>

namespace gb200h
{
namespace
{
// MAT-file data types (miXXX) and the element size of each array class
constexpr uint32_t miINT8 = 1, miINT32 = 5, miUINT32 = 6, miSINGLE = 7, miUINT64 = 13, miMATRIX = 14;

uint32_t mi_of(MatFile5::Class c)
{
    switch (c)
        {
        case MatFile5::SINGLE: return miSINGLE;
        case MatFile5::INT32: return miINT32;
        case MatFile5::UINT32: return miUINT32;
        default: return miUINT64;
        }
}
uint32_t size_of(MatFile5::Class c) { return c == MatFile5::UINT64 ? 8U : 4U; }
uint32_t pad8(uint32_t n) { return (n + 7U) & ~7U; }
}  // namespace

MatFile5::MatFile5(const std::string& filename)
{
    f_ = std::fopen(filename.c_str(), "wb");
    if (!f_) return;
    char header[128];
    std::memset(header, ' ', sizeof header);
    const char* text = "MATLAB 5.0 MAT-file, Platform: GB200, Created by: gnss_sim_receiver_b200 acquisition dump";
    std::memcpy(header, text, std::strlen(text));
    std::memset(header + 116, 0, 8);  // subsystem data offset: none
    const uint16_t version = 0x0100, endian = 0x4D49;  // 'IM' read back on a little-endian machine = no byte swapping
    std::memcpy(header + 124, &version, 2);
    std::memcpy(header + 126, &endian, 2);
    std::fwrite(header, 1, sizeof header, f_);
}

MatFile5::~MatFile5()
{
    if (f_) std::fclose(f_);
}

// one data element: tag (type, byte count), data, zero padding to 8 bytes
void MatFile5::element(uint32_t type, const void* data, uint32_t bytes)
{
    const uint32_t tag[2] = {type, bytes};
    std::fwrite(tag, 4, 2, f_);
    if (bytes) std::fwrite(data, 1, bytes, f_);
    static const char zeros[8] = {0};
    if (pad8(bytes) != bytes) std::fwrite(zeros, 1, pad8(bytes) - bytes, f_);
}

void MatFile5::write(const std::string& name, Class cls, uint32_t rows, uint32_t cols, const void* data)
{
    if (!f_) return;
    const uint32_t name_len = static_cast<uint32_t>(name.size());
    const uint64_t data_bytes64 = static_cast<uint64_t>(rows) * cols * size_of(cls);
    if (data_bytes64 > 0xF0000000ULL) return;  // a level-5 element holds < 4 GiB
    const uint32_t data_bytes = static_cast<uint32_t>(data_bytes64);
    // miMATRIX = array flags (8) + dimensions (8) + name + real part, each with its 8-byte tag
    const uint32_t total = (8 + 8) + (8 + 8) + (8 + pad8(name_len)) + (8 + pad8(data_bytes));
    const uint32_t tag[2] = {miMATRIX, total};
    std::fwrite(tag, 4, 2, f_);
    const uint32_t flags[2] = {static_cast<uint32_t>(cls), 0U};
    element(miUINT32, flags, 8);
    const int32_t dims[2] = {static_cast<int32_t>(rows), static_cast<int32_t>(cols)};
    element(miINT32, dims, 8);
    element(miINT8, name.data(), name_len);
    element(mi_of(cls), data, data_bytes);
}
}  // namespace gb200h
