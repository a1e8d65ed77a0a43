// Minimal MAT-file (level 5, uncompressed) writer for the acquisition dump.
//
// The reference writes its acquisition dump with matio as MAT v7.3 (pcps_acquisition.cc:393-493, Mat_CreateVer(...,
// MAT_FT_MAT73), i.e. HDF5); matio / HDF5 are not part of this tree, so the same variables (names, classes, shapes,
// column-major data) are written as a level-5 file, which MATLAB's `load`, Octave and scipy.io.loadmat — and therefore the
// reference's readers under src/utils/matlab and src/utils/python-skyplot — open the same way.
#ifndef GB200_MAT_WRITER_H
#define GB200_MAT_WRITER_H

#include <cstdint>
#include <cstdio>
#include <string>

namespace gb200h
{
class MatFile5
{
public:
    enum Class : uint8_t
    {
        SINGLE = 7,   // mxSINGLE_CLASS
        INT32 = 12,   // mxINT32_CLASS
        UINT32 = 13,  // mxUINT32_CLASS
        UINT64 = 15   // mxUINT64_CLASS
    };
    explicit MatFile5(const std::string& filename);
    ~MatFile5();
    MatFile5(const MatFile5&) = delete;
    MatFile5& operator=(const MatFile5&) = delete;
    bool ok() const { return f_ != nullptr; }
    // rows x cols, column-major `data` of the given class
    void write(const std::string& name, Class cls, uint32_t rows, uint32_t cols, const void* data);
    template <typename T>
    void scalar(const std::string& name, Class cls, T v)
    {
        write(name, cls, 1, 1, &v);
    }

private:
    void element(uint32_t type, const void* data, uint32_t bytes);
    FILE* f_{nullptr};
};
}  // namespace gb200h
#endif
