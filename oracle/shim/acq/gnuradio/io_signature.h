// Test-infrastructure stub (NOT product code)
#ifndef ORACLE_SHIM_ACQ_GR_IO_SIGNATURE_H
#define ORACLE_SHIM_ACQ_GR_IO_SIGNATURE_H
#include <cstddef>
#include <memory>
namespace gr
{
class io_signature
{
public:
    typedef std::shared_ptr<io_signature> sptr;
    static sptr make(int min_streams, int max_streams, size_t item_size) { return sptr(new io_signature{min_streams, max_streams, item_size}); }
    int min_streams, max_streams;
    size_t item_size;
};
}  // namespace gr
#endif
