// Test-infrastructure stub (NOT product code): GNU Radio is not installed here.  Just enough of gr::sync_interpolator for the
// reference's unpack_2bit_samples (signal_source/gnuradio_blocks/unpack_2bit_samples.{h,cc}) to compile unmodified: its constructor
// forwards (name, input signature, output signature, interpolation) to the base, and work() only uses the vector typedefs.
#ifndef ORACLE_SHIM_GR_SYNC_INTERPOLATOR_H
#define ORACLE_SHIM_GR_SYNC_INTERPOLATOR_H
#include <gnuradio/io_signature.h>
#include <string>
#include <vector>
typedef std::vector<const void*> gr_vector_const_void_star;
typedef std::vector<void*> gr_vector_void_star;
namespace gr
{
class sync_interpolator
{
public:
    sync_interpolator(const std::string& name, io_signature::sptr in, io_signature::sptr out, unsigned interpolation) : d_name(name), d_in(in), d_out(out), d_interpolation(interpolation) {}
    virtual ~sync_interpolator() = default;
    unsigned interpolation() const { return d_interpolation; }

private:
    std::string d_name;
    io_signature::sptr d_in, d_out;
    unsigned d_interpolation;
};
}  // namespace gr
#endif
