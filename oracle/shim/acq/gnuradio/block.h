// Test-infrastructure stub (NOT product code): the members of gr::block that pcps_acquisition uses (constructor, the "events" message
// port, consume_each); published messages and consumed items are kept for the test door to read.
#ifndef ORACLE_SHIM_ACQ_GR_BLOCK_H
#define ORACLE_SHIM_ACQ_GR_BLOCK_H
#include <gnuradio/io_signature.h>
#include <gnuradio/thread/thread.h>
#include <gnuradio/types.h>
#include <pmt/pmt.h>
#include <string>
#include <thread>
#include <vector>
namespace gr
{
namespace thread
{
typedef std::thread thread;
}
class block
{
public:
    block(const std::string& name, io_signature::sptr in, io_signature::sptr out) : d_name(name), d_in(in), d_out(out) {}
    virtual ~block() = default;
    void message_port_register_out(pmt::pmt_t) {}
    void message_port_pub(pmt::pmt_t, pmt::pmt_t msg) { shim_events.push_back(pmt::to_long(msg)); }
    void consume_each(int n) { shim_consumed += n; }
    virtual bool start() { return true; }
    virtual bool stop() { return true; }
    virtual int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) = 0;
    std::vector<long> shim_events;
    long long shim_consumed{0};

protected:
    gr::thread::mutex d_setlock;

private:
    std::string d_name;
    io_signature::sptr d_in, d_out;
};
}  // namespace gr
#endif
