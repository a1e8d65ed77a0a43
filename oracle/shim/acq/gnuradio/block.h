// Test-infrastructure stub (NOT product code): the members of gr::block that pcps_acquisition uses (constructor, the "events" message
// port, consume_each); published messages and consumed items are kept for the test door to read.
#ifndef ORACLE_SHIM_ACQ_GR_BLOCK_H
#define ORACLE_SHIM_ACQ_GR_BLOCK_H
#include <gnuradio/io_signature.h>
#include <gnuradio/thread/thread.h>
#include <gnuradio/types.h>
#include <pmt/pmt.h>
#include <cassert>  // (pulled in transitively by the real header; dll_pll_veml_tracking.cc relies on it)
#include <cstdint>
#include <functional>
#include <string>
#include <thread>
#include <vector>
namespace gr
{
struct tag_t
{
    uint64_t offset{0};
    pmt::pmt_t key, value, srcid;
};
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
    void message_port_register_in(pmt::pmt_t) {}
    template <typename F>
    void set_msg_handler(pmt::pmt_t, F f) { shim_msg_handler = f; }
    void set_relative_rate(uint64_t, uint64_t) {}
    void set_relative_rate(double) {}
    void set_max_noutput_items(int) {}
    enum tag_propagation_policy_t { TPP_DONT = 0, TPP_ALL_TO_ALL = 1, TPP_ONE_TO_ONE = 2 };
    void set_tag_propagation_policy(tag_propagation_policy_t) {}
    // the scheduler's item counters.  As in GNU Radio, consume_each() moves the read pointer AT ONCE (block_detail::consume_each ->
    // buffer_reader::update_read_pointer: d_abs_read_offset += nitems), so nitems_read(0) read after it already counts the items
    uint64_t nitems_read(unsigned) const { return shim_nitems_read; }
    uint64_t nitems_written(unsigned) const { return shim_nitems_written; }
    void get_tags_in_range(std::vector<tag_t>& v, unsigned, uint64_t, uint64_t) { v.clear(); }
    void add_item_tag(unsigned, uint64_t, pmt::pmt_t, pmt::pmt_t) {}
    virtual void forecast(int, gr_vector_int&) {}
    std::function<void(const pmt::pmt_t&)> shim_msg_handler;
    uint64_t shim_nitems_read{0}, shim_nitems_written{0};
    void message_port_pub(pmt::pmt_t, pmt::pmt_t msg) { shim_events.push_back(pmt::to_long(msg)); }
    void consume_each(int n)
    {
        shim_consumed += n;
        shim_pending_consume += n;  // (what the driver advances its own stream position by when general_work returns)
        shim_nitems_read += static_cast<uint64_t>(n);
    }
    long long shim_pending_consume{0};
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
