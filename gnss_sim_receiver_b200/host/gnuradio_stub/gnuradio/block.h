// MINIMAL STAND-IN for <gnuradio/block.h> + <gnuradio/io_signature.h>, used only when this tree is built without GNU Radio
// (it is not installed in this image).  It declares exactly the members of gr::block that b200_gr_blocks.{h,cc} use — same names,
// argument types and meaning as GNU Radio 3.8-3.10 — plus a tiny single-block driver (stub_feed) so that the wrappers can be
// exercised by the tests: one block, one input stream, the scheduler contract of forecast() / general_work() / consume_each().
// With -DGB200_WITH_GNURADIO the real headers are included instead and nothing here is compiled.
#ifndef GB200_GNURADIO_STUB_BLOCK_H
#define GB200_GNURADIO_STUB_BLOCK_H
#include <pmt/pmt.h>
#include <algorithm>
#include <cstdint>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>

typedef std::vector<int> gr_vector_int;
typedef std::vector<const void*> gr_vector_const_void_star;
typedef std::vector<void*> gr_vector_void_star;

namespace gr
{
class io_signature
{
public:
    typedef std::shared_ptr<io_signature> sptr;
    static sptr make(int min_streams, int max_streams, int sizeof_stream_item) { return sptr(new io_signature(min_streams, max_streams, sizeof_stream_item)); }
    int min_streams() const { return d_min; }
    int max_streams() const { return d_max; }
    int sizeof_stream_item(int) const { return d_size; }

private:
    io_signature(int a, int b, int c) : d_min(a), d_max(b), d_size(c) {}
    int d_min, d_max, d_size;
};

class block
{
public:
    block(const std::string& name, io_signature::sptr in, io_signature::sptr out) : d_name(name), d_in(std::move(in)), d_out(std::move(out)) {}
    virtual ~block() = default;
    virtual void forecast(int noutput_items, gr_vector_int& ninput_items_required)
    {
        for (auto& r : ninput_items_required) r = noutput_items;
    }
    virtual int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) = 0;
    virtual bool start() { return true; }
    virtual bool stop() { return true; }
    std::string name() const { return d_name; }
    io_signature::sptr input_signature() const { return d_in; }
    io_signature::sptr output_signature() const { return d_out; }
    uint64_t nitems_read(unsigned) const { return d_nitems_read; }
    void consume_each(int how_many_items) { d_consumed = how_many_items; }
    void set_relative_rate(uint64_t interpolation, uint64_t decimation) { d_interp = interpolation, d_decim = decimation; }
    void set_max_noutput_items(int m) { d_max_noutput = m; }
    void message_port_register_out(pmt::pmt_t port) { d_out_ports[pmt::symbol_to_string(port)]; }
    void message_port_register_in(pmt::pmt_t port) { d_handlers[pmt::symbol_to_string(port)]; }
    void message_port_pub(pmt::pmt_t port, pmt::pmt_t msg) { d_out_ports[pmt::symbol_to_string(port)].push_back(msg); }
    template <typename F>
    void set_msg_handler(pmt::pmt_t port, F handler) { d_handlers[pmt::symbol_to_string(port)] = handler; }

    // ---- test driver (stub only): what the scheduler does for ONE block with one input stream
    // feeds n_items of `items`; calls general_work while forecast() is satisfied; outputs appended to out_bytes (item size from
    // the output signature).  Returns the number of items consumed.
    int64_t stub_feed(const void* items, int64_t n_items, std::vector<char>* out_bytes)
    {
        const int isz = d_in->sizeof_stream_item(0), osz = d_out->max_streams() > 0 ? d_out->sizeof_stream_item(0) : 0;
        int64_t done = 0;
        std::vector<char> obuf(static_cast<size_t>(std::max(1, osz)) * 4);
        for (int guard = 0; guard < (1 << 24); guard++)
            {
                gr_vector_int required(1, 0);
                forecast(1, required);
                const int64_t avail = n_items - done;
                if (avail < std::max(1, required[0])) break;
                gr_vector_int ninput(1, static_cast<int>(std::min<int64_t>(avail, 1 << 30)));
                gr_vector_const_void_star in(1, static_cast<const char*>(items) + done * isz);
                gr_vector_void_star out(1, obuf.data());
                d_consumed = 0;
                const int produced = general_work(std::min(1, d_max_noutput), ninput, in, out);
                if (produced > 0 && out_bytes && osz) out_bytes->insert(out_bytes->end(), obuf.begin(), obuf.begin() + static_cast<size_t>(produced) * osz);
                done += d_consumed;
                d_nitems_read += static_cast<uint64_t>(d_consumed);
                if (produced < 0 || (d_consumed == 0 && produced == 0 && ++d_idle > 8)) break;
                if (d_consumed || produced) d_idle = 0;
            }
        return done;
    }
    void stub_post(const std::string& port, pmt::pmt_t msg)
    {
        auto it = d_handlers.find(port);
        if (it != d_handlers.end() && it->second) it->second(msg);
    }
    std::vector<pmt::pmt_t>& stub_published(const std::string& port) { return d_out_ports[port]; }
    void stub_set_nitems_read(uint64_t n) { d_nitems_read = n; }

private:
    std::string d_name;
    io_signature::sptr d_in, d_out;
    uint64_t d_nitems_read{0}, d_interp{1}, d_decim{1};
    int d_consumed{0}, d_max_noutput{1 << 20}, d_idle{0};
    std::map<std::string, std::vector<pmt::pmt_t>> d_out_ports;
    std::map<std::string, std::function<void(pmt::pmt_t)>> d_handlers;
};
}  // namespace gr
#endif
