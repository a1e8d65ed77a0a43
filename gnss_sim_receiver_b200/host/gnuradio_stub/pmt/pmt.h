// MINIMAL STAND-IN for <pmt/pmt.h> (see gnuradio_stub/gnuradio/block.h): the handful of PMT calls the B200 block wrappers make —
// mp / intern, from_long / to_long, is_integer, symbol_to_string — with GNU Radio's names and meaning.
#ifndef GB200_GNURADIO_STUB_PMT_H
#define GB200_GNURADIO_STUB_PMT_H
#include <memory>
#include <string>

namespace pmt
{
struct pmt_base
{
    bool is_int{false};
    long value{0};
    std::string symbol;
};
typedef std::shared_ptr<pmt_base> pmt_t;
inline pmt_t mp(const std::string& s)
{
    auto p = std::make_shared<pmt_base>();
    p->symbol = s;
    return p;
}
inline pmt_t intern(const std::string& s) { return mp(s); }
inline pmt_t from_long(long v)
{
    auto p = std::make_shared<pmt_base>();
    p->is_int = true;
    p->value = v;
    return p;
}
inline bool is_integer(const pmt_t& p) { return p && p->is_int; }
inline long to_long(const pmt_t& p) { return p->value; }
inline std::string symbol_to_string(const pmt_t& p) { return p->symbol; }
}  // namespace pmt
#endif
