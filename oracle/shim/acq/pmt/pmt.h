// Test-infrastructure stub (NOT product code): the block only makes port names and wraps event numbers
#ifndef ORACLE_SHIM_ACQ_PMT_H
#define ORACLE_SHIM_ACQ_PMT_H
#include <any>
#include <memory>
#include <string>
namespace pmt
{
struct pmt_base
{
    std::string s;
    long v{0};
    std::any a;
};
typedef std::shared_ptr<pmt_base> pmt_t;
inline pmt_t mp(const std::string& s)
{
    auto p = std::make_shared<pmt_base>();
    p->s = s;
    return p;
}
inline pmt_t from_long(long v)
{
    auto p = std::make_shared<pmt_base>();
    p->v = v;
    return p;
}
inline long to_long(const pmt_t& p) { return p->v; }
inline pmt_t make_any(const std::any& a)
{
    auto p = std::make_shared<pmt_base>();
    p->a = a;
    return p;
}
inline const std::any& any_ref(const pmt_t& p) { return p->a; }
}  // namespace pmt
#endif
