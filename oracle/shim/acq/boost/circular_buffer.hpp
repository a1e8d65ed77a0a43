// Test-infrastructure stub (NOT product code): the subset of boost::circular_buffer the tracking block uses (set_capacity, clear,
// push_back dropping the oldest element when full, size, full, operator[], iteration from oldest to newest).
#ifndef ORACLE_SHIM_BOOST_CIRCULAR_BUFFER_H
#define ORACLE_SHIM_BOOST_CIRCULAR_BUFFER_H
#include <cstddef>
#include <deque>
namespace boost
{
template <typename T>
class circular_buffer
{
public:
    typedef typename std::deque<T>::iterator iterator;
    typedef typename std::deque<T>::const_iterator const_iterator;
    circular_buffer() = default;
    explicit circular_buffer(size_t cap) : cap_(cap) {}
    void set_capacity(size_t cap)
    {
        cap_ = cap;
        while (d_.size() > cap_) d_.pop_back();  // boost keeps the first `cap` elements
    }
    size_t capacity() const { return cap_; }
    void clear() { d_.clear(); }
    void push_back(const T& v)
    {
        if (cap_ == 0) return;
        if (d_.size() == cap_) d_.pop_front();
        d_.push_back(v);
    }
    size_t size() const { return d_.size(); }
    bool full() const { return d_.size() == cap_; }
    bool empty() const { return d_.empty(); }
    T& operator[](size_t i) { return d_[i]; }
    const T& operator[](size_t i) const { return d_[i]; }
    T& front() { return d_.front(); }
    T& back() { return d_.back(); }
    const T& front() const { return d_.front(); }
    const T& back() const { return d_.back(); }
    iterator begin() { return d_.begin(); }
    iterator end() { return d_.end(); }
    const_iterator begin() const { return d_.begin(); }
    const_iterator end() const { return d_.end(); }

private:
    size_t cap_{0};
    std::deque<T> d_;
};
}  // namespace boost
#endif
