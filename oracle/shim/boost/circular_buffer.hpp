/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): the subset of boost::circular_buffer the reference's tracking blocks use, so that
 * dll_pll_veml_tracking.cc compiles where Boost is not installed.  Same semantics for that subset: push_back on a full buffer
 * drops the oldest element; operator[] indexes from the oldest. */
#ifndef GB_ORACLE_SHIM_BOOST_CIRCULAR_BUFFER_HPP
#define GB_ORACLE_SHIM_BOOST_CIRCULAR_BUFFER_HPP
#include <cstddef>
#include <deque>
namespace boost
{
template <class T>
class circular_buffer
{
    std::deque<T> d_;
    std::size_t cap_ = 0;

public:
    using const_iterator = typename std::deque<T>::const_iterator;
    circular_buffer() = default;
    explicit circular_buffer(std::size_t c) : cap_(c) {}
    void set_capacity(std::size_t c)
    {
        cap_ = c;
        while (d_.size() > cap_) d_.pop_back();
    }
    void resize(std::size_t n)  // boost: capacity grows to n when needed, size becomes n (value-initialised fill)
    {
        if (n > cap_) cap_ = n;
        d_.resize(n);
    }
    std::size_t capacity() const { return cap_; }
    std::size_t size() const { return d_.size(); }
    bool full() const { return d_.size() == cap_; }
    bool empty() const { return d_.empty(); }
    void clear() { d_.clear(); }
    void push_back(const T& v)
    {
        if (cap_ == 0) return;
        if (d_.size() == cap_) d_.pop_front();
        d_.push_back(v);
    }
    T& operator[](std::size_t i) { return d_[i]; }
    const T& operator[](std::size_t i) const { return d_[i]; }
    const_iterator begin() const { return d_.begin(); }
    const_iterator end() const { return d_.end(); }
};
}  // namespace boost
#endif
