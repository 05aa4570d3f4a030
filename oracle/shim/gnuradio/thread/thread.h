/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): gr::thread::{mutex, scoped_lock, thread} on the standard library.
 * boost::thread detaches when it goes out of scope; so does this one. */
#ifndef GB_ORACLE_SHIM_GR_THREAD_H
#define GB_ORACLE_SHIM_GR_THREAD_H
#include <mutex>
#include <thread>
#include <utility>
namespace gr { namespace thread {
typedef std::recursive_mutex mutex;
typedef std::unique_lock<std::recursive_mutex> scoped_lock;
class thread
{
    std::thread t_;

public:
    template <class F, class... A>
    explicit thread(F&& f, A&&... a) : t_(std::forward<F>(f), std::forward<A>(a)...) {}
    ~thread() { if (t_.joinable()) t_.detach(); }
};
} }
#endif
