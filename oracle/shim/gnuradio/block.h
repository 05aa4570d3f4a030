/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): a stand-in for gr::block that lets the reference's GNU Radio blocks
 * (pcps_acquisition.cc, dll_pll_veml_tracking.cc, compiled unmodified from /root/reference) be driven by a test harness instead
 * of the GNU Radio scheduler.  It implements no signal processing: consume_each() and message_port_pub() only record what the
 * block did so the harness (oracle/ref_blocks.cc) can advance the input pointer and read the events. */
#ifndef GB_ORACLE_SHIM_GR_BLOCK_H
#define GB_ORACLE_SHIM_GR_BLOCK_H
#include <functional>
#include <boost/shared_ptr.hpp>
#include <deque>
#include <fstream>
#include <gnuradio/gr_complex.h>
#include <iostream>
#include <gnuradio/io_signature.h>
#include <gnuradio/thread/thread.h>
#include <gnuradio/types.h>
#include <map>
#include <pmt/pmt.h>
#include <string>
#include <vector>
namespace boost
{
template <class F, class... A>
auto bind(F&& f, A&&... a) -> decltype(std::bind(std::forward<F>(f), std::forward<A>(a)...))
{
    return std::bind(std::forward<F>(f), std::forward<A>(a)...);
}
}  // namespace boost
using std::placeholders::_1;
namespace gr
{
class basic_block
{
public:
    virtual ~basic_block() {}
};
typedef std::shared_ptr<basic_block> basic_block_sptr;

class block : public basic_block
{
public:
    block(const std::string& name, io_signature::sptr in, io_signature::sptr out) : d_name(name), d_in(in), d_out(out) {}
    virtual ~block() {}
    virtual int general_work(int noutput_items, gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items) = 0;
    virtual void forecast(int, gr_vector_int&) {}
    virtual bool start() { return true; }
    void message_port_register_out(pmt::pmt_t) {}
    void message_port_register_in(pmt::pmt_t) {}
    template <class F>
    void set_msg_handler(pmt::pmt_t port, F f)
    {
        d_handlers[port->sym] = f;
    }
    void message_port_pub(pmt::pmt_t port, pmt::pmt_t msg) { gb_events.push_back(pmt::to_long(msg)); (void)port; }
    void set_relative_rate(double) {}
    void consume_each(int n) { gb_consumed += n; }
    long unique_id() const { return 0; }
    // harness side
    long long gb_consumed = 0;
    std::vector<long> gb_events;
    void gb_post(const std::string& port, pmt::pmt_t msg)
    {
        auto it = d_handlers.find(port);
        if (it != d_handlers.end()) it->second(msg);
    }

protected:
    gr::thread::mutex d_setlock;
    std::string d_name;
    io_signature::sptr d_in, d_out;
    std::map<std::string, std::function<void(pmt::pmt_t)>> d_handlers;
};
}  // namespace gr
#endif
