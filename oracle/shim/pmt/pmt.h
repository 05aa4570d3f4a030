/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): the few pmt calls of the acquisition / tracking blocks: port names, the integer
 * event values they publish (1 / 2 acquisition, 3 loss of lock) and the boost::any-carrying telemetry message. */
#ifndef GB_ORACLE_SHIM_PMT_H
#define GB_ORACLE_SHIM_PMT_H
#include <any>
#include <memory>
#include <string>
namespace boost
{
using std::any;
using std::any_cast;
using std::bad_any_cast;
}  // namespace boost
namespace pmt
{
struct pmt_base
{
    std::string sym;
    long lval = 0;
    double dval = 0.0;
    boost::any a;
};
typedef std::shared_ptr<pmt_base> pmt_t;
inline pmt_t mp(const std::string& s)
{
    pmt_t p(new pmt_base);
    p->sym = s;
    return p;
}
inline pmt_t from_long(long v)
{
    pmt_t p(new pmt_base);
    p->lval = v;
    return p;
}
inline long to_long(const pmt_t& p) { return p->lval; }
inline pmt_t from_double(double v)
{
    pmt_t p(new pmt_base);
    p->dval = v;
    return p;
}
inline double to_double(const pmt_t& p) { return p->dval; }
inline pmt_t make_any(const boost::any& a)
{
    pmt_t p(new pmt_base);
    p->a = a;
    return p;
}
inline boost::any& any_ref(const pmt_t& p) { return p->a; }
}  // namespace pmt
#endif
