/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): boost::shared_ptr as std::shared_ptr. */
#ifndef GB_ORACLE_SHIM_BOOST_SHARED_PTR_HPP
#define GB_ORACLE_SHIM_BOOST_SHARED_PTR_HPP
#include <memory>
namespace boost
{
using std::shared_ptr;
}
#endif
