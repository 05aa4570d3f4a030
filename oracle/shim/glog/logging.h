/* TEST INFRASTRUCTURE ONLY: null-stream LOG()/DLOG() so reference scalar libs compile without glog. */
#ifndef GB_ORACLE_SHIM_GLOG_H
#define GB_ORACLE_SHIM_GLOG_H
#include <ostream>
struct gb_null_stream { template <class T> gb_null_stream& operator<<(const T&) { return *this; }
                        gb_null_stream& operator<<(std::ostream& (*)(std::ostream&)) { return *this; } };
namespace google { class LogMessage {}; }  /* some reference files say `using google::LogMessage;` */
#define LOG(x) gb_null_stream()
#define DLOG(x) gb_null_stream()
#endif
