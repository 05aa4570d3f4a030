#ifndef GB_ORACLE_SHIM_GR_MSG_QUEUE_H
#define GB_ORACLE_SHIM_GR_MSG_QUEUE_H
#include <memory>
namespace gr { struct msg_queue { typedef std::shared_ptr<msg_queue> sptr; }; }
#endif
