/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): shadows src/algorithms/channel/libs/channel_fsm.h.  pcps_acquisition only calls
 * Event_valid_acquisition() on it; the harness reads the count. */
#ifndef GB_ORACLE_SHIM_CHANNEL_FSM_H
#define GB_ORACLE_SHIM_CHANNEL_FSM_H
#include <cstdint>
#include <memory>
class ChannelFsm
{
public:
    int valid_acquisitions = 0;
    bool Event_valid_acquisition() { valid_acquisitions++; return true; }
};
#endif
