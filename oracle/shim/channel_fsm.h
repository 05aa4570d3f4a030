/* TEST INFRASTRUCTURE ONLY (oracle/_ref build): shadows src/algorithms/channel/libs/channel_fsm.h.  pcps_acquisition only calls
 * Event_valid_acquisition() on it; the harness reads the count. */
#ifndef GB_ORACLE_SHIM_CHANNEL_FSM_H
#define GB_ORACLE_SHIM_CHANNEL_FSM_H
#include <cstdint>
#include <functional>
#include <memory>
class ChannelFsm
{
public:
    int valid_acquisitions = 0;
    std::function<void()> on_valid_acquisition;  // harness hook: what ChannelFsm::Event_valid_acquisition does next (start_tracking)
    bool Event_valid_acquisition()
    {
        valid_acquisitions++;
        if (on_valid_acquisition) on_valid_acquisition();
        return true;
    }
};
#endif
