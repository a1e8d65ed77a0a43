// Test-infrastructure stub (NOT product code): stands where the reference's channel_fsm.h is included (it needs the whole channel /
// tracking / telemetry interface tree).  pcps_acquisition only raises the three acquisition events on it; they are recorded here.
#ifndef ORACLE_SHIM_ACQ_CHANNEL_FSM_H
#define ORACLE_SHIM_ACQ_CHANNEL_FSM_H
#include "gnss_block_interface.h"
#include <vector>
class ChannelFsm
{
public:
    virtual ~ChannelFsm() = default;
    virtual bool Event_valid_acquisition()
    {
        events.push_back(1);
        return true;
    }
    virtual bool Event_failed_acquisition_repeat()
    {
        events.push_back(2);
        return true;
    }
    virtual bool Event_failed_acquisition_no_repeat()
    {
        events.push_back(3);
        return true;
    }
    std::vector<int> events;
};
#endif
