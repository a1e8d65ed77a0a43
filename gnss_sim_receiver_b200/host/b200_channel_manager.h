// The callers on both sides of the hot paths, batched: what Channel + ChannelFsm + the flowgraph's satellite
// assignment do per channel in the reference, done for all channels of one signal at once.
//
// Reference shape (src/algorithms/channel/adapters/channel.cc:30-304, libs/channel_fsm.cc:44-217,
// src/core/receiver/gnss_flowgraph.cc:1797-1960): every channel owns an acquisition block searching ONE PRN at a time; the
// flowgraph throttles how many channels may be acquiring (`Channels.in_acquisition`), hands a channel the next satellite of
// the `available_..._signals` queue after a failed search (event 0), moves it to tracking after a valid one
// (Event_valid_acquisition -> start_tracking, event 1) and back to acquisition after a loss of lock (event 2 / tracking
// message 3, dll_pll_veml_tracking.cc:1023).
//
// B200 shape: ONE sweep searches every candidate PRN x Doppler bin on the same buffered block (B200AcquisitionSweep) and
// feeds all idle channels, strongest detections first; the channels then run in the lock-step tracking bank.  Channel states
// keep ChannelFsm's numbering: 0 standby (idle, waiting for the next sweep), 1 acquisition (a sweep is buffering / running
// for it), 2 tracking.  Satellites being tracked are not handed out twice; a satellite whose channel lost lock returns to the
// candidates and the channel to state 0.
#ifndef GB200_CHANNEL_MANAGER_H
#define GB200_CHANNEL_MANAGER_H

#include "b200_adapters.h"
#include "b200_pcps_acquisition.h"
#include "b200_tracking_bank.h"
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace gb200h
{
class B200ChannelManager
{
public:
    // `acq_role` / `trk_role`: the .conf roles (e.g. "Acquisition_1C", "Tracking_1C"); `prns`: the satellites to search
    B200ChannelManager(const ConfigurationInterface* configuration, const std::string& acq_role, const std::string& trk_role, int n_channels, const std::vector<unsigned>& prns);

    // Feed the next stretch of the stream (host samples of the configured item type).  Buffers an acquisition block while
    // channels are idle, sweeps, hands detections over to idle channels, advances every tracking channel.
    void push(const void* iq, int64_t n_samples, std::vector<B200TrackingBank::Output>* out);

    struct ChannelInfo
    {
        int state;             // ChannelFsm numbering: 0 standby, 1 acquisition, 2 tracking
        unsigned prn;          // satellite assigned (0 when none)
        int acquisitions;      // valid acquisitions handed to this channel (ChannelFsm::Event_valid_acquisition)
        int losses;            // losses of lock (ChannelFsm::Event_failed_tracking_standby)
        double acq_doppler_hz, acq_delay_samples;
        uint64_t acq_samplestamp_samples;
        float test_statistics;
    };
    ChannelInfo info(int ch) const { return info_[ch]; }
    int size() const { return static_cast<int>(info_.size()); }
    uint64_t sweeps() const { return sweeps_; }
    uint64_t samples() const { return pos_; }
    float threshold() const { return threshold_; }
    B200TrackingBank& bank() { return *bank_; }
    std::shared_ptr<B200CorrelatorEngine> engine() { return engine_; }

private:
    void sweep(uint64_t stamp);
    bool any_idle() const;

    Acq_Conf acq_conf_;
    std::vector<unsigned> prns_;
    std::unique_ptr<B200AcquisitionSweep> sweep_;
    std::shared_ptr<B200CorrelatorEngine> engine_;
    std::unique_ptr<B200TrackingBank> bank_;
    std::vector<std::unique_ptr<Gnss_Synchro>> synchros_;
    std::vector<ChannelInfo> info_;
    std::vector<char> buf_;        // the acquisition data buffer (d_data_buffer, pcps_acquisition.cc:953-985)
    uint32_t consumed_{0}, buffered_{0};
    size_t item_bytes_{8};
    uint64_t pos_{0}, sweeps_{0};
    float threshold_{0.0F};
    std::vector<gb200_acq_result> results_;
};
}  // namespace gb200h
#endif
