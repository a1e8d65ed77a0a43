// Stand-ins for the GNSS-SDR headers the B200 blocks are written against, used ONLY when this tree is built outside
// GNSS-SDR (no GNU Radio here).  Inside GNSS-SDR (-DGB200_WITH_GNSS_SDR) the real headers are included
// instead and nothing below is compiled.  Names, member names and virtual signatures follow
//   src/core/interfaces/gnss_block_interface.h:63-101, tracking_interface.h:45-52, acquisition_interface.h:50-70,
//   configuration_interface.h:44-59, src/core/system_parameters/gnss_synchro.h:36-100,
//   src/algorithms/channel/libs/channel_fsm.h:41-69
// of the reference, so that the adapters compile unchanged against either.
#ifndef GB200_GNSS_SDR_COMPAT_H
#define GB200_GNSS_SDR_COMPAT_H

#ifdef GB200_WITH_GNSS_SDR
#include "acquisition_interface.h"
#include "channel_fsm.h"
#include "configuration_interface.h"
#include "gnss_synchro.h"
#include "tracking_interface.h"
#include <gnuradio/gr_complex.h>
#else

#include <complex>
#include <cstddef>
#include <cstdint>
#include <memory>
#include <string>

using gr_complex = std::complex<float>;

namespace gr
{
// opaque handles: outside GNSS-SDR there is no flowgraph to connect to
using basic_block_sptr = std::shared_ptr<void>;
using top_block_sptr = std::shared_ptr<void>;
}  // namespace gr

class ConfigurationInterface
{
public:
    virtual ~ConfigurationInterface() = default;
    virtual std::string property(std::string property_name, std::string default_value) const = 0;
    virtual bool property(std::string property_name, bool default_value) const = 0;
    virtual int64_t property(std::string property_name, int64_t default_value) const = 0;
    virtual uint64_t property(std::string property_name, uint64_t default_value) const = 0;
    virtual int32_t property(std::string property_name, int32_t default_value) const = 0;
    virtual uint32_t property(std::string property_name, uint32_t default_value) const = 0;
    virtual int16_t property(std::string property_name, int16_t default_value) const = 0;
    virtual uint16_t property(std::string property_name, uint16_t default_value) const = 0;
    virtual float property(std::string property_name, float default_value) const = 0;
    virtual double property(std::string property_name, double default_value) const = 0;
    virtual void set_property(std::string property_name, std::string value) = 0;
};

class Gnss_Synchro
{
public:
    char System{};
    char Signal[3]{};
    uint32_t PRN{};
    int32_t Channel_ID{};
    double Acq_delay_samples{};
    double Acq_doppler_hz{};
    uint64_t Acq_samplestamp_samples{};
    uint32_t Acq_doppler_step{};
    int64_t fs{};
    double Prompt_I{};
    double Prompt_Q{};
    double CN0_dB_hz{};
    double Carrier_Doppler_hz{};
    double Carrier_phase_rads{};
    double Code_phase_samples{};
    uint64_t Tracking_sample_counter{};
    int32_t correlation_length_ms{};
    uint32_t TOW_at_current_symbol_ms{};
    double Pseudorange_m{};
    double RX_time{};
    double interp_TOW_ms{};
    bool Flag_valid_acquisition{};
    bool Flag_valid_symbol_output{};
    bool Flag_valid_word{};
    bool Flag_valid_pseudorange{};
    bool Flag_PLL_180_deg_phase_locked{};
};

class GNSSBlockInterface
{
public:
    virtual ~GNSSBlockInterface() = default;
    virtual std::string role() = 0;
    virtual std::string implementation() = 0;
    virtual size_t item_size() = 0;
    virtual void connect(gr::top_block_sptr top_block) = 0;
    virtual void disconnect(gr::top_block_sptr top_block) = 0;
    virtual gr::basic_block_sptr get_left_block() = 0;
    virtual gr::basic_block_sptr get_right_block() = 0;
    virtual void start() {}
};

class TrackingInterface : public GNSSBlockInterface
{
public:
    virtual void start_tracking() = 0;
    virtual void stop_tracking() = 0;
    virtual void set_gnss_synchro(Gnss_Synchro* gnss_synchro) = 0;
    virtual void set_channel(unsigned int channel) = 0;
};

// the three events the acquisition / tracking blocks raise on their channel (channel_fsm.h:56-62)
class ChannelFsm
{
public:
    virtual ~ChannelFsm() = default;
    virtual bool Event_valid_acquisition() { return true; }
    virtual bool Event_failed_acquisition_repeat() { return true; }
    virtual bool Event_failed_acquisition_no_repeat() { return true; }
    virtual bool Event_failed_tracking_standby() { return true; }
};

class AcquisitionInterface : public GNSSBlockInterface
{
public:
    virtual void set_gnss_synchro(Gnss_Synchro* gnss_synchro) = 0;
    virtual void set_channel(unsigned int channel_id) = 0;
    virtual void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) = 0;
    virtual void set_threshold(float threshold) = 0;
    virtual void set_doppler_max(unsigned int doppler_max) = 0;
    virtual void set_doppler_step(unsigned int doppler_step) = 0;
    virtual void set_doppler_center(int doppler_center __attribute__((unused))) {}
    virtual void init() = 0;
    virtual void set_local_code() = 0;
    virtual void set_state(int state) = 0;
    virtual signed int mag() = 0;
    virtual void reset() = 0;
    virtual void stop_acquisition() = 0;
    virtual void set_resampler_latency(uint32_t latency_samples) = 0;
};

#endif  // GB200_WITH_GNSS_SDR
#endif  // GB200_GNSS_SDR_COMPAT_H
