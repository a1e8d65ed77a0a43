// Adapters: the plugin boundary of the two hot paths.  Each class implements the reference's
// TrackingInterface / AcquisitionInterface with the reference constructor signature
//   (const ConfigurationInterface*, const std::string& role, unsigned in_streams, unsigned out_streams)
// (adapters/gps_l1_ca_dll_pll_tracking.cc:34-37, adapters/gps_l1_ca_pcps_acquisition.cc:39-43) and is selected by the
// `.implementation` string of its role.  GetB200Block() is the string chain INTEGRATION.md adds to
// GNSSBlockFactory::GetBlock / GetTrkBlock / GetAcqBlock (src/core/receiver/gnss_block_factory.cc:1340-1346).
#ifndef GB200_ADAPTERS_H
#define GB200_ADAPTERS_H

#include "b200_dll_pll_veml_tracking.h"
#include "b200_pcps_acquisition.h"
#include "b200_pcps_acquisition_fine_doppler.h"
#include "b200_e5a_noncoherent_iq_acquisition_caf.h"
#include "b200_pcps_quicksync_acquisition.h"
#include "gnss_sdr_compat.h"
#include <memory>
#include <string>
#include <vector>

namespace gb200h
{
// ---- tracking: one adapter class, parameterised by signal, with the thin named subclasses the factory instantiates
class B200DllPllTracking : public TrackingInterface
{
public:
    B200DllPllTracking(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
        char system, const char* signal, const char* implementation, double code_rate_cps, double code_length_chips, int max_extend_symbols);
    std::string role() override { return role_; }
    std::string implementation() override { return implementation_; }
    size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override;
    gr::basic_block_sptr get_right_block() override;
    void start_tracking() override;
    void stop_tracking() override;
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override;
    void set_channel(unsigned int channel) override;
    // the block this adapter wraps (what get_left_block() exposes inside GNSS-SDR)
    std::shared_ptr<b200_dll_pll_veml_tracking> block() { return tracking_; }
    const Dll_Pll_Conf& conf() const { return trk_params_; }

protected:
    Dll_Pll_Conf trk_params_;
    std::shared_ptr<b200_dll_pll_veml_tracking> tracking_;
    std::string role_, implementation_;
    size_t item_size_;
    unsigned int channel_{0}, in_streams_, out_streams_;
};

class GpsL1CaDllPllTrackingB200 : public B200DllPllTracking
{
public:
    GpsL1CaDllPllTrackingB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams)
        : B200DllPllTracking(configuration, role, in_streams, out_streams, 'G', "1C", "GPS_L1_CA_DLL_PLL_Tracking_B200", 1.023e6, 1023.0, 20)
    {
    }
};

class GalileoE1DllPllVemlTrackingB200 : public B200DllPllTracking
{
public:
    GalileoE1DllPllVemlTrackingB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams)
        : B200DllPllTracking(configuration, role, in_streams, out_streams, 'E', "1B", "Galileo_E1_DLL_PLL_VEML_Tracking_B200", 1.023e6, 4092.0, 0)
    {
    }
};

class GpsL5DllPllTrackingB200 : public B200DllPllTracking
{
public:
    GpsL5DllPllTrackingB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams)
        : B200DllPllTracking(configuration, role, in_streams, out_streams, 'G', "L5", "GPS_L5_DLL_PLL_Tracking_B200", 10.23e6, 10230.0, 10)
    {
    }
};

class GalileoE5aDllPllTrackingB200 : public B200DllPllTracking
{
public:
    GalileoE5aDllPllTrackingB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams)
        : B200DllPllTracking(configuration, role, in_streams, out_streams, 'E', "5X", "Galileo_E5a_DLL_PLL_Tracking_B200", 10.23e6, 10230.0, 20)
    {
    }
};

// ---- acquisition: one adapter class parameterised by signal (the four reference adapters differ only in the chip rate, the
// code period, the replica generator and two E1 / E5a options), with the thin named subclasses the factory instantiates
class B200PcpsAcquisition : public AcquisitionInterface
{
public:
    // adapters/gps_l1_ca_pcps_acquisition.cc:39-98, galileo_e1_pcps_ambiguous_acquisition.cc:36-98, gps_l5i_pcps_acquisition.cc:36-95,
    // galileo_e5a_pcps_acquisition.cc:36-90: ms_per_code, SetFromConfiguration(chip rate, optimum resampled rate), code / vector lengths
    B200PcpsAcquisition(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
        const char* implementation, uint32_t ms_per_code, double chip_rate_cps, double opt_acq_fs_sps, double code_length_chips);
    std::string role() override { return role_; }
    std::string implementation() override { return implementation_; }
    size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override;
    gr::basic_block_sptr get_right_block() override;
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override;
    void set_channel(unsigned int channel) override;
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) override;
    void set_threshold(float threshold) override;
    void set_doppler_max(unsigned int doppler_max) override;
    void set_doppler_step(unsigned int doppler_step) override;
    void set_doppler_center(int doppler_center) override;
    void init() override;
    void set_local_code() override;
    void set_state(int state) override;
    signed int mag() override;
    void reset() override;
    void stop_acquisition() override;
    void set_resampler_latency(uint32_t latency_samples) override;
    std::shared_ptr<b200_pcps_acquisition> block() { return acquisition_; }
    const Acq_Conf& conf() const { return acq_parameters_; }
    // the replica of gnss_synchro_'s satellite as set_local_code() builds it, without handing it to the block (the batched
    // all-PRN sweep of B200ChannelManager fills one code-spectrum slot per candidate satellite from it)
    void make_local_code();
    // the replica last built by make_local_code() / set_local_code()
    const std::vector<std::complex<float>>& local_code() const { return code_; }

protected:
    // one code period of the sampled replica of gnss_synchro_->PRN at `fs` into code[0..code_length_)
    virtual void generate_code(std::complex<float>* code, int32_t fs) = 0;
    // how many copies of that period fill the block's local code (sampled_ms for 1 ms codes, sampled_ms / 4 for E1)
    virtual unsigned int code_copies() const { return sampled_ms_; }

    Acq_Conf acq_parameters_;
    const ConfigurationInterface* configuration_;
    std::shared_ptr<b200_pcps_acquisition> acquisition_;
    std::vector<std::complex<float>> code_;
    std::weak_ptr<ChannelFsm> channel_fsm_;
    Gnss_Synchro* gnss_synchro_{nullptr};
    std::string role_, implementation_, item_type_;
    size_t item_size_;
    float threshold_{0.0F};
    int doppler_center_{0};
    unsigned int vector_length_, code_length_, channel_{0}, doppler_max_, doppler_step_, sampled_ms_, in_streams_, out_streams_;
};

// adapters/gps_l1_ca_pcps_acquisition.cc
class GpsL1CaPcpsAcquisitionB200 : public B200PcpsAcquisition
{
public:
    GpsL1CaPcpsAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);

protected:
    void generate_code(std::complex<float>* code, int32_t fs) override;
};

// adapters/galileo_e1_pcps_ambiguous_acquisition.cc: 4 ms codes, `acquire_pilot` (E1-C instead of the channel's signal), `Acquisition<ch>.cboc`
class GalileoE1PcpsAmbiguousAcquisitionB200 : public B200PcpsAcquisition
{
public:
    GalileoE1PcpsAmbiguousAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);

protected:
    void generate_code(std::complex<float>* code, int32_t fs) override;
    unsigned int code_copies() const override { return sampled_ms_ / 4; }
    bool acquire_pilot_;
};

// adapters/gps_l5i_pcps_acquisition.cc
class GpsL5iPcpsAcquisitionB200 : public B200PcpsAcquisition
{
public:
    GpsL5iPcpsAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);

protected:
    void generate_code(std::complex<float>* code, int32_t fs) override;
};

// adapters/galileo_e5a_pcps_acquisition.cc: `acquire_pilot` (5Q), `acquire_iq` (5X, wins over acquire_pilot), else 5I
class GalileoE5aPcpsAcquisitionB200 : public B200PcpsAcquisition
{
public:
    GalileoE5aPcpsAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);

protected:
    void generate_code(std::complex<float>* code, int32_t fs) override;
    bool acq_pilot_, acq_iq_;
};

// adapters/gps_l1_ca_pcps_acquisition_fine_doppler.{h:46-170,cc:30-190}: `GPS_L1_CA_PCPS_Acquisition_Fine_Doppler`; gr_complex items only,
// keys doppler_max (5000), max_dwells (1), coherent_integration_time_ms, blocking_on_standby, dump / dump_filename (the dump is not written here)
class GpsL1CaPcpsAcquisitionFineDopplerB200 : public AcquisitionInterface
{
public:
    GpsL1CaPcpsAcquisitionFineDopplerB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
    std::string role() override { return role_; }
    std::string implementation() override { return "GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200"; }
    size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override { return acquisition_cc_; }
    gr::basic_block_sptr get_right_block() override { return acquisition_cc_; }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override;
    void set_channel(unsigned int channel) override;
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) override { channel_fsm_ = std::move(channel_fsm); }
    void set_threshold(float threshold) override;
    void set_doppler_max(unsigned int doppler_max) override;
    void set_doppler_step(unsigned int doppler_step) override;
    void init() override { acquisition_cc_->init(); }
    void set_local_code() override;
    void set_state(int state) override { acquisition_cc_->set_state(state); }
    signed int mag() override { return static_cast<signed int>(acquisition_cc_->mag()); }
    void reset() override { acquisition_cc_->set_active(true); }
    void stop_acquisition() override;
    void set_resampler_latency(uint32_t latency_samples __attribute__((unused))) override {}
    std::shared_ptr<b200_pcps_acquisition_fine_doppler> block() { return acquisition_cc_; }
    const std::vector<std::complex<float>>& local_code() const { return code_; }

private:
    std::shared_ptr<b200_pcps_acquisition_fine_doppler> acquisition_cc_;
    std::weak_ptr<ChannelFsm> channel_fsm_;
    std::vector<std::complex<float>> code_;
    std::string role_, item_type_, dump_filename_;
    Gnss_Synchro* gnss_synchro_{nullptr};
    int64_t fs_in_;
    size_t item_size_;
    float threshold_{0.0F};
    int doppler_max_, max_dwells_;
    unsigned int vector_length_{0}, channel_{0}, doppler_step_{0}, sampled_ms_;
    bool dump_;
};

// adapters/galileo_e5a_noncoherent_iq_acquisition_caf.{h:50-200,cc:40-326}: `Galileo_E5a_Noncoherent_IQ_Acquisition_CAF`; gr_complex items only;
// keys doppler_max (5000), coherent_integration_time_ms (1, at most 3), max_dwells (1), bit_transition_flag, CAF_window_hz (0 = off),
// Zero_padding (> 0: 1 ms of code + 1 ms of zeros), pfa, and `Channel.signal` ("5X": both components, the default)
class GalileoE5aNoncoherentIQAcquisitionCafB200 : public AcquisitionInterface
{
public:
    GalileoE5aNoncoherentIQAcquisitionCafB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
    std::string role() override { return role_; }
    std::string implementation() override { return "Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200"; }
    size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override { return acquisition_cc_; }
    gr::basic_block_sptr get_right_block() override { return acquisition_cc_; }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override;
    void set_channel(unsigned int channel) override;
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) override { channel_fsm_ = std::move(channel_fsm); }
    void set_threshold(float threshold) override;
    void set_doppler_max(unsigned int doppler_max) override;
    void set_doppler_step(unsigned int doppler_step) override;
    void init() override { acquisition_cc_->init(); }
    void set_local_code() override;
    void set_state(int state) override { acquisition_cc_->set_state(state); }
    signed int mag() override { return static_cast<signed int>(acquisition_cc_->mag()); }
    void reset() override { acquisition_cc_->set_active(true); }
    void stop_acquisition() override;
    void set_resampler_latency(uint32_t latency_samples __attribute__((unused))) override {}
    std::shared_ptr<b200_e5a_noncoherent_iq_acquisition_caf> block() { return acquisition_cc_; }
    float threshold() const { return threshold_; }

private:
    float calculate_threshold(float pfa) const;
    const ConfigurationInterface* configuration_;
    std::shared_ptr<b200_e5a_noncoherent_iq_acquisition_caf> acquisition_cc_;
    std::weak_ptr<ChannelFsm> channel_fsm_;
    std::vector<std::complex<float>> codeI_, codeQ_;
    std::string role_, item_type_;
    Gnss_Synchro* gnss_synchro_{nullptr};
    int64_t fs_in_;
    size_t item_size_;
    float threshold_{0.0F};
    int Zero_padding, CAF_window_hz_, code_length_{0};
    unsigned int vector_length_{0}, channel_{0}, doppler_max_, doppler_step_{0}, sampled_ms_, max_dwells_;
    bool bit_transition_flag_, both_signal_components{false};
};

// adapters/gps_l1_ca_pcps_quicksync_acquisition.{h:50-185,cc:36-312}: `GPS_L1_CA_PCPS_QuickSync_Acquisition`; gr_complex items only; keys doppler_max
// (5000), coherent_integration_time_ms (4; rounded to a multiple of the folding factor), folding_factor (ceil(sqrt(log2(samples per code)))),
// bit_transition_flag (then max_dwells = 2), max_dwells (1), pfa.  The block's input items are vectors of samples-per-code x folding_factor
// samples (the adapter's stream_to_vector, :113-114).
class GpsL1CaPcpsQuickSyncAcquisitionB200 : public AcquisitionInterface
{
public:
    GpsL1CaPcpsQuickSyncAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
    std::string role() override { return role_; }
    std::string implementation() override { return galileo_ ? "Galileo_E1_PCPS_QuickSync_Ambiguous_Acquisition_B200" : "GPS_L1_CA_PCPS_QuickSync_Acquisition_B200"; }
    size_t item_size() override { return item_size_; }
    void connect(gr::top_block_sptr top_block) override;
    void disconnect(gr::top_block_sptr top_block) override;
    gr::basic_block_sptr get_left_block() override { return acquisition_cc_; }  // (the reference returns its stream_to_vector here; the block takes the stream itself)
    gr::basic_block_sptr get_right_block() override { return acquisition_cc_; }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) override;
    void set_channel(unsigned int channel) override;
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) override { channel_fsm_ = std::move(channel_fsm); }
    void set_threshold(float threshold) override;
    void set_doppler_max(unsigned int doppler_max) override;
    void set_doppler_step(unsigned int doppler_step) override;
    void init() override { acquisition_cc_->init(); }
    void set_local_code() override;
    void set_state(int state) override { acquisition_cc_->set_state(state); }
    signed int mag() override { return static_cast<signed int>(acquisition_cc_->mag()); }
    void reset() override { acquisition_cc_->set_active(true); }
    void stop_acquisition() override;
    void set_resampler_latency(uint32_t latency_samples __attribute__((unused))) override {}
    std::shared_ptr<b200_pcps_quicksync_acquisition> block() { return acquisition_cc_; }
    unsigned int folding_factor() const { return folding_factor_; }
    unsigned int sampled_ms() const { return sampled_ms_; }

protected:
    struct GalileoE1Tag
    {
    };
    // adapters/galileo_e1_pcps_quicksync_ambiguous_acquisition.cc:36-150
    GpsL1CaPcpsQuickSyncAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams, GalileoE1Tag);
    bool galileo_{false};

private:
    float calculate_threshold(float pfa) const;
    const ConfigurationInterface* configuration_;
    std::shared_ptr<b200_pcps_quicksync_acquisition> acquisition_cc_;
    std::weak_ptr<ChannelFsm> channel_fsm_;
    std::vector<std::complex<float>> code_;
    std::string role_, item_type_;
    Gnss_Synchro* gnss_synchro_{nullptr};
    int64_t fs_in_;
    size_t item_size_;
    float threshold_{0.0F};
    unsigned int channel_{0}, doppler_max_, doppler_step_{0}, sampled_ms_, max_dwells_{1}, folding_factor_{2}, code_length_{0}, vector_length_{0};
    bool bit_transition_flag_;
};

// adapters/galileo_e1_pcps_quicksync_ambiguous_acquisition.{h,cc}: `Galileo_E1_PCPS_QuickSync_Ambiguous_Acquisition` — the same block over 4 ms E1 codes:
// coherent_integration_time_ms (8) a multiple of 4 x folding_factor (2), `Acquisition<channel>.cboc`, vector items of sampled_ms x samples per ms
class GalileoE1PcpsQuickSyncAmbiguousAcquisitionB200 : public GpsL1CaPcpsQuickSyncAcquisitionB200
{
public:
    GalileoE1PcpsQuickSyncAmbiguousAcquisitionB200(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams)
        : GpsL1CaPcpsQuickSyncAcquisitionB200(configuration, role, in_streams, out_streams, GalileoE1Tag{})
    {
    }
};

// The `.implementation` strings this tree serves; nullptr for any other string (the caller falls through to the
// reference's own chain).
std::unique_ptr<GNSSBlockInterface> GetB200Block(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
std::unique_ptr<TrackingInterface> GetB200TrkBlock(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
std::unique_ptr<AcquisitionInterface> GetB200AcqBlock(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams);
}  // namespace gb200h
#endif
