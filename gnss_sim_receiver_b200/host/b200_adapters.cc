#include "b200_adapters.h"
#include "b200_gnss_codes.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <iostream>

namespace gb200h
{
// ------------------------------------------------------------------ tracking
// Parameter handling of adapters/gps_l1_ca_dll_pll_tracking.cc:42-93 and adapters/galileo_e1_dll_pll_veml_tracking.cc:42-90
B200DllPllTracking::B200DllPllTracking(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
    char system, const char* signal, const char* implementation, double code_rate_cps, double code_length_chips, int max_extend_symbols)
    : role_(role), implementation_(implementation), item_size_(sizeof(gr_complex)), in_streams_(in_streams), out_streams_(out_streams)
{
    trk_params_.SetFromConfiguration(configuration, role_);
    trk_params_.vector_length = static_cast<uint32_t>(std::round(trk_params_.fs_in / (code_rate_cps / code_length_chips)));
    if (trk_params_.extend_correlation_symbols < 1)
        {
            trk_params_.extend_correlation_symbols = 1;
            std::cout << "WARNING: " << implementation << ". extend_correlation_symbols must be bigger than 0. Coherent integration has been set to 1 symbol\n";
        }
    else if (max_extend_symbols > 0 && trk_params_.extend_correlation_symbols > max_extend_symbols)
        {
            trk_params_.extend_correlation_symbols = max_extend_symbols;
            std::cout << "WARNING: " << implementation << ". extend_correlation_symbols must be lower than " << max_extend_symbols + 1 << ". Coherent integration has been set to "
                      << max_extend_symbols << " symbols\n";
        }
    if (system == 'G' && std::string(signal) == "1C")
        {
            trk_params_.track_pilot = configuration->property(role_ + ".track_pilot", false);
            if (trk_params_.track_pilot)
                {
                    trk_params_.track_pilot = false;
                    std::cout << "WARNING: GPS L1 C/A does not have pilot signal. Data tracking has been enabled\n";
                }
        }
    if (system == 'E' && std::string(signal) == "1B" && !trk_params_.track_pilot && trk_params_.extend_correlation_symbols > 1)
        {
            trk_params_.extend_correlation_symbols = 1;
            std::cout << "WARNING: Galileo E1. Extended coherent integration is not allowed when tracking the data component. Coherent integration has been set to 4 ms (1 symbol)\n";
        }
    if (trk_params_.extend_correlation_symbols > 1 && (trk_params_.pll_bw_narrow_hz > trk_params_.pll_bw_hz || trk_params_.dll_bw_narrow_hz > trk_params_.dll_bw_hz))
        std::cout << "WARNING: " << implementation << ". PLL or DLL narrow tracking bandwidth is higher than wide tracking one\n";
    trk_params_.system = system;
    std::memcpy(trk_params_.signal, signal, 3);

    // the engine ingests gr_complex, cshort and cbyte rings; the per-channel block takes the stream's item type
    const int iq = item_type_iq(trk_params_.item_type);
    if (iq >= 0)
        {
            item_size_ = item_type_size(trk_params_.item_type);
            tracking_ = std::make_shared<b200_dll_pll_veml_tracking>(trk_params_);
        }
    else
        {
            item_size_ = 0;  // Channel::connect throws on item_size() == 0 (channel.cc:118-122)
            std::cerr << "WARNING: " << trk_params_.item_type << " unknown tracking item type.\n";
        }
    if (in_streams_ > 1) std::cerr << "ERROR: This implementation only supports one input stream\n";
    if (out_streams_ > 1) std::cerr << "ERROR: This implementation only supports one output stream\n";
}

void B200DllPllTracking::start_tracking()
{
    if (!tracking_) return;
    if (!tracking_->tracking_code().empty() && tracking_->code_table_entries() > 0)
        {
            // attach to the GPU engine of the configured device on first use
            // One engine per (device, item type, input stream).  Channels of different signals on the same stream share it (GPS L1 C/A
            // and Galileo E1 on an L1 front-end), so it is sized for the longest code period of the band, not for this channel's:
            // 4 ms (Galileo E1) on the L1 band, 1 ms on the L5 / E5a band.  `<role>.stream_id` overrides the band default.
            const bool l5_band = std::string(trk_params_.signal) == "L5" || std::string(trk_params_.signal) == "5X";
            const auto longest_period = static_cast<int64_t>(std::llround(trk_params_.fs_in * (l5_band ? 1e-3 : 4e-3)));
            EngineConfig ec;
            ec.device = trk_params_.device;
            ec.iq_dtype = item_type_iq(trk_params_.item_type);
            ec.stream_id = trk_params_.stream_id >= 0 ? trk_params_.stream_id : (l5_band ? 1 : 0);
            ec.max_block = static_cast<int>(2 * std::max<int64_t>(longest_period, trk_params_.vector_length));
            ec.max_code_len = 10230;  // longest table of the four signals (E1: 8184 entries, L5 / E5a: 10230)
            ec.ring_samples = std::max<int64_t>(int64_t{1} << 22, 32LL * ec.max_block);
            tracking_->set_engine(B200CorrelatorEngine::instance(ec));
        }
    tracking_->start_tracking();
}
void B200DllPllTracking::stop_tracking()
{
    if (tracking_) tracking_->stop_tracking();
}
void B200DllPllTracking::set_channel(unsigned int channel)
{
    channel_ = channel;
    if (tracking_) tracking_->set_channel(channel);
}
void B200DllPllTracking::set_gnss_synchro(Gnss_Synchro* p)
{
    if (tracking_) tracking_->set_gnss_synchro(p);
}
void B200DllPllTracking::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect, now the tracking uses gr_sync_decimator (adapters/gps_l1_ca_dll_pll_tracking.cc:133-139) */ }
}
void B200DllPllTracking::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
// Outside GNSS-SDR the "block" handle is the b200 block object itself; inside GNSS-SDR it is the gr::block wrapper of
// INTEGRATION.md that forwards general_work to it.
gr::basic_block_sptr B200DllPllTracking::get_left_block() { return tracking_; }
gr::basic_block_sptr B200DllPllTracking::get_right_block() { return tracking_; }

// ------------------------------------------------------------------ acquisition
// Common constructor body of the four reference adapters (adapters/gps_l1_ca_pcps_acquisition.cc:39-98,
// galileo_e1_pcps_ambiguous_acquisition.cc:36-98, gps_l5i_pcps_acquisition.cc:36-95, galileo_e5a_pcps_acquisition.cc:36-90)
B200PcpsAcquisition::B200PcpsAcquisition(const ConfigurationInterface* configuration, const std::string& role, unsigned int in_streams, unsigned int out_streams,
    const char* implementation, uint32_t ms_per_code, double chip_rate_cps, double opt_acq_fs_sps, double code_length_chips)
    : configuration_(configuration), role_(role), implementation_(implementation), in_streams_(in_streams), out_streams_(out_streams)
{
    acq_parameters_.ms_per_code = ms_per_code;
    acq_parameters_.SetFromConfiguration(configuration, role, chip_rate_cps, opt_acq_fs_sps);
    doppler_max_ = static_cast<unsigned int>(acq_parameters_.doppler_max);
    doppler_step_ = static_cast<unsigned int>(acq_parameters_.doppler_step);
    item_type_ = acq_parameters_.item_type;
    item_size_ = acq_parameters_.it_size;
    code_length_ = static_cast<unsigned int>(std::floor(static_cast<double>(acq_parameters_.resampled_fs) / (chip_rate_cps / code_length_chips)));
    vector_length_ = static_cast<unsigned int>(std::floor(acq_parameters_.sampled_ms * acq_parameters_.samples_per_ms) * (acq_parameters_.bit_transition_flag ? 2.0 : 1.0));
    code_.assign(vector_length_, std::complex<float>(0, 0));
    sampled_ms_ = acq_parameters_.sampled_ms;
    if (item_type_iq(item_type_) >= 0)
        acquisition_ = std::make_shared<b200_pcps_acquisition>(acq_parameters_);
    else
        {
            item_size_ = 0;
            std::cerr << "WARNING: " << item_type_ << " unknown acquisition item type\n";
        }
    if (in_streams_ > 1) std::cerr << "ERROR: This implementation only supports one input stream\n";
    if (out_streams_ > 0) std::cerr << "ERROR: This implementation does not provide an output stream\n";
}

void B200PcpsAcquisition::stop_acquisition() { acquisition_->set_active(false); }
void B200PcpsAcquisition::set_threshold(float threshold)
{
    threshold_ = threshold;
    acquisition_->set_threshold(threshold_);
}
void B200PcpsAcquisition::set_doppler_max(unsigned int doppler_max)
{
    doppler_max_ = doppler_max;
    acquisition_->set_doppler_max(doppler_max_);
}
void B200PcpsAcquisition::set_doppler_step(unsigned int doppler_step)
{
    doppler_step_ = doppler_step;
    acquisition_->set_doppler_step(doppler_step_);
}
void B200PcpsAcquisition::set_doppler_center(int doppler_center)
{
    doppler_center_ = doppler_center;
    acquisition_->set_doppler_center(doppler_center_);
}
void B200PcpsAcquisition::set_gnss_synchro(Gnss_Synchro* p)
{
    gnss_synchro_ = p;
    acquisition_->set_gnss_synchro(gnss_synchro_);
}
void B200PcpsAcquisition::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_->set_channel(channel_);
}
void B200PcpsAcquisition::set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm)
{
    channel_fsm_ = std::move(channel_fsm);
    acquisition_->set_channel_fsm(channel_fsm_);
}
signed int B200PcpsAcquisition::mag() { return static_cast<signed int>(acquisition_->mag()); }
void B200PcpsAcquisition::init() { acquisition_->init(); }

// one period of the sampled replica, repeated to fill the block's local code (gps_l1_ca_pcps_acquisition.cc:151-170,
// galileo_e1_pcps_ambiguous_acquisition.cc:149-199, gps_l5i_pcps_acquisition.cc:151-171, galileo_e5a_pcps_acquisition.cc:145-180)
void B200PcpsAcquisition::set_local_code()
{
    make_local_code();
    acquisition_->set_local_code(code_.data());
}
void B200PcpsAcquisition::make_local_code()
{
    std::vector<std::complex<float>> code(code_length_);
    const auto fs = static_cast<int32_t>(acq_parameters_.use_automatic_resampler ? acq_parameters_.resampled_fs : acq_parameters_.fs_in);
    generate_code(code.data(), fs);
    const unsigned int copies = code_copies();
    for (unsigned int i = 0; i < copies && (i + 1) * code_length_ <= vector_length_; i++) std::copy_n(code.data(), code_length_, code_.data() + i * code_length_);
}
void B200PcpsAcquisition::reset() { acquisition_->set_active(true); }
void B200PcpsAcquisition::set_state(int state) { acquisition_->set_state(state); }
void B200PcpsAcquisition::set_resampler_latency(uint32_t latency_samples) { acquisition_->set_resampler_latency(latency_samples); }
void B200PcpsAcquisition::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* gr_complex, cshort and cbyte are ingested directly: no converter blocks to connect (cf. gps_l1_ca_pcps_acquisition.cc:187-203) */ }
}
void B200PcpsAcquisition::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
gr::basic_block_sptr B200PcpsAcquisition::get_left_block() { return acquisition_; }
gr::basic_block_sptr B200PcpsAcquisition::get_right_block() { return acquisition_; }

// GPS_L1_CA_CODE_RATE_CPS, GPS_L1_CA_OPT_ACQ_FS_SPS, GPS_L1_CA_CODE_LENGTH_CHIPS (GPS_L1_CA.h)
GpsL1CaPcpsAcquisitionB200::GpsL1CaPcpsAcquisitionB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
    : B200PcpsAcquisition(c, role, in, out, "GPS_L1_CA_PCPS_Acquisition_B200", 1, 1.023e6, 2000000.0, 1023.0)
{
}
void GpsL1CaPcpsAcquisitionB200::generate_code(std::complex<float>* code, int32_t fs)
{
    gps_l1_ca_code_gen_complex_sampled(code, static_cast<int>(code_length_), gnss_synchro_->PRN, fs, 0);
}

// GALILEO_E1_CODE_CHIP_RATE_CPS, GALILEO_E1_OPT_ACQ_FS_SPS, GALILEO_E1_B_CODE_LENGTH_CHIPS (Galileo_E1.h:39-53)
GalileoE1PcpsAmbiguousAcquisitionB200::GalileoE1PcpsAmbiguousAcquisitionB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
    : B200PcpsAcquisition(c, role, in, out, "Galileo_E1_PCPS_Ambiguous_Acquisition_B200", 4, 1.023e6, 2000000.0, 4092.0),
      acquire_pilot_(c->property(role + ".acquire_pilot", false))
{
}
void GalileoE1PcpsAmbiguousAcquisitionB200::generate_code(std::complex<float>* code, int32_t fs)
{
    // the reference reads "Acquisition<channel>.cboc" (no signal suffix) at every set_local_code (:151-152)
    const bool cboc = configuration_->property("Acquisition" + std::to_string(channel_) + ".cboc", false);
    char sig[3] = {'1', 'C', '\0'};
    if (!acquire_pilot_)
        {
            sig[0] = gnss_synchro_->Signal[0];
            sig[1] = gnss_synchro_->Signal[1];
        }
    galileo_e1_code_gen_complex_sampled(code, sig, cboc, gnss_synchro_->PRN, fs, 0, false);
}

// GPS_L5I_CODE_RATE_CPS, GPS_L5_OPT_ACQ_FS_SPS, GPS_L5I_CODE_LENGTH_CHIPS (GPS_L5.h)
GpsL5iPcpsAcquisitionB200::GpsL5iPcpsAcquisitionB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
    : B200PcpsAcquisition(c, role, in, out, "GPS_L5i_PCPS_Acquisition_B200", 1, 10.23e6, 10000000.0, 10230.0)
{
}
void GpsL5iPcpsAcquisitionB200::generate_code(std::complex<float>* code, int32_t fs) { gps_l5i_code_gen_complex_sampled(code, gnss_synchro_->PRN, fs); }

// GALILEO_E5A_CODE_CHIP_RATE_CPS, GALILEO_E5A_OPT_ACQ_FS_SPS, GALILEO_E5A_CODE_LENGTH_CHIPS (Galileo_E5a.h)
GalileoE5aPcpsAcquisitionB200::GalileoE5aPcpsAcquisitionB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
    : B200PcpsAcquisition(c, role, in, out, "Galileo_E5a_Pcps_Acquisition_B200", 1, 10.23e6, 10000000.0, 10230.0),
      acq_pilot_(c->property(role + ".acquire_pilot", false)),
      acq_iq_(c->property(role + ".acquire_iq", false))
{
    if (acq_iq_) acq_pilot_ = false;  // :77-80
}
void GalileoE5aPcpsAcquisitionB200::generate_code(std::complex<float>* code, int32_t fs)
{
    const char sig[3] = {'5', acq_iq_ ? 'X' : acq_pilot_ ? 'Q' : 'I', '\0'};
    galileo_e5_a_code_gen_complex_sampled(code, gnss_synchro_->PRN, sig, fs, 0);
}

// ------------------------------------------------------------------ factory chain
std::unique_ptr<TrackingInterface> GetB200TrkBlock(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
{
    const std::string impl = c->property(role + ".implementation", std::string("Pass_Through"));
    if (impl == "GPS_L1_CA_DLL_PLL_Tracking_B200") return std::make_unique<GpsL1CaDllPllTrackingB200>(c, role, in, out);
    if (impl == "Galileo_E1_DLL_PLL_VEML_Tracking_B200") return std::make_unique<GalileoE1DllPllVemlTrackingB200>(c, role, in, out);
    if (impl == "GPS_L5_DLL_PLL_Tracking_B200") return std::make_unique<GpsL5DllPllTrackingB200>(c, role, in, out);
    if (impl == "Galileo_E5a_DLL_PLL_Tracking_B200") return std::make_unique<GalileoE5aDllPllTrackingB200>(c, role, in, out);
    return nullptr;
}

// adapters/gps_l1_ca_pcps_acquisition_fine_doppler.cc:30-100
GpsL1CaPcpsAcquisitionFineDopplerB200::GpsL1CaPcpsAcquisitionFineDopplerB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
    : role_(role),
      item_size_(sizeof(gr_complex)),
      doppler_max_(c->property(role + ".doppler_max", 5000)),
      max_dwells_(c->property(role + ".max_dwells", 1)),
      sampled_ms_(static_cast<unsigned int>(c->property(role + ".coherent_integration_time_ms", 1))),
      dump_(c->property(role + ".dump", false))
{
    Acq_Conf acq_parameters = Acq_Conf();
    item_type_ = c->property(role_ + ".item_type", std::string("gr_complex"));
    const int64_t fs_in_deprecated = c->property("GNSS-SDR.internal_fs_hz", static_cast<int64_t>(2048000));
    fs_in_ = c->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    acq_parameters.fs_in = fs_in_;
    acq_parameters.samples_per_chip = static_cast<unsigned int>(std::ceil((1.0 / 1023000.0) * static_cast<float>(acq_parameters.fs_in)));  // GPS_L1_CA_CHIP_PERIOD_S
    acq_parameters.dump = dump_;
    dump_filename_ = c->property(role_ + ".dump_filename", std::string("./acquisition.mat"));
    acq_parameters.dump_filename = dump_filename_;
    acq_parameters.doppler_max = doppler_max_;
    acq_parameters.sampled_ms = sampled_ms_;
    acq_parameters.max_dwells = static_cast<uint32_t>(max_dwells_);
    acq_parameters.blocking_on_standby = c->property(role + ".blocking_on_standby", false);
    acq_parameters.device = c->property(role + ".device", 0);
    vector_length_ = static_cast<unsigned int>(std::round(static_cast<double>(fs_in_) / (1023000.0 / 1023.0)));
    acq_parameters.samples_per_ms = static_cast<float>(vector_length_);
    code_ = std::vector<std::complex<float>>(vector_length_);
    // the reference block takes gr_complex only; any other item type leaves it without a block (:76-85) — here that is an error at once
    if (item_type_ != "gr_complex") throw std::invalid_argument(item_type_ + " unknown acquisition item type (GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200 takes gr_complex)");
    acquisition_cc_ = std::make_shared<b200_pcps_acquisition_fine_doppler>(acq_parameters);
    if (in > 1) std::cerr << "This implementation only supports one input stream\n";
    if (out > 0) std::cerr << "This implementation does not provide an output stream\n";
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect internally */ }
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::stop_acquisition()
{
    acquisition_cc_->set_state(0);
    acquisition_cc_->set_active(false);
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::set_threshold(float threshold)
{
    threshold_ = threshold;
    acquisition_cc_->set_threshold(threshold_);
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::set_doppler_max(unsigned int doppler_max)
{
    doppler_max_ = static_cast<int>(doppler_max);
    acquisition_cc_->set_doppler_max(doppler_max);
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::set_doppler_step(unsigned int doppler_step)
{
    doppler_step_ = doppler_step;
    acquisition_cc_->set_doppler_step(doppler_step_);
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::set_gnss_synchro(Gnss_Synchro* gnss_synchro)
{
    gnss_synchro_ = gnss_synchro;
    acquisition_cc_->set_gnss_synchro(gnss_synchro_);
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_cc_->set_channel(channel_);
}
void GpsL1CaPcpsAcquisitionFineDopplerB200::set_local_code()
{
    gps_l1_ca_code_gen_complex_sampled(code_.data(), static_cast<int>(code_.size()), gnss_synchro_->PRN, static_cast<int32_t>(fs_in_), 0);
    acquisition_cc_->set_local_code(code_.data());
}

// adapters/galileo_e5a_noncoherent_iq_acquisition_caf.cc:40-128
GalileoE5aNoncoherentIQAcquisitionCafB200::GalileoE5aNoncoherentIQAcquisitionCafB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
    : configuration_(c),
      role_(role),
      item_size_(sizeof(gr_complex)),
      Zero_padding(c->property(role + ".Zero_padding", 0)),
      CAF_window_hz_(c->property(role + ".CAF_window_hz", 0)),
      doppler_max_(static_cast<unsigned int>(c->property(role + ".doppler_max", 5000))),
      sampled_ms_(static_cast<unsigned int>(c->property(role + ".coherent_integration_time_ms", 1))),
      max_dwells_(static_cast<unsigned int>(c->property(role + ".max_dwells", 1))),
      bit_transition_flag_(c->property(role + ".bit_transition_flag", false))
{
    item_type_ = c->property(role_ + ".item_type", std::string("gr_complex"));
    const int64_t fs_in_deprecated = c->property("GNSS-SDR.internal_fs_hz", static_cast<int64_t>(32000000));
    fs_in_ = c->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    if (sampled_ms_ > 3)
        {
            sampled_ms_ = 3;
            std::cout << "Too high coherent integration time. Changing to 3ms\n";
        }
    if (Zero_padding > 0)
        {
            sampled_ms_ = 2;
            std::cout << "Zero padding activated. Changing to 1ms code + 1ms zero padding\n";
        }
    code_length_ = static_cast<int>(std::round(static_cast<double>(fs_in_) / 10230000.0 * 10230.0));  // GALILEO_E5A_CODE_CHIP_RATE_CPS, GALILEO_E5A_CODE_LENGTH_CHIPS
    vector_length_ = static_cast<unsigned int>(code_length_) * sampled_ms_;
    codeI_ = std::vector<std::complex<float>>(vector_length_);
    codeQ_ = std::vector<std::complex<float>>(vector_length_);
    const std::string sig_ = c->property("Channel.signal", std::string("5X"));
    if (sig_.size() >= 2 && sig_.at(0) == '5' && sig_.at(1) == 'X') both_signal_components = true;
    if (item_type_ != "gr_complex") throw std::invalid_argument(item_type_ + " unknown acquisition item type (Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200 takes gr_complex)");
    acquisition_cc_ = std::make_shared<b200_e5a_noncoherent_iq_acquisition_caf>(sampled_ms_, max_dwells_, doppler_max_, fs_in_, code_length_, code_length_, bit_transition_flag_,
        both_signal_components, CAF_window_hz_, Zero_padding, c->property(role + ".device", 0));
    if (in > 1) std::cerr << "This implementation only supports one input stream\n";
    if (out > 0) std::cerr << "This implementation does not provide an output stream\n";
}
void GalileoE5aNoncoherentIQAcquisitionCafB200::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to connect internally */ }
}
void GalileoE5aNoncoherentIQAcquisitionCafB200::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
void GalileoE5aNoncoherentIQAcquisitionCafB200::stop_acquisition()
{
    acquisition_cc_->set_state(0);
    acquisition_cc_->set_active(false);
}
// :134-154
void GalileoE5aNoncoherentIQAcquisitionCafB200::set_threshold(float threshold)
{
    float pfa = configuration_->property(role_ + std::to_string(channel_) + ".pfa", static_cast<float>(0.0));
    if (pfa == 0.0) pfa = configuration_->property(role_ + ".pfa", static_cast<float>(0.0));
    threshold_ = pfa == 0.0 ? threshold : calculate_threshold(pfa);
    acquisition_cc_->set_threshold(threshold_);
}
// :272-289: the (1 - pfa)^(1 / cells) quantile of an exponential distribution of rate vector_length (Boost there: -log1p(-p) / lambda)
float GalileoE5aNoncoherentIQAcquisitionCafB200::calculate_threshold(float pfa) const
{
    unsigned int frequency_bins = 0;
    for (int doppler = static_cast<int>(-doppler_max_); doppler <= static_cast<int>(doppler_max_); doppler += static_cast<int>(doppler_step_)) frequency_bins++;
    const unsigned int ncells = vector_length_ * frequency_bins;
    const double exponent = 1 / static_cast<double>(ncells);
    const double val = std::pow(1.0 - pfa, exponent);
    const auto lambda = static_cast<double>(vector_length_);
    return static_cast<float>(-std::log1p(-val) / lambda);
}
void GalileoE5aNoncoherentIQAcquisitionCafB200::set_doppler_max(unsigned int doppler_max)
{
    doppler_max_ = doppler_max;
    acquisition_cc_->set_doppler_max(doppler_max_);
}
void GalileoE5aNoncoherentIQAcquisitionCafB200::set_doppler_step(unsigned int doppler_step)
{
    doppler_step_ = doppler_step;
    acquisition_cc_->set_doppler_step(doppler_step_);
}
void GalileoE5aNoncoherentIQAcquisitionCafB200::set_gnss_synchro(Gnss_Synchro* gnss_synchro)
{
    gnss_synchro_ = gnss_synchro;
    acquisition_cc_->set_gnss_synchro(gnss_synchro_);
}
void GalileoE5aNoncoherentIQAcquisitionCafB200::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_cc_->set_channel(channel_);
}
// :211-261
void GalileoE5aNoncoherentIQAcquisitionCafB200::set_local_code()
{
    std::vector<std::complex<float>> codeI(static_cast<size_t>(code_length_)), codeQ(static_cast<size_t>(code_length_));
    const bool x = gnss_synchro_->Signal[0] == '5' && gnss_synchro_->Signal[1] == 'X';
    if (x)
        {
            galileo_e5_a_code_gen_complex_sampled(codeI.data(), gnss_synchro_->PRN, "5I", static_cast<int32_t>(fs_in_), 0);
            galileo_e5_a_code_gen_complex_sampled(codeQ.data(), gnss_synchro_->PRN, "5Q", static_cast<int32_t>(fs_in_), 0);
        }
    else
        galileo_e5_a_code_gen_complex_sampled(codeI.data(), gnss_synchro_->PRN, "5X", static_cast<int32_t>(fs_in_), 0);
    const unsigned int copies = Zero_padding == 0 ? sampled_ms_ : 1U;  // zero padding: 1 ms of code, the rest stays zero
    for (unsigned int i = 0; i < copies; i++)
        {
            std::copy_n(codeI.data(), code_length_, codeI_.data() + static_cast<size_t>(i) * code_length_);
            if (x) std::copy_n(codeQ.data(), code_length_, codeQ_.data() + static_cast<size_t>(i) * code_length_);
        }
    acquisition_cc_->set_local_code(codeI_.data(), codeQ_.data());
}

// adapters/gps_l1_ca_pcps_quicksync_acquisition.cc:36-135
GpsL1CaPcpsQuickSyncAcquisitionB200::GpsL1CaPcpsQuickSyncAcquisitionB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
    : configuration_(c),
      role_(role),
      item_size_(sizeof(gr_complex)),
      doppler_max_(static_cast<unsigned int>(c->property(role + ".doppler_max", 5000))),
      sampled_ms_(static_cast<unsigned int>(c->property(role + ".coherent_integration_time_ms", 4))),
      bit_transition_flag_(c->property(role + ".bit_transition_flag", false))
{
    item_type_ = c->property(role_ + ".item_type", std::string("gr_complex"));
    const int64_t fs_in_deprecated = c->property("GNSS-SDR.internal_fs_hz", static_cast<int64_t>(2048000));
    fs_in_ = c->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    code_length_ = static_cast<unsigned int>(std::round(static_cast<double>(fs_in_) / (1023000.0 / 1023.0)));
    const auto temp = static_cast<unsigned int>(std::ceil(std::sqrt(std::log2(code_length_))));
    folding_factor_ = c->property(role_ + ".folding_factor", temp);
    if (folding_factor_ == 0U) throw std::invalid_argument("QuickSync acquisition: folding_factor must be positive");
    if (sampled_ms_ % folding_factor_ != 0)
        {
            std::cout << "QuickSync Algorithm requires a coherent_integration_time multiple of " << folding_factor_ << "ms, Value entered " << sampled_ms_ << " ms\n";
            sampled_ms_ = sampled_ms_ < folding_factor_ ? folding_factor_ : (sampled_ms_ / folding_factor_) * folding_factor_;
            std::cout << " Coherent_integration_time of " << sampled_ms_ << " ms will be used instead.\n";
        }
    vector_length_ = code_length_ * sampled_ms_;
    max_dwells_ = bit_transition_flag_ ? 2U : static_cast<unsigned int>(c->property(role_ + ".max_dwells", 1));
    const int samples_per_ms = static_cast<int>(code_length_);
    code_ = std::vector<std::complex<float>>(code_length_);
    if (item_type_ != "gr_complex") throw std::invalid_argument(item_type_ + " unknown acquisition item type (GPS_L1_CA_PCPS_QuickSync_Acquisition_B200 takes gr_complex)");
    acquisition_cc_ = std::make_shared<b200_pcps_quicksync_acquisition>(folding_factor_, sampled_ms_, max_dwells_, doppler_max_, fs_in_, samples_per_ms, static_cast<int32_t>(code_length_),
        bit_transition_flag_, c->property(role + ".device", 0));
    if (in > 1) std::cerr << "This implementation only supports one input stream\n";
    if (out > 0) std::cerr << "This implementation does not provide an output stream\n";
}
// adapters/galileo_e1_pcps_quicksync_ambiguous_acquisition.cc:36-150
GpsL1CaPcpsQuickSyncAcquisitionB200::GpsL1CaPcpsQuickSyncAcquisitionB200(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out, GalileoE1Tag)
    : galileo_(true),
      configuration_(c),
      role_(role),
      item_size_(sizeof(gr_complex)),
      doppler_max_(static_cast<unsigned int>(c->property(role + ".doppler_max", 5000))),
      sampled_ms_(static_cast<unsigned int>(c->property(role + ".coherent_integration_time_ms", 8))),
      bit_transition_flag_(c->property(role + ".bit_transition_flag", false))
{
    item_type_ = c->property(role_ + ".item_type", std::string("gr_complex"));
    const int64_t fs_in_deprecated = c->property("GNSS-SDR.internal_fs_hz", static_cast<int64_t>(4000000));
    fs_in_ = c->property("GNSS-SDR.internal_fs_sps", fs_in_deprecated);
    code_length_ = static_cast<unsigned int>(std::round(static_cast<double>(fs_in_) / (1023000.0 / 4092.0)));  // GALILEO_E1_CODE_CHIP_RATE_CPS / GALILEO_E1_B_CODE_LENGTH_CHIPS: 4 ms
    const auto samples_per_ms = static_cast<int>(std::round(code_length_ / 4.0));
    folding_factor_ = static_cast<unsigned int>(c->property(role + ".folding_factor", 2));
    if (folding_factor_ == 0U) throw std::invalid_argument("QuickSync acquisition: folding_factor must be positive");
    if (sampled_ms_ % (folding_factor_ * 4) != 0)
        {
            std::cout << "QuickSync Algorithm requires a coherent_integration_time multiple of " << (folding_factor_ * 4) << "ms, Value entered " << sampled_ms_ << " ms\n";
            sampled_ms_ = sampled_ms_ < folding_factor_ * 4 ? folding_factor_ * 4 : (sampled_ms_ / (folding_factor_ * 4)) * (folding_factor_ * 4);
            std::cout << "coherent_integration_time = " << sampled_ms_ << " ms will be used.\n";
        }
    vector_length_ = sampled_ms_ * static_cast<unsigned int>(samples_per_ms);
    max_dwells_ = bit_transition_flag_ ? 2U : static_cast<unsigned int>(c->property(role_ + ".max_dwells", 1));
    code_ = std::vector<std::complex<float>>(code_length_);
    if (item_type_ != "gr_complex") throw std::invalid_argument(item_type_ + " unknown acquisition item type (Galileo_E1_PCPS_QuickSync_Ambiguous_Acquisition_B200 takes gr_complex)");
    acquisition_cc_ = std::make_shared<b200_pcps_quicksync_acquisition>(folding_factor_, sampled_ms_, max_dwells_, doppler_max_, fs_in_, samples_per_ms, static_cast<int32_t>(code_length_),
        bit_transition_flag_, c->property(role + ".device", 0));
    acquisition_cc_->set_item_length(static_cast<int>(vector_length_));  // the reference's stream_to_vector (:129-130)
    if (in > 1) std::cerr << "This implementation only supports one input stream\n";
    if (out > 0) std::cerr << "This implementation does not provide an output stream\n";
}
void GpsL1CaPcpsQuickSyncAcquisitionB200::connect(gr::top_block_sptr top_block)
{
    if (top_block) { /* the block cuts its own vectors out of the stream: no stream_to_vector to connect (cf. :294-300) */ }
}
void GpsL1CaPcpsQuickSyncAcquisitionB200::disconnect(gr::top_block_sptr top_block)
{
    if (top_block) { /* nothing to disconnect */ }
}
void GpsL1CaPcpsQuickSyncAcquisitionB200::stop_acquisition()
{
    acquisition_cc_->set_state(0);
    acquisition_cc_->set_active(false);
}
// :154-176
void GpsL1CaPcpsQuickSyncAcquisitionB200::set_threshold(float threshold)
{
    float pfa = configuration_->property(role_ + std::to_string(channel_) + ".pfa", static_cast<float>(0.0));
    if (pfa == 0.0) pfa = configuration_->property(role_ + ".pfa", static_cast<float>(0.0));
    threshold_ = pfa == 0.0 ? threshold : calculate_threshold(pfa);
    acquisition_cc_->set_threshold(threshold_);
}
// :274-291: the (1 - pfa)^(1 / cells) quantile of an exponential distribution of rate code_length / folding_factor, cells counted on the folded grid
float GpsL1CaPcpsQuickSyncAcquisitionB200::calculate_threshold(float pfa) const
{
    unsigned int frequency_bins = 0;
    for (int doppler = static_cast<int>(-doppler_max_); doppler <= static_cast<int>(doppler_max_); doppler += static_cast<int>(doppler_step_)) frequency_bins++;
    const unsigned int ncells = (code_length_ / folding_factor_) * frequency_bins;
    const double exponent = 1.0 / static_cast<double>(ncells);
    const double val = std::pow(1.0 - pfa, exponent);
    const double lambda = static_cast<double>(code_length_) / static_cast<double>(folding_factor_);
    return static_cast<float>(-std::log1p(-val) / lambda);
}
void GpsL1CaPcpsQuickSyncAcquisitionB200::set_doppler_max(unsigned int doppler_max)
{
    doppler_max_ = doppler_max;
    acquisition_cc_->set_doppler_max(doppler_max_);
}
void GpsL1CaPcpsQuickSyncAcquisitionB200::set_doppler_step(unsigned int doppler_step)
{
    doppler_step_ = doppler_step;
    acquisition_cc_->set_doppler_step(doppler_step_);
}
void GpsL1CaPcpsQuickSyncAcquisitionB200::set_gnss_synchro(Gnss_Synchro* gnss_synchro)
{
    gnss_synchro_ = gnss_synchro;
    acquisition_cc_->set_gnss_synchro(gnss_synchro_);
}
void GpsL1CaPcpsQuickSyncAcquisitionB200::set_channel(unsigned int channel)
{
    channel_ = channel;
    acquisition_cc_->set_channel(channel_);
}
// :228-244.  The reference copies sampled_ms / folding_factor periods into a buffer of ONE period (it overruns it when that quotient
// exceeds 1); the block reads one period only (pcps_quicksync_acquisition_cc.cc:137), which is what is built here.
void GpsL1CaPcpsQuickSyncAcquisitionB200::set_local_code()
{
    if (galileo_)
        {
            // galileo_e1_pcps_quicksync_ambiguous_acquisition.cc:234-258 (one period; the reference's extra copies overrun its one-period buffer)
            const bool cboc = configuration_->property("Acquisition" + std::to_string(channel_) + ".cboc", false);
            const char sig[3] = {gnss_synchro_->Signal[0], gnss_synchro_->Signal[1], '\0'};
            galileo_e1_code_gen_complex_sampled(code_.data(), sig, cboc, gnss_synchro_->PRN, static_cast<int32_t>(fs_in_), 0, false);
            acquisition_cc_->set_local_code(code_.data());
            return;
        }
    gps_l1_ca_code_gen_complex_sampled(code_.data(), static_cast<int>(code_.size()), gnss_synchro_->PRN, static_cast<int32_t>(fs_in_), 0);
    acquisition_cc_->set_local_code(code_.data());
}

std::unique_ptr<AcquisitionInterface> GetB200AcqBlock(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
{
    const std::string impl = c->property(role + ".implementation", std::string("Pass_Through"));
    if (impl == "GPS_L1_CA_PCPS_Acquisition_B200") return std::make_unique<GpsL1CaPcpsAcquisitionB200>(c, role, in, out);
    if (impl == "Galileo_E1_PCPS_Ambiguous_Acquisition_B200") return std::make_unique<GalileoE1PcpsAmbiguousAcquisitionB200>(c, role, in, out);
    if (impl == "GPS_L5i_PCPS_Acquisition_B200") return std::make_unique<GpsL5iPcpsAcquisitionB200>(c, role, in, out);
    if (impl == "Galileo_E5a_Pcps_Acquisition_B200") return std::make_unique<GalileoE5aPcpsAcquisitionB200>(c, role, in, out);
    if (impl == "Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200") return std::make_unique<GalileoE5aNoncoherentIQAcquisitionCafB200>(c, role, in, out);
    if (impl == "Galileo_E1_PCPS_QuickSync_Ambiguous_Acquisition_B200") return std::make_unique<GalileoE1PcpsQuickSyncAmbiguousAcquisitionB200>(c, role, in, out);
    if (impl == "GPS_L1_CA_PCPS_QuickSync_Acquisition_B200") return std::make_unique<GpsL1CaPcpsQuickSyncAcquisitionB200>(c, role, in, out);
    if (impl == "GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200") return std::make_unique<GpsL1CaPcpsAcquisitionFineDopplerB200>(c, role, in, out);
    return nullptr;
}

std::unique_ptr<GNSSBlockInterface> GetB200Block(const ConfigurationInterface* c, const std::string& role, unsigned int in, unsigned int out)
{
    if (auto t = GetB200TrkBlock(c, role, in, out)) return t;
    if (auto a = GetB200AcqBlock(c, role, in, out)) return a;
    return nullptr;
}
}  // namespace gb200h
