// Parameter bundles of the two blocks, read from the same .conf keys with the same defaults and clamps as the reference:
//   Dll_Pll_Conf  src/algorithms/tracking/libs/dll_pll_conf.{h:30-88,cc:36-125}
//   Acq_Conf      src/algorithms/acquisition/libs/acq_conf.{h:34-79,cc:24-118}
// plus keys of our own: `<role>.device` (CUDA ordinal, default 0), `<role>.n_correlators` (extended-correlator taps) and
// `<role>.stream_id` (which input stream the channel reads; default -1 = by band: L1 signals share stream 0, L5 / E5a stream 1).
#ifndef GB200_CONF_H
#define GB200_CONF_H

#include "gnss_sdr_compat.h"
#include <cstdint>
#include <string>

namespace gb200h
{
class Dll_Pll_Conf
{
public:
    Dll_Pll_Conf() = default;
    void SetFromConfiguration(const ConfigurationInterface* configuration, const std::string& role);

    std::string item_type{"gr_complex"};
    std::string dump_filename{"./dll_pll_dump.dat"};
    double fs_in{2000000.0};
    double carrier_lock_th{0.0};
    float pll_pull_in_bw_hz{50.0};
    float dll_pull_in_bw_hz{3.0};
    float fll_bw_hz{35.0};
    float pll_bw_hz{35.0};
    float dll_bw_hz{2.0};
    float pll_bw_narrow_hz{5.0};
    float dll_bw_narrow_hz{0.75};
    float early_late_space_chips{0.25};
    float very_early_late_space_chips{0.5};
    float early_late_space_narrow_chips{0.15};
    float very_early_late_space_narrow_chips{0.5};
    float slope{1.0};
    float spc{0.5};
    float y_intercept{1.0};
    float cn0_smoother_alpha{0.002};
    float carrier_lock_test_smoother_alpha{0.002};
    uint32_t pull_in_time_s{10U};
    uint32_t bit_synchronization_time_limit_s{20U};
    uint32_t vector_length{0U};
    uint32_t smoother_length{10U};
    int32_t fll_filter_order{1};
    int32_t pll_filter_order{3};
    int32_t dll_filter_order{2};
    int32_t extend_correlation_symbols{1};
    int32_t cn0_samples{0};
    int32_t cn0_smoother_samples{200};
    int32_t carrier_lock_test_smoother_samples{25};
    int32_t cn0_min{0};
    int32_t max_code_lock_fail{0};
    int32_t max_carrier_lock_fail{0};
    char signal[3]{};
    char system{'G'};
    bool enable_fll_pull_in{false};
    bool enable_fll_steady_state{false};
    bool track_pilot{true};
    bool enable_doppler_correction{false};
    bool carrier_aiding{true};
    bool high_dyn{false};
    bool dump{false};
    bool dump_mat{true};
    // B200 additions
    int32_t device{0};
    int32_t stream_id{-1};
};

class Acq_Conf
{
public:
    Acq_Conf() = default;
    // throws std::invalid_argument on an unknown item type, like the reference (acq_conf.cc:28-31)
    void SetFromConfiguration(const ConfigurationInterface* configuration, const std::string& role, double chip_rate, double opt_freq);
    void SetDerivedParams();

    std::string item_type{"gr_complex"};
    std::string dump_filename;
    int64_t fs_in{4000000LL};
    int64_t resampled_fs{0LL};
    size_t it_size{sizeof(gr_complex)};
    float doppler_step{250.0};
    float samples_per_ms{0.0};
    float doppler_step2{125.0};
    float pfa{0.0};
    float pfa2{0.0};
    float samples_per_code{0.0};
    float resampler_ratio{1.0};
    uint32_t sampled_ms{1U};
    uint32_t ms_per_code{1U};
    uint32_t samples_per_chip{2U};
    uint32_t chips_per_second{1023000U};
    uint32_t max_dwells{1U};
    uint32_t num_doppler_bins_step2{4U};
    uint32_t resampler_latency_samples{0U};
    uint32_t dump_channel{0U};
    int32_t doppler_max{5000};
    int32_t doppler_min{-5000};
    bool bit_transition_flag{false};
    bool use_CFAR_algorithm_flag{true};
    bool dump{false};
    bool blocking{true};
    bool blocking_on_standby{false};
    bool make_2_steps{false};
    bool use_automatic_resampler{false};
    bool enable_monitor_output{false};
    // B200 additions
    int32_t device{0};

private:
    void ConfigureAutomaticResampler(double opt_freq);
};

// item types the blocks accept (src/algorithms/libs/item_type_helpers.cc): size in bytes, 0 = unknown
size_t item_type_size(const std::string& item_type);
// GB200_IQ_* of an item type, -1 when the engines cannot ingest it
int item_type_iq(const std::string& item_type);
}  // namespace gb200h
#endif
