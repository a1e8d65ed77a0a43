#include "b200_conf.h"
#include "gnss_b200.h"
#include <cmath>
#include <iostream>
#include <stdexcept>

namespace gb200h
{
size_t item_type_size(const std::string& t)
{
    if (t == "byte") return 1;
    if (t == "ibyte" || t == "cbyte") return 2;
    if (t == "short") return 2;
    if (t == "ishort" || t == "cshort") return 4;
    if (t == "float") return 4;
    if (t == "gr_complex") return 8;
    return 0;
}

int item_type_iq(const std::string& t)
{
    if (t == "gr_complex") return GB200_IQ_CF32;
    if (t == "cshort" || t == "ishort") return GB200_IQ_CS16;
    if (t == "cbyte" || t == "ibyte") return GB200_IQ_CS8;
    return -1;
}

namespace
{
void warn(const std::string& m) { std::cerr << "WARNING: " << m << '\n'; }
template <typename T>
T clampi(T v, T lo, T hi, const char* what)
{
    if (v < lo)
        {
            warn(std::string(what) + " below its minimum; set to " + std::to_string(lo));
            return lo;
        }
    if (v > hi)
        {
            warn(std::string(what) + " above its maximum; set to " + std::to_string(hi));
            return hi;
        }
    return v;
}
}  // namespace

void Dll_Pll_Conf::SetFromConfiguration(const ConfigurationInterface* c, const std::string& role)
{
    item_type = c->property(role + ".item_type", item_type);
    if (item_type_size(item_type) == 0)
        {
            warn("Unknown item type: " + item_type + ". Set to gr_complex");
            item_type = "gr_complex";
        }
    const double fs_deprecated = c->property("GNSS-SDR.internal_fs_hz", fs_in);
    fs_in = c->property("GNSS-SDR.internal_fs_sps", fs_deprecated);
    high_dyn = c->property(role + ".high_dyn", high_dyn);
    dump = c->property(role + ".dump", dump);
    dump_filename = c->property(role + ".dump_filename", dump_filename);
    dump_mat = c->property(role + ".dump_mat", dump_mat);
    pll_bw_hz = c->property(role + ".pll_bw_hz", pll_bw_hz);
    pll_bw_narrow_hz = c->property(role + ".pll_bw_narrow_hz", pll_bw_narrow_hz);
    dll_bw_narrow_hz = c->property(role + ".dll_bw_narrow_hz", dll_bw_narrow_hz);
    dll_bw_hz = c->property(role + ".dll_bw_hz", dll_bw_hz);
    dll_filter_order = clampi(c->property(role + ".dll_filter_order", dll_filter_order), 1, 3, "dll_filter_order");
    pll_filter_order = clampi(c->property(role + ".pll_filter_order", pll_filter_order), 2, 3, "pll_filter_order");
    fll_filter_order = pll_filter_order == 2 ? 1 : 2;
    enable_fll_pull_in = c->property(role + ".enable_fll_pull_in", enable_fll_pull_in);
    enable_fll_steady_state = c->property(role + ".enable_fll_steady_state", enable_fll_steady_state);
    fll_bw_hz = c->property(role + ".fll_bw_hz", fll_bw_hz);
    pull_in_time_s = c->property(role + ".pull_in_time_s", pull_in_time_s);
    bit_synchronization_time_limit_s = c->property(role + ".bit_synchronization_time_limit_s", bit_synchronization_time_limit_s);
    early_late_space_chips = c->property(role + ".early_late_space_chips", early_late_space_chips);
    early_late_space_narrow_chips = c->property(role + ".early_late_space_narrow_chips", early_late_space_narrow_chips);
    very_early_late_space_chips = c->property(role + ".very_early_late_space_chips", very_early_late_space_chips);
    very_early_late_space_narrow_chips = c->property(role + ".very_early_late_space_narrow_chips", very_early_late_space_narrow_chips);
    extend_correlation_symbols = c->property(role + ".extend_correlation_symbols", extend_correlation_symbols);
    track_pilot = c->property(role + ".track_pilot", track_pilot);
    cn0_samples = c->property(role + ".cn0_samples", cn0_samples);
    cn0_min = c->property(role + ".cn0_min", cn0_min);
    max_code_lock_fail = c->property(role + ".max_lock_fail", max_code_lock_fail);
    max_carrier_lock_fail = c->property(role + ".max_carrier_lock_fail", max_carrier_lock_fail);
    carrier_lock_th = c->property(role + ".carrier_lock_th", carrier_lock_th);
    carrier_aiding = c->property(role + ".carrier_aiding", carrier_aiding);
    cn0_smoother_samples = c->property(role + ".cn0_smoother_samples", cn0_smoother_samples);
    cn0_smoother_alpha = c->property(role + ".cn0_smoother_alpha", cn0_smoother_alpha);
    smoother_length = c->property(role + ".smoother_length", smoother_length);
    if (smoother_length < 1)
        {
            smoother_length = 1;
            warn("smoother_length must be bigger than 0. It has been set to 1");
        }
    carrier_lock_test_smoother_samples = c->property(role + ".carrier_lock_test_smoother_samples", carrier_lock_test_smoother_samples);
    carrier_lock_test_smoother_alpha = c->property(role + ".carrier_lock_test_smoother_alpha", carrier_lock_test_smoother_alpha);
    device = c->property(role + ".device", device);
    stream_id = c->property(role + ".stream_id", stream_id);
}

void Acq_Conf::SetFromConfiguration(const ConfigurationInterface* c, const std::string& role, double chip_rate, double opt_freq)
{
    item_type = c->property(role + ".item_type", item_type);
    if (item_type_size(item_type) == 0) throw std::invalid_argument("Unknown item type: " + item_type);
    chips_per_second = static_cast<uint32_t>(chip_rate);
    const int64_t fs_deprecated = c->property("GNSS-SDR.internal_fs_hz", fs_in);
    fs_in = c->property("GNSS-SDR.internal_fs_sps", fs_deprecated);
    doppler_max = c->property(role + ".doppler_max", doppler_max);
    sampled_ms = c->property(role + ".coherent_integration_time_ms", sampled_ms);
    bit_transition_flag = c->property(role + ".bit_transition_flag", bit_transition_flag);
    max_dwells = c->property(role + ".max_dwells", max_dwells);
    dump = c->property(role + ".dump", dump);
    dump_channel = c->property(role + ".dump_channel", dump_channel);
    blocking = c->property(role + ".blocking", blocking);
    dump_filename = c->property(role + ".dump_filename", dump_filename);
    use_automatic_resampler = c->property("GNSS-SDR.use_acquisition_resampler", use_automatic_resampler);
    if ((sampled_ms % ms_per_code) != 0)
        {
            warn("Parameter coherent_integration_time_ms should be a multiple of " + std::to_string(ms_per_code) + ". Setting it to " + std::to_string(ms_per_code));
            sampled_ms = ms_per_code;
        }
    resampled_fs = fs_in;
    if (use_automatic_resampler) ConfigureAutomaticResampler(opt_freq);
    it_size = item_type_size(item_type);
    num_doppler_bins_step2 = c->property(role + ".second_nbins", num_doppler_bins_step2);
    doppler_step2 = c->property(role + ".second_doppler_step", doppler_step2);
    doppler_step = c->property(role + ".doppler_step", doppler_step);
    pfa = c->property(role + ".pfa", pfa);
    if (pfa < 0.0F || pfa > 1.0F)
        {
            warn("Parameter pfa should between 0.0 and 1.0. Setting it to 0.0");
            pfa = 0.0F;
        }
    pfa2 = c->property(role + ".pfa_second_step", pfa2);
    if (pfa2 <= 0.0F || pfa2 > 1.0F) pfa2 = pfa;
    make_2_steps = c->property(role + ".make_two_steps", make_2_steps);
    blocking_on_standby = c->property(role + ".blocking_on_standby", blocking_on_standby);
    if (pfa <= 0.0F) use_CFAR_algorithm_flag = false;
    enable_monitor_output = c->property("AcquisitionMonitor.enable_monitor", false);
    device = c->property(role + ".device", device);
    SetDerivedParams();
}

void Acq_Conf::ConfigureAutomaticResampler(double opt_freq)
{
    if (!use_automatic_resampler) return;
    if (fs_in > opt_freq)
        {
            auto decimation = static_cast<uint32_t>(fs_in / opt_freq);
            while (fs_in % decimation > 0) decimation--;
            resampler_ratio = static_cast<float>(decimation);
            resampled_fs = fs_in / static_cast<int>(resampler_ratio);
        }
    SetDerivedParams();
}

void Acq_Conf::SetDerivedParams()
{
    samples_per_ms = static_cast<float>(resampled_fs) * 0.001F;
    samples_per_chip = static_cast<unsigned int>(std::ceil(static_cast<float>(resampled_fs) / chips_per_second));
    samples_per_code = samples_per_ms * ms_per_code;
}
}  // namespace gb200h
