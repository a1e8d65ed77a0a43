#include "b200_dll_pll_veml_tracking.h"
#include "b200_gnss_codes.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <iostream>
#include <numeric>
#include <stdexcept>

namespace gb200h
{
namespace
{
constexpr double TWO_PI_D = 2.0 * 3.1415926535898;  // TWO_PI of MATH_CONSTANTS.h:47-49 (IS-GPS-200 pi), as the reference's loop uses

// GPS L1 C/A telemetry preamble 10001011, each bit held for 20 code periods (GPS_L1_CA.h:73)
std::string gps_ca_preamble_symbols()
{
    std::string s;
    for (char b : std::string("10001011")) s.append(20, b);
    return s;
}

}  // namespace

// Signal table: dll_pll_veml_tracking.cc:171-447 (the signals of BASELINE configs C1-C5; others are rejected)
b200_dll_pll_veml_tracking::b200_dll_pll_veml_tracking(const Dll_Pll_Conf& conf)
    : d_trk_parameters(conf), d_carrier_lock_threshold(conf.carrier_lock_th), d_current_prn_length_samples(static_cast<int32_t>(conf.vector_length)),
      d_extend_correlation_symbols(conf.extend_correlation_symbols), d_dump(conf.dump)
{
    d_signal_type = std::string(d_trk_parameters.signal);
    Dll_Pll_Conf& p = d_trk_parameters;
    const auto plain = [&] {  // BPSK-like signals: unit slope, E-L spacing as configured
        p.slope = 1.0F;
        p.spc = p.early_late_space_chips;
        p.y_intercept = 1.0F;
    };
    if (p.system == 'G' && d_signal_type == "1C")
        {
            d_systemName = "GPS";
            d_signal_carrier_freq = 1575.42e6;
            d_code_period = 0.001;
            d_code_chip_rate = 1.023e6;
            d_correlation_length_ms = 1;
            d_code_samples_per_chip = 1;
            d_code_length_chips = 1023;
            d_secondary = false;
            p.track_pilot = false;  // no pilot component
            plain();
            // bit synchronisation by correlating prompt signs with the preamble pattern (20 symbols per bit)
            d_secondary_code_string = gps_ca_preamble_symbols();
            d_secondary_code_length = 160;
            d_symbols_per_bit = 20;
        }
    else if (p.system == 'G' && d_signal_type == "L5")
        {
            d_systemName = "GPS";
            d_signal_carrier_freq = 1176.45e6;
            d_code_period = 0.001;
            d_code_chip_rate = 10.23e6;
            d_symbols_per_bit = 10;
            d_correlation_length_ms = 1;
            d_code_samples_per_chip = 1;
            d_code_length_chips = 10230;
            d_secondary = true;
            plain();
            if (p.track_pilot)
                {
                    d_secondary_code_length = 20;
                    d_secondary_code_string = "00000100110101001110";  // L5Q Neuman-Hofman
                    d_data_secondary_code_length = 10;
                    d_data_secondary_code_string = "0000110101";  // L5I Neuman-Hofman
                }
            else
                {
                    d_secondary_code_length = 10;
                    d_secondary_code_string = "0000110101";
                    d_interchange_iq = true;
                }
        }
    else if (p.system == 'E' && d_signal_type == "1B")
        {
            d_systemName = "Galileo";
            d_signal_carrier_freq = 1575.42e6;
            d_code_period = 0.004;
            d_code_chip_rate = 1.023e6;
            d_code_length_chips = 4092;
            d_symbols_per_bit = 1;
            d_correlation_length_ms = 4;
            d_code_samples_per_chip = 2;  // sinBOC(1,1) table, 2 entries per chip
            d_veml = true;
            p.spc = p.early_late_space_chips;
            p.slope = static_cast<float>(-sinboc11_slope_abs(p.spc));
            p.y_intercept = static_cast<float>(sinboc11_y_intercept_abs(p.spc));
            if (p.track_pilot)
                {
                    d_secondary = true;
                    d_secondary_code_length = 25;
                    d_secondary_code_string = "0011100000001010110110010";  // E1-C secondary code
                }
            else
                {
                    d_secondary = false;
                }
        }
    else if (p.system == 'E' && d_signal_type == "5X")
        {
            d_systemName = "Galileo";
            d_signal_carrier_freq = 1176.45e6;
            d_code_period = 0.001;
            d_code_chip_rate = 10.23e6;
            d_symbols_per_bit = 20;
            d_correlation_length_ms = 1;
            d_code_samples_per_chip = 1;
            d_code_length_chips = 10230;
            d_secondary = true;
            plain();
            if (p.track_pilot)
                {
                    d_secondary_code_length = 100;  // E5a-Q secondary code is per PRN: supplied through set_secondary_code
                    d_data_secondary_code_length = 20;
                    d_data_secondary_code_string = "10000100001011101001";
                    d_interchange_iq = true;
                }
            else
                {
                    d_secondary_code_length = 20;
                    d_secondary_code_string = "10000100001011101001";
                }
        }
    else
        {
            std::cerr << "Invalid System/Signal argument when instantiating tracking blocks\n";
            d_correlation_length_ms = 1;
            d_secondary = false;
        }

    d_code_freq_chips = d_code_chip_rate;
    d_code_loop_filter = Tracking_loop_filter(static_cast<float>(d_code_period), p.dll_bw_hz, p.dll_filter_order, false);
    d_carrier_loop_filter.set_params(p.fll_bw_hz, p.pll_bw_hz, p.pll_filter_order);

    d_tracking_code.assign(2 * std::max(d_code_length_chips, 1), 0.0F);
    d_n_correlator_taps = d_veml ? 5 : 3;
    d_correlator_outs.assign(d_n_correlator_taps, gr_complex(0, 0));
    d_local_code_shift_chips.assign(d_n_correlator_taps, 0.0F);
    if (d_veml)
        {
            idx_VE_ = 0;
            idx_E_ = 1;
            idx_P_ = 2;
            idx_L_ = 3;
            idx_VL_ = 4;
        }
    set_wide_taps();

    if (p.extend_correlation_symbols > 1)
        d_enable_extended_integration = true;
    else
        {
            d_enable_extended_integration = false;
            p.extend_correlation_symbols = 1;
        }
    if (p.track_pilot) d_data_code.assign(2 * std::max(d_code_length_chips, 1), 0.0F);

    d_Prompt_buffer.assign(std::max(p.cn0_samples, 1), gr_complex(0, 0));
    d_cn0_smoother.set_alpha(p.cn0_smoother_alpha);
    if (d_code_period > 0.0) d_cn0_smoother.set_samples_for_initialization(p.cn0_smoother_samples / static_cast<int>(d_code_period * 1000.0));
    d_carrier_lock_test_smoother.set_alpha(p.carrier_lock_test_smoother_alpha);
    d_carrier_lock_test_smoother.set_min_value(-1.0F);
    d_carrier_lock_test_smoother.set_offset(0.0F);
    d_carrier_lock_test_smoother.set_samples_for_initialization(p.carrier_lock_test_smoother_samples);
    clear_tracking_vars();
}

b200_dll_pll_veml_tracking::~b200_dll_pll_veml_tracking()
{
    if (d_dump_file.is_open()) d_dump_file.close();
}

void b200_dll_pll_veml_tracking::set_wide_taps()
{
    const Dll_Pll_Conf& p = d_trk_parameters;
    const auto spc = static_cast<float>(d_code_samples_per_chip);
    if (d_veml)
        {
            d_local_code_shift_chips[0] = -p.very_early_late_space_chips * spc;
            d_local_code_shift_chips[1] = -p.early_late_space_chips * spc;
            d_local_code_shift_chips[2] = 0.0F;
            d_local_code_shift_chips[3] = p.early_late_space_chips * spc;
            d_local_code_shift_chips[4] = p.very_early_late_space_chips * spc;
        }
    else
        {
            d_local_code_shift_chips[0] = -p.early_late_space_chips * spc;
            d_local_code_shift_chips[1] = 0.0F;
            d_local_code_shift_chips[2] = p.early_late_space_chips * spc;
        }
}

void b200_dll_pll_veml_tracking::set_narrow_taps()
{
    Dll_Pll_Conf& p = d_trk_parameters;
    const auto spc = static_cast<float>(d_code_samples_per_chip);
    if (d_veml)
        {
            d_local_code_shift_chips[0] = -p.very_early_late_space_narrow_chips * spc;
            d_local_code_shift_chips[1] = -p.early_late_space_narrow_chips * spc;
            d_local_code_shift_chips[3] = p.early_late_space_narrow_chips * spc;
            d_local_code_shift_chips[4] = p.very_early_late_space_narrow_chips * spc;
        }
    else
        {
            d_local_code_shift_chips[0] = -p.early_late_space_narrow_chips * spc;
            d_local_code_shift_chips[2] = p.early_late_space_narrow_chips * spc;
        }
    p.spc = p.early_late_space_narrow_chips;
}

// dll_pll_veml_tracking.cc:1675-1705: the dump file is <dump_filename><channel>.dat
void b200_dll_pll_veml_tracking::set_channel(uint32_t channel)
{
    d_channel = channel;
    if (d_dump && !d_dump_file.is_open())
        {
            std::string base = d_trk_parameters.dump_filename;
            const size_t slash = base.find_last_of('/');
            const size_t dot = base.find_last_of('.');
            if (dot != std::string::npos && dot > 0 && (slash == std::string::npos || dot > slash + 1)) base = base.substr(0, dot);  // drop the extension
            d_dump_file.open(base + std::to_string(d_channel) + ".dat", std::ios::out | std::ios::binary);
            if (!d_dump_file.is_open()) std::cerr << "channel " << d_channel << ": cannot open the tracking dump file " << base << '\n';
        }
}
void b200_dll_pll_veml_tracking::set_gnss_synchro(Gnss_Synchro* p) { d_acquisition_gnss_synchro = p; }
void b200_dll_pll_veml_tracking::stop_tracking() { d_state = 0; }

void b200_dll_pll_veml_tracking::set_tracking_code(const float* code, int n)
{
    d_tracking_code.assign(code, code + n);
    tracking_code_supplied_ = true;
}

void b200_dll_pll_veml_tracking::set_data_code(const float* code, int n)
{
    d_data_code.assign(code, code + n);
    data_code_supplied_ = true;
}

void b200_dll_pll_veml_tracking::set_engine(std::shared_ptr<B200CorrelatorEngine> engine)
{
    engine_ = std::move(engine);
    const auto refuse = [&](const char* which) {
        const EngineConfig& ec = engine_->config();
        throw std::runtime_error(std::string("b200_dll_pll_veml_tracking: ") + which + " cannot attach to the correlator engine of device " + std::to_string(ec.device) +
                                 " stream " + std::to_string(ec.stream_id) + " (engine: windows up to " + std::to_string(ec.max_block) + " samples, " +
                                 std::to_string(ec.max_channels) + " channel slots; this channel: 2 x " + std::to_string(d_trk_parameters.vector_length) + " samples, " +
                                 std::to_string(d_n_correlator_taps) + " taps)");
    };
    d_multicorrelator.set_engine(engine_);
    d_multicorrelator.set_high_dynamics_resampler(d_trk_parameters.high_dyn);
    if (!d_multicorrelator.init(static_cast<int>(2 * d_trk_parameters.vector_length), d_n_correlator_taps)) refuse("the multicorrelator");
    if (d_trk_parameters.track_pilot)
        {
            d_correlator_data.set_engine(engine_);
            d_correlator_data.set_high_dynamics_resampler(d_trk_parameters.high_dyn);
            if (!d_correlator_data.init(static_cast<int>(2 * d_trk_parameters.vector_length), 1)) refuse("the data-component correlator");
        }
}

void b200_dll_pll_veml_tracking::upload_codes()
{
    if (!engine_) return;
    bool ok = d_multicorrelator.set_local_code_and_taps(code_table_entries(), d_tracking_code.data(), d_local_code_shift_chips.data());
    if (ok && d_trk_parameters.track_pilot) ok = d_correlator_data.set_local_code_and_taps(code_table_entries(), d_data_code.data(), &d_local_code_shift_chips[idx_P_]);
    if (!ok) throw std::runtime_error("b200_dll_pll_veml_tracking: set_local_code_and_taps before a successful init() (channel not attached to the engine)");
}

// dll_pll_veml_tracking.cc:643-883
void b200_dll_pll_veml_tracking::start_tracking()
{
    const Dll_Pll_Conf& p = d_trk_parameters;
    d_acq_code_phase_samples = d_acquisition_gnss_synchro->Acq_delay_samples;
    d_acq_carrier_doppler_hz = d_acquisition_gnss_synchro->Acq_doppler_hz;
    d_acq_sample_stamp = d_acquisition_gnss_synchro->Acq_samplestamp_samples;

    d_carrier_doppler_hz = d_acq_carrier_doppler_hz;
    d_carrier_phase_step_rad = TWO_PI_D * d_carrier_doppler_hz / p.fs_in;
    d_carrier_phase_rate_step_rad = 0.0;
    d_carr_ph_history.clear();
    d_code_ph_history.clear();
    d_extend_correlation_symbols = p.extend_correlation_symbols;

    if (d_systemName == "GPS" && d_signal_type == "1C" && !tracking_code_supplied_)
        {
            d_tracking_code.assign(2 * d_code_length_chips, 0.0F);
            gps_l1_ca_code_gen_float(d_tracking_code.data(), d_acquisition_gnss_synchro->PRN, 0);
        }
    else if (!tracking_code_supplied_)
        {
            // dll_pll_veml_tracking.cc:670-722: the other signals' tables, same generators (b200_gnss_codes.h)
            const uint32_t prn = d_acquisition_gnss_synchro->PRN;
            const size_t entries = static_cast<size_t>(d_code_samples_per_chip) * d_code_length_chips;
            if (d_systemName == "GPS" && d_signal_type == "L5")
                {
                    d_tracking_code.assign(2 * entries, 0.0F);
                    if (p.track_pilot)
                        {
                            d_data_code.assign(2 * entries, 0.0F);
                            gps_l5q_code_gen_float(d_tracking_code.data(), prn);
                            gps_l5i_code_gen_float(d_data_code.data(), prn);
                        }
                    else
                        gps_l5i_code_gen_float(d_tracking_code.data(), prn);
                }
            else if (d_systemName == "Galileo" && d_signal_type == "1B")
                {
                    d_tracking_code.assign(2 * entries, 0.0F);
                    if (p.track_pilot)
                        {
                            d_data_code.assign(2 * entries, 0.0F);
                            galileo_e1_code_gen_sinboc11_float(d_tracking_code.data(), "1C", prn);
                            galileo_e1_code_gen_sinboc11_float(d_data_code.data(), "1B", prn);
                        }
                    else
                        galileo_e1_code_gen_sinboc11_float(d_tracking_code.data(), "1B", prn);
                }
            else if (d_systemName == "Galileo" && d_signal_type == "5X")
                {
                    std::vector<std::complex<float>> aux(d_code_length_chips);
                    galileo_e5_a_code_gen_complex_primary(aux.data(), static_cast<int32_t>(prn), "5X");
                    d_tracking_code.assign(2 * entries, 0.0F);
                    if (p.track_pilot)
                        {
                            d_secondary_code_string = galileo_e5a_q_secondary_code(prn);
                            d_data_code.assign(2 * entries, 0.0F);
                            for (size_t i = 0; i < entries; i++)
                                {
                                    d_tracking_code[i] = aux[i].imag();
                                    d_data_code[i] = aux[i].real();
                                }
                        }
                    else
                        for (size_t i = 0; i < entries; i++) d_tracking_code[i] = aux[i].real();
                }
        }
    // tables supplied through set_tracking_code / set_data_code take precedence (tests, other signals)
    if (p.track_pilot) d_Prompt_Data = gr_complex(0.0, 0.0);
    upload_codes();
    std::fill(d_correlator_outs.begin(), d_correlator_outs.end(), gr_complex(0.0, 0.0));

    d_carrier_lock_fail_counter = 0;
    d_code_lock_fail_counter = 0;
    d_rem_code_phase_samples = 0.0;
    d_rem_carr_phase_rad = 0.0;
    d_rem_code_phase_chips = 0.0;
    d_acc_carrier_phase_rad = 0.0;
    d_cn0_estimation_counter = 0;
    d_carrier_lock_test = 1.0;
    d_CN0_SNV_dB_Hz = 0.0;
    set_wide_taps();
    d_current_correlation_time_s = d_code_period;

    d_carrier_loop_filter.set_params(p.fll_bw_hz, p.pll_bw_hz, p.pll_filter_order);
    d_code_loop_filter.set_noise_bandwidth(p.dll_bw_hz);
    d_code_loop_filter.set_update_interval(static_cast<float>(d_code_period));
    d_carrier_loop_filter.initialize(static_cast<float>(d_acq_carrier_doppler_hz));
    d_code_loop_filter.initialize();

    d_state = 1;
    pending_skip_ = 0;
    d_cloop = true;
    d_pull_in_transitory = true;
    d_Prompt_circular_buffer.clear();
    d_corrected_doppler = false;
    d_acc_carrier_phase_initialized = false;
}

// dll_pll_veml_tracking.cc:926-969: sign pattern of the last `length` prompts against the secondary code / preamble
bool b200_dll_pll_veml_tracking::acquire_secondary()
{
    int32_t corr_value = 0;
    for (uint32_t i = 0; i < d_secondary_code_length; i++)
        {
            const bool negative = d_Prompt_circular_buffer[i].real() < 0.0;
            const bool zero_chip = d_secondary_code_string[i] == '0';
            corr_value += (negative == zero_chip) ? 1 : -1;
        }
    if (std::abs(corr_value) == static_cast<int32_t>(d_secondary_code_length))
        {
            d_Flag_PLL_180_deg_phase_locked = corr_value < 0;
            return true;
        }
    return false;
}

// dll_pll_veml_tracking.cc:972-1029
bool b200_dll_pll_veml_tracking::cn0_and_tracking_lock_status(double coh_integration_time_s)
{
    const Dll_Pll_Conf& p = d_trk_parameters;
    if (d_cn0_estimation_counter < p.cn0_samples)
        {
            d_Prompt_buffer[d_cn0_estimation_counter] = d_P_accu;  // still filling the estimator window
            d_cn0_estimation_counter++;
            return true;
        }
    d_Prompt_buffer[d_cn0_estimation_counter % p.cn0_samples] = d_P_accu;
    d_cn0_estimation_counter++;
    const float cn0_raw = cn0_m2m4_estimator(d_Prompt_buffer.data(), p.cn0_samples, static_cast<float>(coh_integration_time_s));
    d_CN0_SNV_dB_Hz = d_cn0_smoother.smooth(cn0_raw);
    d_carrier_lock_test = d_carrier_lock_test_smoother.smooth(carrier_lock_detector(d_Prompt_buffer.data(), 1));
    if (!d_pull_in_transitory)
        {
            if (d_carrier_lock_test < d_carrier_lock_threshold)
                d_carrier_lock_fail_counter++;
            else if (d_carrier_lock_fail_counter > 0)
                d_carrier_lock_fail_counter--;
            if (d_CN0_SNV_dB_Hz < p.cn0_min)
                d_code_lock_fail_counter++;
            else if (d_code_lock_fail_counter > 0)
                d_code_lock_fail_counter--;
        }
    if (d_carrier_lock_fail_counter > p.max_carrier_lock_fail || d_code_lock_fail_counter > p.max_code_lock_fail)
        {
            std::cout << "Loss of lock in channel " << d_channel << "!\n";
            if (event_sink_) event_sink_(3);  // 3 -> loss of lock
            d_carrier_lock_fail_counter = 0;
            d_code_lock_fail_counter = 0;
            return false;
        }
    return true;
}

// dll_pll_veml_tracking.cc:1065-1152
void b200_dll_pll_veml_tracking::run_dll_pll()
{
    const Dll_Pll_Conf& p = d_trk_parameters;
    // PLL discriminator: Costas (insensitive to bit flips) until the secondary code / pilot removes the transitions
    d_carr_phase_error_hz = (d_cloop ? pll_cloop_two_quadrant_atan(d_P_accu) : pll_four_quadrant_atan(d_P_accu)) / TWO_PI_D;
    const bool fll_pull_in = d_pull_in_transitory && p.enable_fll_pull_in;
    if (fll_pull_in || p.enable_fll_steady_state)
        {
            d_carr_freq_error_hz = fll_diff_atan(d_P_accu_old, d_P_accu, 0, d_current_correlation_time_s) / TWO_PI_D;
            d_P_accu_old = d_P_accu;
            // pure FLL during pull-in, FLL-aided PLL afterwards
            d_carr_error_filt_hz = d_carrier_loop_filter.get_carrier_error(static_cast<float>(d_carr_freq_error_hz),
                fll_pull_in ? 0.0F : static_cast<float>(d_carr_phase_error_hz), static_cast<float>(d_current_correlation_time_s));
        }
    else
        {
            d_carr_error_filt_hz = d_carrier_loop_filter.get_carrier_error(0, static_cast<float>(d_carr_phase_error_hz), static_cast<float>(d_current_correlation_time_s));
        }
    d_carrier_doppler_hz = d_carr_error_filt_hz;

    // DLL
    if (d_veml)
        d_code_error_chips = dll_nc_vemlp_normalized(d_VE_accu, d_E_accu, d_L_accu, d_VL_accu);
    else
        d_code_error_chips = dll_nc_e_minus_l_normalized(d_E_accu, d_L_accu, p.spc, p.slope, p.y_intercept);
    d_code_error_filt_chips = d_code_loop_filter.apply(static_cast<float>(d_code_error_chips));  // [chips/s]
    d_code_freq_chips = d_code_chip_rate - d_code_error_filt_chips;
    if (p.carrier_aiding) d_code_freq_chips += d_carrier_doppler_hz * d_code_chip_rate / d_signal_carrier_freq;

    // experimental carrier/code Doppler incoherence correction (:1127-1151)
    if (p.enable_doppler_correction && !d_pull_in_transitory && !d_corrected_doppler)
        {
            d_dll_filt_history.push_back(static_cast<float>(d_code_error_filt_chips));
            if (d_dll_filt_history.size() == 1000)
                {
                    const float avg = static_cast<float>(std::accumulate(d_dll_filt_history.begin(), d_dll_filt_history.end(), 0.0)) / 1000.0F;
                    if (std::fabs(avg) > 1.0)
                        {
                            const float err_hz = static_cast<float>(d_signal_carrier_freq) * avg / static_cast<float>(d_code_chip_rate);
                            d_carrier_loop_filter.initialize(static_cast<float>(d_carrier_doppler_hz) - err_hz);
                            d_corrected_doppler = true;
                        }
                    d_dll_filt_history.clear();
                }
        }
}

void b200_dll_pll_veml_tracking::clear_tracking_vars()
{
    std::fill(d_correlator_outs.begin(), d_correlator_outs.end(), gr_complex(0.0, 0.0));
    if (d_trk_parameters.track_pilot)
        {
            d_Prompt_Data = gr_complex(0.0, 0.0);
            d_P_data_accu = gr_complex(0.0, 0.0);
        }
    d_P_accu_old = gr_complex(0.0, 0.0);
    d_carr_phase_error_hz = d_carr_freq_error_hz = d_carr_error_filt_hz = 0.0;
    d_code_error_chips = d_code_error_filt_chips = 0.0;
    d_current_symbol = 0;
    d_current_data_symbol = 0;
    d_Prompt_circular_buffer.clear();
    d_carrier_phase_rate_step_rad = 0.0;
    d_code_phase_rate_step_chips = 0.0;
    d_carr_ph_history.clear();
    d_code_ph_history.clear();
}

namespace
{
// high_dyn rate estimate: difference of the means of the two halves of a 2*L history over the samples of the newer half
// (dll_pll_veml_tracking.cc:1207-1224, 1243-1258)
double history_rate(std::deque<std::pair<double, double>>& hist, double value, double n_samples, unsigned L, double current)
{
    hist.push_back({value, n_samples});
    if (hist.size() > 2 * L) hist.pop_front();
    if (hist.size() < 2 * L) return current;
    double older = 0.0, newer = 0.0, samples = 0.0;
    for (unsigned k = 0; k < L; k++)
        {
            older += hist[k].first;
            newer += hist[2 * L - k - 1].first;
            samples += hist[2 * L - k - 1].second;
        }
    return (newer / L - older / L) / samples;
}
}  // namespace

// dll_pll_veml_tracking.cc:1189-1260
void b200_dll_pll_veml_tracking::update_tracking_vars()
{
    const Dll_Pll_Conf& p = d_trk_parameters;
    d_T_chip_seconds = 1.0 / d_code_freq_chips;
    d_T_prn_seconds = d_T_chip_seconds * static_cast<double>(d_code_length_chips);
    // next block length from the new code period and the residual code phase
    d_T_prn_samples = d_T_prn_seconds * p.fs_in;
    d_K_blk_samples = d_T_prn_samples + d_rem_code_phase_samples;
    d_current_prn_length_samples = static_cast<int32_t>(std::floor(d_K_blk_samples));
    const auto n = static_cast<double>(d_current_prn_length_samples);

    d_carrier_phase_step_rad = TWO_PI_D * d_carrier_doppler_hz / p.fs_in;
    if (p.high_dyn) d_carrier_phase_rate_step_rad = history_rate(d_carr_ph_history, d_carrier_phase_step_rad, n, p.smoother_length, d_carrier_phase_rate_step_rad);
    const double advance = d_carrier_phase_step_rad * n + 0.5 * d_carrier_phase_rate_step_rad * n * n;
    d_rem_carr_phase_rad += static_cast<float>(advance);
    d_rem_carr_phase_rad = std::fmod(d_rem_carr_phase_rad, TWO_PI_D);  // float member, double fmod: as the reference
    d_acc_carrier_phase_rad -= advance;

    d_code_phase_step_chips = d_code_freq_chips / p.fs_in;
    if (p.high_dyn) d_code_phase_rate_step_chips = history_rate(d_code_ph_history, d_code_phase_step_chips, n, p.smoother_length, d_code_phase_rate_step_chips);
    d_rem_code_phase_samples = d_K_blk_samples - n;  // < 1 sample
    d_rem_code_phase_chips = d_code_freq_chips * d_rem_code_phase_samples / p.fs_in;
}

// dll_pll_veml_tracking.cc:1263-1373: accumulate one code period into the extended-integration sums, removing the
// secondary code of the tracked component and of the data component
void b200_dll_pll_veml_tracking::save_correlation_results()
{
    const float sign = (d_secondary && d_secondary_code_string[d_current_symbol] != '0') ? -1.0F : 1.0F;
    if (d_veml)
        {
            d_VE_accu += sign * d_correlator_outs[idx_VE_];
            d_VL_accu += sign * d_correlator_outs[idx_VL_];
        }
    d_E_accu += sign * d_correlator_outs[idx_E_];
    d_P_accu += sign * d_correlator_outs[idx_P_];
    d_L_accu += sign * d_correlator_outs[idx_L_];
    if (d_secondary) d_current_symbol = (d_current_symbol + 1) % d_secondary_code_length;

    const gr_complex data_prompt = d_trk_parameters.track_pilot ? d_Prompt_Data : d_correlator_outs[idx_P_];
    if (d_symbols_per_bit > 1)
        {
            if (d_data_secondary_code_length > 0)
                {
                    d_P_data_accu += (d_data_secondary_code_string[d_current_data_symbol] == '0') ? data_prompt : -data_prompt;
                    d_current_data_symbol = (d_current_data_symbol + 1) % d_data_secondary_code_length;
                }
            else
                {
                    d_P_data_accu += data_prompt;
                    d_current_data_symbol = (d_current_data_symbol + 1) % d_symbols_per_bit;
                }
        }
    else
        {
            d_P_data_accu = data_prompt;
        }
    d_cloop = !d_trk_parameters.track_pilot;  // a pilot has no data transitions: four-quadrant discriminator
}

// One Carrier_wipeoff_multicorrelator_resampler call as do_correlation_step forms it (:1037-1062): doubles cast to
// float, code quantities scaled by the table entries per chip, and — as in this fork — a window of vector_length samples.
void b200_dll_pll_veml_tracking::fill_cmd(gb200_trk_cmd& c, int ch, int n_taps, const float* shifts, uint64_t offset) const
{
    const auto spc = static_cast<float>(d_code_samples_per_chip);
    c = gb200_trk_cmd{};
    c.ch = ch;
    c.n_samples = static_cast<int32_t>(d_trk_parameters.vector_length);
    c.n_taps = n_taps;
    c.flags = d_trk_parameters.high_dyn ? GB200_TRK_HIGH_DYN : 0;
    c.sample_offset = offset;
    c.rem_carrier_phase_rad = d_rem_carr_phase_rad;
    c.phase_step_rad = static_cast<float>(d_carrier_phase_step_rad);
    c.phase_rate_step_rad = static_cast<float>(d_carrier_phase_rate_step_rad);
    c.rem_code_phase_chips = static_cast<float>(d_rem_code_phase_chips) * spc;
    c.code_phase_step_chips = static_cast<float>(d_code_phase_step_chips) * spc;
    c.code_phase_rate_step_chips = static_cast<float>(d_code_phase_rate_step_chips) * spc;
    for (int t = 0; t < n_taps; t++) c.shifts_chips[t] = shifts[t];
}

// First half of general_work (:1728-1790): the pull-in time-out, standby, and the pull-in alignment need no correlation
b200_dll_pll_veml_tracking::Step b200_dll_pll_veml_tracking::plan(uint64_t nitems_read, int engine_channel, int engine_data_channel)
{
    const Dll_Pll_Conf& p = d_trk_parameters;
    Step s;
    if (d_pull_in_transitory && p.pull_in_time_s < (nitems_read - d_acq_sample_stamp) / static_cast<int>(p.fs_in))
        {
            d_pull_in_transitory = false;
            d_carrier_lock_fail_counter = 0;
            d_code_lock_fail_counter = 0;
        }
    switch (d_state)
        {
        case 0:  // standby: the caller consumes whatever it has
            s.consume = -1;
            return s;
        case 1:  // pull-in: skip samples until the incoming code is aligned with the local replica
            {
                const int64_t acq_trk_diff_samples = static_cast<int64_t>(nitems_read) - static_cast<int64_t>(d_acq_sample_stamp);
                const double delta_trk_to_acq_prn_start_samples = static_cast<double>(acq_trk_diff_samples) - d_acq_code_phase_samples;
                d_code_freq_chips = d_code_chip_rate;
                d_code_phase_step_chips = d_code_freq_chips / p.fs_in;
                d_code_phase_rate_step_chips = 0.0;
                const double T_prn_mod_samples = (1.0 / d_code_freq_chips) * static_cast<double>(d_code_length_chips) * p.fs_in;
                d_acq_code_phase_samples = T_prn_mod_samples - std::fmod(delta_trk_to_acq_prn_start_samples, T_prn_mod_samples);
                d_current_prn_length_samples = static_cast<int32_t>(std::round(T_prn_mod_samples));
                const auto samples_offset = static_cast<int32_t>(std::round(d_acq_code_phase_samples));
                d_acc_carrier_phase_rad -= d_carrier_phase_step_rad * static_cast<double>(samples_offset);
                d_state = 2;
                d_cn0_smoother.reset();
                d_carrier_lock_test_smoother.reset();
                s.consume = samples_offset;
                return s;
            }
        default:
            s.correlate = true;
            s.consume = d_current_prn_length_samples;  // provisional: finish() consumes the updated length, as the reference does
            fill_cmd(s.cmd, engine_channel, d_n_correlator_taps, d_local_code_shift_chips.data(), nitems_read);
            if (p.track_pilot)
                {
                    s.has_data = true;
                    fill_cmd(s.data_cmd, engine_data_channel, 1, &d_local_code_shift_chips[idx_P_], nitems_read);
                }
            return s;
        }
}

void b200_dll_pll_veml_tracking::fill_synchro(Gnss_Synchro& s) const
{
    s = *d_acquisition_gnss_synchro;
    const gr_complex pd = d_P_data_accu;
    s.Prompt_I = static_cast<double>(d_interchange_iq ? pd.imag() : pd.real());
    s.Prompt_Q = static_cast<double>(d_interchange_iq ? pd.real() : pd.imag());
    s.Code_phase_samples = d_rem_code_phase_samples;
    s.Carrier_phase_rads = d_acc_carrier_phase_rad;
    s.Carrier_Doppler_hz = d_carrier_doppler_hz;
    s.CN0_dB_hz = d_CN0_SNV_dB_Hz;
    s.correlation_length_ms = d_correlation_length_ms;
    s.Flag_valid_symbol_output = true;
}

// Second half of general_work (:1791-2094), after the correlator outputs of this code period are known.
// On return step-consumption is d_current_prn_length_samples (read it through current_prn_length_samples()).
int b200_dll_pll_veml_tracking::finish(const Step& step, const gb200_cf32* taps, const gb200_cf32* data_tap, uint64_t nitems_read, Gnss_Synchro* out)
{
    const Dll_Pll_Conf& p = d_trk_parameters;
    if (!step.correlate) return 0;
    last_nitems_read_ = nitems_read;
    for (int t = 0; t < d_n_correlator_taps; t++) d_correlator_outs[t] = gr_complex(taps[t].re, taps[t].im);
    if (p.track_pilot && data_tap) d_Prompt_Data = gr_complex(data_tap->re, data_tap->im);

    Gnss_Synchro current{};
    current.Flag_valid_symbol_output = false;
    bool loss_of_lock = false;
    const auto reset_accumulators = [&](bool with_data) {
        d_VE_accu = d_E_accu = d_P_accu = d_L_accu = d_VL_accu = gr_complex(0.0, 0.0);
        if (with_data) d_P_data_accu = gr_complex(0.0, 0.0);
    };
    switch (d_state)
        {
        case 2:  // wide tracking and symbol synchronisation
            {
                if (d_veml)
                    {
                        d_VE_accu = d_correlator_outs[idx_VE_];
                        d_VL_accu = d_correlator_outs[idx_VL_];
                    }
                d_E_accu = d_correlator_outs[idx_E_];
                d_P_accu = d_correlator_outs[idx_P_];
                d_L_accu = d_correlator_outs[idx_L_];
                d_trk_parameters.spc = p.early_late_space_chips;
                // fail-safe: no bit / secondary-code synchronisation within the time limit forces a loss of lock
                if (p.bit_synchronization_time_limit_s < (nitems_read - d_acq_sample_stamp) / static_cast<int>(p.fs_in)) d_carrier_lock_fail_counter = 300000;
                if (!cn0_and_tracking_lock_status(d_code_period))
                    {
                        clear_tracking_vars();
                        d_state = 0;
                        loss_of_lock = true;
                        current = *d_acquisition_gnss_synchro;
                        break;
                    }
                run_dll_pll();
                update_tracking_vars();
                log_data();
                bool next_state = false;
                if (!d_pull_in_transitory)
                    {
                        if (d_secondary || d_symbols_per_bit > 1)
                            {
                                // secondary-code lock, or (no secondary code) bit edges against the preamble pattern
                                d_Prompt_circular_buffer.push_back(d_correlator_outs[idx_P_]);
                                if (d_Prompt_circular_buffer.size() > d_secondary_code_length) d_Prompt_circular_buffer.pop_front();
                                if (d_Prompt_circular_buffer.size() == d_secondary_code_length)
                                    {
                                        next_state = acquire_secondary();
                                        if (next_state)
                                            std::cout << d_systemName << " " << d_signal_type << (d_secondary ? " secondary code locked in channel " : " tracking bit synchronization locked in channel ")
                                                      << d_channel << " for satellite PRN " << d_acquisition_gnss_synchro->PRN << '\n';
                                    }
                            }
                        else
                            {
                                next_state = true;
                            }
                    }
                if (next_state)
                    {
                        reset_accumulators(true);
                        d_Prompt_circular_buffer.clear();
                        d_current_symbol = 0;
                        d_current_data_symbol = 0;
                        if (d_enable_extended_integration)
                            {
                                d_extend_correlation_symbols_count = 0;
                                d_current_correlation_time_s = static_cast<float>(d_extend_correlation_symbols) * static_cast<float>(d_code_period);
                                d_state = 3;
                                std::cout << "Enabled " << d_extend_correlation_symbols * static_cast<int32_t>(d_code_period * 1000.0) << " ms extended correlator in channel " << d_channel
                                          << " for satellite PRN " << d_acquisition_gnss_synchro->PRN << '\n';
                                d_code_loop_filter.set_update_interval(static_cast<float>(d_current_correlation_time_s));
                                d_code_loop_filter.set_noise_bandwidth(p.dll_bw_narrow_hz);
                                d_carrier_loop_filter.set_params(p.fll_bw_hz, p.pll_bw_narrow_hz, p.pll_filter_order);
                                set_narrow_taps();
                            }
                        else
                            {
                                d_state = 4;
                            }
                    }
                break;
            }
        case 3:  // coherent integration (correlation time extension)
            {
                save_correlation_results();
                update_tracking_vars();
                if (d_current_data_symbol == 0)
                    {
                        log_data();
                        fill_synchro(current);
                        d_P_data_accu = gr_complex(0.0, 0.0);
                    }
                d_extend_correlation_symbols_count++;
                if (d_extend_correlation_symbols_count == (d_extend_correlation_symbols - 1))
                    {
                        d_extend_correlation_symbols_count = 0;
                        d_state = 4;
                    }
                break;
            }
        case 4:  // narrow tracking
            {
                save_correlation_results();
                if (!cn0_and_tracking_lock_status(d_code_period * static_cast<double>(d_extend_correlation_symbols)))
                    {
                        clear_tracking_vars();
                        d_state = 0;
                        loss_of_lock = true;
                        current = *d_acquisition_gnss_synchro;
                        break;
                    }
                run_dll_pll();
                update_tracking_vars();
                if (!d_acc_carrier_phase_initialized)  // check_carrier_phase_coherent_initialization (:1155-1162)
                    {
                        d_acc_carrier_phase_rad = -d_rem_carr_phase_rad;
                        d_acc_carrier_phase_initialized = true;
                    }
                if (d_current_data_symbol == 0)
                    {
                        log_data();
                        fill_synchro(current);
                        d_P_data_accu = gr_complex(0.0, 0.0);
                    }
                reset_accumulators(false);
                if (d_enable_extended_integration) d_state = 3;
                break;
            }
        default:
            break;
        }
    if (current.Flag_valid_symbol_output || loss_of_lock)
        {
            current.fs = static_cast<int64_t>(p.fs_in);
            // nitems_read(0) after consume_each(d_current_prn_length_samples)
            current.Tracking_sample_counter = nitems_read + static_cast<uint64_t>(d_current_prn_length_samples);
            current.Flag_valid_symbol_output = !loss_of_lock;
            current.Flag_PLL_180_deg_phase_locked = d_Flag_PLL_180_deg_phase_locked;
            if (out) *out = current;
            return 1;
        }
    return 0;
}

// Binary dump record, field for field the reference's (:1376-1470), so tracking_dump_reader
// (src/tests/unit-tests/signal-processing-blocks/libs/tracking_dump_reader.cc) and the MATLAB/Python readers keep working:
// 5 x float |VE|,|E|,|P|,|L|,|VL|; 2 x float prompt I,Q; uint64 PRN start sample; 12 x float (acc. carrier phase, Doppler,
// Doppler rate, code freq, code rate, PLL error, PLL filtered, DLL error, DLL filtered, CN0, lock test, rem code phase);
// double sample counter; uint32 PRN.
void b200_dll_pll_veml_tracking::log_data()
{
    if (!d_dump || !d_dump_file.is_open()) return;
    const Dll_Pll_Conf& p = d_trk_parameters;
    const gr_complex prompt = p.track_pilot ? d_Prompt_Data : d_correlator_outs[idx_P_];
    const uint64_t stamp = last_nitems_read_ + static_cast<uint64_t>(d_current_prn_length_samples);
    const float f[19] = {d_veml ? std::abs(d_VE_accu) : 0.0F, std::abs(d_E_accu), std::abs(d_P_accu), std::abs(d_L_accu), d_veml ? std::abs(d_VL_accu) : 0.0F, prompt.real(),
        prompt.imag(), static_cast<float>(d_acc_carrier_phase_rad), static_cast<float>(d_carrier_doppler_hz),
        static_cast<float>(d_carrier_phase_rate_step_rad * p.fs_in * p.fs_in / TWO_PI_D), static_cast<float>(d_code_freq_chips),
        static_cast<float>(d_code_phase_rate_step_chips * p.fs_in * p.fs_in), static_cast<float>(d_carr_phase_error_hz), static_cast<float>(d_carr_error_filt_hz),
        static_cast<float>(d_code_error_chips), static_cast<float>(d_code_error_filt_chips), static_cast<float>(d_CN0_SNV_dB_Hz), static_cast<float>(d_carrier_lock_test),
        static_cast<float>(d_rem_code_phase_samples)};
    d_dump_file.write(reinterpret_cast<const char*>(f), 7 * sizeof(float));
    d_dump_file.write(reinterpret_cast<const char*>(&stamp), sizeof(uint64_t));
    d_dump_file.write(reinterpret_cast<const char*>(f + 7), 12 * sizeof(float));
    const auto counter = static_cast<double>(stamp);
    d_dump_file.write(reinterpret_cast<const char*>(&counter), sizeof(double));
    const uint32_t prn = d_acquisition_gnss_synchro->PRN;
    d_dump_file.write(reinterpret_cast<const char*>(&prn), sizeof(uint32_t));
}

// Per-channel blocking form of general_work: what a gr::block wrapper calls with its input buffer.
// The correlation goes through the shared engine (one upload per sample block, requests of concurrent channels batched).
int b200_dll_pll_veml_tracking::general_work(const void* input_items, int ninput_items, uint64_t nitems_read, Gnss_Synchro* out, int* consumed)
{
    if (pending_skip_ > 0)
        {
            // rest of a pull-in alignment skip that did not fit into the previous call's input (see below)
            const int c = static_cast<int>(std::min<int64_t>(pending_skip_, ninput_items));
            pending_skip_ -= c;
            *consumed = c;
            return 0;
        }
    Step s = plan(nitems_read, d_multicorrelator.channel(), d_correlator_data.channel());
    if (!s.correlate)
        {
            // The reference's forecast() asks for 2 * vector_length items, so its consume_each(samples_offset) (dll_pll_veml_tracking.cc:1786)
            // always fits; a caller of this blocking form may hand over less: plan() has already left state 1 and charged the full
            // offset to d_acc_carrier_phase_rad, so the part that does not fit is skipped at the start of the following calls.
            *consumed = s.consume < 0 ? ninput_items : std::min(s.consume, ninput_items);
            if (s.consume > ninput_items) pending_skip_ = static_cast<int64_t>(s.consume) - ninput_items;
            return 0;
        }
    if (ninput_items < s.cmd.n_samples)
        throw std::runtime_error("b200_dll_pll_veml_tracking::general_work: " + std::to_string(ninput_items) + " input items, the correlation needs " +
                                 std::to_string(s.cmd.n_samples) + " (the reference's forecast() guarantees 2 * vector_length)");
    if (!engine_) throw std::runtime_error("b200_dll_pll_veml_tracking: no correlator engine (set_engine)");
    gr_complex data_out(0, 0);
    d_multicorrelator.set_input_output_vectors(d_correlator_outs.data(), input_items, nitems_read, ninput_items);
    bool ok = d_multicorrelator.Carrier_wipeoff_multicorrelator_resampler(s.cmd.rem_carrier_phase_rad, s.cmd.phase_step_rad, s.cmd.phase_rate_step_rad,
        s.cmd.rem_code_phase_chips, s.cmd.code_phase_step_chips, s.cmd.code_phase_rate_step_chips, s.cmd.n_samples);
    if (ok && s.has_data)
        {
            d_correlator_data.set_input_output_vectors(&data_out, input_items, nitems_read, ninput_items);
            ok = d_correlator_data.Carrier_wipeoff_multicorrelator_resampler(s.data_cmd.rem_carrier_phase_rad, s.data_cmd.phase_step_rad, s.data_cmd.phase_rate_step_rad,
                s.data_cmd.rem_code_phase_chips, s.data_cmd.code_phase_step_chips, s.data_cmd.code_phase_rate_step_chips, s.data_cmd.n_samples);
        }
    if (!ok) throw std::runtime_error("b200_dll_pll_veml_tracking: correlation request rejected (window not in the device ring)");
    std::vector<gb200_cf32> taps(d_n_correlator_taps);
    for (int t = 0; t < d_n_correlator_taps; t++) taps[t] = {d_correlator_outs[t].real(), d_correlator_outs[t].imag()};
    const gb200_cf32 dt{data_out.real(), data_out.imag()};
    const int produced = finish(s, taps.data(), s.has_data ? &dt : nullptr, nitems_read, out);
    *consumed = d_current_prn_length_samples;
    return produced;
}
}  // namespace gb200h
