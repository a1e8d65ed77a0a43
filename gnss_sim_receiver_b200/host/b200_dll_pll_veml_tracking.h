// Loop closure of the DLL/PLL (VEML) tracking block around the GPU multicorrelator.
//
// Mirrors dll_pll_veml_tracking (src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.{h:58-230,cc:88-2094}):
// same states (0 standby, 1 pull-in, 2 wide tracking + symbol sync, 3 extended integration, 4 narrow tracking), same
// double/float state variables, same order of operations per code period.  What differs is the shape of general_work:
// it is split into plan() -> [correlation on the GPU, batched over channels] -> finish(), so that B200TrackingBank can
// step thousands of channels with one kernel launch per code period.  general_work() is the per-channel blocking form
// (plan + B200_Multicorrelator_Real_Codes + finish) a GNU Radio block wrapper calls.
#ifndef GB200_DLL_PLL_VEML_TRACKING_H
#define GB200_DLL_PLL_VEML_TRACKING_H

#include "b200_conf.h"
#include "b200_correlator_engine.h"
#include "b200_tracking_libs.h"
#include "gnss_sdr_compat.h"
#include <deque>
#include <fstream>
#include <functional>
#include <string>
#include <vector>

namespace gb200h
{
class b200_dll_pll_veml_tracking
{
public:
    explicit b200_dll_pll_veml_tracking(const Dll_Pll_Conf& conf);
    ~b200_dll_pll_veml_tracking();

    // ---- TrackingInterface-facing controls (dll_pll_veml_tracking.h:67-75)
    void set_channel(uint32_t channel);
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro);
    void start_tracking();
    void stop_tracking();
    // "events" message port: 3 = loss of lock (dll_pll_veml_tracking.cc:1023)
    void set_event_sink(std::function<void(int)> sink) { event_sink_ = std::move(sink); }
    // "telemetry_to_trk" message port (msg_handler_telemetry_to_trk, dll_pll_veml_tracking.cc:617-641): event 1 = the telemetry decoder
    // reports a fault -> the loss-of-lock condition is forced (d_carrier_lock_fail_counter = 200000); other events are ignored
    void telemetry_to_trk(int tlm_event)
    {
        if (tlm_event == 1) d_carrier_lock_fail_counter = 200000;
    }

    // Code tables for signals whose generators are not part of this tree (Galileo E1/E5a, GPS L5...): the adapter hands
    // them over before start_tracking().  GPS L1 C/A tables are generated here when none was supplied.
    void set_tracking_code(const float* code, int n_entries);
    void set_data_code(const float* code, int n_entries);

    // ---- one general_work call, split
    struct Step
    {
        int consume{0};          // samples this call consumes (consume_each)
        bool correlate{false};   // states 2-4: the command(s) below must be evaluated before finish()
        gb200_trk_cmd cmd{};     // main correlator (n_taps = 3, 5 or n_correlators)
        bool has_data{false};    // track_pilot: 1-tap data-component correlator with the same NCO arguments
        gb200_trk_cmd data_cmd{};
    };
    // samples general_work needs in its input buffer (forecast, :607-614)
    int required_samples() const { return static_cast<int>(d_trk_parameters.vector_length) * 2; }
    // `nitems_read` = absolute index of the first unread input sample; `engine_channel`/`engine_data_channel` = code-table
    // slots of this block in the engine
    Step plan(uint64_t nitems_read, int engine_channel, int engine_data_channel);
    // taps: outputs of step.cmd (n_taps values), data_tap: output of step.data_cmd.  Returns 1 when *out was filled.
    int finish(const Step& step, const gb200_cf32* taps, const gb200_cf32* data_tap, uint64_t nitems_read, Gnss_Synchro* out);

    // per-channel blocking form: returns the number of Gnss_Synchro produced (0/1); *consumed = consume_each
    int general_work(const void* input_items, int ninput_items, uint64_t nitems_read, Gnss_Synchro* out, int* consumed);
    void set_engine(std::shared_ptr<B200CorrelatorEngine> engine);
    std::shared_ptr<B200CorrelatorEngine> engine() const { return engine_; }
    int engine_channel() const { return d_multicorrelator.channel(); }
    // (re)upload the code tables of this block to its engine slots; start_tracking does it when an engine is attached
    void upload_codes();

    // ---- observers
    int state() const { return d_state; }
    int n_correlator_taps() const { return d_n_correlator_taps; }
    double carrier_doppler_hz() const { return d_carrier_doppler_hz; }
    double code_freq_chips() const { return d_code_freq_chips; }
    double cn0_db_hz() const { return d_CN0_SNV_dB_Hz; }
    double carrier_lock_test() const { return d_carrier_lock_test; }
    double rem_code_phase_samples() const { return d_rem_code_phase_samples; }
    double acc_carrier_phase_rad() const { return d_acc_carrier_phase_rad; }
    int current_prn_length_samples() const { return d_current_prn_length_samples; }
    const std::vector<gr_complex>& correlator_outs() const { return d_correlator_outs; }
    const std::vector<float>& local_code_shift_chips() const { return d_local_code_shift_chips; }
    const std::vector<float>& tracking_code() const { return d_tracking_code; }
    const std::vector<float>& data_code() const { return d_data_code; }
    const Dll_Pll_Conf& parameters() const { return d_trk_parameters; }
    bool track_pilot() const { return d_trk_parameters.track_pilot; }
    int code_table_entries() const { return d_code_samples_per_chip * d_code_length_chips; }

private:
    bool cn0_and_tracking_lock_status(double coh_integration_time_s);
    bool acquire_secondary();
    void run_dll_pll();
    void update_tracking_vars();
    void clear_tracking_vars();
    void save_correlation_results();
    void log_data();
    void fill_cmd(gb200_trk_cmd& c, int ch, int n_taps, const float* shifts, uint64_t offset) const;
    void fill_synchro(Gnss_Synchro& s) const;
    void set_wide_taps();
    void set_narrow_taps();

    Dll_Pll_Conf d_trk_parameters;
    Gnss_Synchro* d_acquisition_gnss_synchro{nullptr};
    std::function<void(int)> event_sink_;
    std::shared_ptr<B200CorrelatorEngine> engine_;
    B200_Multicorrelator_Real_Codes d_multicorrelator, d_correlator_data;

    Tracking_loop_filter d_code_loop_filter;
    Tracking_FLL_PLL_filter d_carrier_loop_filter;
    Exponential_Smoother d_cn0_smoother, d_carrier_lock_test_smoother;

    std::vector<float> d_tracking_code, d_data_code, d_local_code_shift_chips;
    bool tracking_code_supplied_{false}, data_code_supplied_{false};
    std::vector<gr_complex> d_correlator_outs, d_Prompt_buffer;
    std::deque<gr_complex> d_Prompt_circular_buffer;
    std::deque<float> d_dll_filt_history;
    std::deque<std::pair<double, double>> d_carr_ph_history, d_code_ph_history;
    gr_complex d_Prompt_Data{0, 0};
    gr_complex d_VE_accu{0, 0}, d_E_accu{0, 0}, d_P_accu{0, 0}, d_P_accu_old{0, 0}, d_L_accu{0, 0}, d_VL_accu{0, 0}, d_P_data_accu{0, 0};
    int idx_VE_{-1}, idx_E_{0}, idx_P_{1}, idx_L_{2}, idx_VL_{-1};

    std::string d_secondary_code_string, d_data_secondary_code_string, d_systemName, d_signal_type;
    std::ofstream d_dump_file;

    double d_signal_carrier_freq{0}, d_code_period{0}, d_code_chip_rate{0}, d_acq_code_phase_samples{0}, d_acq_carrier_doppler_hz{0};
    double d_current_correlation_time_s{0}, d_carr_phase_error_hz{0}, d_carr_freq_error_hz{0}, d_carr_error_filt_hz{0};
    double d_code_error_chips{0}, d_code_error_filt_chips{0}, d_code_freq_chips{0}, d_carrier_doppler_hz{0}, d_acc_carrier_phase_rad{0};
    double d_rem_code_phase_chips{0}, d_T_chip_seconds{0}, d_T_prn_seconds{0}, d_T_prn_samples{0}, d_K_blk_samples{0};
    double d_carrier_lock_test{1.0}, d_CN0_SNV_dB_Hz{0}, d_carrier_lock_threshold{0};
    double d_carrier_phase_step_rad{0}, d_carrier_phase_rate_step_rad{0}, d_code_phase_step_chips{0}, d_code_phase_rate_step_chips{0};
    double d_rem_code_phase_samples{0};
    uint64_t d_acq_sample_stamp{0}, last_nitems_read_{0};
    int64_t pending_skip_{0};  // general_work: rest of the pull-in alignment skip that did not fit into one call
    float d_rem_carr_phase_rad{0};
    int32_t d_symbols_per_bit{0}, d_state{0}, d_correlation_length_ms{0}, d_n_correlator_taps{3}, d_current_prn_length_samples{0};
    int32_t d_extend_correlation_symbols_count{0}, d_extend_correlation_symbols{1}, d_cn0_estimation_counter{0};
    int32_t d_carrier_lock_fail_counter{0}, d_code_lock_fail_counter{0}, d_code_samples_per_chip{1}, d_code_length_chips{0};
    uint32_t d_channel{0}, d_secondary_code_length{0}, d_data_secondary_code_length{0}, d_current_symbol{0}, d_current_data_symbol{0};
    bool d_pull_in_transitory{true}, d_corrected_doppler{false}, d_interchange_iq{false}, d_veml{false}, d_cloop{true}, d_secondary{false};
    bool d_dump{false}, d_acc_carrier_phase_initialized{false}, d_enable_extended_integration{false}, d_Flag_PLL_180_deg_phase_locked{false};
};
}  // namespace gb200h
#endif
