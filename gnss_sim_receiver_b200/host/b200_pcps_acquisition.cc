#include "b200_pcps_acquisition.h"
#include "b200_mat_writer.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <iostream>
#include <stdexcept>

namespace gb200h
{
namespace
{
[[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + gb200_last_error()); }

gb200_acq_config engine_config(const Acq_Conf& c, uint32_t fft_size, uint32_t consumed, int iq, int max_prns)
{
    gb200_acq_config g{};
    g.device = c.device;
    g.iq_dtype = iq;
    g.fft_size = static_cast<int32_t>(fft_size);
    g.consumed_samples = static_cast<int32_t>(consumed);
    g.bit_transition = c.bit_transition_flag;
    g.use_cfar = c.use_CFAR_algorithm_flag;
    g.samples_per_chip = static_cast<int32_t>(c.samples_per_chip);
    g.max_dwells = static_cast<int32_t>(c.max_dwells);
    g.max_prns = max_prns;
    g.samples_per_code = c.samples_per_code;
    g.fs_in = static_cast<double>(c.resampled_fs > 0 ? c.resampled_fs : c.fs_in);
    return g;
}

void sizes(const Acq_Conf& c, uint32_t* consumed, uint32_t* fft_size)
{
    // pcps_acquisition.cc:71 and :84-91
    *consumed = static_cast<uint32_t>(c.sampled_ms * c.samples_per_ms * (c.bit_transition_flag ? 2.0 : 1.0));
    *fft_size = (c.sampled_ms == c.ms_per_code) ? *consumed : *consumed * 2;
}
}  // namespace

b200_pcps_acquisition::b200_pcps_acquisition(const Acq_Conf& conf)
    : d_acq_parameters(conf), d_doppler_step(static_cast<uint32_t>(conf.doppler_step)), d_use_CFAR_algorithm_flag(conf.use_CFAR_algorithm_flag)
{
    sizes(conf, &d_consumed_samples, &d_fft_size);
    iq_dtype_ = item_type_iq(conf.item_type);
    if (iq_dtype_ < 0) throw std::invalid_argument("Unknown item type: " + conf.item_type);
    const gb200_acq_config g = engine_config(conf, d_fft_size, d_consumed_samples, iq_dtype_, 1);
    if (gb200_acq_create(&g, &h_) != GB200_OK) fail("gb200_acq_create");
    d_dump = conf.dump;
    if (d_dump)
        {
            // :134-165: the dump name without its extension; the grid must stay on the device after a sweep
            d_dump_filename = conf.dump_filename.empty() ? std::string("./acquisition") : conf.dump_filename;
            const size_t slash = d_dump_filename.find_last_of('/');
            const size_t dot = d_dump_filename.find_last_of('.');
            if (dot != std::string::npos && dot != 0 && (slash == std::string::npos || dot > slash + 1)) d_dump_filename.erase(dot);
            if (gb200_acq_keep_grid(h_, 1) != GB200_OK) fail("gb200_acq_keep_grid");
        }
    if (conf.make_2_steps)
        {
            if (conf.num_doppler_bins_step2 == 0U) throw std::invalid_argument("make_two_steps needs second_nbins > 0");
            if (gb200_acq_create(&g, &h2_) != GB200_OK) fail("gb200_acq_create (second step)");
            if (d_dump && gb200_acq_keep_grid(h2_, 1) != GB200_OK) fail("gb200_acq_keep_grid");
            d_doppler_center_step_two = 0.0F;
            update_grid_doppler_wipeoffs_step2();  // sizes the fine grid; re-centred after every positive first step
        }
    d_data_buffer.assign(static_cast<size_t>(d_consumed_samples) * conf.it_size, 0);
}

b200_pcps_acquisition::~b200_pcps_acquisition()
{
    gb200_acq_destroy(h_);
    gb200_acq_destroy(h2_);
}

// :304-312: fine bins at  centre + (k - floor(n2 / 2)) * second_doppler_step,  float arithmetic
void b200_pcps_acquisition::update_grid_doppler_wipeoffs_step2()
{
    const uint32_t n2 = d_acq_parameters.num_doppler_bins_step2;
    std::vector<float> f(n2);
    for (uint32_t k = 0; k < n2; k++)
        {
            const float doppler = (static_cast<float>(k) - static_cast<float>(std::floor(n2 / 2.0))) * d_acq_parameters.doppler_step2;
            f[k] = d_doppler_center_step_two + doppler;
        }
    if (gb200_acq_set_doppler_bins(h2_, f.data(), static_cast<int>(n2)) != GB200_OK) fail("gb200_acq_set_doppler_bins");
}

void b200_pcps_acquisition::apply_grid()
{
    if (gb200_acq_set_doppler_grid(h_, d_acq_parameters.doppler_max, static_cast<int>(d_doppler_step), d_doppler_center) != GB200_OK) fail("gb200_acq_set_doppler_grid");
    d_num_doppler_bins = static_cast<uint32_t>(gb200_acq_num_doppler_bins(h_));
    grid_dirty_ = false;
}

// :248-292
void b200_pcps_acquisition::init()
{
    d_gnss_synchro->Flag_valid_acquisition = false;
    d_gnss_synchro->Flag_valid_symbol_output = false;
    d_gnss_synchro->Flag_valid_pseudorange = false;
    d_gnss_synchro->Flag_valid_word = false;
    d_gnss_synchro->Acq_doppler_step = 0U;
    d_gnss_synchro->Acq_delay_samples = 0.0;
    d_gnss_synchro->Acq_doppler_hz = 0.0;
    d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
    d_mag = 0.0;
    d_input_power = 0.0;
    apply_grid();  // d_num_doppler_bins = ceil(2 * doppler_max / doppler_step) and the wipe-off grid (:261, :295-302)
    d_worker_active = false;
}

void b200_pcps_acquisition::set_doppler_center(int32_t doppler_center)
{
    std::lock_guard<std::mutex> lk(d_setlock);
    if (doppler_center != d_doppler_center)
        {
            d_doppler_center = doppler_center;
            apply_grid();  // update_grid_doppler_wipeoffs (:222-229)
        }
}

// :175-208: the device zero-pads, transforms and conjugates
void b200_pcps_acquisition::set_local_code(std::complex<float>* code)
{
    std::lock_guard<std::mutex> lk(d_setlock);
    const int n = d_acq_parameters.bit_transition_flag ? static_cast<int>(d_fft_size / 2) : static_cast<int>(d_consumed_samples);
    if (gb200_acq_set_local_code(h_, 0, reinterpret_cast<const gb200_cf32*>(code), n) != GB200_OK) fail("gb200_acq_set_local_code");
    if (h2_ && gb200_acq_set_local_code(h2_, 0, reinterpret_cast<const gb200_cf32*>(code), n) != GB200_OK) fail("gb200_acq_set_local_code");
}

// :315-336
void b200_pcps_acquisition::set_state(int32_t state)
{
    std::lock_guard<std::mutex> lk(d_setlock);
    d_state = state;
    if (d_state == 1)
        {
            d_gnss_synchro->Acq_delay_samples = 0.0;
            d_gnss_synchro->Acq_doppler_hz = 0.0;
            d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
            d_gnss_synchro->Acq_doppler_step = 0U;
            d_mag = 0.0;
            d_test_statistics = 0.0;
            d_active = true;
        }
}

void b200_pcps_acquisition::send_positive_acquisition()
{
    d_positive_acq = 1;
    if (auto fsm = d_channel_fsm.lock())
        fsm->Event_valid_acquisition();
    else if (event_sink_)
        event_sink_(1);
}

void b200_pcps_acquisition::send_negative_acquisition()
{
    d_positive_acq = 0;
    if (event_sink_) event_sink_(2);
}

bool b200_pcps_acquisition::start()
{
    std::lock_guard<std::mutex> lk(d_setlock);
    d_sample_counter = 0ULL;
    calculate_threshold();
    return true;
}

// :884-899 (gb200_acq_threshold evaluates 2 * gamma_p_inv(2 * dwells, (1 - pfa)^(1 / (N_eff * bins))))
void b200_pcps_acquisition::calculate_threshold()
{
    const float pfa = d_step_two ? d_acq_parameters.pfa2 : d_acq_parameters.pfa;  // :886
    if (pfa <= 0.0F) return;
    if (grid_dirty_) apply_grid();
    d_threshold = gb200_acq_threshold(d_step_two ? h2_ : h_, pfa);  // :894: the step's own number of Doppler bins
}

// :600-871
void b200_pcps_acquisition::acquisition_core(uint64_t samp_count)
{
    std::unique_lock<std::mutex> lk(d_setlock);
    d_mag = 0.0;
    d_num_noncoherent_integrations_counter++;
    if (grid_dirty_) apply_grid();
    const int32_t slot = 0;
    if (gb200_acq_run(d_step_two ? h2_ : h_, d_data_buffer.data(), static_cast<int>(d_num_noncoherent_integrations_counter) - 1, &slot, 1, &last_) != GB200_OK) fail("gb200_acq_run");
    if (d_step_two)
        {
            // :524 / :561: the fine bin's frequency, float arithmetic then truncated to int32
            const float n2 = static_cast<float>(d_acq_parameters.num_doppler_bins_step2);
            last_.doppler_hz = static_cast<int32_t>(d_doppler_center_step_two + (static_cast<float>(last_.doppler_index) - static_cast<float>(std::floor(n2 / 2.0))) * d_acq_parameters.doppler_step2);
            last_.acq_doppler_hz = static_cast<float>(last_.doppler_hz);
        }
    d_mag = last_.peak;
    if (d_step_two && d_use_CFAR_algorithm_flag)
        {
            // :517-528: max_to_input_power_statistic estimates the input power on the coarse grid only; the fine step divides its own
            // maximum by the power the first step left behind
            last_.input_power = d_input_power;
            last_.test_statistics = last_.peak / d_input_power;
        }
    d_input_power = last_.input_power;
    d_test_statistics = last_.test_statistics;
    const Acq_Conf& p = d_acq_parameters;
    if (p.use_automatic_resampler)
        {
            d_gnss_synchro->Acq_delay_samples = static_cast<double>(last_.acq_delay_samples) * p.resampler_ratio;
            d_gnss_synchro->Acq_delay_samples -= static_cast<double>(p.resampler_latency_samples);
            d_gnss_synchro->Acq_doppler_hz = static_cast<double>(last_.doppler_hz);
            d_gnss_synchro->Acq_samplestamp_samples = static_cast<uint64_t>(std::rint(static_cast<double>(samp_count) * p.resampler_ratio));
        }
    else
        {
            d_gnss_synchro->Acq_delay_samples = static_cast<double>(last_.acq_delay_samples);
            d_gnss_synchro->Acq_doppler_hz = static_cast<double>(last_.doppler_hz);
            d_gnss_synchro->Acq_samplestamp_samples = samp_count;
        }
    d_gnss_synchro->Acq_doppler_step = d_step_two ? static_cast<uint32_t>(d_acq_parameters.doppler_step2) : d_doppler_step;  // :749,:756

    // a positive search: declare it, or (make_two_steps, first step) arm the fine Doppler grid around it (:770-787, :818-838)
    const auto positive = [&](bool reset_counters) {
        if (p.make_2_steps)
            {
                if (d_step_two)
                    {
                        send_positive_acquisition();
                        d_step_two = false;
                        d_state = 0;
                    }
                else
                    {
                        d_step_two = true;  // next: a fresh buffer searched on the small grid
                        d_doppler_center_step_two = static_cast<float>(d_gnss_synchro->Acq_doppler_hz);
                        update_grid_doppler_wipeoffs_step2();
                        d_num_noncoherent_integrations_counter = 0U;
                        if (reset_counters) d_positive_acq = 0;
                        d_state = 0;
                    }
                calculate_threshold();
            }
        else
            {
                send_positive_acquisition();
                d_state = 0;
            }
    };
    const auto leave_step_two = [&] {
        const bool was_step_two = d_step_two;
        d_step_two = false;
        if (was_step_two) calculate_threshold();
    };
    if (!p.bit_transition_flag)
        {
            if (d_test_statistics > d_threshold)
                {
                    d_active = false;
                    positive(true);
                }
            else
                {
                    d_buffer_count = 0;
                    d_state = 1;  // another dwell
                }
            if (d_num_noncoherent_integrations_counter == p.max_dwells)
                {
                    if (d_state != 0) send_negative_acquisition();
                    d_state = 0;
                    d_active = false;
                    leave_step_two();
                }
        }
    else
        {
            d_active = false;
            if (d_test_statistics > d_threshold)
                positive(false);
            else
                {
                    d_state = 0;
                    leave_step_two();
                    send_negative_acquisition();
                }
        }
    d_worker_active = false;
    if (d_num_noncoherent_integrations_counter == p.max_dwells || d_positive_acq == 1 || p.bit_transition_flag)
        {
            if (d_dump && d_positive_acq) dump_results(static_cast<int32_t>(p.bit_transition_flag ? d_fft_size / 2 : d_fft_size));  // :862-866
            d_num_noncoherent_integrations_counter = 0U;
            d_positive_acq = 0;
        }
}

// :902-1031
int b200_pcps_acquisition::general_work(const void* input_items, int ninput_items)
{
    std::unique_lock<std::mutex> lk(d_setlock);
    const Acq_Conf& p = d_acq_parameters;
    int consumed = 0;
    if (!d_active || d_worker_active)
        {
            const bool consume_samples = !d_active || (d_worker_active && d_num_noncoherent_integrations_counter == p.max_dwells);
            if (!p.blocking_on_standby && consume_samples)
                {
                    d_sample_counter += static_cast<uint64_t>(ninput_items);
                    consumed = ninput_items;
                }
            if (d_step_two)  // :927-931: the fine search starts on its own
                {
                    d_state = 0;
                    d_active = true;
                }
            return consumed;
        }
    switch (d_state)
        {
        case 0:
            d_gnss_synchro->Acq_delay_samples = 0.0;
            d_gnss_synchro->Acq_doppler_hz = 0.0;
            d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
            d_gnss_synchro->Acq_doppler_step = 0U;
            d_mag = 0.0;
            d_state = 1;
            d_buffer_count = 0U;
            if (!p.blocking_on_standby)
                {
                    d_sample_counter += static_cast<uint64_t>(ninput_items);
                    consumed = ninput_items;
                }
            break;
        case 1:
            {
                // fill the data buffer; the full-buffer test comes BEFORE the increment, as in the reference, so the
                // transition to state 2 happens on the call after the one that completed the buffer
                const uint32_t n = static_cast<uint32_t>(ninput_items);
                const uint32_t incr = (n + d_buffer_count <= d_consumed_samples) ? n : d_consumed_samples - d_buffer_count;
                std::memcpy(d_data_buffer.data() + static_cast<size_t>(d_buffer_count) * p.it_size, input_items, static_cast<size_t>(incr) * p.it_size);
                if (d_buffer_count >= d_consumed_samples) d_state = 2;
                d_buffer_count += incr;
                d_sample_counter += static_cast<uint64_t>(incr);
                consumed = static_cast<int>(incr);
                break;
            }
        case 2:
            {
                const uint64_t stamp = d_sample_counter;
                lk.unlock();
                acquisition_core(stamp);  // blocking form; the GPU sweep takes well under a millisecond
                lk.lock();
                d_buffer_count = 0U;
                break;
            }
        default:
            break;
        }
    return consumed;
}

// :393-493 — same file name scheme, variable names, classes and shapes; MAT level 5 instead of v7.3 (b200_mat_writer.h)
void b200_pcps_acquisition::dump_results(int32_t effective_fft_size)
{
    d_dump_number++;
    std::string filename = d_dump_filename;
    filename.append("_");
    filename.append(1, d_gnss_synchro->System);
    filename.append("_");
    filename.append(1, d_gnss_synchro->Signal[0]);
    filename.append(1, d_gnss_synchro->Signal[1]);
    filename.append("_ch_");
    filename.append(std::to_string(d_channel));
    filename.append("_");
    filename.append(std::to_string(d_dump_number));
    filename.append("_sat_");
    filename.append(std::to_string(d_gnss_synchro->PRN));
    filename.append(".mat");
    MatFile5 mat(filename);
    if (!mat.ok())
        {
            std::cout << "Unable to create or open Acquisition dump file\n";
            return;
        }
    const auto neff = static_cast<uint32_t>(effective_fft_size);
    // d_grid: effective_fft_size x num_doppler_bins, column-major = one Doppler bin after the other
    std::vector<float> grid(static_cast<size_t>(d_num_doppler_bins) * neff);
    if (gb200_acq_read_grid(h_, 0, grid.data()) != GB200_OK) fail("gb200_acq_read_grid");
    mat.write("acq_grid", MatFile5::SINGLE, neff, d_num_doppler_bins, grid.data());
    mat.scalar("doppler_max", MatFile5::INT32, static_cast<int32_t>(d_acq_parameters.doppler_max));
    mat.scalar("doppler_step", MatFile5::INT32, static_cast<int32_t>(d_doppler_step));
    mat.scalar("d_positive_acq", MatFile5::INT32, static_cast<int32_t>(d_positive_acq));
    mat.scalar("acq_doppler_hz", MatFile5::SINGLE, static_cast<float>(d_gnss_synchro->Acq_doppler_hz));
    mat.scalar("acq_delay_samples", MatFile5::SINGLE, static_cast<float>(d_gnss_synchro->Acq_delay_samples));
    mat.scalar("test_statistic", MatFile5::SINGLE, d_test_statistics);
    mat.scalar("threshold", MatFile5::SINGLE, d_threshold);
    mat.scalar("input_power", MatFile5::SINGLE, d_input_power);
    mat.scalar("sample_counter", MatFile5::UINT64, static_cast<uint64_t>(d_sample_counter));
    mat.scalar("PRN", MatFile5::UINT32, static_cast<uint32_t>(d_gnss_synchro->PRN));
    mat.scalar("num_dwells", MatFile5::INT32, static_cast<int32_t>(d_num_noncoherent_integrations_counter));
    if (d_acq_parameters.make_2_steps)
        {
            const uint32_t n2 = d_acq_parameters.num_doppler_bins_step2;
            std::vector<float> narrow(static_cast<size_t>(n2) * neff);
            if (gb200_acq_read_grid(h2_, 0, narrow.data()) != GB200_OK) fail("gb200_acq_read_grid");
            mat.write("acq_grid_narrow", MatFile5::SINGLE, neff, n2, narrow.data());
            mat.scalar("doppler_step_narrow", MatFile5::SINGLE, d_acq_parameters.doppler_step2);
            mat.scalar("doppler_grid_narrow_min", MatFile5::SINGLE, d_doppler_center_step_two - static_cast<float>(std::floor(n2 / 2.0)) * d_acq_parameters.doppler_step2);
        }
}

std::vector<float> b200_pcps_acquisition::magnitude_grid()
{
    const uint32_t neff = static_cast<uint32_t>(gb200_acq_neff(h_));
    std::vector<float> g(static_cast<size_t>(d_num_doppler_bins) * neff);
    if (gb200_acq_read_grid(h_, 0, g.data()) != GB200_OK) fail("gb200_acq_read_grid");
    return g;
}

// ------------------------------------------------------------------ all PRNs in one sweep
B200AcquisitionSweep::B200AcquisitionSweep(const Acq_Conf& conf, int n_prn) : n_prn_(n_prn)
{
    uint32_t consumed, fft_size;
    sizes(conf, &consumed, &fft_size);
    const int iq = item_type_iq(conf.item_type);
    if (iq < 0) throw std::invalid_argument("Unknown item type: " + conf.item_type);
    const gb200_acq_config g = engine_config(conf, fft_size, consumed, iq, n_prn);
    if (gb200_acq_create(&g, &h_) != GB200_OK) fail("gb200_acq_create");
}

B200AcquisitionSweep::~B200AcquisitionSweep() { gb200_acq_destroy(h_); }

void B200AcquisitionSweep::set_local_code(int slot, const std::complex<float>* code, int n_code)
{
    if (gb200_acq_set_local_code(h_, slot, reinterpret_cast<const gb200_cf32*>(code), n_code) != GB200_OK) fail("gb200_acq_set_local_code");
}

void B200AcquisitionSweep::set_doppler_grid(int doppler_max, int doppler_step, int doppler_center)
{
    if (gb200_acq_set_doppler_grid(h_, doppler_max, doppler_step, doppler_center) != GB200_OK) fail("gb200_acq_set_doppler_grid");
}

void B200AcquisitionSweep::run(const void* iq_host, int dwell_index, std::vector<gb200_acq_result>* results)
{
    results->resize(n_prn_);
    if (gb200_acq_run(h_, iq_host, dwell_index, nullptr, n_prn_, results->data()) != GB200_OK) fail("gb200_acq_run");
}
}  // namespace gb200h
