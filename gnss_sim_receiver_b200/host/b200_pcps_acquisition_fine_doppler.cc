#include "b200_pcps_acquisition_fine_doppler.h"
#include "b200_gnss_codes.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <stdexcept>

namespace gb200h
{
namespace
{
[[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + gb200_last_error()); }
constexpr int kPrnReplicas = 10;       // pcps_acquisition_fine_doppler_cc.cc:341
constexpr int kZeroPaddingFactor = 8;  // :340
}  // namespace

b200_pcps_acquisition_fine_doppler::b200_pcps_acquisition_fine_doppler(const Acq_Conf& conf)
    : acq_parameters(conf),
      d_fs_in(conf.fs_in),
      d_samples_per_ms(static_cast<int>(conf.samples_per_ms)),
      d_max_dwells(static_cast<int>(conf.max_dwells)),
      d_config_doppler_max(conf.doppler_max),
      d_fft_size(static_cast<int>(conf.samples_per_ms))
{
    // the reference's 50 ms buffer (:72) takes one block per dwell; more dwells than that write past its end there
    if (d_max_dwells < 1 || d_max_dwells > 50) throw std::invalid_argument("Acquisition max_dwells must be 1 ... 50 for the fine Doppler implementation");
    if (d_fft_size < 8) throw std::invalid_argument("Acquisition: less than 8 samples per ms");
    gb200_acq_config g{};
    g.device = conf.device;
    g.iq_dtype = GB200_IQ_CF32;  // the block takes gr_complex only (adapters/gps_l1_ca_pcps_acquisition_fine_doppler.cc:76-85)
    g.fft_size = d_fft_size;
    g.consumed_samples = d_fft_size;
    g.bit_transition = 0;
    g.use_cfar = 0;  // compute_CAF: first peak over second peak (:211-283)
    g.samples_per_chip = static_cast<int32_t>(conf.samples_per_chip);
    g.max_dwells = std::max(2, d_max_dwells);  // the grid is kept: dwells accumulate (:330)
    g.max_prns = 1;
    g.samples_per_code = static_cast<float>(d_fft_size);  // Acq_delay_samples = index_time (:277): no reduction below one block
    g.fs_in = static_cast<double>(conf.fs_in);
    if (gb200_acq_create(&g, &h_) != GB200_OK) fail("gb200_acq_create");
    // the fine estimate runs kZeroPaddingFactor * fold_ transforms of 10 * fft_size / fold_ points: the smallest fold the engine plans
    const float zero_hz = 0.0F;
    for (const int fold : {1, 2, 4, 5, 8, 10})
        {
            gb200_acq_config f = g;
            f.fft_size = f.consumed_samples = kPrnReplicas * d_fft_size / fold;
            f.max_dwells = 1;
            f.samples_per_code = static_cast<float>(f.fft_size);
            gb200_acq* h = nullptr;
            if ((kPrnReplicas * d_fft_size) % fold == 0 && gb200_acq_create(&f, &h) == GB200_OK)
                {
                    // a handle whose size the plan only serves through a padded transform has no fft_size-point transform: next fold
                    int64_t idx = 0;
                    float val = 0.0F;
                    std::vector<std::complex<float>> probe(static_cast<size_t>(f.fft_size), std::complex<float>(0.0F, 0.0F));
                    const std::complex<float> one(1.0F, 0.0F);
                    if (gb200_acq_set_doppler_bins(h, &zero_hz, 1) == GB200_OK &&
                        gb200_acq_spectrum_peak(h, reinterpret_cast<const gb200_cf32*>(probe.data()), f.fft_size, reinterpret_cast<const gb200_cf32*>(&one), 1, 1, &idx, &val) == GB200_OK)
                        {
                            h10_ = h;
                            fold_ = fold;
                            break;
                        }
                    gb200_acq_destroy(h);
                }
        }
    if (!h10_)
        {
            gb200_acq_destroy(h_);
            h_ = nullptr;
            fail("fine Doppler estimate: no transform plan for 10 code periods");
        }
    d_10_ms_buffer.assign(static_cast<size_t>(50) * d_samples_per_ms, std::complex<float>(0.0F, 0.0F));  // :72
}

b200_pcps_acquisition_fine_doppler::~b200_pcps_acquisition_fine_doppler()
{
    gb200_acq_destroy(h_);
    gb200_acq_destroy(h10_);
}

// :126-140 with update_carrier_wipeoff (:192-208): doppler_hz = doppler_step * index - doppler_max (integers), handed on as a float
void b200_pcps_acquisition_fine_doppler::set_doppler_step(unsigned int doppler_step)
{
    d_doppler_step = doppler_step;
    if (doppler_step == 0U) throw std::invalid_argument("Acquisition doppler_step must be positive");
    d_num_doppler_points = static_cast<int>(std::floor(std::abs(2 * d_config_doppler_max) / d_doppler_step));
    if (d_num_doppler_points < 1) throw std::invalid_argument("Acquisition: doppler_max / doppler_step leave no Doppler bin");
    std::vector<float> f(static_cast<size_t>(d_num_doppler_points));
    for (int i = 0; i < d_num_doppler_points; i++) f[static_cast<size_t>(i)] = static_cast<float>(static_cast<int>(d_doppler_step) * i - d_config_doppler_max);
    if (gb200_acq_set_doppler_bins(h_, f.data(), d_num_doppler_points) != GB200_OK) fail("gb200_acq_set_doppler_bins");
}

// :143-149: fft_size samples, transformed and conjugated on the device
void b200_pcps_acquisition_fine_doppler::set_local_code(std::complex<float>* code)
{
    if (gb200_acq_set_local_code(h_, 0, reinterpret_cast<const gb200_cf32*>(code), d_fft_size) != GB200_OK) fail("gb200_acq_set_local_code");
}

// :152-163
void b200_pcps_acquisition_fine_doppler::init()
{
    d_gnss_synchro->Flag_valid_acquisition = false;
    d_gnss_synchro->Flag_valid_symbol_output = false;
    d_gnss_synchro->Flag_valid_pseudorange = false;
    d_gnss_synchro->Flag_valid_word = false;
    d_gnss_synchro->Acq_doppler_step = 0U;
    d_gnss_synchro->Acq_delay_samples = 0.0;
    d_gnss_synchro->Acq_doppler_hz = 0.0;
    d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
    d_state = 0;
}

bool b200_pcps_acquisition_fine_doppler::start()
{
    d_sample_counter = 0ULL;  // :405-409
    return true;
}

// :412-436
void b200_pcps_acquisition_fine_doppler::set_state(int state)
{
    d_state = state;
    if (d_state == 1)
        {
            d_gnss_synchro->Acq_delay_samples = 0.0;
            d_gnss_synchro->Acq_doppler_hz = 0.0;
            d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
            d_gnss_synchro->Acq_doppler_step = 0U;
            d_well_count = 0;
            d_test_statistics = 0.0;
            d_active = true;
            reset_grid();
        }
}

// :300-335: one 1 ms block searched over the whole grid and added to the kept one; the engine also leaves the statistic of the grid so far
void b200_pcps_acquisition_fine_doppler::compute_and_accumulate_grid(const std::complex<float>* in)
{
    if (d_num_doppler_points < 1) throw std::logic_error("Acquisition: set_doppler_step was not called");
    const int32_t slot = 0;
    if (gb200_acq_run(h_, in, d_well_count, &slot, 1, &last_) != GB200_OK) fail("gb200_acq_run");
}

// :211-283 — the search of the accumulated grid ran on the device with the last dwell
float b200_pcps_acquisition_fine_doppler::compute_CAF()
{
    d_test_statistics = last_.test_statistics;
    d_gnss_synchro->Acq_delay_samples = static_cast<double>(last_.indext);
    // :278 as written there: int * unsigned int - int is UNSIGNED arithmetic, so a bin below zero Doppler comes out as 2^32 - |doppler|
    // (e.g. 4294964796 for -2500 Hz) and the fine estimate's 1 kHz plausibility test (:393) then never replaces it.  Kept: the block
    // is a drop-in, and the reference's own compiled block is the referee of the tests (tests/test_acq_fine_doppler_gpu.py).
    const int index_doppler = last_.doppler_index;
    d_gnss_synchro->Acq_doppler_hz = static_cast<double>(index_doppler * d_doppler_step - d_config_doppler_max);
    d_gnss_synchro->Acq_samplestamp_samples = d_sample_counter;
    d_gnss_synchro->Acq_doppler_step = d_doppler_step;
    return d_test_statistics;
}

// :338-402
void b200_pcps_acquisition_fine_doppler::estimate_Doppler()
{
    const int signal_samples = kPrnReplicas * d_fft_size;
    const int fft_size_extended = signal_samples * kZeroPaddingFactor;
    // one period of the sampled code (gps_l1_ca_code_gen_complex_sampled fills samples_per_code entries of the span, :349-351)
    std::vector<std::complex<float>> code_replica(static_cast<size_t>(d_fft_size), std::complex<float>(0.0F, 0.0F));
    const int spc = static_cast<int>(static_cast<double>(d_fs_in) / (1023000.0 / 1023.0));
    gps_l1_ca_code_gen_complex_sampled(code_replica.data(), std::min(spc, d_fft_size), d_gnss_synchro->PRN, static_cast<int32_t>(d_fs_in), 0);
    const int shift_index = static_cast<int>(d_gnss_synchro->Acq_delay_samples);
    // :355-358: std::rotate over [0, fft_size - 1) — the last sample of the period keeps its place
    if (shift_index != 0) std::rotate(code_replica.data(), code_replica.data() + (d_fft_size - shift_index), code_replica.data() + d_fft_size - 1);
    int64_t idx = 0;
    float val = 0.0F;
    if (gb200_acq_spectrum_peak(h10_, reinterpret_cast<const gb200_cf32*>(d_10_ms_buffer.data()), signal_samples, reinterpret_cast<const gb200_cf32*>(code_replica.data()), d_fft_size,
            kZeroPaddingFactor, &idx, &val) != GB200_OK)
        fail("gb200_acq_spectrum_peak");
    fine_index_ = idx;
    // :376-391: the bin's frequency, double arithmetic rounded to float (the table entry of fftFreqBins)
    const int half = fft_size_extended / 2;
    float freq;
    if (idx < half)
        freq = static_cast<float>(((static_cast<float>(d_fs_in) / 2.0) * static_cast<float>(idx)) / (static_cast<float>(fft_size_extended) / 2.0));
    else
        freq = static_cast<float>(((-static_cast<float>(d_fs_in) / 2.0) * static_cast<float>(fft_size_extended - idx)) / (static_cast<float>(fft_size_extended) / 2.0));
    // :393-401
    if (std::abs(freq - d_gnss_synchro->Acq_doppler_hz) < 1000) d_gnss_synchro->Acq_doppler_hz = static_cast<double>(freq);
}

// :450-563
int b200_pcps_acquisition_fine_doppler::general_work(int noutput_items, const std::complex<float>* in, int ninput_items)
{
    int consumed = 0;
    const auto consume_each = [&](int n) {
        if (n > ninput_items) throw std::logic_error("Acquisition: the scheduler handed over fewer items than the block consumes");
        consumed = n;
    };
    switch (d_state)
        {
        case 0:  // stand-by
            if (d_active)
                {
                    reset_grid();
                    d_n_samples_in_buffer = 0;
                    d_state = 1;
                }
            if (!acq_parameters.blocking_on_standby)
                {
                    d_sample_counter += static_cast<uint64_t>(d_fft_size);
                    consume_each(d_fft_size);
                }
            break;
        case 1:  // grid of one block
            if (ninput_items < d_fft_size) throw std::logic_error("Acquisition: forecast() asks for fft_size items");
            compute_and_accumulate_grid(in);
            std::copy(in, in + d_fft_size, &d_10_ms_buffer[static_cast<size_t>(d_n_samples_in_buffer)]);
            d_n_samples_in_buffer += d_fft_size;
            d_well_count++;
            if (d_well_count >= d_max_dwells) d_state = 2;
            d_sample_counter += static_cast<uint64_t>(d_fft_size);
            consume_each(d_fft_size);
            break;
        case 2:  // decision
            d_test_statistics = compute_CAF();
            if (d_test_statistics > d_threshold)
                d_state = 3;
            else
                {
                    d_state = 5;
                    d_n_samples_in_buffer = 0;
                }
            break;
        case 3:  // the rest of the 10 ms, then the fine estimate
            {
                const int samples_remaining = 10 * d_samples_per_ms - d_n_samples_in_buffer;
                if (samples_remaining > noutput_items)
                    {
                        if (noutput_items > ninput_items) throw std::logic_error("Acquisition: noutput_items exceeds the items at the input");
                        std::copy(in, in + noutput_items, &d_10_ms_buffer[static_cast<size_t>(d_n_samples_in_buffer)]);
                        d_n_samples_in_buffer += noutput_items;
                        d_sample_counter += static_cast<uint64_t>(noutput_items);
                        consume_each(noutput_items);
                    }
                else
                    {
                        if (samples_remaining > 0)
                            {
                                if (samples_remaining > ninput_items) throw std::logic_error("Acquisition: fewer items at the input than the 10 ms buffer needs");
                                std::copy(in, in + samples_remaining, &d_10_ms_buffer[static_cast<size_t>(d_n_samples_in_buffer)]);
                                d_sample_counter += static_cast<uint64_t>(samples_remaining);
                                consume_each(samples_remaining);
                            }
                        estimate_Doppler();
                        d_n_samples_in_buffer = 0;
                        d_state = 4;
                    }
            }
            break;
        case 4:  // positive
            d_positive_acq = 1;
            d_active = false;
            if (event_sink_) event_sink_(1);
            d_state = 0;
            if (!acq_parameters.blocking_on_standby)
                {
                    d_sample_counter += static_cast<uint64_t>(noutput_items);
                    consume_each(noutput_items);
                }
            break;
        case 5:  // negative
            d_positive_acq = 0;
            d_active = false;
            if (event_sink_) event_sink_(2);
            d_state = 0;
            if (!acq_parameters.blocking_on_standby)
                {
                    d_sample_counter += static_cast<uint64_t>(noutput_items);
                    consume_each(noutput_items);
                }
            break;
        default:
            d_state = 0;
            if (!acq_parameters.blocking_on_standby)
                {
                    d_sample_counter += static_cast<uint64_t>(noutput_items);
                    consume_each(noutput_items);
                }
            break;
        }
    return consumed;
}
}  // namespace gb200h
