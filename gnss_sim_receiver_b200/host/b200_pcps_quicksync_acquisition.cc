#include "b200_pcps_quicksync_acquisition.h"
#include <algorithm>
#include <array>
#include <cmath>
#include <stdexcept>

namespace gb200h
{
namespace
{
[[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + gb200_last_error()); }
}  // namespace

// pcps_quicksync_acquisition_cc.cc:54-112
b200_pcps_quicksync_acquisition::b200_pcps_quicksync_acquisition(uint32_t folding_factor, uint32_t sampled_ms, uint32_t max_dwells, uint32_t doppler_max, int64_t fs_in,
    int32_t samples_per_ms, int32_t samples_per_code, bool bit_transition_flag, int device)
    : d_fs_in(fs_in),
      d_samples_per_ms(samples_per_ms),
      d_samples_per_code(samples_per_code),
      d_folding_factor(folding_factor),
      d_doppler_max(doppler_max),
      d_sampled_ms(sampled_ms),
      d_max_dwells(max_dwells),
      d_fft_size(folding_factor ? static_cast<uint32_t>(samples_per_code) / folding_factor : 0U),
      device_(device),
      d_bit_transition_flag(bit_transition_flag)
{
    // folding_factor 1 makes the reference read past its wiped block when it resolves the delay (:339), and its candidate accumulator
    // holds 100 entries (:330)
    if (folding_factor < 2 || folding_factor > 100) throw std::invalid_argument("QuickSync acquisition: folding_factor must be 2 ... 100");
    if (d_fft_size < 8) throw std::invalid_argument("QuickSync acquisition: samples_per_code / folding_factor below 8");
    gb200_acq_config g{};
    g.device = device;
    g.iq_dtype = GB200_IQ_CF32;
    g.fft_size = static_cast<int32_t>(d_fft_size);
    g.consumed_samples = g.fft_size;
    g.bit_transition = 0;
    g.use_cfar = 0;
    g.samples_per_chip = 1;
    g.max_dwells = 1;
    g.max_prns = 1;
    g.samples_per_code = static_cast<float>(d_fft_size);
    g.fs_in = static_cast<double>(fs_in);
    if (gb200_acq_create(&g, &h_) != GB200_OK) fail("gb200_acq_create");
    d_code.assign(static_cast<size_t>(d_samples_per_code), std::complex<float>(0.0F, 0.0F));
}

void b200_pcps_quicksync_acquisition::set_item_length(int n)
{
    if (n < d_samples_per_code * static_cast<int>(d_folding_factor)) throw std::invalid_argument("QuickSync acquisition: vector items shorter than samples_per_code * folding_factor");
    item_length_ = n;
}

b200_pcps_quicksync_acquisition::~b200_pcps_quicksync_acquisition() { gb200_acq_destroy(h_); }

// :133-156: the unfolded replica is kept for the delay resolution; the transform gets the replica folded folding_factor times
void b200_pcps_quicksync_acquisition::set_local_code(std::complex<float>* code)
{
    std::copy(code, code + d_samples_per_code, d_code.data());
    std::vector<std::complex<float>> folded(d_fft_size, std::complex<float>(0.0F, 0.0F));
    for (uint32_t i = 0; i < d_folding_factor; i++)
        for (uint32_t k = 0; k < d_fft_size; k++) folded[k] += code[i * d_fft_size + k];
    if (gb200_acq_set_local_code(h_, 0, reinterpret_cast<const gb200_cf32*>(folded.data()), static_cast<int>(d_fft_size)) != GB200_OK) fail("gb200_acq_set_local_code");
}

// :159-194
void b200_pcps_quicksync_acquisition::init()
{
    d_gnss_synchro->Flag_valid_acquisition = false;
    d_gnss_synchro->Flag_valid_symbol_output = false;
    d_gnss_synchro->Flag_valid_pseudorange = false;
    d_gnss_synchro->Flag_valid_word = false;
    d_gnss_synchro->Acq_delay_samples = 0.0;
    d_gnss_synchro->Acq_doppler_hz = 0.0;
    d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
    d_gnss_synchro->Acq_doppler_step = 0U;
    d_mag = 0.0;
    d_input_power = 0.0;
    if (d_doppler_step == 0) d_doppler_step = 250;
    d_num_doppler_bins = 0;
    for (auto doppler = static_cast<int32_t>(-d_doppler_max); doppler <= static_cast<int32_t>(d_doppler_max); doppler += d_doppler_step) d_num_doppler_bins++;
    std::vector<float> f(d_num_doppler_bins);
    // :188-189: phase_step_rad = float(TWO_PI) * doppler / float(fs) with an int doppler: the engine forms the same product from the float
    for (uint32_t k = 0; k < d_num_doppler_bins; k++) f[k] = static_cast<float>(-static_cast<int32_t>(d_doppler_max) + static_cast<int32_t>(d_doppler_step * k));
    if (gb200_acq_set_doppler_bins(h_, f.data(), static_cast<int>(d_num_doppler_bins)) != GB200_OK) fail("gb200_acq_set_doppler_bins");
    if (gb200_acq_set_input_folding(h_, static_cast<int>(d_folding_factor * d_folding_factor), d_samples_per_code * static_cast<int>(d_folding_factor)) != GB200_OK) fail("gb200_acq_set_input_folding");
}

// :197-219
void b200_pcps_quicksync_acquisition::set_state(int32_t state)
{
    d_state = state;
    if (d_state == 1)
        {
            d_gnss_synchro->Acq_delay_samples = 0.0;
            d_gnss_synchro->Acq_doppler_hz = 0.0;
            d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
            d_gnss_synchro->Acq_doppler_step = 0U;
            d_well_count = 0;
            d_mag = 0.0;
            d_input_power = 0.0;
            d_test_statistics = 0.0;
            d_active = true;
        }
}

// state 1 (:262-430): one item
void b200_pcps_quicksync_acquisition::dwell(const std::complex<float>* in)
{
    if (d_num_doppler_bins < 1) throw std::logic_error("QuickSync acquisition: init() was not called");
    const float fft_normalization_factor = static_cast<float>(d_fft_size) * static_cast<float>(d_fft_size);
    d_input_power = 0.0;
    d_mag = 0.0;
    d_test_statistics = 0.0;
    d_sample_counter += static_cast<uint64_t>(d_sampled_ms) * d_samples_per_ms;
    d_well_count++;
    // :291-296: mean |x|^2 over the item, summed in sample order like volk_32f_accumulator_s32f
    const int n = d_samples_per_code * static_cast<int>(d_folding_factor);
    for (int i = 0; i < n; i++)
        {
            const float re = in[i].real(), im = in[i].imag();
            d_input_power += (re * re) + (im * im);
        }
    d_input_power /= static_cast<float>(n);

    const int32_t slot = 0;
    gb200_acq_result res{};
    if (gb200_acq_run(h_, in, 0, &slot, 1, &res) != GB200_OK) fail("gb200_acq_run");
    const uint32_t nb = d_num_doppler_bins;
    std::vector<float> vmax(nb), sum(nb);
    std::vector<int32_t> imax(nb);
    if (gb200_acq_read_cells(h_, 1, vmax.data(), imax.data(), sum.data()) != GB200_OK) fail("gb200_acq_read_cells");
    // :298-365: bins in ascending order; every strictly larger peak re-resolves the delay and overwrites the results, so the last such
    // bin — the first bin that holds the largest peak — is the one whose candidates are correlated
    int win = -1;
    for (uint32_t b = 0; b < nb; b++)
        {
            const float magt = vmax[b] / (fft_normalization_factor * fft_normalization_factor);
            if (d_mag < magt)
                {
                    d_mag = magt;
                    if (d_test_statistics < (d_mag / d_input_power) || !d_bit_transition_flag)
                        {
                            win = static_cast<int>(b);
                            d_test_statistics = d_mag / d_input_power;
                        }
                }
        }
    if (win < 0) return;
    const uint32_t detected_delay_samples_folded = static_cast<uint32_t>(imax[static_cast<size_t>(win)]) % static_cast<uint32_t>(d_samples_per_code);
    std::vector<std::complex<float>> acc(d_folding_factor);
    if (gb200_acq_folded_candidates(h_, win, static_cast<int>(detected_delay_samples_folded), reinterpret_cast<const gb200_cf32*>(d_code.data()), d_samples_per_code,
            static_cast<int>(d_folding_factor), reinterpret_cast<gb200_cf32*>(acc.data())) != GB200_OK)
        fail("gb200_acq_folded_candidates");
    uint32_t best = 0;
    float best_v = (acc[0].real() * acc[0].real()) + (acc[0].imag() * acc[0].imag());
    for (uint32_t i = 1; i < d_folding_factor; i++)
        {
            const float v = (acc[i].real() * acc[i].real()) + (acc[i].imag() * acc[i].imag());
            if (v > best_v)
                {
                    best_v = v;
                    best = i;
                }
        }
    d_gnss_synchro->Acq_delay_samples = static_cast<double>(detected_delay_samples_folded + best * d_fft_size);
    d_gnss_synchro->Acq_doppler_hz = static_cast<double>(-static_cast<int32_t>(d_doppler_max) + static_cast<int32_t>(d_doppler_step * static_cast<uint32_t>(win)));
    d_gnss_synchro->Acq_samplestamp_samples = d_sample_counter;
    d_gnss_synchro->Acq_doppler_step = d_doppler_step;
}

// :214-550
int b200_pcps_quicksync_acquisition::general_work(const std::complex<float>* in, int ninput_items)
{
    int consumed = 0;
    switch (d_state)
        {
        case 0:
            if (d_active)
                {
                    d_gnss_synchro->Acq_delay_samples = 0.0;
                    d_gnss_synchro->Acq_doppler_hz = 0.0;
                    d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
                    d_gnss_synchro->Acq_doppler_step = 0U;
                    d_well_count = 0;
                    d_mag = 0.0;
                    d_input_power = 0.0;
                    d_test_statistics = 0.0;
                    d_state = 1;
                }
            d_sample_counter += static_cast<uint64_t>(d_sampled_ms) * d_samples_per_ms * ninput_items;
            consumed = ninput_items;
            break;
        case 1:
            if (ninput_items < 1) throw std::logic_error("QuickSync acquisition: no item at the input");
            dwell(in);
            if (!d_bit_transition_flag)
                {
                    if (d_test_statistics > d_threshold)
                        d_state = 2;
                    else if (d_well_count == d_max_dwells)
                        d_state = 3;
                }
            else if (d_well_count == d_max_dwells)
                d_state = d_test_statistics > d_threshold ? 2 : 3;
            consumed = 1;
            break;
        case 2:
            d_active = false;
            d_state = 0;
            d_sample_counter += static_cast<uint64_t>(d_sampled_ms) * d_samples_per_ms * ninput_items;
            consumed = ninput_items;
            if (event_sink_) event_sink_(1);
            break;
        case 3:
            d_active = false;
            d_state = 0;
            d_sample_counter += static_cast<uint64_t>(d_sampled_ms) * d_samples_per_ms * ninput_items;
            consumed = ninput_items;
            if (event_sink_) event_sink_(2);
            break;
        default:
            break;
        }
    return consumed;
}
}  // namespace gb200h
