#include "b200_e5a_noncoherent_iq_acquisition_caf.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <stdexcept>

namespace gb200h
{
namespace
{
[[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + gb200_last_error()); }
}  // namespace

// galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc:55-134
b200_e5a_noncoherent_iq_acquisition_caf::b200_e5a_noncoherent_iq_acquisition_caf(unsigned int sampled_ms, unsigned int max_dwells, unsigned int doppler_max, int64_t fs_in,
    int samples_per_ms, int samples_per_code, bool bit_transition_flag, bool both_signal_components, int CAF_window_hz, int Zero_padding, int device)
    : d_fs_in(fs_in),
      d_samples_per_ms(samples_per_ms),
      d_samples_per_code(samples_per_code),
      d_CAF_window_hz(CAF_window_hz),
      d_doppler_max(static_cast<int>(doppler_max)),
      d_fft_size(static_cast<int>(sampled_ms) * samples_per_ms),
      d_max_dwells(max_dwells),
      d_sampled_ms(Zero_padding > 0 ? 1U : sampled_ms),  // :107-114
      d_bit_transition_flag(bit_transition_flag),
      d_both_signal_components(both_signal_components)
{
    if (d_fft_size < 8 || samples_per_code < 1 || samples_per_code > d_fft_size) throw std::invalid_argument("E5a CAF acquisition: bad block / code length");
    // slots of the engine sweep: I_A always; Q_A with both components; I_B / Q_B when more than one code period is integrated (:116-131)
    if (d_both_signal_components) slot_qa_ = n_slots_++;
    if (d_sampled_ms > 1)
        {
            slot_ib_ = n_slots_++;
            if (d_both_signal_components) slot_qb_ = n_slots_++;
        }
    gb200_acq_config g{};
    g.device = device;
    g.iq_dtype = GB200_IQ_CF32;  // gr_complex only (adapters/galileo_e5a_noncoherent_iq_acquisition_caf.cc:94-106)
    g.fft_size = d_fft_size;
    g.consumed_samples = d_fft_size;
    g.bit_transition = 0;
    g.use_cfar = 0;
    g.samples_per_chip = 1;
    g.max_dwells = 1;
    g.max_prns = n_slots_;
    g.samples_per_code = static_cast<float>(d_fft_size);
    g.fs_in = static_cast<double>(fs_in);
    if (gb200_acq_create(&g, &h_) != GB200_OK) fail("gb200_acq_create");
    if (gb200_acq_keep_grid(h_, 1) != GB200_OK)
        {
            gb200_acq_destroy(h_);
            h_ = nullptr;
            fail("gb200_acq_keep_grid");
        }
    d_inbuffer.assign(static_cast<size_t>(d_fft_size), std::complex<float>(0.0F, 0.0F));
}

b200_e5a_noncoherent_iq_acquisition_caf::~b200_e5a_noncoherent_iq_acquisition_caf() { gb200_acq_destroy(h_); }

// :160-206.  The reference builds every replica in the SAME transform input buffer: the B variants overwrite its first samples_per_code
// entries with the sign-flipped code and transform whatever the buffer held behind them — codeQ when both components are searched (the
// Q_A copy came last), codeI otherwise.  Reproduced as built there.
void b200_e5a_noncoherent_iq_acquisition_caf::set_local_code(std::complex<float>* codeI, std::complex<float>* codeQ)
{
    const auto n = static_cast<size_t>(d_fft_size);
    std::vector<std::complex<float>> inbuf(codeI, codeI + n);
    const auto install = [&](int slot) {
        if (gb200_acq_set_local_code(h_, slot, reinterpret_cast<const gb200_cf32*>(inbuf.data()), d_fft_size) != GB200_OK) fail("gb200_acq_set_local_code");
    };
    install(0);
    if (d_both_signal_components)
        {
            std::copy(codeQ, codeQ + n, inbuf.begin());
            install(slot_qa_);
        }
    if (d_sampled_ms > 1)
        {
            for (int i = 0; i < d_samples_per_code; i++) inbuf[static_cast<size_t>(i)] = codeI[i] * std::complex<float>(-1.0F, 0.0F);
            install(slot_ib_);
            if (d_both_signal_components)
                {
                    for (int i = 0; i < d_samples_per_code; i++) inbuf[static_cast<size_t>(i)] = codeQ[i] * std::complex<float>(-1.0F, 0.0F);
                    install(slot_qb_);
                }
        }
}

// :209-254
void b200_e5a_noncoherent_iq_acquisition_caf::init()
{
    d_gnss_synchro->Flag_valid_acquisition = false;
    d_gnss_synchro->Flag_valid_symbol_output = false;
    d_gnss_synchro->Flag_valid_pseudorange = false;
    d_gnss_synchro->Flag_valid_word = false;
    d_gnss_synchro->Acq_delay_samples = 0.0;
    d_gnss_synchro->Acq_doppler_hz = 0.0;
    d_gnss_synchro->Acq_doppler_step = 0U;
    d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
    d_mag = 0.0;
    d_input_power = 0.0;
    if (d_doppler_step <= 0) throw std::invalid_argument("E5a CAF acquisition: doppler_step must be positive");
    d_num_doppler_bins = 0;
    for (int doppler = -d_doppler_max; doppler <= d_doppler_max; doppler += d_doppler_step) d_num_doppler_bins++;
    std::vector<float> f(static_cast<size_t>(d_num_doppler_bins));
    for (int k = 0; k < d_num_doppler_bins; k++) f[static_cast<size_t>(k)] = static_cast<float>(-d_doppler_max + d_doppler_step * k);  // :236-239
    if (gb200_acq_set_doppler_bins(h_, f.data(), d_num_doppler_bins) != GB200_OK) fail("gb200_acq_set_doppler_bins");
    if (d_CAF_window_hz > 0)
        {
            if (d_CAF_window_hz / (2 * d_doppler_step) < 1) throw std::invalid_argument("E5a CAF acquisition: CAF_window_hz below two Doppler steps (the reference divides by zero there)");
            if (2 * (d_CAF_window_hz / (2 * d_doppler_step)) > d_num_doppler_bins) throw std::invalid_argument("E5a CAF acquisition: CAF window wider than the Doppler grid");
            d_CAF_vector.assign(static_cast<size_t>(d_num_doppler_bins), 0.0F);
            d_CAF_vector_I.assign(static_cast<size_t>(d_num_doppler_bins), 0.0F);
            if (d_both_signal_components) d_CAF_vector_Q.assign(static_cast<size_t>(d_num_doppler_bins), 0.0F);
        }
}

// :257-280
void b200_e5a_noncoherent_iq_acquisition_caf::set_state(int state)
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
        }
}

void b200_e5a_noncoherent_iq_acquisition_caf::search()
{
    const int nb = d_num_doppler_bins;
    if (nb < 1) throw std::logic_error("E5a CAF acquisition: init() was not called");
    const float fft_normalization_factor = static_cast<float>(d_fft_size) * static_cast<float>(d_fft_size);
    const float norm2 = fft_normalization_factor * fft_normalization_factor;
    d_input_power = 0.0;
    d_mag = 0.0;
    d_well_count++;
    // :378-381: mean |x|^2 of the buffered block, summed in sample order like volk_32f_accumulator_s32f (fft_size additions: host)
    for (int i = 0; i < d_fft_size; i++)
        {
            const float re = d_inbuffer[static_cast<size_t>(i)].real(), im = d_inbuffer[static_cast<size_t>(i)].imag();
            d_input_power += (re * re) + (im * im);
        }
    d_input_power /= static_cast<float>(d_fft_size);

    // every replica x every Doppler bin in one sweep; per (replica, bin): maximum and its first index (:383-437)
    std::vector<int32_t> slots(static_cast<size_t>(n_slots_));
    for (int s = 0; s < n_slots_; s++) slots[static_cast<size_t>(s)] = s;
    std::vector<gb200_acq_result> res(static_cast<size_t>(n_slots_));
    if (gb200_acq_run(h_, d_inbuffer.data(), 0, slots.data(), n_slots_, res.data()) != GB200_OK) fail("gb200_acq_run");
    const auto cells = static_cast<size_t>(n_slots_) * nb;
    std::vector<float> vmax(cells), sum(cells);
    std::vector<int32_t> imax(cells);
    if (gb200_acq_read_cells(h_, n_slots_, vmax.data(), imax.data(), sum.data()) != GB200_OK) fail("gb200_acq_read_cells");
    const auto v_of = [&](int slot, int bin) { return vmax[static_cast<size_t>(slot) * nb + bin]; };
    const auto i_of = [&](int slot, int bin) { return imax[static_cast<size_t>(slot) * nb + bin]; };
    // :435: magt_QB is read from the I_B grid at Q_B's index
    std::vector<float> ib_at_qb;
    if (slot_qb_ >= 0)
        {
            std::vector<int32_t> ps(static_cast<size_t>(nb), slot_ib_), pb(static_cast<size_t>(nb)), pi(static_cast<size_t>(nb));
            for (int b = 0; b < nb; b++) pb[static_cast<size_t>(b)] = b, pi[static_cast<size_t>(b)] = i_of(slot_qb_, b);
            ib_at_qb.resize(static_cast<size_t>(nb));
            if (gb200_acq_grid_probe(h_, nb, ps.data(), pb.data(), pi.data(), ib_at_qb.data()) != GB200_OK) fail("gb200_acq_grid_probe");
        }
    // per bin: the I replica and the Q replica that enter the non-coherent sum (:443-531)
    std::vector<int32_t> sel_i(static_cast<size_t>(nb), 0), sel_q(static_cast<size_t>(nb), -1);
    for (int b = 0; b < nb; b++)
        {
            const float magt_IA = v_of(0, b) / norm2;
            if (d_sampled_ms > 1)
                {
                    const float magt_IB = v_of(slot_ib_, b) / norm2;
                    sel_i[static_cast<size_t>(b)] = magt_IA >= magt_IB ? 0 : slot_ib_;
                    if (d_both_signal_components)
                        {
                            const float magt_QA = v_of(slot_qa_, b) / norm2;
                            const float magt_QB = ib_at_qb[static_cast<size_t>(b)] / norm2;
                            sel_q[static_cast<size_t>(b)] = magt_QA >= magt_QB ? slot_qa_ : slot_qb_;
                        }
                }
            else if (d_both_signal_components)
                sel_q[static_cast<size_t>(b)] = slot_qa_;
            if (d_CAF_window_hz > 0)
                {
                    d_CAF_vector_I[static_cast<size_t>(b)] = v_of(sel_i[static_cast<size_t>(b)], b);
                    if (d_both_signal_components) d_CAF_vector_Q[static_cast<size_t>(b)] = v_of(sel_q[static_cast<size_t>(b)], b);
                }
        }
    std::vector<float> pv(static_cast<size_t>(nb));
    std::vector<int32_t> pi(static_cast<size_t>(nb));
    if (d_both_signal_components)
        {
            if (gb200_acq_grid_pair_max(h_, sel_i.data(), sel_q.data(), pv.data(), pi.data()) != GB200_OK) fail("gb200_acq_grid_pair_max");
        }
    else
        for (int b = 0; b < nb; b++) pv[static_cast<size_t>(b)] = v_of(sel_i[static_cast<size_t>(b)], b), pi[static_cast<size_t>(b)] = i_of(sel_i[static_cast<size_t>(b)], b);
    // :533-545, bins in ascending order
    for (int b = 0; b < nb; b++)
        {
            const int doppler = -d_doppler_max + d_doppler_step * b;
            const float magt = pv[static_cast<size_t>(b)] / norm2;
            if (d_mag < magt)
                {
                    d_mag = magt;
                    if (d_test_statistics < (d_mag / d_input_power) || !d_bit_transition_flag)
                        {
                            d_gnss_synchro->Acq_delay_samples = static_cast<double>(static_cast<uint32_t>(pi[static_cast<size_t>(b)]) % d_samples_per_code);
                            d_gnss_synchro->Acq_doppler_hz = static_cast<double>(doppler);
                            d_gnss_synchro->Acq_samplestamp_samples = d_sample_counter;
                            d_gnss_synchro->Acq_doppler_step = static_cast<uint32_t>(d_doppler_step);
                            d_test_statistics = d_mag / d_input_power;
                        }
                }
        }
    if (d_CAF_window_hz <= 0) return;

    // CAF filter over the Doppler bins (:590-690): a triangular window of half-width h bins over the per-bin maxima of the I (and Q)
    // component, in three stretches (lower edge, interior, upper edge) each with its own normaliser.  The weights are the reference's, sign
    // conventions included (the I sums of the first two stretches use (bin - i) without the absolute value).
    const int h = d_CAF_window_hz / (2 * d_doppler_step);
    const float w = 0.5F / static_cast<float>(h);
    const auto wsum = [&](const std::vector<float>& v, int b, int lo, int hi, bool use_abs) {
        float acc = 0;
        for (int i = lo; i < hi; i++) acc += v[static_cast<size_t>(i)] * (1.0F - w * static_cast<float>(use_abs ? std::abs(b - i) : (b - i)));
        return acc;
    };
    for (int b = 0; b < h; b++)
        {
            float c = wsum(d_CAF_vector_I, b, 0, h + b + 1, false);
            c /= 1.0F + static_cast<float>(h + b) - w * static_cast<float>(h) * ((static_cast<float>(h) + 1.0F) / 2.0F) - w * static_cast<float>(b) * (static_cast<float>(b) + 1.0F) / 2.0F;
            if (d_both_signal_components)
                {
                    float a = wsum(d_CAF_vector_Q, b, 0, h + b + 1, true);
                    a /= 1.0F + static_cast<float>(h + b) - w * static_cast<float>(h) * static_cast<float>(h + 1) / 2.0F - w * static_cast<float>(b) * static_cast<float>(b + 1) / 2.0F;
                    c += a;
                }
            d_CAF_vector[static_cast<size_t>(b)] = c;
        }
    for (int b = h; b < nb - h; b++)
        {
            float c = wsum(d_CAF_vector_I, b, b - h, b + h + 1, false);
            const float den = 1.0F + 2.0F * static_cast<float>(h) - 2.0F * w * static_cast<float>(h) * static_cast<float>(h + 1) / 2.0F;
            c /= den;
            if (d_both_signal_components)
                {
                    float a = wsum(d_CAF_vector_Q, b, b - h, b + h + 1, false);
                    a /= den;
                    c += a;
                }
            d_CAF_vector[static_cast<size_t>(b)] = c;
        }
    for (int b = nb - h; b < nb; b++)
        {
            float c = wsum(d_CAF_vector_I, b, b - h, nb, true);
            c /= 1.0F + static_cast<float>(h) + static_cast<float>(nb - b - 1) - w * static_cast<float>(h) * (static_cast<float>(h) + 1.0F) / 2.0F - w * (nb - b - 1) * static_cast<float>(nb - b) / 2.0F;
            if (d_both_signal_components)
                {
                    float a = wsum(d_CAF_vector_Q, b, b - h, nb, true);
                    // (this one normaliser is evaluated in double there: the literals 1.0 and 2.0 of :680)
                    a /= static_cast<float>(1.0F + static_cast<float>(h) + static_cast<float>(nb - b - 1) - w * static_cast<float>(h) * static_cast<float>(h + 1.0) / 2.0 - w * static_cast<float>(nb - b - 1) * static_cast<float>(nb - b) / 2.0);
                    c += a;
                }
            d_CAF_vector[static_cast<size_t>(b)] = c;
        }
    // :687-689: the Doppler of the largest filtered value replaces the one of the peak (first maximum)
    int best = 0;
    for (int b = 1; b < nb; b++)
        if (d_CAF_vector[static_cast<size_t>(b)] > d_CAF_vector[static_cast<size_t>(best)]) best = b;
    d_gnss_synchro->Acq_doppler_hz = static_cast<double>(-d_doppler_max + d_doppler_step * best);
}

// :295-777
int b200_e5a_noncoherent_iq_acquisition_caf::general_work(const std::complex<float>* in, int ninput_items)
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
            d_sample_counter += static_cast<uint64_t>(ninput_items);
            consumed = ninput_items;
            break;
        case 1:
            {
                const int buff_increment = (ninput_items + d_buffer_count) <= d_fft_size ? ninput_items : d_fft_size - d_buffer_count;
                std::copy(in, in + buff_increment, d_inbuffer.begin() + d_buffer_count);
                if (d_buffer_count >= d_fft_size - d_gr_stream_buffer) d_state = 2;  // (tested BEFORE the count moves: the search starts with the next call)
                d_buffer_count += buff_increment;
                d_sample_counter += static_cast<uint64_t>(buff_increment);
                consumed = buff_increment;
                break;
            }
        case 2:
            {
                const int missing = d_fft_size - d_buffer_count;
                if (missing > 0)
                    {
                        if (missing > ninput_items) throw std::logic_error("E5a CAF acquisition: fewer items at the input than the buffer misses");
                        std::copy(in, in + missing, d_inbuffer.begin() + d_buffer_count);
                    }
                d_sample_counter += static_cast<uint64_t>(missing);
                search();
                if (d_well_count == d_max_dwells)
                    d_state = d_test_statistics > d_threshold ? 3 : 4;
                else
                    d_state = 1;
                consumed = missing;
                d_buffer_count = 0;
                break;
            }
        case 3:
            d_active = false;
            d_state = 0;
            if (event_sink_) event_sink_(1);
            d_sample_counter += static_cast<uint64_t>(ninput_items);
            consumed = ninput_items;
            break;
        case 4:
            d_active = false;
            d_state = 0;
            d_sample_counter += static_cast<uint64_t>(ninput_items);
            consumed = ninput_items;
            if (event_sink_) event_sink_(2);
            break;
        default:
            break;
        }
    return consumed;
}
}  // namespace gb200h
