// TEST INFRASTRUCTURE — not part of the product.
// C door to the reference's OWN Galileo E5a non-coherent I/Q acquisition block with CAF filtering:
// src/algorithms/acquisition/gnuradio_blocks/galileo_e5a_noncoherent_iq_acquisition_caf_cc.{h,cc}, compiled UNMODIFIED from /root/reference
// against the stand-in declarations of oracle/shim/acq/ (README.txt there; the DFT behind gr::fft is oracle/oracle_fft.c).  The block is
// made, fed its replicas and driven the way adapters/galileo_e5a_noncoherent_iq_acquisition_caf.cc:40-261 does (coherent time capped at
// 3 ms, Zero_padding -> 2 ms with 1 ms of code, "5X" -> separate 5I / 5Q replicas, else one 5X-composed replica), the replicas coming
// from the reference's own galileo_e5_a_code_gen_complex_sampled.
#include <algorithm>
#include <any>
#include <array>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <deque>
#include <fstream>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <queue>
#include <span>
#include <sstream>
#include <string>
#include <thread>
#include <vector>
#define private public
#define protected public
#include "galileo_e5a_noncoherent_iq_acquisition_caf_cc.h"
#undef private
#undef protected
#include "Galileo_E5a.h"
#include "galileo_e5_signal_replica.h"
#include "gnss_synchro.h"

namespace
{
struct RefAcqCaf
{
    galileo_e5a_noncoherentIQ_acquisition_caf_cc_sptr block;
    Gnss_Synchro synchro{};
    std::vector<std::complex<float>> codeI, codeQ;
    int64_t fs_in{0};
    int code_length{0}, zero_padding{0};
    unsigned sampled_ms{1};
};
}  // namespace

extern "C" {
struct ref_acqcaf_status
{
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    int64_t consumed_total;
    float test_statistics, threshold, mag, input_power;
    int32_t state, active, num_doppler_bins, fft_size, well_count, n_events, events[8];
    uint32_t acq_doppler_step;
};

void* ref_acqcaf_new(long long fs_in, unsigned sampled_ms, unsigned max_dwells, unsigned doppler_max, int bit_transition, int caf_window_hz, int zero_padding, unsigned prn,
    const char* signal, const char* channel_signal)
{
    auto* r = new RefAcqCaf;
    r->fs_in = fs_in;
    if (sampled_ms > 3) sampled_ms = 3;
    if (zero_padding > 0) sampled_ms = 2;
    r->sampled_ms = sampled_ms;
    r->zero_padding = zero_padding;
    r->code_length = static_cast<int>(std::round(static_cast<double>(fs_in) / GALILEO_E5A_CODE_CHIP_RATE_CPS * static_cast<double>(GALILEO_E5A_CODE_LENGTH_CHIPS)));
    const unsigned vector_length = static_cast<unsigned>(r->code_length) * sampled_ms;
    r->codeI.assign(vector_length, std::complex<float>(0.0F, 0.0F));
    r->codeQ.assign(vector_length, std::complex<float>(0.0F, 0.0F));
    const bool both = channel_signal[0] == '5' && channel_signal[1] == 'X';
    r->block = galileo_e5a_noncoherentIQ_make_acquisition_caf_cc(sampled_ms, max_dwells, doppler_max, fs_in, r->code_length, r->code_length, bit_transition != 0, false, "", both,
        caf_window_hz, zero_padding, false);
    r->synchro.PRN = prn;
    r->synchro.System = 'E';
    std::memset(r->synchro.Signal, 0, sizeof r->synchro.Signal);
    r->synchro.Signal[0] = signal[0];
    r->synchro.Signal[1] = signal[1];
    r->block->set_gnss_synchro(&r->synchro);
    r->block->set_channel(0);
    return r;
}
void ref_acqcaf_free(void* h) { delete static_cast<RefAcqCaf*>(h); }

void ref_acqcaf_arm(void* h, unsigned doppler_max, unsigned doppler_step, float threshold)
{
    auto* r = static_cast<RefAcqCaf*>(h);
    r->block->set_doppler_max(doppler_max);
    r->block->set_doppler_step(doppler_step);
    r->block->set_threshold(threshold);
    r->block->init();
    // the adapter's set_local_code (:211-261)
    std::vector<std::complex<float>> codeI(r->code_length), codeQ(r->code_length);
    const bool x = r->synchro.Signal[0] == '5' && r->synchro.Signal[1] == 'X';
    if (x)
        {
            galileo_e5_a_code_gen_complex_sampled(codeI, r->synchro.PRN, std::array<char, 3>{{'5', 'I', '\0'}}, static_cast<int32_t>(r->fs_in), 0);
            galileo_e5_a_code_gen_complex_sampled(codeQ, r->synchro.PRN, std::array<char, 3>{{'5', 'Q', '\0'}}, static_cast<int32_t>(r->fs_in), 0);
        }
    else
        galileo_e5_a_code_gen_complex_sampled(codeI, r->synchro.PRN, std::array<char, 3>{{'5', 'X', '\0'}}, static_cast<int32_t>(r->fs_in), 0);
    const unsigned copies = r->zero_padding == 0 ? r->sampled_ms : 1U;
    for (unsigned i = 0; i < copies; i++)
        {
            std::copy_n(codeI.data(), r->code_length, r->codeI.data() + static_cast<size_t>(i) * r->code_length);
            if (x) std::copy_n(codeQ.data(), r->code_length, r->codeQ.data() + static_cast<size_t>(i) * r->code_length);
        }
    r->block->set_local_code(r->codeI.data(), r->codeQ.data());
    r->block->set_state(1);
    r->block->set_active(true);
}

// plays the scheduler like gb200h_acqcaf_work: general_work with at most `chunk` items per call until the block has raised its event and
// gone idle or the buffer is used up; returns the items consumed
long long ref_acqcaf_work(void* h, const void* items, long long n_items, int chunk)
{
    auto* r = static_cast<RefAcqCaf*>(h);
    auto& b = *r->block;
    const auto* in = static_cast<const std::complex<float>*>(items);
    long long pos = 0;
    int idle = 0;
    while (pos < n_items && idle < 3)
        {
            const int before_state = b.d_state;
            gr_vector_int ninput{static_cast<int>(std::min<long long>(n_items - pos, chunk))};
            gr_vector_const_void_star inv{in + pos};
            Gnss_Synchro out_item{};
            Gnss_Synchro* out_ptr = &out_item;
            gr_vector_void_star out{&out_ptr};
            const long long before = b.shim_consumed;
            b.general_work(1, ninput, inv, out);
            const long long c = b.shim_consumed - before;
            pos += c;
            if (b.d_state == 0 && !b.d_active) break;
            idle = (c == 0 && b.d_state == before_state) ? idle + 1 : 0;
        }
    return pos;
}

void ref_acqcaf_get_status(void* h, ref_acqcaf_status* s)
{
    auto* r = static_cast<RefAcqCaf*>(h);
    auto& b = *r->block;
    std::memset(s, 0, sizeof *s);
    s->acq_delay_samples = r->synchro.Acq_delay_samples, s->acq_doppler_hz = r->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = r->synchro.Acq_samplestamp_samples, s->sample_counter = b.d_sample_counter;
    s->acq_doppler_step = r->synchro.Acq_doppler_step;
    s->consumed_total = b.shim_consumed;
    s->test_statistics = b.d_test_statistics, s->threshold = b.d_threshold, s->mag = b.d_mag, s->input_power = b.d_input_power;
    s->state = b.d_state, s->active = b.d_active, s->num_doppler_bins = b.d_num_doppler_bins, s->fft_size = b.d_fft_size, s->well_count = static_cast<int32_t>(b.d_well_count);
    s->n_events = static_cast<int>(std::min<size_t>(8, b.shim_events.size()));
    for (int i = 0; i < s->n_events; i++) s->events[i] = static_cast<int>(b.shim_events[b.shim_events.size() - s->n_events + i]);
}

int ref_acqcaf_caf_vector(void* h, float* dst, int max_bins)
{
    auto& v = static_cast<RefAcqCaf*>(h)->block->d_CAF_vector;
    const int n = std::min<int>(max_bins, static_cast<int>(v.size()));
    if (n > 0 && dst) std::copy(v.begin(), v.begin() + n, dst);
    return static_cast<int>(v.size());
}
}
