// TEST INFRASTRUCTURE — not part of the product.
// C door to the reference's OWN QuickSync acquisition block: src/algorithms/acquisition/gnuradio_blocks/pcps_quicksync_acquisition_cc.{h,cc},
// compiled UNMODIFIED from /root/reference against the stand-in declarations of oracle/shim/acq/ (README.txt there; the DFT behind gr::fft is
// oracle/oracle_fft.c).  The block is made and driven the way adapters/gps_l1_ca_pcps_quicksync_acquisition.cc:36-244 does (folding factor,
// coherent time rounded to a multiple of it, max_dwells = 2 with the bit-transition flag, one period of gps_l1_ca_code_gen_complex_sampled as
// the replica); the door plays stream_to_vector (items of samples-per-code x folding_factor samples, :113-114) and the scheduler.
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
#include "pcps_quicksync_acquisition_cc.h"
#undef private
#undef protected
#include "GPS_L1_CA.h"
#include "Galileo_E1.h"
#include "galileo_e1_signal_replica.h"
#include "gnss_synchro.h"
#include "gps_sdr_signal_replica.h"

namespace
{
struct RefAcqQs
{
    pcps_quicksync_acquisition_cc_sptr block;
    Gnss_Synchro synchro{};
    std::vector<std::complex<float>> code;
    int64_t fs_in{0};
    unsigned code_length{0}, folding{0}, sampled_ms{0}, item_length{0};
    bool galileo{false}, cboc{false};
};
}  // namespace

extern "C" {
struct ref_acqqs_status
{
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    int64_t consumed_total;
    float test_statistics, threshold, mag, input_power;
    int32_t state, active, num_doppler_bins, fft_size, well_count, item_length, folding_factor, sampled_ms, n_events, events[8];
    uint32_t acq_doppler_step;
};

// folding_factor 0: the adapter's default ceil(sqrt(log2(samples per code)))
void* ref_acqqs_new(long long fs_in, unsigned folding_factor, unsigned sampled_ms, unsigned max_dwells, unsigned doppler_max, int bit_transition, unsigned prn)
{
    auto* r = new RefAcqQs;
    r->fs_in = fs_in;
    r->code_length = static_cast<unsigned int>(std::round(fs_in / (GPS_L1_CA_CODE_RATE_CPS / GPS_L1_CA_CODE_LENGTH_CHIPS)));
    if (folding_factor == 0) folding_factor = static_cast<unsigned int>(std::ceil(std::sqrt(std::log2(r->code_length))));
    if (sampled_ms % folding_factor != 0) sampled_ms = sampled_ms < folding_factor ? folding_factor : (sampled_ms / folding_factor) * folding_factor;
    if (bit_transition) max_dwells = 2;
    r->folding = folding_factor;
    r->sampled_ms = sampled_ms;
    r->code.assign(r->code_length, std::complex<float>(0.0F, 0.0F));
    r->block = pcps_quicksync_make_acquisition_cc(folding_factor, sampled_ms, max_dwells, doppler_max, fs_in, static_cast<int>(r->code_length), static_cast<int>(r->code_length),
        bit_transition != 0, false, "", false);
    r->synchro.PRN = prn;
    r->synchro.System = 'G';
    std::memset(r->synchro.Signal, 0, sizeof r->synchro.Signal);
    r->synchro.Signal[0] = '1';
    r->synchro.Signal[1] = 'C';
    r->block->set_gnss_synchro(&r->synchro);
    r->block->set_channel(0);
    return r;
}
// the Galileo E1 adapter (adapters/galileo_e1_pcps_quicksync_ambiguous_acquisition.cc:36-150): 4 ms codes, folding_factor 0 = its default 2, coherent
// time a multiple of 4 x folding_factor, vector items of sampled_ms x samples per ms
void* ref_acqqs_new_e1(long long fs_in, unsigned folding_factor, unsigned sampled_ms, unsigned max_dwells, unsigned doppler_max, int bit_transition, unsigned prn, const char* signal, int cboc)
{
    auto* r = new RefAcqQs;
    r->fs_in = fs_in;
    r->galileo = true;
    r->cboc = cboc != 0;
    r->code_length = static_cast<unsigned int>(std::round(fs_in / (GALILEO_E1_CODE_CHIP_RATE_CPS / GALILEO_E1_B_CODE_LENGTH_CHIPS)));
    const auto samples_per_ms = static_cast<int>(std::round(r->code_length / 4.0));
    if (folding_factor == 0) folding_factor = 2;
    if (sampled_ms % (folding_factor * 4) != 0) sampled_ms = sampled_ms < folding_factor * 4 ? folding_factor * 4 : (sampled_ms / (folding_factor * 4)) * (folding_factor * 4);
    if (bit_transition) max_dwells = 2;
    r->folding = folding_factor;
    r->sampled_ms = sampled_ms;
    r->item_length = sampled_ms * static_cast<unsigned>(samples_per_ms);
    r->code.assign(r->code_length, std::complex<float>(0.0F, 0.0F));
    r->block = pcps_quicksync_make_acquisition_cc(folding_factor, sampled_ms, max_dwells, doppler_max, fs_in, samples_per_ms, static_cast<int>(r->code_length), bit_transition != 0, false, "",
        false);
    r->synchro.PRN = prn;
    r->synchro.System = 'E';
    std::memset(r->synchro.Signal, 0, sizeof r->synchro.Signal);
    r->synchro.Signal[0] = signal[0];
    r->synchro.Signal[1] = signal[1];
    r->block->set_gnss_synchro(&r->synchro);
    r->block->set_channel(0);
    return r;
}
void ref_acqqs_free(void* h) { delete static_cast<RefAcqQs*>(h); }

void ref_acqqs_arm(void* h, unsigned doppler_max, unsigned doppler_step, float threshold)
{
    auto* r = static_cast<RefAcqQs*>(h);
    r->block->set_doppler_max(doppler_max);
    r->block->set_doppler_step(doppler_step);
    r->block->set_threshold(threshold);
    r->block->init();
    if (r->galileo)
        {
            const std::array<char, 3> sig{{r->synchro.Signal[0], r->synchro.Signal[1], '\0'}};
            galileo_e1_code_gen_complex_sampled(r->code, sig, r->cboc, r->synchro.PRN, static_cast<int32_t>(r->fs_in), 0, false);
        }
    else
        gps_l1_ca_code_gen_complex_sampled(r->code, r->synchro.PRN, static_cast<int32_t>(r->fs_in), 0);
    r->block->set_local_code(r->code.data());
    r->block->set_state(1);
}

// the stream cut into vector items of code_length * folding samples; general_work with at most max_items of them per call until the block has
// raised its event and gone idle or the items are used up; returns the ITEMS consumed
long long ref_acqqs_work(void* h, const void* samples, long long n_samples, int max_items)
{
    auto* r = static_cast<RefAcqQs*>(h);
    auto& b = *r->block;
    const auto* in = static_cast<const std::complex<float>*>(samples);
    const long long len = r->item_length ? r->item_length : static_cast<long long>(r->code_length) * r->folding, n_items = n_samples / len;
    long long pos = 0;
    while (pos < n_items)
        {
            gr_vector_int ninput{static_cast<int>(std::min<long long>(n_items - pos, max_items))};
            gr_vector_const_void_star inv{in + pos * len};
            Gnss_Synchro out_item{};
            Gnss_Synchro* out_ptr = &out_item;
            gr_vector_void_star out{&out_ptr};
            const long long before = b.shim_consumed;
            b.general_work(1, ninput, inv, out);
            pos += b.shim_consumed - before;
            if (b.d_state == 0 && !b.d_active) break;
        }
    return pos;
}

void ref_acqqs_get_status(void* h, ref_acqqs_status* s)
{
    auto* r = static_cast<RefAcqQs*>(h);
    auto& b = *r->block;
    std::memset(s, 0, sizeof *s);
    s->acq_delay_samples = r->synchro.Acq_delay_samples, s->acq_doppler_hz = r->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = r->synchro.Acq_samplestamp_samples, s->sample_counter = b.d_sample_counter;
    s->acq_doppler_step = r->synchro.Acq_doppler_step;
    s->consumed_total = b.shim_consumed;
    s->test_statistics = b.d_test_statistics, s->threshold = b.d_threshold, s->mag = b.d_mag, s->input_power = b.d_input_power;
    s->state = b.d_state, s->active = b.d_active, s->num_doppler_bins = static_cast<int32_t>(b.d_num_doppler_bins), s->fft_size = static_cast<int32_t>(b.d_fft_size);
    s->well_count = static_cast<int32_t>(b.d_well_count);
    s->item_length = static_cast<int32_t>(r->item_length ? r->item_length : r->code_length * r->folding), s->folding_factor = static_cast<int32_t>(r->folding), s->sampled_ms = static_cast<int32_t>(r->sampled_ms);
    s->n_events = static_cast<int>(std::min<size_t>(8, b.shim_events.size()));
    for (int i = 0; i < s->n_events; i++) s->events[i] = static_cast<int>(b.shim_events[b.shim_events.size() - s->n_events + i]);
}
}
