// TEST INFRASTRUCTURE — not part of the product.
// C door to the reference's OWN fine-Doppler acquisition block: src/algorithms/acquisition/gnuradio_blocks/
// pcps_acquisition_fine_doppler_cc.{h,cc}, compiled UNMODIFIED from /root/reference against the stand-in declarations of
// oracle/shim/acq/ (README.txt there; the DFT behind gr::fft is oracle/oracle_fft.c — also for the 80 x samples-per-ms point transform of
// estimate_Doppler).  The block is filled and driven the way adapters/gps_l1_ca_pcps_acquisition_fine_doppler.cc:30-150 does: Acq_Conf
// fields by hand (fs_in, samples_per_chip, doppler_max, sampled_ms, max_dwells, blocking_on_standby, samples_per_ms), set_gnss_synchro,
// set_channel, set_doppler_max, set_doppler_step, set_threshold, init, set_local_code (gps_l1_ca_code_gen_complex_sampled), set_state(1).
#include <algorithm>
#include <any>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <deque>
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
#include "pcps_acquisition_fine_doppler_cc.h"
#undef private
#undef protected
#include "GPS_L1_CA.h"
#include "gnss_synchro.h"
#include "gps_sdr_signal_replica.h"

namespace
{
struct RefAcqFd
{
    Acq_Conf conf;
    pcps_acquisition_fine_doppler_cc_sptr block;
    Gnss_Synchro synchro{};
    std::vector<std::complex<float>> code;
    int64_t fs_in{0};
};
}  // namespace

extern "C" {
struct ref_acqfd_status
{
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    int64_t consumed_total;
    float test_statistics, threshold;
    int32_t state, active, positive_acq, num_doppler_points, fft_size, well_count, n_events, events[8];
    uint32_t acq_doppler_step;
};

void* ref_acqfd_new(long long fs_in, int doppler_max, int max_dwells, int blocking_on_standby, unsigned prn)
{
    auto* r = new RefAcqFd;
    r->fs_in = fs_in;
    Acq_Conf& p = r->conf;
    p.fs_in = fs_in;
    p.samples_per_chip = static_cast<unsigned int>(std::ceil(GPS_L1_CA_CHIP_PERIOD_S * static_cast<float>(p.fs_in)));
    p.dump = false;
    p.doppler_max = doppler_max;
    p.sampled_ms = 1;
    p.max_dwells = static_cast<uint32_t>(max_dwells);
    p.blocking_on_standby = blocking_on_standby != 0;
    const auto vector_length = static_cast<unsigned int>(std::round(fs_in / (GPS_L1_CA_CODE_RATE_CPS / GPS_L1_CA_CODE_LENGTH_CHIPS)));
    p.samples_per_ms = static_cast<float>(vector_length);
    r->code.assign(vector_length, std::complex<float>(0.0F, 0.0F));
    r->block = pcps_make_acquisition_fine_doppler_cc(p);
    r->synchro.PRN = prn;
    r->synchro.System = 'G';
    std::memset(r->synchro.Signal, 0, sizeof r->synchro.Signal);
    r->synchro.Signal[0] = '1';
    r->synchro.Signal[1] = 'C';
    r->block->set_gnss_synchro(&r->synchro);
    r->block->set_channel(0);
    return r;
}
void ref_acqfd_free(void* h) { delete static_cast<RefAcqFd*>(h); }

void ref_acqfd_arm(void* h, unsigned doppler_max, unsigned doppler_step, float threshold)
{
    auto* r = static_cast<RefAcqFd*>(h);
    r->block->set_doppler_max(doppler_max);
    r->block->set_doppler_step(doppler_step);
    r->block->set_threshold(threshold);
    r->block->init();
    gps_l1_ca_code_gen_complex_sampled(r->code, r->synchro.PRN, static_cast<int32_t>(r->fs_in), 0);
    r->block->set_local_code(r->code.data());
    r->block->start();
    r->block->set_state(1);
}

// plays the scheduler like gb200h_acqfd_work: general_work(noutput_items) while max(fft_size, noutput_items) items are left, until the
// block has raised its event and gone idle; returns the items consumed
long long ref_acqfd_work(void* h, const void* items, long long n_items, int noutput_items)
{
    auto* r = static_cast<RefAcqFd*>(h);
    auto& b = *r->block;
    const auto* in = static_cast<const std::complex<float>*>(items);
    const long long need = std::max(b.d_fft_size, noutput_items);
    long long pos = 0;
    while (n_items - pos >= need)
        {
            gr_vector_int ninput{static_cast<int>(std::min<long long>(n_items - pos, 1 << 30))};
            gr_vector_const_void_star inv{in + pos};
            Gnss_Synchro out_item{};
            Gnss_Synchro* out_ptr = &out_item;
            gr_vector_void_star out{&out_ptr};
            const long long before = b.shim_consumed;
            b.general_work(noutput_items, ninput, inv, out);
            pos += b.shim_consumed - before;
            if (b.d_state == 0 && !b.d_active) break;
        }
    return pos;
}

void ref_acqfd_get_status(void* h, ref_acqfd_status* s)
{
    auto* r = static_cast<RefAcqFd*>(h);
    auto& b = *r->block;
    std::memset(s, 0, sizeof *s);
    s->acq_delay_samples = r->synchro.Acq_delay_samples, s->acq_doppler_hz = r->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = r->synchro.Acq_samplestamp_samples, s->sample_counter = b.d_sample_counter;
    s->acq_doppler_step = r->synchro.Acq_doppler_step;
    s->consumed_total = b.shim_consumed;
    s->test_statistics = b.d_test_statistics, s->threshold = b.d_threshold;
    s->state = b.d_state, s->active = b.d_active, s->positive_acq = b.d_positive_acq, s->num_doppler_points = b.d_num_doppler_points;
    s->fft_size = b.d_fft_size, s->well_count = b.d_well_count;
    s->n_events = static_cast<int>(std::min<size_t>(8, b.shim_events.size()));
    for (int i = 0; i < s->n_events; i++) s->events[i] = static_cast<int>(b.shim_events[b.shim_events.size() - s->n_events + i]);
}
}
