// TEST INFRASTRUCTURE — not part of the product.
// C door to the reference's OWN acquisition block: src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.{h,cc} and
// src/algorithms/acquisition/libs/acq_conf.{h,cc}, compiled UNMODIFIED from /root/reference against the stand-in declarations of
// oracle/shim/acq/ (see its README.txt for what is real and what is a stand-in; the one arithmetic stand-in that matters is the DFT
// behind gr::fft, which is oracle/oracle_fft.c in float32 or float64).  The block is driven the way Channel / the adapters drive it:
// Acq_Conf::SetFromConfiguration from "key=value" text, set_gnss_synchro / set_channel_fsm / set_doppler_max / set_doppler_step /
// set_threshold, init(), set_local_code(), set_state(1), set_active(true), general_work() over the input stream.
#include <algorithm>
#include <any>
#include <deque>
#include <functional>
#include <complex>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <queue>
#include <sstream>
#include <string>
#include <vector>
#include <span>
#include <mutex>
#include <thread>
#include <iostream>
// the door reads the block's private results (d_test_statistics, d_mag, d_input_power, ...) and calls calculate_threshold()
#define private public
#define protected public
#include "pcps_acquisition.h"
#undef private
#undef protected
#include "gnss_synchro.h"
#include "oracle.h"

extern "C" {
int oracle_shim_fft_f64 = 0;
void oracle_shim_fft(const std::complex<float>* in, std::complex<float>* out, int n, int inverse)
{
    if (!oracle_shim_fft_f64)
        {
            orc_fft_c32(reinterpret_cast<const float*>(in), reinterpret_cast<float*>(out), n, inverse);
            return;
        }
    std::vector<double> a(2 * static_cast<size_t>(n)), b(2 * static_cast<size_t>(n));
    for (int i = 0; i < n; i++) a[2 * i] = in[i].real(), a[2 * i + 1] = in[i].imag();
    orc_fft_c64(a.data(), b.data(), n, inverse);
    for (int i = 0; i < n; i++) out[i] = std::complex<float>(static_cast<float>(b[2 * i]), static_cast<float>(b[2 * i + 1]));
}
}

namespace
{
// ConfigurationInterface over "key=value" lines (what InMemoryConfiguration is in the reference's tests)
class MapConfiguration : public ConfigurationInterface
{
public:
    explicit MapConfiguration(const char* text)
    {
        std::istringstream is(text ? text : "");
        std::string line;
        while (std::getline(is, line))
            {
                const size_t eq = line.find('=');
                if (eq == std::string::npos) continue;
                kv_[line.substr(0, eq)] = line.substr(eq + 1);
            }
    }
    std::string property(std::string k, std::string d) const override { return has(k) ? kv_.at(k) : d; }
    bool property(std::string k, bool d) const override { return has(k) ? (kv_.at(k) == "true" || kv_.at(k) == "1") : d; }
    int64_t property(std::string k, int64_t d) const override { return has(k) ? static_cast<int64_t>(std::stod(kv_.at(k))) : d; }
    uint64_t property(std::string k, uint64_t d) const override { return has(k) ? static_cast<uint64_t>(std::stod(kv_.at(k))) : d; }
    int32_t property(std::string k, int32_t d) const override { return has(k) ? static_cast<int32_t>(std::stod(kv_.at(k))) : d; }
    uint32_t property(std::string k, uint32_t d) const override { return has(k) ? static_cast<uint32_t>(std::stod(kv_.at(k))) : d; }
    int16_t property(std::string k, int16_t d) const override { return has(k) ? static_cast<int16_t>(std::stod(kv_.at(k))) : d; }
    uint16_t property(std::string k, uint16_t d) const override { return has(k) ? static_cast<uint16_t>(std::stod(kv_.at(k))) : d; }
    float property(std::string k, float d) const override { return has(k) ? std::stof(kv_.at(k)) : d; }
    double property(std::string k, double d) const override { return has(k) ? std::stod(kv_.at(k)) : d; }
    void set_property(std::string k, std::string v) override { kv_[k] = v; }

private:
    bool has(const std::string& k) const { return kv_.count(k) != 0; }
    std::map<std::string, std::string> kv_;
};

struct RefAcq
{
    Acq_Conf conf;
    pcps_acquisition_sptr block;
    std::shared_ptr<ChannelFsm> fsm;
    Gnss_Synchro synchro{};
};
}  // namespace

extern "C" {
struct ref_acq_status
{
    int32_t state, active, positive_acq, n_events, events[8], fsm_events[8], n_fsm_events;
    uint32_t fft_size, consumed_samples, num_doppler_bins, dwells_counter, samples_per_chip;
    float threshold, mag, input_power, test_statistics, samples_per_code, samples_per_ms;
    double acq_delay_samples, acq_doppler_hz;
    uint64_t acq_samplestamp_samples, sample_counter;
    int64_t consumed_total;
    int32_t doppler_step, doppler_max, use_cfar, bit_transition, it_size;
    int32_t step_two, pad_;  // d_step_two: the coarse step was positive, the fine step (make_two_steps) starts with the next call
};

// conf_text: "role.key=value" lines as in a .conf; ms_per_code / chip_rate / opt_freq: what the adapter passes (e.g. 1, 1.023e6, 2e6)
void* ref_acq_new(const char* conf_text, const char* role, unsigned ms_per_code, double chip_rate, double opt_freq, unsigned prn, const char* system, const char* signal)
{
    auto* r = new RefAcq;
    MapConfiguration cfg(conf_text);
    r->conf.ms_per_code = ms_per_code;
    r->conf.SetFromConfiguration(&cfg, role, chip_rate, opt_freq);
    r->block = pcps_make_acquisition(r->conf);
    r->fsm = std::make_shared<ChannelFsm>();
    r->synchro.PRN = prn;
    r->synchro.System = system[0];
    std::memset(r->synchro.Signal, 0, sizeof r->synchro.Signal);
    std::strncpy(r->synchro.Signal, signal, 2);
    r->block->set_gnss_synchro(&r->synchro);
    r->block->set_channel(0);
    r->block->set_channel_fsm(r->fsm);
    return r;
}
void ref_acq_free(void* h) { delete static_cast<RefAcq*>(h); }

// the adapter's set_doppler_max / set_doppler_step / set_threshold / init (gps_l1_ca_pcps_acquisition.cc:110-190): with pfa > 0 in the
// configuration the block computes its own threshold in init(), otherwise `threshold` is installed
void ref_acq_init(void* h, unsigned doppler_max, unsigned doppler_step, float threshold)
{
    auto* r = static_cast<RefAcq*>(h);
    r->block->set_doppler_max(doppler_max);
    r->block->set_doppler_step(doppler_step);
    r->block->set_threshold(threshold);
    r->block->init();
    r->block->start();  // what the GNU Radio scheduler calls when the flowgraph starts: resets the sample counter, computes the pfa threshold (:875-881)
}
// after init(): replaces the threshold init() computed from pfa (e.g. an unreachable one, so that every dwell of max_dwells is taken)
void ref_acq_set_threshold(void* h, float threshold) { static_cast<RefAcq*>(h)->block->set_threshold(threshold); }
void ref_acq_set_doppler_center(void* h, int center) { static_cast<RefAcq*>(h)->block->set_doppler_center(center); }
// code: vector_length complex samples as the adapter prepares them (set_local_code of the adapter ends in acquisition_->set_local_code(code_.data()))
void ref_acq_set_local_code(void* h, const float* code_ri)
{
    static_cast<RefAcq*>(h)->block->set_local_code(reinterpret_cast<std::complex<float>*>(const_cast<float*>(code_ri)));
}
// Channel::start_acquisition -> acq_->reset(): set_active(true) after set_state(1)
void ref_acq_arm(void* h)
{
    auto* r = static_cast<RefAcq*>(h);
    r->block->set_state(1);
    r->block->set_active(true);
}
// one scheduler call with n_items input items available; returns the items consumed by this call
long ref_acq_work(void* h, const void* items, int n_items)
{
    auto* r = static_cast<RefAcq*>(h);
    gr_vector_int ninput{n_items};
    gr_vector_const_void_star in{items};
    Gnss_Synchro out_item{};
    gr_vector_void_star out{&out_item};
    const long long before = r->block->shim_consumed;
    r->block->general_work(1, ninput, in, out);
    return static_cast<long>(r->block->shim_consumed - before);
}
void ref_acq_get_status(void* h, ref_acq_status* s)
{
    auto* r = static_cast<RefAcq*>(h);
    auto& b = *r->block;
    std::memset(s, 0, sizeof *s);
    s->state = b.d_state, s->active = b.d_active, s->positive_acq = b.d_positive_acq;
    s->n_events = static_cast<int>(std::min<size_t>(8, b.shim_events.size()));
    for (int i = 0; i < s->n_events; i++) s->events[i] = static_cast<int>(b.shim_events[b.shim_events.size() - s->n_events + i]);
    s->n_fsm_events = static_cast<int>(std::min<size_t>(8, r->fsm->events.size()));
    for (int i = 0; i < s->n_fsm_events; i++) s->fsm_events[i] = r->fsm->events[r->fsm->events.size() - s->n_fsm_events + i];
    s->fft_size = b.d_fft_size, s->consumed_samples = b.d_consumed_samples, s->num_doppler_bins = b.d_num_doppler_bins;
    s->dwells_counter = b.d_num_noncoherent_integrations_counter, s->samples_per_chip = b.d_samplesPerChip;
    s->threshold = b.d_threshold, s->mag = b.d_mag, s->input_power = b.d_input_power, s->test_statistics = b.d_test_statistics;
    s->samples_per_code = b.d_acq_parameters.samples_per_code, s->samples_per_ms = b.d_acq_parameters.samples_per_ms;
    s->acq_delay_samples = r->synchro.Acq_delay_samples, s->acq_doppler_hz = r->synchro.Acq_doppler_hz;
    s->acq_samplestamp_samples = r->synchro.Acq_samplestamp_samples, s->sample_counter = b.d_sample_counter;
    s->consumed_total = b.shim_consumed;
    s->doppler_step = static_cast<int32_t>(b.d_doppler_step), s->doppler_max = b.d_acq_parameters.doppler_max;
    s->step_two = b.d_step_two ? 1 : 0;
    s->use_cfar = b.d_use_CFAR_algorithm_flag, s->bit_transition = b.d_acq_parameters.bit_transition_flag, s->it_size = static_cast<int32_t>(b.d_acq_parameters.it_size);
}
// the magnitude grid of the last dwell: out[bin][neff]
int ref_acq_read_grid(void* h, float* out, int max_bins, int neff)
{
    auto& b = *static_cast<RefAcq*>(h)->block;
    const int bins = std::min<int>(max_bins, static_cast<int>(b.d_magnitude_grid.size()));
    for (int k = 0; k < bins; k++) std::memcpy(out + static_cast<size_t>(k) * neff, b.d_magnitude_grid[k].data(), sizeof(float) * std::min<size_t>(neff, b.d_magnitude_grid[k].size()));
    return bins;
}
void ref_acq_fft_precision(int f64) { oracle_shim_fft_f64 = f64 ? 1 : 0; }
}
