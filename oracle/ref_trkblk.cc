// TEST INFRASTRUCTURE — not part of the product.
// C door to the reference's OWN tracking block: src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.{h,cc} and
// src/algorithms/tracking/libs/dll_pll_conf.{h,cc}, compiled UNMODIFIED from /root/reference against the stand-in declarations of
// oracle/shim/acq/ (gr::block, pmt, boost::circular_buffer, gflags variables, matio, glog: scheduler plumbing and containers, no
// arithmetic).  Everything numerical is the reference's: Cpu_Multicorrelator_Real_Codes over the volk_gnsssdr protokernels, the
// discriminators, loop filters, lock detectors, smoothers, the code generators and the whole general_work state machine.
// The door plays the GNU Radio scheduler for one block: it keeps the unread part of the stream, asks forecast() how much the block
// wants, calls general_work(), applies consume_each() to nitems_read and collects the Gnss_Synchro items produced.
#include <algorithm>
#include <any>
#include <complex>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <span>
#include <fstream>
#include <iostream>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#define private public
#define protected public
#include "dll_pll_veml_tracking.h"
#undef private
#undef protected
#include "gnss_synchro.h"

namespace
{
class MapConfigurationT : public ConfigurationInterface
{
public:
    explicit MapConfigurationT(const char* text)
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

struct RefTrkBlk
{
    Dll_Pll_Conf conf;
    dll_pll_veml_tracking_sptr block;
    Gnss_Synchro synchro{};
    std::vector<std::complex<float>> pending;  // unread samples of the stream
};
}  // namespace

extern "C" {
// same record as oracle/ref_loop.cc (tests/_oracle.py REF_LOOP_REC)
struct ref_trkblk_rec
{
    int32_t channel;
    uint32_t prn;
    int32_t valid, correlation_length_ms;
    double prompt_i, prompt_q, cn0_db_hz, carrier_doppler_hz, carrier_phase_rads, code_phase_samples;
    uint64_t tracking_sample_counter;
};

// conf_text: "role.key=value" lines; vector_length_chips_rate: the adapter's vector_length = round(fs_in / (chip_rate / code_length)) is
// computed here from chip_rate and code_length (gps_l1_ca_dll_pll_tracking.cc:46-47 and its siblings); track_pilot as the adapter reads it
void* ref_trkblk_new(const char* conf_text, const char* role, const char* system, const char* signal, double chip_rate, double code_length_chips, unsigned prn,
    double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp)
{
    auto* r = new RefTrkBlk;
    MapConfigurationT cfg(conf_text);
    r->conf = Dll_Pll_Conf();
    r->conf.SetFromConfiguration(&cfg, role);
    r->conf.vector_length = static_cast<int>(std::round(r->conf.fs_in / (chip_rate / code_length_chips)));
    if (r->conf.extend_correlation_symbols < 1) r->conf.extend_correlation_symbols = 1;
    r->conf.system = system[0];
    std::memset(r->conf.signal, 0, sizeof r->conf.signal);
    std::strncpy(r->conf.signal, signal, 2);
    r->block = dll_pll_veml_make_tracking(r->conf);
    r->synchro.PRN = prn;
    r->synchro.System = system[0];
    std::memset(r->synchro.Signal, 0, sizeof r->synchro.Signal);
    std::strncpy(r->synchro.Signal, signal, 2);
    r->synchro.Acq_delay_samples = acq_delay_samples;
    r->synchro.Acq_doppler_hz = acq_doppler_hz;
    r->synchro.Acq_samplestamp_samples = acq_samplestamp;
    r->block->set_channel(0);
    r->block->set_gnss_synchro(&r->synchro);
    r->block->start_tracking();
    return r;
}
void ref_trkblk_free(void* h) { delete static_cast<RefTrkBlk*>(h); }

// append n samples to the stream and run the block while its forecast is met; returns the number of records written to out
int ref_trkblk_feed(void* h, const float* iq_ri, int64_t n, ref_trkblk_rec* out, int max_records)
{
    auto* r = static_cast<RefTrkBlk*>(h);
    auto& b = *r->block;
    const auto* x = reinterpret_cast<const std::complex<float>*>(iq_ri);
    r->pending.insert(r->pending.end(), x, x + n);
    size_t pos = 0;
    int n_out = 0;
    for (;;)
        {
            gr_vector_int required{0};
            b.forecast(1, required);
            const size_t avail = r->pending.size() - pos;
            if (avail < static_cast<size_t>(std::max(1, required[0])) || n_out >= max_records) break;
            gr_vector_int ninput{static_cast<int>(std::min<size_t>(avail, 1u << 30))};
            gr_vector_const_void_star in{r->pending.data() + pos};
            Gnss_Synchro item{};
            gr_vector_void_star ov{&item};
            b.shim_pending_consume = 0;
            const int produced = b.general_work(1, ninput, in, ov);
            pos += static_cast<size_t>(b.shim_pending_consume);
            if (produced > 0)
                {
                    b.shim_nitems_written += 1;
                    ref_trkblk_rec& o = out[n_out++];
                    o.channel = item.Channel_ID, o.prn = item.PRN, o.valid = item.Flag_valid_symbol_output ? 1 : 0;
                    o.correlation_length_ms = item.correlation_length_ms;
                    o.prompt_i = item.Prompt_I, o.prompt_q = item.Prompt_Q, o.cn0_db_hz = item.CN0_dB_hz, o.carrier_doppler_hz = item.Carrier_Doppler_hz;
                    o.carrier_phase_rads = item.Carrier_phase_rads, o.code_phase_samples = item.Code_phase_samples;
                    o.tracking_sample_counter = item.Tracking_sample_counter;
                }
            else if (b.shim_pending_consume == 0)
                break;  // nothing consumed, nothing produced: the block is idle (standby) and would spin
        }
    r->pending.erase(r->pending.begin(), r->pending.begin() + static_cast<long>(pos));
    return n_out;
}

// out: state, cn0_db_hz, carrier_doppler_hz, code_freq_chips, carrier_lock_test, events (last message-port value, 0 if none)
void ref_trkblk_state(void* h, double* out)
{
    auto& b = *static_cast<RefTrkBlk*>(h)->block;
    out[0] = b.d_state, out[1] = b.d_CN0_SNV_dB_Hz, out[2] = b.d_carrier_doppler_hz, out[3] = b.d_code_freq_chips, out[4] = b.d_carrier_lock_test;
    out[5] = b.shim_events.empty() ? 0.0 : static_cast<double>(b.shim_events.back());
}
}
