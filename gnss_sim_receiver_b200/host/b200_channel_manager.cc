#include "b200_channel_manager.h"
#include "b200_gnss_codes.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace gb200h
{
B200ChannelManager::B200ChannelManager(const ConfigurationInterface* configuration, const std::string& acq_role, const std::string& trk_role, int n_channels,
    const std::vector<unsigned>& prns)
    : prns_(prns)
{
    if (n_channels <= 0 || prns.empty()) throw std::invalid_argument("channel manager needs channels and satellites");
    // ---- acquisition side: one sweep engine with a code-spectrum slot per candidate satellite
    auto acq = GetB200AcqBlock(configuration, acq_role, 1, 0);
    if (!acq) throw std::runtime_error(acq_role + ".implementation is not a B200 acquisition block");
    auto* acq_adapter = static_cast<B200PcpsAcquisition*>(acq.get());
    acq_conf_ = acq_adapter->conf();
    if (acq_conf_.pfa <= 0.0F) throw std::invalid_argument(acq_role + ".pfa must be set: the batched sweep judges every satellite against the CFAR threshold");
    item_bytes_ = acq_conf_.it_size;
    consumed_ = static_cast<uint32_t>(acq_conf_.sampled_ms * acq_conf_.samples_per_ms * (acq_conf_.bit_transition_flag ? 2.0 : 1.0));  // pcps_acquisition.cc:71
    sweep_ = std::make_unique<B200AcquisitionSweep>(acq_conf_, static_cast<int>(prns_.size()));
    sweep_->set_doppler_grid(acq_conf_.doppler_max, static_cast<int>(acq_conf_.doppler_step), 0);
    threshold_ = sweep_->threshold(acq_conf_.pfa);
    {
        // the sampled replica of every candidate satellite, built by the signal's own adapter (GPS L1 C/A, Galileo E1, GPS L5I,
        // Galileo E5a: adapters/*_pcps_acquisition.cc set_local_code); the channel's system / signal come from the tracking role
        auto trk_probe = GetB200TrkBlock(configuration, trk_role, 1, 1);
        if (!trk_probe) throw std::runtime_error(trk_role + ".implementation is not a B200 tracking block");
        const Dll_Pll_Conf& tp = static_cast<B200DllPllTracking*>(trk_probe.get())->conf();
        Gnss_Synchro probe{};
        probe.System = tp.system;
        std::memcpy(probe.Signal, tp.signal, 3);
        acq_adapter->set_gnss_synchro(&probe);
        const unsigned n_code = acq_conf_.bit_transition_flag ? consumed_ / 2 : consumed_;
        for (size_t s = 0; s < prns_.size(); s++)
            {
                probe.PRN = prns_[s];
                acq_adapter->make_local_code();
                const auto& code = acq_adapter->local_code();
                if (code.size() < n_code) throw std::runtime_error("acquisition adapter produced a short local code");
                sweep_->set_local_code(static_cast<int>(s), code.data(), static_cast<int>(n_code));
            }
    }
    buf_.assign(static_cast<size_t>(consumed_) * item_bytes_, 0);
    // ---- tracking side: the lock-step bank over one correlator engine
    for (int i = 0; i < n_channels; i++)
        {
            auto adapter = GetB200TrkBlock(configuration, trk_role, 1, 1);
            if (!adapter) throw std::runtime_error(trk_role + ".implementation is not a B200 tracking block");
            auto* a = static_cast<B200DllPllTracking*>(adapter.get());
            if (a->item_size() == 0) throw std::runtime_error("unsupported tracking item type");
            if (a->item_size() != item_bytes_) throw std::invalid_argument("acquisition and tracking must read the same item type");
            if (!engine_)
                {
                    EngineConfig ec;
                    ec.device = a->conf().device;
                    ec.iq_dtype = item_type_iq(a->conf().item_type);
                    ec.max_channels = 4 * n_channels;
                    ec.max_code_len = 2 * a->block()->code_table_entries();
                    ec.max_block = static_cast<int>(2 * a->conf().vector_length);
                    ec.ring_samples = std::max<int64_t>(int64_t{1} << 21, 64LL * a->conf().vector_length);
                    engine_ = std::make_shared<B200CorrelatorEngine>(ec);
                    bank_ = std::make_unique<B200TrackingBank>(engine_);
                }
            synchros_.push_back(std::make_unique<Gnss_Synchro>());
            auto blk = std::make_unique<b200_dll_pll_veml_tracking>(a->conf());
            blk->set_gnss_synchro(synchros_.back().get());
            blk->set_channel(static_cast<uint32_t>(i));
            bank_->add_channel(std::move(blk));
            info_.push_back(ChannelInfo{0, 0U, 0, 0, 0.0, 0.0, 0ULL, 0.0F});
        }
}

bool B200ChannelManager::any_idle() const
{
    return std::any_of(info_.begin(), info_.end(), [](const ChannelInfo& c) { return c.state != 2; });
}

// One sweep = what every acquiring channel's acquisition_core does in the reference (pcps_acquisition.cc:600-871), for all
// candidate satellites at once; then the flowgraph's part: valid detections go to idle channels (Event_valid_acquisition ->
// start_tracking), strongest first, satellites already in tracking are skipped.
void B200ChannelManager::sweep(uint64_t stamp)
{
    sweep_->run(buf_.data(), 0, &results_);
    sweeps_++;
    std::vector<int> order;
    for (size_t s = 0; s < prns_.size(); s++)
        {
            if (!(results_[s].test_statistics > threshold_)) continue;  // :770 strict '>'
            const bool tracked = std::any_of(info_.begin(), info_.end(), [&](const ChannelInfo& c) { return c.state == 2 && c.prn == prns_[s]; });
            if (!tracked) order.push_back(static_cast<int>(s));
        }
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return results_[a].test_statistics > results_[b].test_statistics; });
    size_t next = 0;
    for (int ch = 0; ch < size() && next < order.size(); ch++)
        {
            ChannelInfo& c = info_[ch];
            if (c.state == 2) continue;
            const gb200_acq_result& r = results_[order[next]];
            Gnss_Synchro& s = *synchros_[ch];
            const Dll_Pll_Conf& p = bank_->block(ch).parameters();
            s = Gnss_Synchro{};
            s.System = p.system;
            std::memcpy(s.Signal, p.signal, 3);
            s.PRN = prns_[order[next]];
            s.Channel_ID = ch;
            // what acquisition_core leaves in the Gnss_Synchro (:751-757)
            s.Acq_delay_samples = static_cast<double>(r.acq_delay_samples);
            s.Acq_doppler_hz = static_cast<double>(r.acq_doppler_hz);
            s.Acq_samplestamp_samples = stamp;
            s.Acq_doppler_step = static_cast<uint32_t>(acq_conf_.doppler_step);
            s.Flag_valid_acquisition = true;
            bank_->start_tracking(ch, stamp);
            c.state = 2;
            c.prn = s.PRN;
            c.acquisitions++;
            c.acq_doppler_hz = s.Acq_doppler_hz;
            c.acq_delay_samples = s.Acq_delay_samples;
            c.acq_samplestamp_samples = stamp;
            c.test_statistics = r.test_statistics;
            next++;
        }
    // channels left without a detection stay in acquisition: the next buffer is searched again (Event_failed_acquisition_repeat)
}

void B200ChannelManager::push(const void* iq, int64_t n_samples, std::vector<B200TrackingBank::Output>* out)
{
    const char* src = static_cast<const char*>(iq);
    int64_t done = 0;
    // the device ring takes the samples in pieces it can hold; the acquisition buffer fills from the same samples
    const int64_t piece = std::max<int64_t>(consumed_, engine_->config().ring_samples / 4);
    while (done < n_samples)
        {
            const int64_t n = std::min<int64_t>(piece, n_samples - done);
            bank_->push(src + done * item_bytes_, n);
            int64_t used = 0;
            while (used < n)
                {
                    if (!any_idle())
                        {
                            buffered_ = 0;  // nobody is acquiring: like a block in standby, samples pass by
                            break;
                        }
                    for (auto& c : info_)
                        if (c.state == 0) c.state = 1;  // Event_start_acquisition
                    const uint32_t take = static_cast<uint32_t>(std::min<int64_t>(consumed_ - buffered_, n - used));
                    std::memcpy(buf_.data() + static_cast<size_t>(buffered_) * item_bytes_, src + (done + used) * item_bytes_, static_cast<size_t>(take) * item_bytes_);
                    buffered_ += take;
                    used += take;
                    if (buffered_ == consumed_)
                        {
                            sweep(pos_ + static_cast<uint64_t>(used));  // d_sample_counter at the end of the buffered block
                            buffered_ = 0;
                        }
                }
            pos_ += static_cast<uint64_t>(n);
            done += n;
            const size_t first = out->size();
            bank_->step(out);
            (void)first;
            // loss of lock: the block went back to standby (dll_pll_veml_tracking.cc:1018-1028, message 3) ->
            // ChannelFsm::Event_failed_tracking_standby; the satellite returns to the candidates
            for (int ch = 0; ch < size(); ch++)
                if (info_[ch].state == 2 && bank_->block(ch).state() == 0)
                    {
                        info_[ch].state = 0;
                        info_[ch].losses++;
                    }
        }
}
}  // namespace gb200h
