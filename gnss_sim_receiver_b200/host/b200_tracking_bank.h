// Lock-step driver of many tracking blocks over one correlator engine: the B200-native replacement for "one GNU Radio
// thread per tracking block".  Every channel keeps its own read pointer into the shared stream (its nitems_read);
// step() advances every channel that has 2 * vector_length samples available (the reference's forecast,
// dll_pll_veml_tracking.cc:607-614) by one general_work call, with ALL correlations of the round in one kernel launch.
#ifndef GB200_TRACKING_BANK_H
#define GB200_TRACKING_BANK_H

#include "b200_dll_pll_veml_tracking.h"
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace gb200h
{
class B200TrackingBank
{
public:
    B200TrackingBank(std::shared_ptr<B200CorrelatorEngine> engine) : engine_(std::move(engine)) {}
    // adds a channel; returns its index.  The block's Gnss_Synchro must have been set.
    int add_channel(std::unique_ptr<b200_dll_pll_veml_tracking> block);
    int size() const { return static_cast<int>(ch_.size()); }
    b200_dll_pll_veml_tracking& block(int i) { return *ch_[i].block; }
    uint64_t nitems_read(int i) const { return ch_[i].read; }
    // start_tracking on channel i: uploads its code tables and puts the read pointer at `nitems_read`
    void start_tracking(int i, uint64_t nitems_read);
    // append stream samples (engine's iq_dtype) to the device ring
    void push(const void* iq, int64_t n_samples) { engine_->push(iq, n_samples); }
    int64_t push_raw(const void* raw, int64_t n_bytes, const gb200_ingest_format& fmt)
    {
        int64_t n = 0;
        engine_->push_raw(raw, n_bytes, fmt, &n);
        return n;
    }
    // Run rounds until no channel can advance.  Every produced Gnss_Synchro is appended to `out` with its channel index.
    // Returns the number of kernel launches.
    struct Output
    {
        int channel;
        Gnss_Synchro synchro;
    };
    int step(std::vector<Output>* out);

private:
    struct Chan
    {
        std::unique_ptr<b200_dll_pll_veml_tracking> block;
        int slot{-1}, data_slot{-1};
        uint64_t read{0};
    };
    std::shared_ptr<B200CorrelatorEngine> engine_;
    std::vector<Chan> ch_;
    std::vector<gb200_trk_cmd> cmds_;
    // correlator outputs of one round in page-locked memory (gb200_host_alloc): the engine DMAs results straight into it
    struct PinnedTaps
    {
        gb200_cf32* p{nullptr};
        size_t cap{0};
        PinnedTaps() = default;
        PinnedTaps(const PinnedTaps&) = delete;
        PinnedTaps& operator=(const PinnedTaps&) = delete;
        ~PinnedTaps() { gb200_host_free(p); }
        void resize(size_t n)
        {
            if (n <= cap) return;
            gb200_host_free(p);
            p = nullptr;
            cap = 0;
            void* q = nullptr;
            const size_t want = n + n / 2;
            if (gb200_host_alloc(&q, want * sizeof(gb200_cf32)) != GB200_OK) throw std::runtime_error(std::string("gb200_host_alloc: ") + gb200_last_error());
            p = static_cast<gb200_cf32*>(q);
            cap = want;
        }
        gb200_cf32* data() { return p; }
        const gb200_cf32& operator[](size_t i) const { return p[i]; }
    };
    PinnedTaps res_;
    std::vector<int> who_, data_idx_;
    std::vector<b200_dll_pll_veml_tracking::Step> steps_;
};
}  // namespace gb200h
#endif
