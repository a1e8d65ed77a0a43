// QuickSync PCPS acquisition over the GPU engine.
//
// b200_pcps_quicksync_acquisition mirrors pcps_quicksync_acquisition_cc
// (src/algorithms/acquisition/gnuradio_blocks/pcps_quicksync_acquisition_cc.{h:80-260,cc:54-550}): the members the adapter calls, the state
// machine of general_work over VECTOR items (0 stand-by / 1 one item = one dwell / 2 positive / 3 negative) and the results written into the
// shared Gnss_Synchro.  An item is samples_per_code * folding_factor samples (adapters/gps_l1_ca_pcps_quicksync_acquisition.cc:113-114).
//
// On the GPU: the block is wiped off over its whole length per Doppler bin and folded folding_factor^2 times into one
// samples_per_code / folding_factor point transform (gb200_acq_set_input_folding), correlated with the folded replica by the PCPS engine;
// the folding_factor candidate delays of the winning bin are resolved by time-domain correlations against the unfolded replica
// (gb200_acq_folded_candidates).  The mean input power is one pass over the item in the reference's summation order on the host.
#ifndef GB200_PCPS_QUICKSYNC_ACQUISITION_H
#define GB200_PCPS_QUICKSYNC_ACQUISITION_H

#include "gnss_b200.h"
#include "gnss_sdr_compat.h"
#include <complex>
#include <functional>
#include <vector>

namespace gb200h
{
class b200_pcps_quicksync_acquisition
{
public:
    // the argument list of pcps_quicksync_make_acquisition_cc (.h:60-70) without the dump / monitor ones, plus the device
    b200_pcps_quicksync_acquisition(uint32_t folding_factor, uint32_t sampled_ms, uint32_t max_dwells, uint32_t doppler_max, int64_t fs_in, int32_t samples_per_ms,
        int32_t samples_per_code, bool bit_transition_flag, int device);
    ~b200_pcps_quicksync_acquisition();
    b200_pcps_quicksync_acquisition(const b200_pcps_quicksync_acquisition&) = delete;
    b200_pcps_quicksync_acquisition& operator=(const b200_pcps_quicksync_acquisition&) = delete;

    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) { d_gnss_synchro = p_gnss_synchro; }
    uint32_t mag() const { return static_cast<uint32_t>(d_mag); }
    void init();
    void set_local_code(std::complex<float>* code);
    void set_active(bool active) { d_active = active; }
    void set_state(int32_t state);
    void set_channel(uint32_t channel) { d_channel = channel; }
    void set_threshold(float threshold) { d_threshold = threshold; }
    void set_doppler_max(uint32_t doppler_max) { d_doppler_max = doppler_max; }
    void set_doppler_step(uint32_t doppler_step) { d_doppler_step = doppler_step; }
    void set_event_sink(std::function<void(int)> sink) { event_sink_ = std::move(sink); }

    // general_work (:214-550): `in` holds ninput_items vector items of item_length() samples each; returns the ITEMS consumed
    int general_work(const std::complex<float>* in, int ninput_items);
    // samples per vector item: samples_per_code * folding_factor are read of each; the Galileo adapter's stream_to_vector makes items of
    // sampled_ms * samples_per_ms samples (galileo_e1_pcps_quicksync_ambiguous_acquisition.cc:129-130), which is longer when the coherent time
    // exceeds folding_factor code periods
    int item_length() const { return item_length_ ? item_length_ : d_samples_per_code * static_cast<int>(d_folding_factor); }
    void set_item_length(int n);

    int state() const { return d_state; }
    bool active() const { return d_active; }
    float threshold() const { return d_threshold; }
    float test_statistics() const { return d_test_statistics; }
    float input_power() const { return d_input_power; }
    float mag_f() const { return d_mag; }
    uint32_t fft_size() const { return d_fft_size; }
    uint32_t num_doppler_bins() const { return d_num_doppler_bins; }
    uint32_t well_count() const { return d_well_count; }
    uint64_t sample_counter() const { return d_sample_counter; }

private:
    void dwell(const std::complex<float>* in);

    gb200_acq* h_{nullptr};
    Gnss_Synchro* d_gnss_synchro{nullptr};
    std::function<void(int)> event_sink_;
    std::vector<std::complex<float>> d_code;
    int64_t d_fs_in;
    uint64_t d_sample_counter{0};
    float d_threshold{0.0F}, d_mag{0.0F}, d_input_power{0.0F}, d_test_statistics{0.0F};
    int32_t d_samples_per_ms, d_samples_per_code, d_state{0};
    uint32_t d_channel{0}, d_folding_factor, d_doppler_max, d_doppler_step{0}, d_sampled_ms, d_max_dwells, d_well_count{0}, d_fft_size, d_num_doppler_bins{0};
    int device_, item_length_{0};
    bool d_bit_transition_flag, d_active{false};
};
}  // namespace gb200h
#endif
