// Galileo E5a non-coherent I/Q acquisition with CAF filtering over the GPU engine.
//
// b200_e5a_noncoherent_iq_acquisition_caf mirrors galileo_e5a_noncoherentIQ_acquisition_caf_cc
// (src/algorithms/acquisition/gnuradio_blocks/galileo_e5a_noncoherent_iq_acquisition_caf_cc.{h:70-250,cc:55-777}): the members the adapter
// calls, the buffering state machine of general_work (0 stand-by / 1 fill / 2 search / 3 positive / 4 negative) and the results written
// into the shared Gnss_Synchro.
//
// On the GPU: the up to four replicas of the block (I_A, Q_A and — for more than one code period — the sign-flipped I_B, Q_B, :160-206)
// are four code-spectrum slots of ONE engine sweep per buffer (the forward transform of every Doppler bin shared by the four); the
// per-bin choice between A and B and the non-coherent sum of the chosen I and Q grids followed by the first maximum (:443-531) run on the
// kept grid on the device (gb200_acq_grid_pair_max); the CAF smoothing over Doppler bins (:590-690) is a few dozen numbers on the host.
// There is no CPU path for the search.
#ifndef GB200_E5A_NONCOHERENT_IQ_ACQUISITION_CAF_H
#define GB200_E5A_NONCOHERENT_IQ_ACQUISITION_CAF_H

#include "gnss_b200.h"
#include "gnss_sdr_compat.h"
#include <complex>
#include <functional>
#include <string>
#include <vector>

namespace gb200h
{
class b200_e5a_noncoherent_iq_acquisition_caf
{
public:
    // the argument list of galileo_e5a_noncoherentIQ_make_acquisition_caf_cc (.h:52-66) without the dump / monitor ones, plus the device
    b200_e5a_noncoherent_iq_acquisition_caf(unsigned int sampled_ms, unsigned int max_dwells, unsigned int doppler_max, int64_t fs_in, int samples_per_ms, int samples_per_code,
        bool bit_transition_flag, bool both_signal_components, int CAF_window_hz, int Zero_padding, int device);
    ~b200_e5a_noncoherent_iq_acquisition_caf();
    b200_e5a_noncoherent_iq_acquisition_caf(const b200_e5a_noncoherent_iq_acquisition_caf&) = delete;
    b200_e5a_noncoherent_iq_acquisition_caf& operator=(const b200_e5a_noncoherent_iq_acquisition_caf&) = delete;

    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) { d_gnss_synchro = p_gnss_synchro; }
    unsigned int mag() const { return static_cast<unsigned int>(d_mag); }
    void init();
    void set_local_code(std::complex<float>* codeI, std::complex<float>* codeQ);
    void set_active(bool active) { d_active = active; }
    void set_state(int state);
    void set_channel(unsigned int channel) { d_channel = channel; }
    void set_threshold(float threshold) { d_threshold = threshold; }
    void set_doppler_max(unsigned int doppler_max) { d_doppler_max = static_cast<int>(doppler_max); }
    void set_doppler_step(unsigned int doppler_step) { d_doppler_step = static_cast<int>(doppler_step); }
    void set_event_sink(std::function<void(int)> sink) { event_sink_ = std::move(sink); }

    // general_work (:295-777): ninput_items gr_complex items at `in`; returns the items consumed
    int general_work(const std::complex<float>* in, int ninput_items);

    int state() const { return d_state; }
    bool active() const { return d_active; }
    float threshold() const { return d_threshold; }
    float test_statistics() const { return d_test_statistics; }
    float input_power() const { return d_input_power; }
    float mag_f() const { return d_mag; }
    int fft_size() const { return d_fft_size; }
    int num_doppler_bins() const { return d_num_doppler_bins; }
    int well_count() const { return d_well_count; }
    uint64_t sample_counter() const { return d_sample_counter; }
    const std::vector<float>& caf_vector() const { return d_CAF_vector; }

private:
    void search();  // the Doppler loop of state 2 (:383-560) and the CAF filter (:590-700)

    gb200_acq* h_{nullptr};
    Gnss_Synchro* d_gnss_synchro{nullptr};
    std::function<void(int)> event_sink_;
    std::vector<std::complex<float>> d_inbuffer;
    std::vector<float> d_CAF_vector, d_CAF_vector_I, d_CAF_vector_Q;
    int64_t d_fs_in;
    uint64_t d_sample_counter{0};
    float d_threshold{0.0F}, d_mag{0.0F}, d_input_power{0.0F}, d_test_statistics{0.0F};
    int d_state{0}, d_samples_per_ms, d_samples_per_code, d_CAF_window_hz, d_buffer_count{0}, d_doppler_max, d_doppler_step{250}, d_fft_size, d_num_doppler_bins{0};
    int d_gr_stream_buffer{0};
    unsigned int d_channel{0}, d_max_dwells, d_well_count{0}, d_sampled_ms;
    int n_slots_{1}, slot_qa_{-1}, slot_ib_{-1}, slot_qb_{-1};
    bool d_bit_transition_flag, d_active{false}, d_both_signal_components;
};
}  // namespace gb200h
#endif
