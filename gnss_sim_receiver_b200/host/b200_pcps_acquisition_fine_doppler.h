// Fine-Doppler PCPS acquisition over the GPU engine.
//
// b200_pcps_acquisition_fine_doppler mirrors pcps_acquisition_fine_doppler_cc
// (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition_fine_doppler_cc.{h:70-240,cc:40-563}): the members a channel /
// adapter calls (set_doppler_step, set_local_code, init, set_state, set_active, set_threshold, set_doppler_max), the state machine
// of general_work (0 stand-by / 1 grid per 1 ms block, max_dwells of them / 2 decision / 3 buffering 10 ms + fine estimate /
// 4 positive / 5 negative) and the results written in place into the shared Gnss_Synchro.
//
// On the GPU: the coarse search is the PCPS engine with the block's own grid (floor(2 doppler_max / doppler_step) bins from
// -doppler_max, :128-137,:192-208), non-coherent accumulation over the dwells (:330) and the first-to-second-peak statistic (:211-283);
// the fine estimate (estimate_Doppler, :338-402) is gb200_acq_spectrum_peak on a second handle.  There is no CPU path.
#ifndef GB200_PCPS_ACQUISITION_FINE_DOPPLER_H
#define GB200_PCPS_ACQUISITION_FINE_DOPPLER_H

#include "b200_conf.h"
#include "gnss_b200.h"
#include "gnss_sdr_compat.h"
#include <complex>
#include <functional>
#include <string>
#include <vector>

namespace gb200h
{
class b200_pcps_acquisition_fine_doppler
{
public:
    // conf: fs_in, samples_per_ms, samples_per_chip, doppler_max, max_dwells, blocking_on_standby (the fields the reference's adapter
    // fills, adapters/gps_l1_ca_pcps_acquisition_fine_doppler.cc:51-69).  Throws when the engines cannot be created.
    explicit b200_pcps_acquisition_fine_doppler(const Acq_Conf& conf);
    ~b200_pcps_acquisition_fine_doppler();
    b200_pcps_acquisition_fine_doppler(const b200_pcps_acquisition_fine_doppler&) = delete;
    b200_pcps_acquisition_fine_doppler& operator=(const b200_pcps_acquisition_fine_doppler&) = delete;

    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) { d_gnss_synchro = p_gnss_synchro; }
    unsigned int mag() const { return static_cast<unsigned int>(d_test_statistics); }  // .h:104-107
    void init();
    void set_local_code(std::complex<float>* code);
    void set_active(bool active) { d_active = active; }
    void set_state(int state);
    void set_channel(unsigned int channel) { d_channel = channel; }
    void set_threshold(float threshold) { d_threshold = threshold; }
    void set_doppler_max(unsigned int doppler_max) { d_config_doppler_max = static_cast<int>(doppler_max); }
    void set_doppler_step(unsigned int doppler_step);
    void set_event_sink(std::function<void(int)> sink) { event_sink_ = std::move(sink); }  // the "events" port: 1 positive, 2 negative
    bool start();

    // general_work (:450-563): `in` holds ninput_items gr_complex items (forecast asks for fft_size of them, :171-178), noutput_items as
    // the scheduler gives it; returns the number of items consumed (consume_each)
    int general_work(int noutput_items, const std::complex<float>* in, int ninput_items);

    int state() const { return d_state; }
    bool active() const { return d_active; }
    float threshold() const { return d_threshold; }
    float test_statistics() const { return d_test_statistics; }
    int fft_size() const { return d_fft_size; }
    int num_doppler_points() const { return d_num_doppler_points; }
    int positive_acq() const { return d_positive_acq; }
    uint64_t sample_counter() const { return d_sample_counter; }
    int64_t fine_index() const { return fine_index_; }  // index of the padded-spectrum maximum of the last estimate_Doppler

private:
    void reset_grid() { d_well_count = 0; }  // the engine overwrites its grid on dwell 0 (:181-190)
    void compute_and_accumulate_grid(const std::complex<float>* in);
    float compute_CAF();
    void estimate_Doppler();

    Acq_Conf acq_parameters;
    gb200_acq* h_{nullptr};    // coarse search, fft_size = samples per ms
    gb200_acq* h10_{nullptr};  // fine estimate: transforms of 10 * fft_size / fold points
    int fold_{1};
    Gnss_Synchro* d_gnss_synchro{nullptr};
    std::function<void(int)> event_sink_;
    std::vector<std::complex<float>> d_10_ms_buffer;
    gb200_acq_result last_{};
    int64_t d_fs_in;
    uint64_t d_sample_counter{0};
    int64_t fine_index_{-1};
    float d_threshold{0.0F}, d_test_statistics{0.0F};
    int d_positive_acq{0}, d_state{0}, d_samples_per_ms, d_max_dwells, d_config_doppler_max, d_num_doppler_points{0}, d_well_count{0};
    int d_n_samples_in_buffer{0}, d_fft_size;
    unsigned int d_doppler_step{0}, d_channel{0};
    bool d_active{false};
};
}  // namespace gb200h
#endif
