// PCPS acquisition block over the GPU engine.
//
// b200_pcps_acquisition mirrors pcps_acquisition (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.{h:92-290,
// cc:53-1031}): same members a channel/adapter calls (init, set_local_code, set_state, set_threshold, set_doppler_*,
// mag, set_active), the same buffering FSM in general_work (0 idle / 1 fill / 2 run), the same dwell bookkeeping and
// events (1 positive, 2 negative), results written in place into the shared Gnss_Synchro (:691-696).  acquisition_core
// hands the whole Doppler search to gb200_acq_run.
//
// B200AcquisitionSweep is the batched form the GPU makes possible: all PRNs x all Doppler bins of one signal on one
// sample block in a single sweep (the forward FFT of each Doppler bin is shared by every PRN).
#ifndef GB200_PCPS_ACQUISITION_H
#define GB200_PCPS_ACQUISITION_H

#include "b200_conf.h"
#include "gnss_b200.h"
#include "gnss_sdr_compat.h"
#include <complex>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

namespace gb200h
{
class b200_pcps_acquisition
{
public:
    // throws std::runtime_error when the GPU engine cannot be created (no CPU fallback)
    explicit b200_pcps_acquisition(const Acq_Conf& conf);
    ~b200_pcps_acquisition();

    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro) { d_gnss_synchro = p_gnss_synchro; }
    uint32_t mag() const { return static_cast<uint32_t>(d_mag); }
    void init();
    // `code`: the sampled replica the adapter builds (vector_length_ samples, adapters/gps_l1_ca_pcps_acquisition.cc:151-170)
    void set_local_code(std::complex<float>* code);
    void set_active(bool active) { d_active = active; }
    void set_state(int32_t state);
    void set_channel(uint32_t channel) { d_channel = channel; }
    void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) { d_channel_fsm = std::move(channel_fsm); }
    void set_threshold(float threshold) { d_threshold = threshold; }
    void set_doppler_max(uint32_t doppler_max) { d_acq_parameters.doppler_max = static_cast<int32_t>(doppler_max); }
    void set_doppler_step(uint32_t doppler_step) { d_doppler_step = doppler_step; }
    void set_doppler_center(int32_t doppler_center);
    void set_resampler_latency(uint32_t latency_samples) { d_acq_parameters.resampler_latency_samples = latency_samples; }
    // "events" message port: 1 positive, 2 negative acquisition (:362,389), used when no ChannelFsm is attached
    void set_event_sink(std::function<void(int)> sink) { event_sink_ = std::move(sink); }
    bool start();  // resets the sample counter and computes the threshold from pfa (:873-881)

    // general_work (:902-1031): returns the number of input items consumed
    int general_work(const void* input_items, int ninput_items);
    void acquisition_core(uint64_t samp_count);

    // observers
    int state() const { return d_state; }
    bool active() const { return d_active; }
    bool step_two() const { return d_step_two; }  // armed for (or running) the fine-Doppler step of make_two_steps
    float threshold() const { return d_threshold; }
    float test_statistics() const { return d_test_statistics; }
    float input_power() const { return d_input_power; }
    uint32_t fft_size() const { return d_fft_size; }
    uint32_t consumed_samples() const { return d_consumed_samples; }
    uint32_t num_doppler_bins() const { return d_num_doppler_bins; }
    uint64_t sample_counter() const { return d_sample_counter; }
    const gb200_acq_result& last_result() const { return last_; }
    // magnitude grid of the last run (bins x effective_fft_size), for dumps (:393-493)
    std::vector<float> magnitude_grid();

private:
    void calculate_threshold();
    void send_positive_acquisition();
    void send_negative_acquisition();
    void apply_grid();

    Acq_Conf d_acq_parameters;
    gb200_acq* h_{nullptr};
    gb200_acq* h2_{nullptr};  // second (fine Doppler) step of make_two_steps: its own grid of second_nbins bins
    bool d_step_two{false};
    float d_doppler_center_step_two{0.0F};
    void update_grid_doppler_wipeoffs_step2();
    void dump_results(int32_t effective_fft_size);  // :393-493
    bool d_dump{false};
    long long d_dump_number{0};
    std::string d_dump_filename;
    Gnss_Synchro* d_gnss_synchro{nullptr};
    std::weak_ptr<ChannelFsm> d_channel_fsm;
    std::function<void(int)> event_sink_;
    std::mutex d_setlock;
    std::vector<unsigned char> d_data_buffer;  // consumed_samples items of it_size bytes
    gb200_acq_result last_{};
    uint64_t d_sample_counter{0};
    float d_threshold{0.0F}, d_mag{0.0F}, d_input_power{0.0F}, d_test_statistics{0.0F};
    int32_t d_state{0}, d_positive_acq{0}, d_doppler_center{0};
    uint32_t d_channel{0}, d_doppler_step{0}, d_num_noncoherent_integrations_counter{0}, d_fft_size{0}, d_consumed_samples{0};
    uint32_t d_num_doppler_bins{0}, d_buffer_count{0};
    int iq_dtype_{GB200_IQ_CF32};
    bool d_active{false}, d_worker_active{false}, d_use_CFAR_algorithm_flag{true}, grid_dirty_{true};
};

// All PRNs of one signal in one sweep.
class B200AcquisitionSweep
{
public:
    B200AcquisitionSweep(const Acq_Conf& conf, int n_prn);
    ~B200AcquisitionSweep();
    void set_local_code(int slot, const std::complex<float>* code, int n_code);
    void set_doppler_grid(int doppler_max, int doppler_step, int doppler_center);
    float threshold(float pfa) const { return gb200_acq_threshold(h_, pfa); }
    int num_doppler_bins() const { return gb200_acq_num_doppler_bins(h_); }
    // one dwell over `consumed_samples` host samples for every slot; results[i] for slot i
    void run(const void* iq_host, int dwell_index, std::vector<gb200_acq_result>* results);

private:
    gb200_acq* h_{nullptr};
    int n_prn_;
};
}  // namespace gb200h
#endif
