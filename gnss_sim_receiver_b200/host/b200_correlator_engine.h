// Per-GPU correlator engine and the per-channel facade the tracking blocks hold.
//
// The reference gives every tracking block its own Cpu_Multicorrelator_Real_Codes
// (src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.{h:37-61,cc:36-160}) and every block thread correlates its
// own copy of the stream.  Here all blocks of one GPU share ONE engine: one device IQ ring (each sample block is uploaded
// once, by whichever channel asks for it first) and one kernel launch per batch of channel requests.
//   * B200_Multicorrelator_Real_Codes — same methods and argument meaning as the reference class, so a tracking block's
//     do_correlation_step (dll_pll_veml_tracking.cc:1037-1062) is unchanged apart from passing the absolute sample index;
//   * B200CorrelatorEngine::correlate_batch — the lock-step form used by B200TrackingBank and the benches.
#ifndef GB200_CORRELATOR_ENGINE_H
#define GB200_CORRELATOR_ENGINE_H

#include "gnss_b200.h"
#include <complex>
#include <condition_variable>
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

namespace gb200h
{
struct EngineConfig
{
    int device{0};
    int iq_dtype{GB200_IQ_CF32};
    int stream_id{0};                // which input stream (RF band / antenna) the ring holds: one engine per (device, iq_dtype, stream)
    int max_channels{256};
    int max_code_len{10230};
    int max_block{1 << 17};          // longest correlation window, samples
    int64_t ring_samples{1 << 22};   // device ring capacity, samples
    int batch_window_us{200};        // how long a batch leader waits for the other channels' requests
};

class B200CorrelatorEngine
{
public:
    // one engine per (device, iq_dtype, stream_id): the ring is ONE stream's samples addressed by nitems_read, so two streams of
    // the same item type (an L1 band and an L5 band, two antennas) never share it.  Created on first use with `cfg`; a later call
    // whose channel needs a longer window, a longer code table or a deeper ring than the existing engine was sized for throws
    // std::invalid_argument (size the first one from the receiver-wide maximum, as the adapters do).
    // Throws std::runtime_error (message from gb200_last_error) when the GPU engine cannot be created: no CPU fallback.
    static std::shared_ptr<B200CorrelatorEngine> instance(const EngineConfig& cfg);
    // drop the process-wide instances (tests)
    static void reset_instances();
    explicit B200CorrelatorEngine(const EngineConfig& cfg);
    ~B200CorrelatorEngine();
    B200CorrelatorEngine(const B200CorrelatorEngine&) = delete;
    B200CorrelatorEngine& operator=(const B200CorrelatorEngine&) = delete;

    const EngineConfig& config() const { return cfg_; }
    int attach_channel();              // returns a channel slot, -1 when full
    void detach_channel(int ch);
    void set_code(int ch, const float* code, int code_len);
    void set_code(int ch, const std::complex<float>* code, int code_len);

    // Append samples to the ring; returns the absolute index of the first one.
    uint64_t push(const void* iq, int64_t n_samples);
    // raw file items (File_Signal_Source item types, 2-bit packed): converted to the ring type on the device (gb200_trk_push_raw);
    // *n_samples receives how many samples the bytes held
    uint64_t push_raw(const void* raw, int64_t n_bytes, const gb200_ingest_format& fmt, int64_t* n_samples);
    // Make sure the ring holds stream samples [abs_index, abs_index + n): uploads only the part beyond the ring's head.
    // `iq` points at the sample with absolute index abs_index.  Returns false when abs_index is older than the ring.
    bool ensure(uint64_t abs_index, const void* iq, int64_t n_samples);
    uint64_t head() const;

    // lock-step form: n commands in one launch; out[i * GB200_MAX_TAPS + t]
    void correlate_batch(const gb200_trk_cmd* cmds, int n, gb200_cf32* out);
    // per-channel blocking form: requests arriving from several threads within batch_window_us share one launch
    void correlate_one(const gb200_trk_cmd& cmd, gb200_cf32* out_taps);
    // number of channels that will take part in the next batches (a leader stops waiting once all have arrived)
    void set_expected_batch(int n) { expected_ = n; }
    uint64_t launches() const { return launches_; }
    uint64_t batched_requests() const { return batched_; }

private:
    EngineConfig cfg_;
    gb200_trk* h_{nullptr};
    mutable std::mutex mu_;       // engine state (ring, codes, launches)
    std::vector<std::pair<uint64_t, uint64_t>> gaps_;  // stretches [a, b) behind the head that were skipped, not uploaded
    std::vector<char> used_;
    // batching of per-channel requests: the first request opens a batch and leads it
    struct Batch;
    std::mutex bmu_;
    std::condition_variable bcv_;
    std::shared_ptr<Batch> open_;
    int expected_{0};
    uint64_t launches_{0}, batched_{0};
};

// Drop-in for Cpu_Multicorrelator_Real_Codes: same call sequence
//   init -> set_local_code_and_taps -> [set_input_output_vectors -> Carrier_wipeoff_multicorrelator_resampler]* -> free
class B200_Multicorrelator_Real_Codes
{
public:
    B200_Multicorrelator_Real_Codes() = default;
    ~B200_Multicorrelator_Real_Codes();
    void set_engine(std::shared_ptr<B200CorrelatorEngine> engine) { engine_ = std::move(engine); }
    void set_high_dynamics_resampler(bool use_high_dynamics_resampler) { high_dyn_ = use_high_dynamics_resampler; }
    bool init(int max_signal_length_samples, int n_correlators);
    bool set_local_code_and_taps(int code_length_chips, const float* local_code_in, float* shifts_chips);
    // `abs_sample_index` = absolute stream index of sig_in[0] (nitems_read(0) in a GNU Radio block); `n_available` =
    // samples readable at sig_in (the block's forecast guarantees 2 * vector_length, dll_pll_veml_tracking.cc:607-614)
    bool set_input_output_vectors(std::complex<float>* corr_out, const void* sig_in, uint64_t abs_sample_index, int n_available);
    void update_local_code(int correlator_length_samples, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips = 0.0);
    bool Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad, float phase_rate_step_rad,
        float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples);
    bool free();
    int channel() const { return ch_; }

private:
    std::shared_ptr<B200CorrelatorEngine> engine_;
    std::complex<float>* corr_out_{nullptr};
    const void* sig_in_{nullptr};
    const float* shifts_{nullptr};
    uint64_t abs_index_{0};
    int n_available_{0};
    int ch_{-1}, n_correlators_{0}, max_len_{0};
    bool high_dyn_{false};
};
}  // namespace gb200h
#endif
