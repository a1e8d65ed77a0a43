#include "b200_correlator_engine.h"
#include <algorithm>
#include <chrono>
#include <cstring>
#include <map>
#include <tuple>
#include <stdexcept>

namespace gb200h
{
namespace
{
[[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string(what) + ": " + gb200_last_error()); }
std::mutex g_inst_mu;
std::map<std::tuple<int, int, int>, std::shared_ptr<B200CorrelatorEngine>> g_instances;
int elem_bytes(int iq) { return iq == GB200_IQ_CF32 ? 8 : iq == GB200_IQ_CS16 ? 4 : 2; }
}  // namespace

struct B200CorrelatorEngine::Batch
{
    std::vector<gb200_trk_cmd> cmds;
    std::vector<gb200_cf32*> outs;
    bool closed{false}, done{false};
    std::string error;
};

std::shared_ptr<B200CorrelatorEngine> B200CorrelatorEngine::instance(const EngineConfig& cfg)
{
    std::lock_guard<std::mutex> lk(g_inst_mu);
    auto key = std::make_tuple(cfg.device, cfg.iq_dtype, cfg.stream_id);
    auto it = g_instances.find(key);
    if (it != g_instances.end())
        {
            const EngineConfig& have = it->second->config();
            if (cfg.max_block > have.max_block || cfg.max_code_len > have.max_code_len)
                throw std::invalid_argument("B200CorrelatorEngine: the engine of device " + std::to_string(cfg.device) + " stream " + std::to_string(cfg.stream_id) +
                                            " was created for windows of " + std::to_string(have.max_block) + " samples and code tables of " + std::to_string(have.max_code_len) +
                                            " entries; this channel needs " + std::to_string(cfg.max_block) + " / " + std::to_string(cfg.max_code_len));
            return it->second;
        }
    auto e = std::make_shared<B200CorrelatorEngine>(cfg);
    g_instances[key] = e;
    return e;
}

void B200CorrelatorEngine::reset_instances()
{
    std::lock_guard<std::mutex> lk(g_inst_mu);
    g_instances.clear();
}

B200CorrelatorEngine::B200CorrelatorEngine(const EngineConfig& cfg) : cfg_(cfg)
{
    gb200_trk_config c{};
    c.device = cfg.device;
    c.max_channels = cfg.max_channels;
    c.max_code_len = cfg.max_code_len;
    c.iq_dtype = cfg.iq_dtype;
    c.ring_samples = std::max<int64_t>(cfg.ring_samples, 2LL * cfg.max_block);
    c.max_block = cfg.max_block;
    c.max_cmds = cfg.max_channels;
    cfg_.ring_samples = c.ring_samples;
    if (gb200_trk_create(&c, &h_) != GB200_OK) fail("gb200_trk_create");
    used_.assign(cfg.max_channels, 0);
}

B200CorrelatorEngine::~B200CorrelatorEngine() { gb200_trk_destroy(h_); }

int B200CorrelatorEngine::attach_channel()
{
    std::lock_guard<std::mutex> lk(mu_);
    for (size_t i = 0; i < used_.size(); i++)
        if (!used_[i])
            {
                used_[i] = 1;
                return static_cast<int>(i);
            }
    return -1;
}

void B200CorrelatorEngine::detach_channel(int ch)
{
    std::lock_guard<std::mutex> lk(mu_);
    if (ch >= 0 && ch < static_cast<int>(used_.size())) used_[ch] = 0;
}

void B200CorrelatorEngine::set_code(int ch, const float* code, int code_len)
{
    std::lock_guard<std::mutex> lk(mu_);
    if (gb200_trk_set_code(h_, ch, code, code_len) != GB200_OK) fail("gb200_trk_set_code");
}

void B200CorrelatorEngine::set_code(int ch, const std::complex<float>* code, int code_len)
{
    std::lock_guard<std::mutex> lk(mu_);
    if (gb200_trk_set_code_complex(h_, ch, reinterpret_cast<const gb200_cf32*>(code), code_len) != GB200_OK) fail("gb200_trk_set_code_complex");
}

uint64_t B200CorrelatorEngine::push(const void* iq, int64_t n)
{
    std::lock_guard<std::mutex> lk(mu_);
    uint64_t first = 0;
    if (gb200_trk_push_iq(h_, iq, n, &first) != GB200_OK) fail("gb200_trk_push_iq");
    return first;
}

uint64_t B200CorrelatorEngine::push_raw(const void* raw, int64_t n_bytes, const gb200_ingest_format& fmt, int64_t* n_samples)
{
    std::lock_guard<std::mutex> lk(mu_);
    uint64_t first = 0;
    if (gb200_trk_push_raw(h_, raw, n_bytes, &fmt, &first, n_samples) != GB200_OK) fail("gb200_trk_push_raw");
    return first;
}

uint64_t B200CorrelatorEngine::head() const
{
    std::lock_guard<std::mutex> lk(mu_);
    return gb200_trk_head(h_);
}

bool B200CorrelatorEngine::ensure(uint64_t abs_index, const void* iq, int64_t n)
{
    std::lock_guard<std::mutex> lk(mu_);
    const uint64_t head = gb200_trk_head(h_);
    const uint64_t end = abs_index + static_cast<uint64_t>(n);
    const size_t eb = elem_bytes(cfg_.iq_dtype);
    if (end > head)
        {
            uint64_t from = head;
            if (abs_index > head)  // nobody has asked for [head, abs_index) yet: skip it, remember it (see gb200_trk_skip)
                {
                    if (gb200_trk_skip(h_, static_cast<int64_t>(abs_index - head)) != GB200_OK) fail("gb200_trk_skip");
                    gaps_.emplace_back(head, abs_index);
                    from = abs_index;
                }
            const char* src = static_cast<const char*>(iq) + (from - abs_index) * eb;
            if (gb200_trk_push_iq(h_, src, static_cast<int64_t>(end - from), nullptr) != GB200_OK) fail("gb200_trk_push_iq");
        }
    const uint64_t new_head = gb200_trk_head(h_);
    const uint64_t oldest = new_head > static_cast<uint64_t>(cfg_.ring_samples) ? new_head - static_cast<uint64_t>(cfg_.ring_samples) : 0;
    if (abs_index < oldest) return false;  // the ring has wrapped past the start of this stretch
    // a channel that runs behind the one that skipped may need part of a skipped stretch: back-fill the overlap
    for (size_t g = 0; g < gaps_.size();)
        {
            const uint64_t a = gaps_[g].first, b = gaps_[g].second;
            if (b <= oldest)
                {
                    gaps_.erase(gaps_.begin() + static_cast<std::ptrdiff_t>(g));
                    continue;
                }
            const uint64_t lo = std::max(a, abs_index), hi = std::min(b, end);
            if (lo < hi)
                {
                    const char* src = static_cast<const char*>(iq) + (lo - abs_index) * eb;
                    if (gb200_trk_write_iq(h_, lo, src, static_cast<int64_t>(hi - lo)) != GB200_OK) fail("gb200_trk_write_iq");
                    gaps_.erase(gaps_.begin() + static_cast<std::ptrdiff_t>(g));
                    if (a < lo) gaps_.emplace_back(a, lo);
                    if (hi < b) gaps_.emplace_back(hi, b);
                    g = 0;  // the list changed; it holds a handful of entries at most
                    continue;
                }
            g++;
        }
    return true;
}

void B200CorrelatorEngine::correlate_batch(const gb200_trk_cmd* cmds, int n, gb200_cf32* out)
{
    std::lock_guard<std::mutex> lk(mu_);
    if (gb200_trk_correlate(h_, cmds, n, out) != GB200_OK) fail("gb200_trk_correlate");
    launches_++;
    batched_ += static_cast<uint64_t>(n);
}

void B200CorrelatorEngine::correlate_one(const gb200_trk_cmd& cmd, gb200_cf32* out_taps)
{
    std::unique_lock<std::mutex> lk(bmu_);
    if (!open_) open_ = std::make_shared<Batch>();
    std::shared_ptr<Batch> b = open_;
    b->cmds.push_back(cmd);
    b->outs.push_back(out_taps);
    const bool leader = b->cmds.size() == 1;
    const auto full = [&] { return (expected_ > 0 && static_cast<int>(b->cmds.size()) >= expected_) || static_cast<int>(b->cmds.size()) >= cfg_.max_channels; };
    if (leader)
        {
            bcv_.wait_for(lk, std::chrono::microseconds(cfg_.batch_window_us), full);
            b->closed = true;
            open_.reset();  // later requests open the next batch
            lk.unlock();
            std::vector<gb200_cf32> res(b->cmds.size() * GB200_MAX_TAPS);
            try
                {
                    correlate_batch(b->cmds.data(), static_cast<int>(b->cmds.size()), res.data());
                    for (size_t i = 0; i < b->cmds.size(); i++) std::memcpy(b->outs[i], &res[i * GB200_MAX_TAPS], sizeof(gb200_cf32) * b->cmds[i].n_taps);
                }
            catch (const std::exception& e)
                {
                    b->error = e.what();
                }
            lk.lock();
            b->done = true;
            bcv_.notify_all();
        }
    else
        {
            if (full()) bcv_.notify_all();  // wake the leader early
            bcv_.wait(lk, [&] { return b->done; });
        }
    if (!b->error.empty()) throw std::runtime_error(b->error);
}

// ------------------------------------------------------------------ per-channel facade
B200_Multicorrelator_Real_Codes::~B200_Multicorrelator_Real_Codes() { free(); }

bool B200_Multicorrelator_Real_Codes::init(int max_signal_length_samples, int n_correlators)
{
    if (!engine_) return false;
    if (n_correlators < 1 || n_correlators > GB200_MAX_TAPS) return false;
    if (max_signal_length_samples > engine_->config().max_block) return false;  // the engine's longest window (prepare_cmd refuses longer ones)
    if (ch_ < 0) ch_ = engine_->attach_channel();
    n_correlators_ = n_correlators;
    max_len_ = max_signal_length_samples;
    return ch_ >= 0;
}

bool B200_Multicorrelator_Real_Codes::set_local_code_and_taps(int code_length_chips, const float* local_code_in, float* shifts_chips)
{
    if (ch_ < 0) return false;
    engine_->set_code(ch_, local_code_in, code_length_chips);  // copied to the device; the reference borrows the pointer
    shifts_ = shifts_chips;                                    // borrowed and re-read on every call, like the reference
    return true;
}

bool B200_Multicorrelator_Real_Codes::set_input_output_vectors(std::complex<float>* corr_out, const void* sig_in, uint64_t abs_sample_index, int n_available)
{
    corr_out_ = corr_out;
    sig_in_ = sig_in;
    abs_index_ = abs_sample_index;
    n_available_ = n_available;
    return true;
}

void B200_Multicorrelator_Real_Codes::update_local_code(int, float, float, float)
{
    // the resampled replicas are never materialised on the GPU (trk_kernel2.cuh): nothing to do
}

bool B200_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler(float rem_carrier_phase_in_rad, float phase_step_rad,
    float phase_rate_step_rad, float rem_code_phase_chips, float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples)
{
    if (ch_ < 0 || !corr_out_ || !shifts_) return false;
    if (signal_length_samples > n_available_) return false;
    if (!engine_->ensure(abs_index_, sig_in_, n_available_)) return false;
    gb200_trk_cmd c{};
    c.ch = ch_;
    c.n_samples = signal_length_samples;
    c.n_taps = n_correlators_;
    c.flags = high_dyn_ ? GB200_TRK_HIGH_DYN : 0;
    c.sample_offset = abs_index_;
    c.rem_carrier_phase_rad = rem_carrier_phase_in_rad;
    c.phase_step_rad = phase_step_rad;
    c.phase_rate_step_rad = phase_rate_step_rad;
    c.rem_code_phase_chips = rem_code_phase_chips;
    c.code_phase_step_chips = code_phase_step_chips;
    c.code_phase_rate_step_chips = code_phase_rate_step_chips;
    for (int t = 0; t < n_correlators_; t++) c.shifts_chips[t] = shifts_[t];
    engine_->correlate_one(c, reinterpret_cast<gb200_cf32*>(corr_out_));
    return true;
}

bool B200_Multicorrelator_Real_Codes::free()
{
    if (engine_ && ch_ >= 0) engine_->detach_channel(ch_);
    ch_ = -1;
    corr_out_ = nullptr;
    return true;
}
}  // namespace gb200h
