#include "b200_tracking_bank.h"
#include <stdexcept>

namespace gb200h
{
int B200TrackingBank::add_channel(std::unique_ptr<b200_dll_pll_veml_tracking> block)
{
    Chan c;
    c.slot = engine_->attach_channel();
    if (c.slot < 0) throw std::runtime_error("B200TrackingBank: the engine has no free channel slot");
    if (block->track_pilot())
        {
            c.data_slot = engine_->attach_channel();
            if (c.data_slot < 0) throw std::runtime_error("B200TrackingBank: the engine has no free channel slot");
        }
    c.block = std::move(block);
    ch_.push_back(std::move(c));
    return static_cast<int>(ch_.size()) - 1;
}

void B200TrackingBank::start_tracking(int i, uint64_t nitems_read)
{
    Chan& c = ch_[i];
    c.block->start_tracking();
    engine_->set_code(c.slot, c.block->tracking_code().data(), c.block->code_table_entries());
    if (c.data_slot >= 0) engine_->set_code(c.data_slot, c.block->data_code().data(), c.block->code_table_entries());
    c.read = nitems_read;
}

int B200TrackingBank::step(std::vector<Output>* out)
{
    int launches = 0;
    for (;;)
        {
            const uint64_t head = engine_->head();
            cmds_.clear();
            who_.clear();
            data_idx_.clear();
            steps_.clear();
            for (int i = 0; i < size(); i++)
                {
                    Chan& c = ch_[i];
                    if (c.block->state() == 0)  // standby consumes everything it is given
                        {
                            c.read = head;
                            continue;
                        }
                    if (c.read + static_cast<uint64_t>(c.block->required_samples()) > head) continue;  // forecast not met yet
                    auto s = c.block->plan(c.read, c.slot, c.data_slot);
                    if (!s.correlate)
                        {
                            c.read += static_cast<uint64_t>(s.consume < 0 ? 0 : s.consume);  // pull-in alignment
                            i--;                                                             // the same channel may go on in this round
                            continue;
                        }
                    who_.push_back(i);
                    steps_.push_back(s);
                    cmds_.push_back(s.cmd);
                    data_idx_.push_back(-1);
                    if (s.has_data)
                        {
                            data_idx_.back() = static_cast<int>(cmds_.size());
                            cmds_.push_back(s.data_cmd);
                        }
                }
            if (cmds_.empty()) return launches;
            // one launch for the whole round: main and data commands go in the same call (a launch is specialised on the padded tap
            // count of the widest command, narrower ones are padded); split only by the engine's command capacity
            res_.resize(cmds_.size() * GB200_MAX_TAPS);
            const int cap = engine_->config().max_channels;
            for (size_t b = 0; b < cmds_.size(); b += cap)
                {
                    const int n = static_cast<int>(std::min(cmds_.size() - b, static_cast<size_t>(cap)));
                    engine_->correlate_batch(cmds_.data() + b, n, res_.data() + b * GB200_MAX_TAPS);
                    launches++;
                }
            size_t k = 0;
            for (size_t j = 0; j < who_.size(); j++)
                {
                    Chan& c = ch_[who_[j]];
                    const gb200_cf32* taps = &res_[k * GB200_MAX_TAPS];
                    const gb200_cf32* data = data_idx_[j] >= 0 ? &res_[static_cast<size_t>(data_idx_[j]) * GB200_MAX_TAPS] : nullptr;
                    k += data ? 2 : 1;
                    Gnss_Synchro s{};
                    const int produced = c.block->finish(steps_[j], taps, data, c.read, &s);
                    c.read += static_cast<uint64_t>(c.block->current_prn_length_samples());
                    if (produced && out) out->push_back({who_[j], s});
                }
        }
}
}  // namespace gb200h
