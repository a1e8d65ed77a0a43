// Small persistent worker pool for the host-side command preprocessing (one pool per process, started on first use).
// parallel_for(n, f) runs f(begin, end) over [0, n) split into one contiguous slice per worker plus the caller.
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace gb200
{
class HostPool
{
public:
    static HostPool& get()
    {
        static HostPool p;
        return p;
    }
    int workers() const { return (int)threads_.size(); }

    // f(begin, end); returns after every slice has run.  Calls from several threads are serialised.
    void parallel_for(int n, const std::function<void(int, int)>& f, int min_per_slice = 512)
    {
        const int slices = std::max(1, std::min((int)threads_.size() + 1, n / std::max(1, min_per_slice)));
        if (slices == 1)
            {
                f(0, n);
                return;
            }
        std::lock_guard<std::mutex> call_lock(call_mu_);
        {
            std::lock_guard<std::mutex> lk(mu_);
            fn_ = &f;
            n_ = n;
            slices_ = slices;
            pending_.store(slices - 1, std::memory_order_relaxed);
            generation_++;
            gen_hint_.store(generation_, std::memory_order_release);
        }
        cv_.notify_all();
        run_slice(slices - 1);  // the caller takes the last slice
        // short spin, then sleep: slices take ~100 us
        for (int spin = 0; pending_.load(std::memory_order_acquire) != 0; spin++)
            if (spin > 2000) std::this_thread::yield();
        std::lock_guard<std::mutex> lk(mu_);
        fn_ = nullptr;
    }

private:
    HostPool()
    {
        int want = (int)std::thread::hardware_concurrency() / 2 - 1;
        want = std::min(want, 7);
        if (const char* e = std::getenv("GB200_HOST_THREADS")) want = std::atoi(e) - 1;  // explicit: up to 32 threads
        want = std::max(0, std::min(want, 31));
        if (const char* e = std::getenv("GB200_HOST_SPIN_US")) spin_us_ = std::max(0, std::atoi(e));
        for (int i = 0; i < want; i++) threads_.emplace_back([this, i] { loop(i); });
    }
    ~HostPool()
    {
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : threads_) t.join();
    }
    void run_slice(int s)
    {
        const int per = (n_ + slices_ - 1) / slices_;
        const int b = std::min(n_, s * per), e = std::min(n_, b + per);
        if (b < e) (*fn_)(b, e);
    }
    void loop(int id)
    {
        unsigned long seen = 0;
        for (;;)
            {
                // calls come in bursts (preprocess, then plan, every code period): poll for a short while before sleeping, a
                // condition-variable wake-up costs 50-100 us per worker
                const auto t0 = std::chrono::steady_clock::now();
                while (gen_hint_.load(std::memory_order_acquire) == seen && std::chrono::steady_clock::now() - t0 < std::chrono::microseconds(spin_us_))
                    {
                    }
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || generation_ != seen; });
                if (stop_) return;
                seen = generation_;
                const bool mine = id < slices_ - 1;
                lk.unlock();
                if (mine)
                    {
                        run_slice(id);
                        pending_.fetch_sub(1, std::memory_order_release);
                    }
            }
    }
    std::vector<std::thread> threads_;
    std::mutex mu_, call_mu_;
    std::condition_variable cv_;
    const std::function<void(int, int)>* fn_{nullptr};
    int n_{0}, slices_{1};
    std::atomic<int> pending_{0};
    unsigned long generation_{0};
    std::atomic<unsigned long> gen_hint_{0};
    int spin_us_{200};
    bool stop_{false};
};
}  // namespace gb200
