// A small persistent pool of host threads for the per-segment table builders of the alignment pipeline
// (gmm_align.cu): at a few hundred million frames per second the GPU side of gmm_score_align_host takes less time
// than one host thread needs to turn the caller's automata into kernel tables.
#pragma once
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>
#include <stdint.h>

class HostPool {
public:
    typedef std::function<void(uint32_t begin, uint32_t end, int worker)> Fn;
    explicit HostPool(int workers) {
        for (int w = 0; w < workers; ++w)
            threads_.emplace_back([this, w] { loop(w + 1); });
    }
    ~HostPool() {
        {
            std::unique_lock<std::mutex> l(m_);
            stop_ = true;
        }
        wake_.notify_all();
        for (auto& t : threads_)
            t.join();
    }
    int size() const { return (int)threads_.size() + 1; }  // the calling thread works too (worker 0)
    // fn(begin, end, worker) over [0, n) in chunks handed out dynamically; returns when all of it is done
    void run(uint32_t n, uint32_t chunk, const Fn& fn) {
        if (n == 0)
            return;
        chunk = chunk ? chunk : 1;
        if (threads_.empty() || n <= chunk) {
            fn(0, n, 0);
            return;
        }
        {
            std::unique_lock<std::mutex> l(m_);
            fn_ = &fn, n_ = n, chunk_ = chunk;
            next_.store(0);
            active_ = (int)threads_.size();
            ++epoch_;
        }
        wake_.notify_all();
        work(0);
        std::unique_lock<std::mutex> l(m_);
        done_.wait(l, [this] { return active_ == 0; });
        fn_ = nullptr;
    }

private:
    void work(int worker) {
        for (;;) {
            const uint32_t b = next_.fetch_add(chunk_);
            if (b >= n_)
                return;
            (*fn_)(b, b + chunk_ < n_ ? b + chunk_ : n_, worker);
        }
    }
    void loop(int worker) {
        uint64_t seen = 0;
        for (;;) {
            {
                std::unique_lock<std::mutex> l(m_);
                wake_.wait(l, [&] { return stop_ || epoch_ != seen; });
                if (stop_)
                    return;
                seen = epoch_;
            }
            work(worker);
            std::unique_lock<std::mutex> l(m_);
            if (--active_ == 0)
                done_.notify_one();
        }
    }
    std::vector<std::thread> threads_;
    std::mutex               m_;
    std::condition_variable  wake_, done_;
    const Fn*                fn_ = nullptr;
    uint32_t                 n_ = 0, chunk_ = 1;
    std::atomic<uint32_t>    next_{0};
    int                      active_ = 0;
    uint64_t                 epoch_ = 0;
    bool                     stop_ = false;
};
