// walk_pool.h -- a few persistent worker threads for walking long tensor lists on the host.
// Header-only, no torch / Python dependency (stress-tested stand-alone by tests/test_walk_pool.py).
//
// Creating and joining 8 std::threads per call costs about as much as the walk itself at 50k tensors.  Workers only
// read tensor metadata; they never call into Python.  The process-wide pool is leaked on purpose (no static destructor
// racing interpreter shutdown) and rebuilt in a forked child (threads do not survive fork()).
#pragma once
#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace mt_host {

class WalkPool {
  public:
    static WalkPool& get() {
        static WalkPool* pool = nullptr;
        static pid_t owner = 0;
        static std::mutex mu;
        std::lock_guard<std::mutex> lk(mu);
        if (!pool || owner != getpid()) {
            pool = new WalkPool(std::min(7u, std::max(1u, std::thread::hardware_concurrency() / 2) - 1));
            owner = getpid();
        }
        return *pool;
    }
    explicit WalkPool(unsigned nworkers) {
        for (unsigned i = 0; i < nworkers; i++) {
            threads_.emplace_back([this] { loop(); });
            threads_.back().detach();
        }
    }
    // Never destroyed: the detached workers wait on this object's condition variable for the life of the process.
    ~WalkPool() = delete;
    WalkPool(const WalkPool&) = delete;
    WalkPool& operator=(const WalkPool&) = delete;

    unsigned workers() const { return (unsigned)threads_.size(); }

    // Run fn(part) for every part in [0, parts): part 0 on the calling thread, the others on whichever workers pick
    // them up (a worker may take several).  Returns when all parts are done.  One run at a time.
    void run(unsigned parts, const std::function<void(unsigned)>& fn) {
        if (parts == 0) return;
        std::lock_guard<std::mutex> serial(run_mu_);
        {
            std::lock_guard<std::mutex> lk(mu_);
            fn_ = &fn;
            parts_ = parts;
            next_ = 1;
            pending_ = parts - 1;
        }
        if (parts > 1) cv_.notify_all();
        fn(0);
        std::unique_lock<std::mutex> lk(mu_);
        if (threads_.empty()) {  // no workers: the caller does everything
            while (next_ < parts_) {
                const unsigned part = next_++;
                lk.unlock();
                fn(part);
                lk.lock();
                --pending_;
            }
        }
        done_.wait(lk, [&] { return pending_ == 0; });
        fn_ = nullptr;
        parts_ = next_ = 0;
    }

  private:
    void loop() {
        for (;;) {
            unsigned part;
            const std::function<void(unsigned)>* fn;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return next_ < parts_; });
                part = next_++;
                fn = fn_;
            }
            (*fn)(part);
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    std::vector<std::thread> threads_;
    std::mutex mu_, run_mu_;
    std::condition_variable cv_, done_;
    const std::function<void(unsigned)>* fn_ = nullptr;
    unsigned parts_ = 0, next_ = 0, pending_ = 0;
};

}  // namespace mt_host
