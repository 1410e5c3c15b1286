// fxcu_hostpar.h -- host-thread helper: split [0, n) over the host cores.  Used by the cache builders that need the HOST's libm
// (coordinate search, fxcu_kdtree.cu and fxcu_nn.cu) and by the host entry points' pageable <-> pinned copies (fxcu_apply.cu).
// The workers are started once and kept: creating and joining threads for every level chunk of a slab cost ~0.4 ms per chunk on
// the GPU box's host (profiles/exp/hostcopy_probe.cu), a quarter of the chunk's own copy time.
#pragma once

#include <algorithm>
#include <condition_variable>
#include <cstddef>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <unistd.h>
#include <vector>

namespace fxcu {

inline unsigned host_threads()
{
    unsigned hw = std::thread::hardware_concurrency();
    return hw == 0 ? 4u : std::min(hw, 64u);
}

namespace hostpar_detail {

inline bool& in_worker()
{
    thread_local bool v = false;
    return v;
}

// A queue of closures served by detached threads.  The object is never destroyed (its threads sleep in it until the process ends).
class Workers {
public:
    static Workers& get()
    {
        static Workers* w = new Workers();
        return *w;
    }
    // returns false if no worker could be started at all (the caller then runs the task itself)
    bool submit(std::function<void()> task)
    {
        std::unique_lock<std::mutex> lock(m_);
        if (pid_ != getpid()) { // a forked child inherits the bookkeeping but none of the threads: start over
            pid_ = getpid();
            started_ = 0;
            idle_ = 0;
            queue_.clear();
        }
        if (idle_ <= queue_.size() && started_ < host_threads()) { // every idle worker already has a task waiting for it
            try {
                std::thread([this]() { serve(); }).detach();
                ++started_;
            } catch (...) {
            }
        }
        if (started_ == 0)
            return false;
        queue_.push_back(std::move(task));
        lock.unlock();
        cv_.notify_one();
        return true;
    }

private:
    void serve()
    {
        in_worker() = true;
        for (;;) {
            std::function<void()> task;
            {
                std::unique_lock<std::mutex> lock(m_);
                ++idle_;
                cv_.wait(lock, [this]() { return !queue_.empty(); });
                --idle_;
                task = std::move(queue_.front());
                queue_.pop_front();
            }
            task();
        }
    }
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<std::function<void()>> queue_;
    unsigned started_ = 0;
    size_t idle_ = 0;
    pid_t pid_ = getpid();
};

struct Latch {
    std::mutex m;
    std::condition_variable cv;
    size_t left = 0;
    void done()
    {
        std::lock_guard<std::mutex> lock(m);
        if (--left == 0)
            cv.notify_all(); // under the lock: the waiter cannot destroy the latch before this returns
    }
    void wait()
    {
        std::unique_lock<std::mutex> lock(m);
        cv.wait(lock, [this]() { return left == 0; });
    }
};

} // namespace hostpar_detail

// fn(a, b) over disjoint ranges covering [0, n): at most max_parts of them (and at most one per host core), none smaller than
// `grain` except the last.  The calling thread takes the first range and returns when all are done.  fn must not throw.
template <class F> void parallel_ranges(size_t n, size_t grain, F&& fn, size_t max_parts = 64)
{
    using namespace hostpar_detail;
    const size_t parts = std::min<size_t>(std::min<size_t>(host_threads(), std::max<size_t>(max_parts, 1)), (n + std::max<size_t>(grain, 1) - 1) / std::max<size_t>(grain, 1));
    if (parts <= 1 || in_worker()) { // a task that fans out again runs its share alone: workers never wait for workers
        fn(size_t(0), n);
        return;
    }
    const size_t step = (n + parts - 1) / parts;
    Latch latch;
    latch.left = parts - 1;
    for (size_t k = 1; k < parts; ++k) {
        const size_t a = std::min(n, k * step), b = std::min(n, (k + 1) * step);
        auto task = [&fn, &latch, a, b]() {
            fn(a, b);
            latch.done();
        };
        if (!Workers::get().submit(task))
            task();
    }
    fn(size_t(0), std::min(n, step));
    latch.wait();
}

} // namespace fxcu
