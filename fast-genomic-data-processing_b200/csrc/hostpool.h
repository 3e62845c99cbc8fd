// Process-wide pool of host worker threads for the packer's parallel loops (plan / bin / copy / scatter over regions).
// The reference spreads host work over its T Mutect2 worker threads (M2/main.cpp:199,570-574) and shares one region's
// testcases between idle threads (PairHMMConcurrentControl.h:15-53, IntelPairHmm.cc:296-330); here the host work of a
// call is spread over persistent pool threads instead of std::threads created and joined on every call (that cost
// ~40 us of a 0.27 ms per-region call), and several handles -- the devices of a multi-GPU handle, or callers on
// different threads -- share the same pool, so the box's cores go to whoever has packing to do.
#pragma once
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstddef>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace phmm {

class HostPool {
public:
    static HostPool &instance() {
        static HostPool *p = new HostPool();   // leaked on purpose: worker threads may outlive static destruction
        return *p;
    }
    int size() const { return (int)threads_.size(); }

    // fn(i) for i in [0, n), in chunks of `grain`, on the calling thread plus up to max_threads - 1 pool threads.
    template <class F> void parallel_for(size_t n, int max_threads, size_t grain, F &&fn) {
        if (grain == 0) grain = 1;
        if (max_threads <= 1 || n <= grain || threads_.empty()) { for (size_t i = 0; i < n; ++i) fn(i); return; }
        auto job = std::make_shared<Job>();
        job->n = n; job->grain = grain;
        job->helpers_left.store((int)std::min<size_t>((size_t)max_threads - 1, (n + grain - 1) / grain - 1));
        job->body = [&fn](size_t b, size_t e) { for (size_t i = b; i < e; ++i) fn(i); };
        {
            std::lock_guard<std::mutex> lk(m_);
            jobs_.push_back(job);
        }
        cv_.notify_all();
        work_on(*job);
        {   // wait for the helpers' last chunks; the job is taken off the list first so that nobody else joins it
            std::unique_lock<std::mutex> lk(m_);
            for (auto it = jobs_.begin(); it != jobs_.end(); ++it) if (it->get() == job.get()) { jobs_.erase(it); break; }
            done_cv_.wait(lk, [&] { return job->done.load(std::memory_order_acquire) >= n; });
        }
    }

private:
    struct Job {
        size_t n = 0, grain = 1;
        std::atomic<size_t> next{0}, done{0};
        std::atomic<int> helpers_left{0};
        std::function<void(size_t, size_t)> body;
    };
    HostPool() {
        unsigned hc = std::thread::hardware_concurrency();
        const int nt = (int)std::max(1u, std::min(hc ? hc : 1u, 64u)) - 1;     // the caller is one of the workers
        for (int t = 0; t < nt; ++t) threads_.emplace_back([this] { worker(); }), threads_.back().detach();
    }
    void work_on(Job &j) {
        for (;;) {
            const size_t b = j.next.fetch_add(j.grain, std::memory_order_relaxed);
            if (b >= j.n) break;
            const size_t e = std::min(j.n, b + j.grain);
            j.body(b, e);
            if (j.done.fetch_add(e - b, std::memory_order_acq_rel) + (e - b) >= j.n) {
                std::lock_guard<std::mutex> lk(m_);
                done_cv_.notify_all();
            }
        }
    }
    void worker() {
        std::unique_lock<std::mutex> lk(m_);
        for (;;) {
            std::shared_ptr<Job> job;
            for (auto &j : jobs_)
                if (j->next.load(std::memory_order_relaxed) < j->n && j->helpers_left.load(std::memory_order_relaxed) > 0 &&
                    j->helpers_left.fetch_sub(1, std::memory_order_acq_rel) > 0) { job = j; break; }
            if (!job) { cv_.wait(lk); continue; }
            lk.unlock();
            work_on(*job);
            lk.lock();
        }
    }
    std::mutex m_;
    std::condition_variable cv_, done_cv_;
    std::deque<std::shared_ptr<Job>> jobs_;
    std::vector<std::thread> threads_;
};

}  // namespace phmm
