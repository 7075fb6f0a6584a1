// Host-side helper threads for the staging copy of lo_odometry_process (pageable scan -> pinned ring): one core reads
// cold DRAM at ~12 GB/s, which makes the 0.9 MB copy of a VLP16 scan (~77 us) the largest host item of the blocking call;
// three helpers bring it to a third. A helper spins (pause) for 1 ms after its last job — a caller in a tight loop
// finds it awake — and sleeps on a condition variable otherwise, so an idle or 10 Hz caller pays nothing.
// Not thread-safe: one pool per lo_odometry handle, used by the thread that calls into the handle.
#pragma once

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace lo {

class CopyPool {
public:
    static constexpr int kSpinMicros = 1000;   // longer than one blocking call (~0.3 ms), 1 % of a core for a 10 Hz caller

    explicit CopyPool(int helpers) {
        for (int k = 0; k < helpers; k++) threads_.emplace_back([this, k] { run(k); });
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_.store(true, std::memory_order_release);
            gen_.fetch_add(1, std::memory_order_release);
        }
        cv_.notify_all();
        for (std::thread& t : threads_) t.join();
    }
    CopyPool(const CopyPool&) = delete;
    CopyPool& operator=(const CopyPool&) = delete;

    int helpers() const { return static_cast<int>(threads_.size()); }

    // memcpy(dst, src, bytes) with the helpers; returns when every byte is in place
    void copy(void* dst, const void* src, size_t bytes) {
        const int parts = helpers() + 1;
        if (parts == 1 || bytes < 64 * 1024) {
            std::memcpy(dst, src, bytes);
            return;
        }
        dst_ = static_cast<char*>(dst);
        src_ = static_cast<const char*>(src);
        bytes_ = bytes;
        part_ = ((bytes + parts - 1) / parts + 4095) & ~static_cast<size_t>(4095);
        done_.store(0, std::memory_order_relaxed);
        // sequentially consistent on both sides (here: publish the job, then look for sleepers; helper: announce the sleep, then
        // look for a job): one of the two sees the other
        gen_.fetch_add(1);
        if (sleepers_.load() > 0) {
            std::lock_guard<std::mutex> lk(m_);   // a helper between its last check and its wait holds the mutex: no lost wake-up
            cv_.notify_all();
        }
        copy_part(parts - 1);
        while (done_.load(std::memory_order_acquire) != helpers()) cpu_relax();
    }

private:
    static void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    }
    void copy_part(int k) const {
        const size_t begin = static_cast<size_t>(k) * part_;
        if (begin >= bytes_) return;
        const size_t len = bytes_ - begin < part_ ? bytes_ - begin : part_;
        std::memcpy(dst_ + begin, src_ + begin, len);
    }
    void run(int k) {
        uint64_t seen = 0;
        while (true) {
            // wait for the next job: spin for a while, then sleep
            const auto spin_until = std::chrono::steady_clock::now() + std::chrono::microseconds(kSpinMicros);
            uint64_t g;
            int polls = 0;
            while ((g = gen_.load(std::memory_order_acquire)) == seen) {
                cpu_relax();
                if ((++polls & 255) == 0 && std::chrono::steady_clock::now() > spin_until) {
                    std::unique_lock<std::mutex> lk(m_);
                    sleepers_.fetch_add(1);
                    cv_.wait(lk, [&] { return gen_.load() != seen; });
                    sleepers_.fetch_sub(1);
                }
            }
            seen = g;
            if (stop_.load(std::memory_order_acquire)) return;
            copy_part(k);
            done_.fetch_add(1, std::memory_order_release);
        }
    }

    std::vector<std::thread> threads_;
    std::mutex m_;
    std::condition_variable cv_;
    std::atomic<uint64_t> gen_{0};
    std::atomic<int> done_{0}, sleepers_{0};
    std::atomic<bool> stop_{false};
    char* dst_ = nullptr;
    const char* src_ = nullptr;
    size_t bytes_ = 0, part_ = 0;
};

}  // namespace lo
