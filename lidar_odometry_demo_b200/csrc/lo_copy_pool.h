// Host-side helper threads for the staging copy of lo_odometry_process (pageable scan -> pinned ring): one core reads
// cold DRAM at ~12 GB/s, which makes the 0.9 MB copy of a VLP16 scan (~77 us) the largest host item of the blocking call.
// A job is cut into chunks that the helpers — and the caller, whenever it has nothing else to do — take from a shared
// counter; the caller watches the completed prefix and hands it to the DMA engine piece by piece while the rest is still
// being copied. A helper spins (pause) for 1 ms after its last chunk — a caller in a tight loop finds it awake — and
// sleeps on a condition variable otherwise, so an idle or 10 Hz caller pays nothing.
// Not thread-safe: one pool per lo_odometry handle, used by the thread that calls into the handle.
#pragma once

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace lo {

class CopyPool {
public:
    static constexpr int kSpinMicros = 1000;        // longer than one blocking call (~0.3 ms), 1 % of a core for a 10 Hz caller
    static constexpr size_t kChunk = 64 * 1024;     // what one thread copies at a time
    static constexpr size_t kMaxChunks = 1024;      // larger jobs use larger chunks

    explicit CopyPool(int helpers) : done_(new std::atomic<uint64_t>[kMaxChunks]) {
        for (size_t k = 0; k < kMaxChunks; k++) done_[k].store(0, std::memory_order_relaxed);
        for (int k = 0; k < helpers; k++) threads_.emplace_back([this] { run(); });
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_.store(true);
            gen_.fetch_add(1);
        }
        cv_.notify_all();
        for (std::thread& t : threads_) t.join();
    }
    CopyPool(const CopyPool&) = delete;
    CopyPool& operator=(const CopyPool&) = delete;

    int helpers() const { return static_cast<int>(threads_.size()); }

    // Start copying src[0, bytes) to dst. The previous job must have been finished (finish(), or ready_bytes() == its size).
    void begin(void* dst, const void* src, size_t bytes) {
        dst_ = static_cast<char*>(dst);
        src_ = static_cast<const char*>(src);
        bytes_ = bytes;
        chunk_ = kChunk;
        while ((bytes + chunk_ - 1) / chunk_ > kMaxChunks) chunk_ *= 2;
        chunks_ = (bytes + chunk_ - 1) / chunk_;
        cursor_ = 0;
        job_ = gen_.load() + 1;
        work_.store(pack(job_, chunks_, 0), std::memory_order_release);
        // sequentially consistent on both sides (here: publish the job, then look for sleepers; helper: announce the sleep, then
        // look for a job): one of the two sees the other
        gen_.store(job_);
        if (chunks_ > 1 && sleepers_.load() > 0) {
            std::lock_guard<std::mutex> lk(m_);   // a helper between its last check and its wait holds the mutex: no lost wake-up
            cv_.notify_all();
        }
    }
    // The caller copies one chunk itself; false when every chunk has been taken.
    bool help_one() { return take_chunk(job_); }
    // Bytes of the completed prefix: dst[0, ready_bytes()) is in place.
    size_t ready_bytes() {
        while (cursor_ < chunks_ && done_[cursor_].load(std::memory_order_acquire) == job_) cursor_++;
        return cursor_ == chunks_ ? bytes_ : cursor_ * chunk_;
    }
    // Copy what is left and wait for the chunks the helpers hold.
    void finish() {
        while (help_one()) {
        }
        while (ready_bytes() != bytes_) cpu_relax();
    }
    // memcpy(dst, src, bytes) with the helpers; returns when every byte is in place
    void copy(void* dst, const void* src, size_t bytes) {
        begin(dst, src, bytes);
        finish();
    }

    static void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    }

private:
    // One word holds what a thread needs to claim a chunk: the job (low 24 bits of its number), its chunk count, the next
    // chunk. A claim is a compare-exchange on that word, so a helper that is late for job J can neither take a ticket of job
    // J + 1 nor act on J's fields after J is complete: a successful claim means the job still had that chunk open, and the
    // caller starts the next job (rewrites the fields) only after every chunk of this one is marked done.
    static uint64_t pack(uint64_t job, size_t chunks, size_t next) {
        return ((job & 0xFFFFFFull) << 40) | (static_cast<uint64_t>(chunks) << 20) | static_cast<uint64_t>(next);
    }
    bool take_chunk(uint64_t job) {
        uint64_t v = work_.load(std::memory_order_acquire);
        while (true) {
            const size_t c = static_cast<size_t>(v & 0xFFFFFull), n = static_cast<size_t>((v >> 20) & 0xFFFFFull);
            if ((v >> 40) != (job & 0xFFFFFFull) || c >= n) return false;
            if (work_.compare_exchange_weak(v, v + 1, std::memory_order_acq_rel, std::memory_order_acquire)) {
                const size_t begin = c * chunk_;
                const size_t len = bytes_ - begin < chunk_ ? bytes_ - begin : chunk_;
                std::memcpy(dst_ + begin, src_ + begin, len);
                done_[c].store(job, std::memory_order_release);
                return true;
            }
        }
    }
    void run() {
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
            if (stop_.load()) return;
            while (take_chunk(g)) {
            }
        }
    }

    std::vector<std::thread> threads_;
    std::mutex m_;
    std::condition_variable cv_;
    std::atomic<uint64_t> gen_{0};    // number of the current job; a change wakes the helpers
    std::atomic<uint64_t> work_{0};   // pack(job, chunks, next chunk)
    std::atomic<int> sleepers_{0};
    std::atomic<bool> stop_{false};
    std::unique_ptr<std::atomic<uint64_t>[]> done_;   // [chunk] = job number once the chunk is in place
    uint64_t job_ = 0;
    char* dst_ = nullptr;
    const char* src_ = nullptr;
    size_t bytes_ = 0, chunk_ = kChunk, chunks_ = 0, cursor_ = 0;
};

}  // namespace lo
