// hostcopy.hpp -- moving PAGEABLE host buffers (what a Rust `Vec<Fr>` is) to and from the device at PCIe speed.
//
// `cudaMemcpyAsync` on pageable memory is staged by the driver through one thread (~8 GB/s measured on the B200 box, and it
// blocks the calling thread, which serialises the devices of a multi-GPU context).  The library instead copies pageable
// buffers through its own ring of pinned slots: a small pool of host threads memcpy's slot-sized pieces in parallel while the
// previous piece is in flight on the DMA engine.  Pinned / registered caller buffers bypass all of this.
#pragma once
#include <cuda_runtime.h>

#include <condition_variable>
#include <cstdint>
#include <cstring>
#include "streamcopy.hpp"
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

namespace h2b {

class CopyPool {
  public:
    static CopyPool& instance() {
        static CopyPool p;
        return p;
    }
    // dst <- src, split across the pool; returns when every byte has been copied.  Safe to call from several threads.
    void copy(void* dst, const void* src, size_t bytes) {
        if (bytes < (size_t)(2u << 20) || workers_.empty()) { std::memcpy(dst, src, bytes); return; }
        Call call;
        const size_t parts = workers_.size();
        const size_t step = ((bytes / parts) + 4095) & ~(size_t)4095;
        {
            std::lock_guard<std::mutex> lk(mu_);
            for (size_t off = 0; off < bytes; off += step) {
                size_t len = bytes - off < step ? bytes - off : step;
                queue_.push_back(Job{static_cast<char*>(dst) + off, static_cast<const char*>(src) + off, len, &call});
                call.remaining++;
            }
        }
        cv_.notify_all();
        std::unique_lock<std::mutex> lk(call.mu);
        call.cv.wait(lk, [&] { return call.remaining == 0; });
    }

    // Several processes on one host (one rank per GPU) each own a pool: `share` = how many of them there are, so that together
    // they start about as many copy threads as one process would.  Call between copies only (h2b_peer_attach does).
    void set_share(unsigned share) {
        if (share == 0) share = 1;
        if (std::getenv("H2B_COPY_THREADS")) return;                  // an explicit setting wins
        // one process: the default; `share` processes: the hardware threads divided between them (at most the default each)
        unsigned n = std::thread::hardware_concurrency() / share;
        if (n > default_threads()) n = default_threads();
        if (n < 2) n = default_threads() ? 2 : 0;
        if (n == workers_.size()) return;
        stop_workers();
        start_workers(n);
    }
    size_t threads() const { return workers_.size(); }

  private:
    struct Call { std::mutex mu; std::condition_variable cv; size_t remaining = 0; };
    struct Job { char* dst; const char* src; size_t len; Call* call; };
    static unsigned default_threads() {
        unsigned hw = std::thread::hardware_concurrency();
        return hw >= 64 ? 16 : (hw >= 32 ? 12 : (hw >= 16 ? 8 : (hw >= 8 ? 4 : (hw >= 4 ? 2 : 0))));
    }
    void start_workers(unsigned n) {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = false; }
        for (unsigned i = 0; i < n; i++) workers_.emplace_back([this] { run(); });
    }
    void stop_workers() {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
        workers_.clear();
    }
    CopyPool() {
        unsigned n = default_threads();
        if (const char* env = std::getenv("H2B_COPY_THREADS")) { int v = std::atoi(env); if (v >= 0 && v <= 32) n = (unsigned)v; }
        start_workers(n);
    }
    ~CopyPool() { stop_workers(); }
    void run() {
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !queue_.empty(); });
                if (queue_.empty()) return;
                j = queue_.front();
                queue_.pop_front();
            }
            stream_copy(j.dst, j.src, j.len);
            bool last;
            { std::lock_guard<std::mutex> lk(j.call->mu); last = --j.call->remaining == 0; if (last) j.call->cv.notify_one(); }
        }
    }
    std::vector<std::thread> workers_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Job> queue_;
    bool stop_ = false;
};

// true when `p` is ordinary pageable host memory (not pinned, registered, managed or device memory)
inline bool host_is_pageable(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

constexpr size_t STAGE_SLOT_BYTES = (size_t)16 << 20;
constexpr int STAGE_SLOTS = 4;

// Per-device ring of pinned slots.
struct StageRing {
    uint8_t* base = nullptr;
    cudaEvent_t ev[STAGE_SLOTS] = {};
    bool busy[STAGE_SLOTS] = {};
    int next = 0;
    cudaError_t ensure() {
        if (base) return cudaSuccess;
        cudaError_t e = cudaMallocHost(reinterpret_cast<void**>(&base), STAGE_SLOT_BYTES * STAGE_SLOTS);
        if (e != cudaSuccess) { base = nullptr; return e; }
        for (int i = 0; i < STAGE_SLOTS; i++) {
            e = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    }
    void release() {
        if (!base) return;
        for (int i = 0; i < STAGE_SLOTS; i++) if (ev[i]) cudaEventDestroy(ev[i]);
        cudaFreeHost(base);
        base = nullptr;
    }
    uint8_t* slot(int i) const { return base + (size_t)i * STAGE_SLOT_BYTES; }

    // host (any kind) -> device on stream `st`.  Pageable sources go through the ring; the call returns once the last piece has
    // been handed to the DMA engine (the source buffer may then be reused by the caller).
    cudaError_t h2d(void* dev_dst, const void* src, size_t bytes, cudaStream_t st) {
        if (bytes == 0) return cudaSuccess;
        if (bytes < ((size_t)1 << 20) || !host_is_pageable(src)) return cudaMemcpyAsync(dev_dst, src, bytes, cudaMemcpyHostToDevice, st);
        cudaError_t e = ensure();
        if (e != cudaSuccess) return e;
        for (size_t off = 0; off < bytes; off += STAGE_SLOT_BYTES) {
            size_t len = bytes - off < STAGE_SLOT_BYTES ? bytes - off : STAGE_SLOT_BYTES;
            int s = next;
            next = (next + 1) % STAGE_SLOTS;
            if (busy[s]) { e = cudaEventSynchronize(ev[s]); if (e != cudaSuccess) return e; }
            CopyPool::instance().copy(slot(s), static_cast<const uint8_t*>(src) + off, len);
            e = cudaMemcpyAsync(static_cast<uint8_t*>(dev_dst) + off, slot(s), len, cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess) return e;
            e = cudaEventRecord(ev[s], st);
            if (e != cudaSuccess) return e;
            busy[s] = true;
        }
        return cudaSuccess;
    }
    // device -> host (any kind), ordered after the work already enqueued on `st`.  For pageable destinations the call returns
    // when the bytes are in `dst`; for pinned destinations it only enqueues the copy.
    cudaError_t d2h(void* dst, const void* dev_src, size_t bytes, cudaStream_t st) {
        if (bytes == 0) return cudaSuccess;
        if (bytes < ((size_t)1 << 20) || !host_is_pageable(dst)) return cudaMemcpyAsync(dst, dev_src, bytes, cudaMemcpyDeviceToHost, st);
        cudaError_t e = ensure();
        if (e != cudaSuccess) return e;
        for (int i = 0; i < STAGE_SLOTS; i++)                        // drain earlier h2d traffic still using the slots
            if (busy[i]) { e = cudaEventSynchronize(ev[i]); if (e != cudaSuccess) return e; busy[i] = false; }
        const size_t pieces = (bytes + STAGE_SLOT_BYTES - 1) / STAGE_SLOT_BYTES;
        auto len_of = [&](size_t i) { size_t off = i * STAGE_SLOT_BYTES; return bytes - off < STAGE_SLOT_BYTES ? bytes - off : STAGE_SLOT_BYTES; };
        auto finish = [&](size_t i) -> cudaError_t {
            int s = (int)(i % STAGE_SLOTS);
            cudaError_t e2 = cudaEventSynchronize(ev[s]);
            if (e2 != cudaSuccess) return e2;
            CopyPool::instance().copy(static_cast<uint8_t*>(dst) + i * STAGE_SLOT_BYTES, slot(s), len_of(i));
            return cudaSuccess;
        };
        for (size_t i = 0; i < pieces; i++) {
            if (i >= (size_t)STAGE_SLOTS) { e = finish(i - STAGE_SLOTS); if (e != cudaSuccess) return e; }
            int s = (int)(i % STAGE_SLOTS);
            e = cudaMemcpyAsync(slot(s), static_cast<const uint8_t*>(dev_src) + i * STAGE_SLOT_BYTES, len_of(i), cudaMemcpyDeviceToHost, st);
            if (e != cudaSuccess) return e;
            e = cudaEventRecord(ev[s], st);
            if (e != cudaSuccess) return e;
        }
        for (size_t i = pieces > (size_t)STAGE_SLOTS ? pieces - STAGE_SLOTS : 0; i < pieces; i++) { e = finish(i); if (e != cudaSuccess) return e; }
        next = 0;
        return cudaSuccess;
    }
};

}  // namespace h2b
