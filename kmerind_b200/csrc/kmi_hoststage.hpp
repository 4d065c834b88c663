// kmi_hoststage.hpp — PAGEABLE host memory <-> device through a ring of pinned slices that host threads fill / drain.
//
// The reference's callers hold their input and output in ordinary heap memory (std::vector<unsigned char> file data,
// std::vector<Kmer> queries, std::vector<std::pair<Kmer, T>> answers: kmer_index.hpp:132-225, file.hpp:160-221). A
// cudaMemcpy from such memory goes through the driver's own staging buffer on ONE host thread (measured on the bench
// box: 9.4 GB/s, 223 ms for the 2.1 GB of the 1 Gbp configuration, against 38 ms from pinned memory). Here the copy is
// cut into slices of a few MiB; T host threads memcpy slices into a ring of pinned slots, one issuer enqueues the device
// copies in order and records an event at every byte offset the consumer named ("marks": the rounds of the pipelined
// builds), so the build of round r still runs under the transfer of round r + 1.
//
// Plain C++ over a small backend interface so that the threading can be exercised without a GPU
// (tests/cpp/stage_check.cpp); the CUDA backend is in kmi_api.cu.
//
//   struct Backend {
//     void thread_init();                                        // once per worker thread (cudaSetDevice)
//     bool copy_in(void* dev, const void* pinned, size_t n);      // enqueue pinned -> device on the copy stream
//     bool copy_out(void* pinned, const void* dev, size_t n);     // enqueue device -> pinned on the copy stream
//     bool record_slot(int slot);                                 // event after the copy just enqueued for this slot
//     bool wait_slot(int slot);                                   // host waits for the slot's last recorded event
//     bool record_mark(size_t i);                                 // event i: every byte before marks[i] is enqueued
//   };
#pragma once
#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace kmi {

template <class Backend> class HostStageJob {
 public:
  HostStageJob(Backend& b, uint8_t* ring, size_t slice_bytes, int n_slots, int n_threads)
      : b_(b), ring_(ring), slice_(slice_bytes), R_((size_t)std::max(n_slots, 1)), T_(std::max(n_threads, 1)) {}
  HostStageJob(const HostStageJob&) = delete;
  HostStageJob& operator=(const HostStageJob&) = delete;
  ~HostStageJob() { finish(true); }

  // where the bytes of a host -> device job come from when it is not memory: fill(dst, offset, len) puts bytes [offset,
  // offset + len) of the source into dst (a pread of a file: kmi_build_file) and returns false on failure. Set before start_in.
  void set_fill(std::function<bool(uint8_t*, uint64_t, size_t)> f) { fill_ = std::move(f); }

  // host -> device, asynchronous: range i is [ends[i-1], ends[i]) (ends ascending, empty ranges allowed). Returns at once;
  // wait_mark(i) blocks the caller until range i has been enqueued and record_mark(i) has run.
  void start_in(uint8_t* dev, const uint8_t* host, const std::vector<uint64_t>& ends) {
    dev_ = dev; hsrc_ = host;
    uint64_t at = 0;
    for (size_t i = 0; i < ends.size(); ++i) {
      const uint64_t e = std::max(ends[i], at);
      do {   // at least one (possibly empty) slice per range: it carries the mark
        const uint64_t len = std::min<uint64_t>(slice_, e - at);
        slices_.push_back(Slice{at, len, -1});
        at += len;
      } while (at < e);
      slices_.back().mark = (long)i;
    }
    state_.assign(slices_.size(), 0);
    n_marks_ = ends.size();
    for (int t = 0; t < T_; ++t) threads_.emplace_back([this] { fill_loop(); });
    threads_.emplace_back([this] { issue_in_loop(); });
  }
  bool wait_mark(size_t i) {
    if (i >= n_marks_) return false;
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return marks_done_ > i || failed_ || cancel_; });
    return marks_done_ > i && !failed_;
  }

  // device -> host, blocking: the caller issues the device copies, the threads drain the slots into `host`.
  // The backend's copy stream must already be ordered behind whatever produced `dev`.
  bool run_out(uint8_t* host, const uint8_t* dev, uint64_t n) {
    hdst_ = host; dsrc_ = dev;
    for (uint64_t at = 0; at < n; at += slice_) slices_.push_back(Slice{at, std::min<uint64_t>(slice_, n - at), -1});
    state_.assign(slices_.size(), 0);
    for (int t = 0; t < T_; ++t) threads_.emplace_back([this] { drain_loop(); });
    for (size_t s = 0; s < slices_.size(); ++s) {
      if (s >= R_ && !wait_state(s - R_, 2)) break;          // the slot's previous slice has left the ring
      const int slot = (int)(s % R_);
      if (!b_.copy_out(ring_ + (size_t)slot * slice_, dsrc_ + slices_[s].off, (size_t)slices_[s].len) || !b_.record_slot(slot)) { fail(); break; }
      set_state(s, 1);
    }
    return finish(false);
  }

  // joins the workers; with cancel they stop at the next slice. true when every slice went through without an error.
  bool finish(bool cancel) {
    if (cancel) {
      std::lock_guard<std::mutex> lk(mu_);
      cancel_ = true;
    }
    cv_.notify_all();
    for (std::thread& t : threads_) if (t.joinable()) t.join();
    threads_.clear();
    return !failed_ && !cancel_;
  }

 private:
  struct Slice { uint64_t off, len; long mark; };
  // state_[s]: H2D 0 waiting, 1 filled (in the ring), 2 enqueued; D2H 0 waiting, 1 enqueued, 2 drained
  bool wait_state(size_t s, int want) {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return state_[s] >= want || cancel_ || failed_; });
    return state_[s] >= want && !failed_ && !cancel_;
  }
  void set_state(size_t s, int v) {
    { std::lock_guard<std::mutex> lk(mu_); state_[s] = (char)v; }
    cv_.notify_all();
  }
  void fail() {
    { std::lock_guard<std::mutex> lk(mu_); failed_ = true; }
    cv_.notify_all();
  }
  bool take(size_t* s) {
    std::lock_guard<std::mutex> lk(mu_);
    if (cancel_ || failed_ || next_ >= slices_.size()) return false;
    *s = next_++;
    return true;
  }
  void fill_loop() {
    b_.thread_init();
    size_t s;
    while (take(&s)) {
      const Slice& sl = slices_[s];
      if (sl.len) {
        const int slot = (int)(s % R_);
        // the slot is free once the device copy of the slice that used it last has finished
        if (s >= R_ && !wait_state(s - R_, 2)) return;
        if (!b_.wait_slot(slot)) { fail(); return; }
        uint8_t* dst = ring_ + (size_t)slot * slice_;
        if (!fill_) std::memcpy(dst, hsrc_ + sl.off, (size_t)sl.len);
        else if (!fill_(dst, sl.off, (size_t)sl.len)) { fail(); return; }
      }
      set_state(s, 1);
    }
  }
  void issue_in_loop() {
    b_.thread_init();
    for (size_t s = 0; s < slices_.size(); ++s) {
      if (!wait_state(s, 1)) return;
      const Slice& sl = slices_[s];
      const int slot = (int)(s % R_);
      if (sl.len && (!b_.copy_in(dev_ + sl.off, ring_ + (size_t)slot * slice_, (size_t)sl.len) || !b_.record_slot(slot))) { fail(); return; }
      if (sl.mark >= 0 && !b_.record_mark((size_t)sl.mark)) { fail(); return; }
      {
        std::lock_guard<std::mutex> lk(mu_);
        state_[s] = 2;
        if (sl.mark >= 0) marks_done_ = (size_t)sl.mark + 1;
      }
      cv_.notify_all();
    }
  }
  void drain_loop() {
    b_.thread_init();
    size_t s;
    while (take(&s)) {
      const Slice& sl = slices_[s];
      const int slot = (int)(s % R_);
      if (!wait_state(s, 1)) return;
      if (!b_.wait_slot(slot)) { fail(); return; }
      std::memcpy(hdst_ + sl.off, ring_ + (size_t)slot * slice_, (size_t)sl.len);
      set_state(s, 2);
    }
  }

  Backend& b_;
  uint8_t* ring_;
  const uint64_t slice_;
  const size_t R_;
  const int T_;
  uint8_t* dev_ = nullptr; const uint8_t* hsrc_ = nullptr;
  uint8_t* hdst_ = nullptr; const uint8_t* dsrc_ = nullptr;
  std::function<bool(uint8_t*, uint64_t, size_t)> fill_;
  std::vector<Slice> slices_;
  std::vector<char> state_;
  std::vector<std::thread> threads_;
  std::mutex mu_;
  std::condition_variable cv_;
  size_t next_ = 0, marks_done_ = 0, n_marks_ = 0;
  bool cancel_ = false, failed_ = false;
};

}  // namespace kmi
