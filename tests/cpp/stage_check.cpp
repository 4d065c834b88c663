// stage_check.cpp — the threading of kmerind_b200/csrc/kmi_hoststage.hpp without a GPU: a fake "device" whose copy stream
// is a worker thread that executes the enqueued copies LATER and in order (with small random delays), so a ring slot that is
// refilled before its device copy has run, a mark recorded early, or a lost wake-up shows up as wrong bytes or a hang.
// Built and run by tests/test_hoststage.py (also under -fsanitize=thread).
#include "../../kmerind_b200/csrc/kmi_hoststage.hpp"

#include <cstdio>
#include <cstdlib>
#include <deque>
#include <functional>
#include <random>

namespace {
struct FakeStream {
  std::mutex mu;
  std::condition_variable cv;
  std::deque<std::function<void()>> q;
  uint64_t enq = 0, done = 0;
  bool stop = false;
  std::thread worker;
  FakeStream() : worker([this] { run(); }) {}
  ~FakeStream() {
    { std::lock_guard<std::mutex> lk(mu); stop = true; }
    cv.notify_all();
    worker.join();
  }
  void run() {
    std::minstd_rand rng(7);
    for (;;) {
      std::function<void()> f;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return stop || !q.empty(); });
        if (q.empty()) return;
        f = std::move(q.front());
        q.pop_front();
      }
      if (rng() % 4 == 0) std::this_thread::sleep_for(std::chrono::microseconds(rng() % 200));
      f();
      { std::lock_guard<std::mutex> lk(mu); ++done; }
      cv.notify_all();
    }
  }
  uint64_t push(std::function<void()> f) {
    uint64_t seq;
    { std::lock_guard<std::mutex> lk(mu); q.push_back(std::move(f)); seq = ++enq; }
    cv.notify_all();
    return seq;
  }
  uint64_t now() { std::lock_guard<std::mutex> lk(mu); return enq; }
  void wait(uint64_t seq) {
    std::unique_lock<std::mutex> lk(mu);
    cv.wait(lk, [&] { return done >= seq; });
  }
};

struct FakeBackend {
  FakeStream& st;
  std::vector<uint64_t> slot_seq, mark_seq;
  std::mutex mu;
  long fail_at = -1;   // copy number that reports an error
  long copies = 0;
  FakeBackend(FakeStream& s, int slots, size_t marks) : st(s), slot_seq((size_t)slots, 0), mark_seq(marks, 0) {}
  void thread_init() {}
  bool copy(void* dst, const void* src, size_t n) {
    { std::lock_guard<std::mutex> lk(mu); if (copies++ == fail_at) return false; }
    st.push([=] { std::memcpy(dst, src, n); });
    return true;
  }
  bool copy_in(void* dev, const void* pinned, size_t n) { return copy(dev, pinned, n); }
  bool copy_out(void* pinned, const void* dev, size_t n) { return copy(pinned, dev, n); }
  bool record_slot(int slot) { uint64_t s = st.now(); std::lock_guard<std::mutex> lk(mu); slot_seq[(size_t)slot] = s; return true; }
  bool wait_slot(int slot) { uint64_t s; { std::lock_guard<std::mutex> lk(mu); s = slot_seq[(size_t)slot]; } st.wait(s); return true; }
  bool record_mark(size_t i) { uint64_t s = st.now(); std::lock_guard<std::mutex> lk(mu); mark_seq[i] = s; return true; }
  uint64_t mark(size_t i) { std::lock_guard<std::mutex> lk(mu); return mark_seq[i]; }
};

int check(bool ok, const char* what) {
  if (!ok) { std::fprintf(stderr, "stage_check FAILED: %s\n", what); std::exit(1); }
  return 0;
}
}  // namespace

int main() {
  std::mt19937_64 rng(12345);
  for (int trial = 0; trial < 40; ++trial) {
    const size_t n = trial == 0 ? 0 : (size_t)(rng() % 200000);
    const size_t slice = 1 + (size_t)(rng() % 4096);
    const int slots = 1 + (int)(rng() % 7), threads = 1 + (int)(rng() % 6);
    std::vector<uint8_t> host(n + 1), dev(n + 1, 0), ring(slice * (size_t)slots), back(n + 1, 0);
    for (size_t i = 0; i < n; ++i) host[i] = (uint8_t)rng();
    // ascending range ends with repeats (empty ranges); the last one is n
    std::vector<uint64_t> ends;
    const int n_marks = 1 + (int)(rng() % 9);
    for (int i = 0; i + 1 < n_marks; ++i) ends.push_back(n ? rng() % (n + 1) : 0);
    std::sort(ends.begin(), ends.end());
    if (trial % 3 == 0 && ends.size() > 1) ends[1] = ends[0];
    ends.push_back(n);
    FakeStream st;
    {
      FakeBackend b(st, slots, ends.size());
      kmi::HostStageJob<FakeBackend> job(b, ring.data(), slice, slots, threads);
      job.start_in(dev.data(), host.data(), ends);
      for (size_t i = 0; i < ends.size(); ++i) {
        check(job.wait_mark(i), "wait_mark");
        st.wait(b.mark(i));   // what cudaStreamWaitEvent(stream, copy_events[i]) gives the consumer
        check(std::memcmp(dev.data(), host.data(), (size_t)ends[i]) == 0, "bytes before a mark differ after its event");
      }
      check(!job.wait_mark(ends.size()), "a mark that does not exist must not block");
      check(job.finish(false), "finish after start_in");
    }
    {
      FakeBackend b(st, slots, 0);
      kmi::HostStageJob<FakeBackend> job(b, ring.data(), slice, slots, threads);
      check(job.run_out(back.data(), dev.data(), n), "run_out");
      check(std::memcmp(back.data(), host.data(), n) == 0, "device -> host bytes differ");
    }
  }
  // an error in the middle of a job: nobody hangs, the consumer is told
  for (int dir = 0; dir < 2; ++dir) {
    const size_t n = 100000, slice = 1000;
    std::vector<uint8_t> host(n, 1), dev(n, 2), ring(slice * 4);
    FakeStream st;
    FakeBackend b(st, 4, 1);
    b.fail_at = 37;
    kmi::HostStageJob<FakeBackend> job(b, ring.data(), slice, 4, 3);
    if (dir == 0) {
      job.start_in(dev.data(), host.data(), std::vector<uint64_t>{n});
      check(!job.wait_mark(0), "wait_mark must report the failed copy");
      check(!job.finish(false), "finish must report the failed copy");
    } else {
      check(!job.run_out(host.data(), dev.data(), n), "run_out must report the failed copy");
    }
  }
  // a source that is not memory (the pread of kmi_build_file): bytes come from the fill function; a failing read is reported
  for (int bad = 0; bad < 2; ++bad) {
    const size_t n = 300000, slice = 4096;
    std::vector<uint8_t> file(n), dev(n, 0), ring(slice * 5);
    for (size_t i = 0; i < n; ++i) file[i] = (uint8_t)(i * 2654435761u >> 13);
    FakeStream st;
    FakeBackend b(st, 5, 3);
    kmi::HostStageJob<FakeBackend> job(b, ring.data(), slice, 5, 4);
    job.set_fill([&](uint8_t* dst, uint64_t off, size_t len) {
      if (bad && off >= 200000) return false;
      std::memcpy(dst, file.data() + off, len);
      return true;
    });
    job.start_in(dev.data(), nullptr, std::vector<uint64_t>{100000, 100000, n});
    if (!bad) {   // (once a read has failed every wait reports it, also for marks that were reached before)
      check(job.wait_mark(0) && job.wait_mark(1), "marks of a healthy job");
      st.wait(b.mark(1));
      check(std::memcmp(dev.data(), file.data(), 100000) == 0, "bytes from the fill function");
    }
    check(job.wait_mark(2) == !bad, "a failed read must fail the mark");
    check(job.finish(false) == !bad, "a failed read must fail the job");
    if (!bad) { st.wait(b.mark(2)); check(std::memcmp(dev.data(), file.data(), n) == 0, "all bytes from the fill function"); }
  }
  // a job abandoned half way (early return of the caller): the destructor stops the workers
  {
    const size_t n = 4000000, slice = 1000;
    std::vector<uint8_t> host(n, 1), dev(n, 2), ring(slice * 4);
    FakeStream st;
    FakeBackend b(st, 4, 2);
    kmi::HostStageJob<FakeBackend> job(b, ring.data(), slice, 4, 3);
    job.start_in(dev.data(), host.data(), std::vector<uint64_t>{1000, n});
    check(job.wait_mark(0), "first mark");
  }
  std::puts("stage_check ok");
  return 0;
}
