// kmi_api.cu — the C ABI of include/kmi.h: host-side orchestration of the sm_100a kernels.
// No CPU fallback: every compute entry point needs a CUDA device and fails with KMI_ERR_CUDA
// otherwise. Reference interfaces replaced are cited in include/kmi.h next to each function.
#define KMI_SHARED_KERNELS
#include "kmi_internal.cuh"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace kmi {
const Ops* kmi_ops_k15(int);
const Ops* kmi_ops_k21(int);
const Ops* kmi_ops_k31(int);
const Ops* kmi_ops_k35(int);
const Ops* kmi_ops_k47(int);
const Ops* kmi_ops_k63(int);
}  // namespace kmi

using namespace kmi;

static thread_local char g_err[512] = "";
namespace kmi {
int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
}  // namespace kmi

// any other K: its kernels live in kmods/libkmi_k<K>.so next to this library (python -m kmerind_b200.build --k K ...)
#include <dlfcn.h>
#include <map>
#include <mutex>
#include <string>
static const Ops* find_ops_module(int k, int hash) {
  static std::mutex mu;
  static std::map<int, void*> loaded;
  std::lock_guard<std::mutex> lock(mu);
  void* h = nullptr;
  auto it = loaded.find(k);
  if (it != loaded.end()) h = it->second;
  else {
    std::string dir;
    if (const char* e = getenv("KMI_KMOD_DIR")) dir = e;
    else {
      Dl_info info;
      if (dladdr(reinterpret_cast<const void*>(&find_ops_module), &info) && info.dli_fname) {
        dir = info.dli_fname;
        const size_t slash = dir.rfind('/');
        dir = (slash == std::string::npos ? std::string(".") : dir.substr(0, slash)) + "/kmods";
      }
    }
    const std::string path = dir + "/libkmi_k" + std::to_string(k) + ".so";
    h = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (h) {
      typedef unsigned long (*abi_fn)(void);
      typedef int (*k_fn)(void);
      abi_fn abi = reinterpret_cast<abi_fn>(dlsym(h, "kmi_ops_module_abi"));
      k_fn kk = reinterpret_cast<k_fn>(dlsym(h, "kmi_ops_module_k"));
      if (!abi || !kk || abi() != sizeof(Ops) || kk() != k) { dlclose(h); h = nullptr; }   // built from other headers
    }
    loaded[k] = h;
  }
  if (!h) return nullptr;
  typedef const Ops* (*ops_fn)(int);
  ops_fn f = reinterpret_cast<ops_fn>(dlsym(h, "kmi_ops_module"));
  return f ? f(hash) : nullptr;
}

static const Ops* find_ops(int k, int hash) {
  switch (k) {
    case 15: return kmi_ops_k15(hash);
    case 21: return kmi_ops_k21(hash);
    case 31: return kmi_ops_k31(hash);
    case 35: return kmi_ops_k35(hash);
    case 47: return kmi_ops_k47(hash);
    case 63: return kmi_ops_k63(hash);
    default: return find_ops_module(k, hash);
  }
}

static u64 pow2_at_least(u64 x) {
  u64 p = 1024;
  while (p < x) p <<= 1;
  return p;
}
static const double DEFAULT_MAX_LOAD = 0.7;  // dense_hash_map max load, reference densehash_map.hpp:969 (KMI_PARAM_MAX_LOAD changes it per index)
#define MAX_LOAD (idx->max_load)

int kmi::read_ctr(kmi_index* idx, int which, u64* out) {
  CU(cudaMemcpyAsync(out, idx->ctr + which, sizeof(u64), cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  return KMI_OK;
}
int kmi::zero_ctr(kmi_index* idx, int which, int count) {
  CU(cudaMemsetAsync(idx->ctr + which, 0, sizeof(u64) * count, idx->stream));
  return KMI_OK;
}
// CTR_ERR holds the sticky TERR_* bits (cleared by kmi_clear only)
int kmi::table_error(u64 e) {
  if (e & TERR_SPILL) return fail(KMI_ERR_CAPACITY, "segment spill list exhausted (keys homed past their table segment's end): index contents are incomplete");
  if (e) return fail(KMI_ERR_CAPACITY, "hash table overflow (probe limit reached)");
  return KMI_OK;
}
int kmi::check_table_error(kmi_index* idx) {
  u64 e = 0;
  OK(read_ctr(idx, CTR_ERR, &e));
  return table_error(e);
}

// ---- pageable host memory <-> device (kmi_hoststage.hpp) ----------------------------------------------------------------
// The reference's callers hold file bytes, query vectors and answers in ordinary heap memory; cudaMemcpy from there runs on
// one driver thread (~9 GB/s on the bench box). Large copies are cut into slices that KMI_STAGE_THREADS host threads (default
// min(8, cores); 0 = off: plain cudaMemcpyAsync) move through a ring of pinned slots of KMI_STAGE_SLICE_BYTES (4 MiB),
// 3 slots per thread, kept with the index; copies below KMI_STAGE_MIN_BYTES (16 MiB) and pinned / registered / managed memory
// go straight to cudaMemcpyAsync.
#include "kmi_hoststage.hpp"
#include <thread>
#include <unistd.h>
namespace kmi {
struct HostRing {
  uint8_t* ring = nullptr;
  size_t slice = 0, min_bytes = 0;
  int slots = 0, threads = 0;
  std::vector<cudaEvent_t> ev;   // per slot: the last device copy that touched it
  cudaEvent_t one = nullptr;     // the single mark of copy_to_device / the hand-over from idx->stream
  bool busy = false;             // a HostFeed is running on it
};
struct CudaStageBackend {
  int device; cudaStream_t stream; cudaEvent_t* slot_ev; cudaEvent_t* mark_ev; size_t n_marks;
  void thread_init() { cudaSetDevice(device); }
  bool copy_in(void* dev, const void* pinned, size_t n) { return cudaMemcpyAsync(dev, pinned, n, cudaMemcpyHostToDevice, stream) == cudaSuccess; }
  bool copy_out(void* pinned, const void* dev, size_t n) { return cudaMemcpyAsync(pinned, dev, n, cudaMemcpyDeviceToHost, stream) == cudaSuccess; }
  bool record_slot(int s) { return cudaEventRecord(slot_ev[s], stream) == cudaSuccess; }
  bool wait_slot(int s) { return cudaEventSynchronize(slot_ev[s]) == cudaSuccess; }
  bool record_mark(size_t i) { return i < n_marks && cudaEventRecord(mark_ev[i], stream) == cudaSuccess; }
};
static size_t env_size(const char* name, size_t dflt) {
  const char* e = getenv(name);
  return (e && *e) ? (size_t)strtoull(e, nullptr, 10) : dflt;
}
// the index's ring, or nullptr when staging is off or the pinned memory cannot be had (callers then copy directly)
static HostRing* host_ring(kmi_index* idx) {
  if (idx->hring || idx->hring_tried) return idx->hring;
  idx->hring_tried = true;
  // default: up to 8 threads (the bench box's memcpy rate peaks there: 39 GB/s), and the cores are shared with the other
  // ranks of a sharded index (one process per GPU on one host)
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  const unsigned share = (unsigned)std::max(1, idx->cfg.nranks);
  const int threads = (int)std::min<size_t>(env_size("KMI_STAGE_THREADS", std::min(8u, std::max(1u, hw / share))), 64);
  if (threads <= 0) return nullptr;
  HostRing* r = new (std::nothrow) HostRing();
  if (!r) return nullptr;
  r->threads = threads;
  r->slots = std::max(4, 3 * threads);
  r->slice = std::max<size_t>(env_size("KMI_STAGE_SLICE_BYTES", 4u << 20), 256);
  r->min_bytes = env_size("KMI_STAGE_MIN_BYTES", 16u << 20);
  bool ok = cudaHostAlloc(reinterpret_cast<void**>(&r->ring), r->slice * (size_t)r->slots, cudaHostAllocDefault) == cudaSuccess;
  if (!ok) r->ring = nullptr;
  for (int i = 0; ok && i < r->slots; ++i) {
    cudaEvent_t e = nullptr;
    ok = cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess;
    if (ok) r->ev.push_back(e);
  }
  ok = ok && cudaEventCreateWithFlags(&r->one, cudaEventDisableTiming) == cudaSuccess;
  ok = ok && (idx->copy_stream || cudaStreamCreateWithFlags(&idx->copy_stream, cudaStreamNonBlocking) == cudaSuccess);
  idx->hring = r;
  if (!ok) { cudaGetLastError(); host_ring_release(idx); }
  return idx->hring;
}
void host_ring_release(kmi_index* idx) {
  HostRing* r = idx->hring;
  if (!r) return;
  for (cudaEvent_t e : r->ev) cudaEventDestroy(e);
  if (r->one) cudaEventDestroy(r->one);
  if (r->ring) cudaFreeHost(r->ring);
  delete r;
  idx->hring = nullptr;
}
// ordinary heap / stack / mmap memory: neither pinned nor registered nor managed nor device
static bool host_is_pageable(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
  return a.type == cudaMemoryTypeUnregistered;
}
bool host_ring_ready(kmi_index* idx) {
  HostRing* r = host_ring(idx);
  return r && !r->busy;
}
// `want` bytes of the file at offset `at` (short only at the end of the file), or -1
static long long pread_full(int fd, uint8_t* dst, size_t want, unsigned long long at) {
  size_t got = 0;
  while (got < want) {
    const ssize_t r = pread(fd, dst + got, want - got, (off_t)(at + got));
    if (r < 0) return -1;
    if (r == 0) break;
    got += (size_t)r;
  }
  return (long long)got;
}
static HostRing* ring_for(kmi_index* idx, const void* host, size_t bytes) {
  if (bytes == 0) return nullptr;
  HostRing* r = host_ring(idx);
  if (!r || r->busy || bytes < r->min_bytes || !host_is_pageable(host)) return nullptr;
  return r;
}

int copy_to_device(kmi_index* idx, void* d_dst, const void* src, size_t bytes) {
  if (bytes == 0) return KMI_OK;
  HostRing* r = ring_for(idx, src, bytes);
  if (!r) { CU(cudaMemcpyAsync(d_dst, src, bytes, cudaMemcpyHostToDevice, idx->stream)); return KMI_OK; }
  CU(cudaEventRecord(r->one, idx->stream));                 // d_dst may still be in use by work queued earlier
  CU(cudaStreamWaitEvent(idx->copy_stream, r->one, 0));
  CudaStageBackend b{idx->cfg.device, idx->copy_stream, r->ev.data(), &r->one, 1};
  bool ok;
  {
    HostStageJob<CudaStageBackend> job(b, r->ring, r->slice, r->slots, r->threads);
    job.start_in(static_cast<uint8_t*>(d_dst), static_cast<const uint8_t*>(src), std::vector<uint64_t>{(uint64_t)bytes});
    ok = job.wait_mark(0);          // every slice has left the caller's memory and is enqueued
    ok = job.finish(false) && ok;
  }
  if (!ok) { cudaError_t e = cudaGetLastError(); return fail(KMI_ERR_CUDA, "staged host -> device copy of %zu bytes: %s", bytes, cudaGetErrorString(e)); }
  CU(cudaStreamWaitEvent(idx->stream, r->one, 0));
  return KMI_OK;
}
int copy_to_host(kmi_index* idx, void* dst, const void* d_src, size_t bytes) {
  if (bytes == 0) return KMI_OK;
  HostRing* r = ring_for(idx, dst, bytes);
  if (!r) { CU(cudaMemcpyAsync(dst, d_src, bytes, cudaMemcpyDeviceToHost, idx->stream)); return KMI_OK; }
  CU(cudaEventRecord(r->one, idx->stream));                 // whatever produces d_src
  CU(cudaStreamWaitEvent(idx->copy_stream, r->one, 0));
  CudaStageBackend b{idx->cfg.device, idx->copy_stream, r->ev.data(), nullptr, 0};
  HostStageJob<CudaStageBackend> job(b, r->ring, r->slice, r->slots, r->threads);
  if (!job.run_out(static_cast<uint8_t*>(dst), static_cast<const uint8_t*>(d_src), bytes)) {
    cudaError_t e = cudaGetLastError();
    return fail(KMI_ERR_CUDA, "staged device -> host copy of %zu bytes: %s", bytes, cudaGetErrorString(e));
  }
  return KMI_OK;
}

struct HostFeed::Impl {
  CudaStageBackend b;
  HostStageJob<CudaStageBackend> job;
  Impl(const CudaStageBackend& b_, HostRing* r) : b(b_), job(b, r->ring, r->slice, r->slots, r->threads) {}
};
int copy_to_device(kmi_index* idx, void* d_dst, const HostSrc& src, size_t bytes) {
  if (src.fd < 0) return copy_to_device(idx, d_dst, static_cast<const void*>(src.mem), bytes);
  if (bytes == 0) return KMI_OK;
  HostFeed feed;
  OK(feed.start(idx, static_cast<uint8_t*>(d_dst), src, std::vector<u64>{(u64)bytes}, idx->stream));
  return feed.wait(0, idx->stream);   // (the feed's destructor then waits for the copies themselves)
}

int HostFeed::start(kmi_index* idx, uint8_t* d_dst, const HostSrc& hs, const std::vector<u64>& ends, cudaStream_t after) {
  idx_ = idx;
  const uint8_t* src = hs.mem;
  if (!idx->copy_stream) CU(cudaStreamCreateWithFlags(&idx->copy_stream, cudaStreamNonBlocking));
  CU(cudaEventRecord(idx->ev[3], after));                   // the destination may still be in use by earlier work
  CU(cudaStreamWaitEvent(idx->copy_stream, idx->ev[3], 0));
  while (idx->copy_events.size() < ends.size()) {
    cudaEvent_t e;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    idx->copy_events.push_back(e);
  }
  const u64 total = ends.empty() ? 0 : ends.back();
  HostRing* r = nullptr;
  if (hs.fd >= 0) {
    if (!host_ring_ready(idx)) return fail(KMI_ERR_UNSUPPORTED, "reading a file into the device needs the pinned slice ring (KMI_STAGE_THREADS > 0)");
    r = idx->hring;
  } else {
    r = ring_for(idx, src, (size_t)total);
  }
  if (!r) {   // pinned (or small): cudaMemcpyAsync does not block, everything is enqueued now
    u64 at = 0;
    for (size_t i = 0; i < ends.size(); ++i) {
      if (ends[i] > at) CU(cudaMemcpyAsync(d_dst + at, src + at, ends[i] - at, cudaMemcpyHostToDevice, idx->copy_stream));
      at = std::max(at, ends[i]);
      CU(cudaEventRecord(idx->copy_events[i], idx->copy_stream));
    }
    return KMI_OK;
  }
  CudaStageBackend b{idx->cfg.device, idx->copy_stream, r->ev.data(), idx->copy_events.data(), ends.size()};
  impl_ = new (std::nothrow) Impl(b, r);
  if (!impl_) return fail(KMI_ERR_NOMEM, "out of host memory");
  r->busy = true;
  std::vector<uint64_t> e64(ends.begin(), ends.end());
  if (hs.fd >= 0) {
    const int fd = hs.fd;
    const u64 base = hs.fd_off;
    impl_->job.set_fill([fd, base](uint8_t* dst, uint64_t off, size_t len) { return pread_full(fd, dst, len, base + off) == (long long)len; });
  }
  impl_->job.start_in(d_dst, src, e64);
  return KMI_OK;
}
int HostFeed::wait(size_t i, cudaStream_t consumer) {
  if (impl_ && !impl_->job.wait_mark(i)) {
    cudaError_t e = cudaGetLastError();
    return fail(KMI_ERR_CUDA, "staged host -> device copy (range %zu): %s", i, cudaGetErrorString(e));
  }
  CU(cudaStreamWaitEvent(consumer, idx_->copy_events[i], 0));
  return KMI_OK;
}
HostFeed::~HostFeed() {
  if (impl_) {
    delete impl_;   // stops and joins the workers
    if (idx_->hring) idx_->hring->busy = false;
  }
  if (idx_ && idx_->copy_stream) cudaStreamSynchronize(idx_->copy_stream);
}
}  // namespace kmi

// allocate a table of `cap` slots (contents undefined: the caller clears it or marks the index virgin)
static int table_alloc(kmi_index* idx, u64 cap, void** out) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, cap * idx->slot_bytes);
  if (e != cudaSuccess) return fail(KMI_ERR_NOMEM, "cudaMalloc table %llu slots: %s", (unsigned long long)cap, cudaGetErrorString(e));
  *out = p;
  return KMI_OK;
}

// grow (rehash) to `new_cap` slots
static int table_grow(kmi_index* idx, u64 new_cap) {
  if (idx->pend_keys) OK(flush_pending(idx));            // staged fine buckets are cut for the current capacity
  void* fresh = nullptr;
  OK(table_alloc(idx, new_cap, &fresh));
  void* old = idx->slots;
  u64 old_cap = idx->cap;
  const bool old_has_content = old && !idx->virgin;
  idx->slots = fresh;
  idx->cap = new_cap;
  idx->virgin = true;                                   // nothing written yet
  OK(zero_ctr(idx, CTR_SIZE));
  if (old_has_content) {
    idx->ops->rehash(idx->stream, old, old_cap, idx->view());   // view() clears the new table first
    idx->launches++;
  }
  if (old) {
    CU(cudaStreamSynchronize(idx->stream));
    CU(cudaFree(old));
  }
  OK(read_ctr(idx, CTR_SIZE, &idx->size_exact));
  idx->size_upper = idx->size_exact;
  idx->tombstones = 0;                                  // the rehash copies live keys only
  return KMI_OK;
}

// Make room for up to `incoming` more distinct keys. Returns in *grant how many keys the caller
// may issue before calling again (>= 1 unless incoming == 0). The table never exceeds MAX_LOAD:
// the bound is tracked on the host and refreshed from the device counter only when it runs out.
static int table_room(kmi_index* idx, u64 incoming, u64* grant) {
  if (incoming == 0) { *grant = 0; return KMI_OK; }
  if (!idx->slots) OK(table_grow(idx, pow2_at_least((u64)(std::min<u64>(incoming, 1u << 20) / MAX_LOAD) + 1)));
  for (;;) {
    u64 limit = (u64)(idx->cap * MAX_LOAD);
    const u64 occupied = idx->size_upper + idx->tombstones;
    u64 slack = limit > occupied ? limit - occupied : 0;
    if (slack >= incoming) { *grant = incoming; break; }
    // refresh the bound
    if (idx->size_upper != idx->size_exact) {
      OK(read_ctr(idx, CTR_SIZE, &idx->size_exact));
      idx->size_upper = idx->size_exact;
      continue;
    }
    u64 min_chunk = std::max<u64>(idx->cap / 8, 1u << 12);
    if (slack >= min_chunk) { *grant = std::min(slack, incoming); break; }
    // tombstones alone may be what fills the table: then a rehash at the same capacity purges them
    const bool purge = idx->tombstones && idx->size_upper + std::min(incoming, idx->cap / 4) <= limit;
    OK(table_grow(idx, purge ? idx->cap : idx->cap * 2));
  }
  idx->size_upper += *grant;
  return KMI_OK;
}


extern "C" const char* kmi_last_error(void) { return g_err; }
extern "C" int kmi_device_count(int* n) {
  if (!n) return fail(KMI_ERR_ARG, "null argument");
  *n = 0;
  cudaError_t e = cudaGetDeviceCount(n);
  if (e != cudaSuccess) { *n = 0; return fail(KMI_ERR_CUDA, "no CUDA device (%s)", cudaGetErrorString(e)); }
  return KMI_OK;
}

extern "C" int kmi_create(const kmi_config* cfg, kmi_index** out) {
  if (!cfg || !out) return fail(KMI_ERR_ARG, "null argument");
  *out = nullptr;
  if (cfg->k < 2 || cfg->k > 63 || cfg->k == 32) return fail(KMI_ERR_ARG, "k=%d needs >= 2 pad bits (k <= 31 or 33..63)", cfg->k);
  if (cfg->mode != KMI_MODE_COUNT && cfg->mode != KMI_MODE_POSITION) return fail(KMI_ERR_ARG, "bad mode %d", cfg->mode);
  if (cfg->hash < 0 || cfg->hash > 4) return fail(KMI_ERR_ARG, "bad hash kind %d", cfg->hash);
  if (cfg->nranks < 1 || cfg->nranks > MAX_RANKS || cfg->rank < 0 || cfg->rank >= cfg->nranks)
    return fail(KMI_ERR_ARG, "bad rank %d / nranks %d (max %d)", cfg->rank, cfg->nranks, MAX_RANKS);
  const Ops* ops = find_ops(cfg->k, cfg->hash);
  if (!ops) return fail(KMI_ERR_UNSUPPORTED, "k=%d: libkmi.so carries 15, 21, 31, 35, 47, 63 and found no kmods/libkmi_k%d.so next to it "
                                 "(python -m kmerind_b200.build --k %d builds it)", cfg->k, cfg->k, cfg->k);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(KMI_ERR_CUDA, "no CUDA device (%s): this library has no CPU path", cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= ndev) return fail(KMI_ERR_ARG, "device %d of %d", cfg->device, ndev);
  CU(cudaSetDevice(cfg->device));
  kmi_index* idx = new (std::nothrow) kmi_index();
  if (!idx) return fail(KMI_ERR_NOMEM, "host allocation");
  idx->cfg = *cfg;
  idx->ops = ops;
  idx->nw = ops->NW;
  idx->slot_bytes = ops->NW == 1 ? sizeof(Slot<1>) : sizeof(Slot<2>);
  e = cudaMalloc(&idx->ctr, sizeof(u64) * CTR_N);
  if (e != cudaSuccess) { delete idx; return fail(KMI_ERR_CUDA, "cudaMalloc counters: %s", cudaGetErrorString(e)); }
  cudaMemset(idx->ctr, 0, sizeof(u64) * CTR_N);
  for (int i = 0; i < 6; ++i) cudaEventCreate(&idx->ev[i]);
  for (int i = 0; i < 4; ++i) cudaEventCreate(&idx->kev[i]);
  *out = idx;
  if (cfg->reserve) {
    int r = kmi_reserve(idx, cfg->reserve);
    if (r != KMI_OK) { kmi_destroy(idx); *out = nullptr; return r; }
  }
  return KMI_OK;
}

extern "C" void kmi_destroy(kmi_index* idx) {
  if (!idx) return;
  cudaSetDevice(idx->cfg.device);
  cudaStreamSynchronize(idx->stream);
  if (idx->slots) cudaFree(idx->slots);
  if (idx->ctr) cudaFree(idx->ctr);
  DevBuf* bufs[] = {&idx->tile_lines, &idx->block_sums, &idx->stage_in, &idx->stage_keys, &idx->stage_vals,
                    &idx->stage_out, &idx->stage_found, &idx->hist, &idx->hist_sums, &idx->log_keys, &idx->log_ids,
                    &idx->log_keys2, &idx->log_ids2, &idx->offsets, &idx->csr_sums, &idx->values, &idx->bkt_keys, &idx->bkt_meta, &idx->fine_keys, &idx->fine_cur, &idx->parse_keys, &idx->parse_ids, &idx->tile_rec, &idx->mid_keys, &idx->mid_cur,
                    &idx->pend_ovf, &idx->pend_ovf_cur, &idx->pend_cur_save, &idx->scan_tmp};
  for (DevBuf* b : bufs) b->release();
  for (int i = 0; i < 6; ++i) if (idx->ev[i]) cudaEventDestroy(idx->ev[i]);
  for (int i = 0; i < 4; ++i) if (idx->kev[i]) cudaEventDestroy(idx->kev[i]);
  if (idx->copy_stream) cudaStreamSynchronize(idx->copy_stream);
  for (cudaEvent_t e : idx->copy_events) cudaEventDestroy(e);
  if (idx->copy_stream) cudaStreamDestroy(idx->copy_stream);
  for (int i = 0; i < 2; ++i) if (idx->file_buf[i]) cudaFreeHost(idx->file_buf[i]);
  host_ring_release(idx);
  if (idx->dist) kmi_dist_release(idx);
  delete idx;
}

extern "C" int kmi_set_stream(kmi_index* idx, void* cuda_stream) {
  if (!idx) return fail(KMI_ERR_ARG, "null index");
  idx->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
  return KMI_OK;
}

extern "C" int kmi_set_param(kmi_index* idx, int param, int64_t value) {
  if (!idx) return fail(KMI_ERR_ARG, "null index");
  if (param == KMI_PARAM_BUILD_PATH) {
    if (value < 0 || value > 2) return fail(KMI_ERR_ARG, "build path must be 0 (auto), 1 (direct) or 2 (bucketed)");
    if (idx->saturate && value != 1) return fail(KMI_ERR_ARG, "a saturating index inserts with direct atomics (build path 1)");
    idx->build_path = (int)value;
    return KMI_OK;
  }
  if (param == KMI_PARAM_SEGMENT_INSERT) {
    idx->seg_path = value != 0;
    return KMI_OK;
  }
  if (param == KMI_PARAM_SATURATE) {
    if (idx->cfg.mode != KMI_MODE_COUNT) return fail(KMI_ERR_ARG, "saturating counts belong to a COUNT index");
    idx->saturate = value != 0;
    if (idx->saturate) idx->build_path = 1;              // the shared-memory segment insert adds with plain wrap-around
    return KMI_OK;
  }
  if (param == KMI_PARAM_DEFER_INSERT) {
    if (value < 0) return fail(KMI_ERR_ARG, "staging budget must be >= 0 keys");
    if (idx->cfg.mode != KMI_MODE_COUNT) return fail(KMI_ERR_ARG, "bulk-load staging belongs to a COUNT index");
    if (!value) OK(flush_pending(idx));
    idx->defer_budget = (u64)value;
    return KMI_OK;
  }
  if (param == KMI_PARAM_MAX_LOAD) {
    if (value < 10 || value > 70) return fail(KMI_ERR_ARG, "max load is a percentage between 10 and 70");
    idx->max_load = (double)value / 100.0;               // takes effect at the next sizing decision (reserve / insert / build)
    return KMI_OK;
  }
  if (param == KMI_PARAM_DIST_TRANS) {
    if (value < 0 || value > 2) return fail(KMI_ERR_ARG, "dist transform must be 0 (identity), 1 (lex_less) or 2 (xor_rev_comp)");
    if (idx->cfg.canonical && value == 2)
      return fail(KMI_ERR_ARG, "a canonical index distributes its lex_less keys as they are (CanonicalHashMapParams fixes DistTrans)");
    idx->dist_trans = (u32)value;
    return KMI_OK;
  }
  return fail(KMI_ERR_ARG, "unknown parameter %d", param);
}

extern "C" int kmi_flush(kmi_index* idx) {
  if (!idx) return fail(KMI_ERR_ARG, "null index");
  CU(cudaSetDevice(idx->cfg.device));
  OK(flush_pending(idx));
  CU(cudaStreamSynchronize(idx->stream));
  harvest_commit_times(idx);
  return check_table_error(idx);
}

extern "C" int kmi_bulk_stats(kmi_index* idx, uint64_t* staged_keys, uint64_t* flushes) {
  if (!idx) return fail(KMI_ERR_ARG, "null index");
  if (staged_keys) *staged_keys = idx->pend_keys;
  if (flushes) *flushes = idx->flushes;
  return KMI_OK;
}

extern "C" int kmi_reserve(kmi_index* idx, uint64_t n_distinct) {
  if (!idx) return fail(KMI_ERR_ARG, "null index");
  CU(cudaSetDevice(idx->cfg.device));
  u64 want = pow2_at_least((u64)(n_distinct / MAX_LOAD) + 1);
  if (want > idx->cap) OK(table_grow(idx, want));
  return KMI_OK;
}

extern "C" int kmi_clear(kmi_index* idx) {
  if (!idx) return fail(KMI_ERR_ARG, "null index");
  CU(cudaSetDevice(idx->cfg.device));
  idx->virgin = idx->slots != nullptr;                  // cleared lazily (kmi_index::touch) or never (segment insert)
  if (idx->pend_keys) {                                 // staged keys are dropped with everything else
    idx->pend_keys = 0;
    CU(cudaMemsetAsync(idx->fine_cur.p, 0, ((u64)1 << idx->pend_log_segs) * sizeof(u32), idx->stream));
  }
  CU(cudaMemsetAsync(idx->ctr, 0, sizeof(u64) * CTR_N, idx->stream));
  idx->size_exact = idx->size_upper = 0;
  idx->tombstones = 0;
  idx->log_n = 0;
  idx->csr_dirty = false;
  idx->non_hll_upper = 0;
  if (idx->hll_ready) CU(cudaMemsetAsync(idx->bkt_meta.as<u32>() + META_HLL, 0, sizeof(u32) * HLL_M, idx->stream));
  return KMI_OK;
}


// ---- hash-prefix buckets (kmi_bucket.cuh) -------------------------------------------------------------

int kmi::bucket_meta_prepare(kmi_index* idx) {
  if (!idx->hll_ready) {
    OK(idx->bkt_meta.ensure(sizeof(u32) * META_WORDS));
    CU(cudaMemsetAsync(idx->bkt_meta.p, 0, sizeof(u32) * META_WORDS, idx->stream));
    idx->hll_ready = true;
  } else {
    CU(cudaMemsetAsync(idx->bkt_meta.p, 0, sizeof(u32) * META_HLL, idx->stream));
  }
  return KMI_OK;
}

// (re)size the bucket regions for up to `max_keys` keys and zero the cursors
int kmi::bucket_prepare(kmi_index* idx, u64 max_keys, BucketSink* bs) {
  if (max_keys > BUCKET_MAX_KEYS) return fail(KMI_ERR_ARG, "bucket round of %llu keys", (unsigned long long)max_keys);
  // uniform hash: a bucket receives max_keys/256 +- a few sigma; 6 % + 1024 slack, the rest goes to the overflow region
  u64 capb = (max_keys / N_BUCKETS + max_keys / (N_BUCKETS * 16) + 1024 + 7) & ~7ULL;   // multiple of 8: flush chunks stay 64-byte aligned
  u64 ovf = max_keys / 8 + 4096;
  OK(idx->bkt_keys.ensure((capb * N_BUCKETS + ovf) * sizeof(u64) * idx->nw));
  OK(bucket_meta_prepare(idx));
  bs->keys = idx->bkt_keys.as<u64>();
  bs->cursor = idx->bkt_meta.as<u32>();
  bs->hll = idx->bkt_meta.as<u32>() + META_HLL;
  bs->capb = (u32)capb;
  bs->ovf_cap = (u32)ovf;
  bs->bases = nullptr;
  return KMI_OK;
}

// HyperLogLog estimate (Flajolet et al. 2007, with the small-range linear-counting correction)
static double hll_estimate(const u32* reg) {
  double sum = 0;
  int zeros = 0;
  for (int i = 0; i < HLL_M; ++i) { sum += std::ldexp(1.0, -(int)reg[i]); zeros += reg[i] == 0; }
  const double m = HLL_M, alpha = 0.7213 / (1.0 + 1.079 / m);
  double e = alpha * m * m / sum;
  if (e <= 2.5 * m && zeros) e = m * std::log(m / zeros);
  return e;
}

// add the kernel times of the last bucket_commit (its events have completed: call after a sync)
void kmi::harvest_commit_times(kmi_index* idx) {
  if (!idx->kev_pending) return;
  float a = 0, b = 0, c = 0;
  cudaEventElapsedTime(&a, idx->kev[0], idx->kev[1]);
  cudaEventElapsedTime(&b, idx->kev[1], idx->kev[2]);
  cudaEventElapsedTime(&c, idx->kev[2], idx->kev[3]);
  idx->kernel_ms[2] += a; idx->kernel_ms[3] += b; idx->kernel_ms[4] += c;
  idx->kev_pending = false;
}

// one synchronisation per round: counters + HLL registers
int kmi::bucket_readback(kmi_index* idx, const BucketSink& bs, u64* host_ctr, std::vector<u32>& reg) {
  reg.resize(HLL_M);
  idx->ops->hll_buckets(idx->stream, bs);
  idx->launches++;
  CU(cudaMemcpyAsync(reg.data(), bs.hll, sizeof(u32) * HLL_M, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaMemcpyAsync(host_ctr, idx->ctr, sizeof(u64) * CTR_N, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaMemcpyAsync(&idx->h_ovf_cursor, ovf_cur(bs), sizeof(u32), cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  harvest_commit_times(idx);
  return KMI_OK;
}

// ---- geometry of the segment path for the current table capacity and a sink's prefix layout --------------------------------
namespace {
struct SegPlan {
  bool ok = false;        // table segments fit shared memory: fine scatter + k_insert_segments can run
  bool two_pass = false;  // very large tables: mid buckets first, then 256 fine buckets per group of 256 mid regions
  int log_pre = 0, log_segs = 0, log_seg = 0;
  u32 log_f = 0;
  u64 n_regions = 0;
};
SegPlan seg_plan(const kmi_index* idx, const BucketSink& bs) {
  SegPlan pl;
  // second level: table segments of <= 4096 slots (8192 beyond 2^31 slots) updated in shared memory
  const int log_cap = __builtin_ctzll(idx->cap);
  // (regions of the exchange buffer share only log_pre < 8 prefix bits: their fine scatter is wider, never skipped)
  pl.log_pre = (int)bs.log_pre;
  const int log_seg_slots = idx->nw == 1 ? KMI_LOG_SEG_SLOTS : LOG_SEG_SLOTS2;
  // (a receive buffer has `sub` regions per prefix: they always need the fine pass to become one region per segment)
  pl.log_segs = std::max(std::max(LOG_BUCKETS, pl.log_pre + ((pl.log_pre < LOG_BUCKETS || bs.sub > 1) ? 1 : 0)), log_cap - log_seg_slots);
  // beyond 2^9 fine buckets per coarse bucket the scatter runs shrink to one or two keys: very large tables take TWO
  // fine passes instead (2^(log_f - 8) mid buckets, then 256 fine buckets inside every group of 256 mid regions)
  pl.two_pass = pl.log_pre == LOG_BUCKETS && pl.log_segs - pl.log_pre >= 10 && pl.log_segs - pl.log_pre <= 16;
  if (!pl.two_pass) pl.log_segs = std::min(pl.log_segs, pl.log_pre + MAX_LOG_F);
  pl.log_seg = log_cap - pl.log_segs;
  pl.ok = idx->seg_path && idx->ops->insert_segments && pl.log_seg <= (idx->nw == 1 ? 13 : 12) && pl.log_seg >= 8;
  pl.log_f = (u32)(pl.log_segs - pl.log_pre);
  pl.n_regions = (u64)1 << pl.log_segs;
  return pl;
}

// coarse regions of `bs` -> the fine regions `fs` (one per table segment), in one or two passes. The fine cursors are NOT
// reset here: a regular commit zeroes them first, a deferred one appends to what earlier rounds staged.
int fine_scatter_launch(kmi_index* idx, const BucketSink& bs, const SegPlan& pl, FineSink fs, u64 n_keys) {
  if (!pl.two_pass) {
    idx->ops->scatter_fine(idx->stream, bs, fs, reinterpret_cast<u32*>(idx->ctr + CTR_PERR), n_keys);
    idx->launches++;
  } else {
    FineSink mid;
    mid.log_f = pl.log_f - LOG_BUCKETS;
    const u64 n_mid = (u64)N_BUCKETS << mid.log_f;
    const double mean1 = (double)n_keys / (double)n_mid;
    const u64 capm = (u64)(mean1 + 8.0 * std::sqrt(mean1) + 32.0);
    OK(idx->mid_keys.ensure(n_mid * capm * sizeof(u64) * idx->nw));
    OK(idx->mid_cur.ensure(n_mid * sizeof(u32)));
    CU(cudaMemsetAsync(idx->mid_cur.p, 0, n_mid * sizeof(u32), idx->stream));
    mid.keys = idx->mid_keys.as<u64>(); mid.cursor = idx->mid_cur.as<u32>(); mid.capf = (u32)capm;
    idx->ops->scatter_fine(idx->stream, bs, mid, reinterpret_cast<u32*>(idx->ctr + CTR_PERR), n_keys);
    idx->launches++;
    fs.log_f = LOG_BUCKETS;                            // per group of 256 mid regions
    for (u64 g = 0; g < n_mid / N_BUCKETS; ++g) {
      BucketSink bg;
      bg.keys = mid.keys + g * N_BUCKETS * capm * idx->nw;
      bg.cursor = mid.cursor + g * N_BUCKETS;
      bg.capb = (u32)capm; bg.ovf_cap = bs.ovf_cap; bg.hll = nullptr; bg.bases = nullptr;
      bg.log_pre = (u32)LOG_BUCKETS + mid.log_f; bg.pre_mask = N_BUCKETS - 1; bg.pre_base = (u32)(g * N_BUCKETS);
      bg.ovf_keys = idx->nw == 1 ? ovf_base<1>(bs) : ovf_base<2>(bs); bg.ovf_cursor = ovf_cur(bs);
      idx->ops->scatter_fine(idx->stream, bg, fs, reinterpret_cast<u32*>(idx->ctr + CTR_PERR), n_keys / (n_mid / N_BUCKETS) + 1);
      idx->launches++;
    }
  }
  CU(cudaGetLastError());                              // a scatter that did not launch must not pass for an empty bucket
  return KMI_OK;
}

// distinct keys after an insert of n_keys more: at most size + n_keys, and (6 sigma) at most 1.05 * HLL + keys the HLL never saw
// (the registers see a 1/HLL_SAMPLE hash-sampled subset: scale back, add 6 sigma of the sampling error)
u64 need_after(const kmi_index* idx, u64 n_keys, u64 size_now, const std::vector<u32>& reg) {
  const double est = hll_estimate(reg.data()) * (double)HLL_SAMPLE;
  const u64 by_hll = (u64)(est * 1.05 + 6.0 * std::sqrt(est * (double)HLL_SAMPLE)) + 4096 + idx->non_hll_upper;
  return std::min(size_now + n_keys, std::max(by_hll, size_now));
}
}  // namespace

// ---- KMI_PARAM_DEFER_INSERT (bulk load): the fine buckets of several rounds / calls are STAGED and inserted together ---------
// A build far larger than one bucket round used to sweep the whole table once per round (the segment insert reads and writes
// every segment): at the 100 Gbp configuration's per-GPU share that was 5-9 sweeps of a 68.7 GB table, all of the insert
// time. In bulk-load mode a round stops after the fine scatter, which APPENDS to fine regions sized for `defer_budget` keys;
// the one segment insert runs when the staging area is full, when the table has to grow, or when anything looks at the table
// (kmi_index::touch, the size readers, kmi_flush). Invariant: staged keys always fit the current table (checked with the HLL
// bound when they are staged), so a flush never grows the table and the regions never outlive the geometry they were cut for.
int kmi::flush_pending(kmi_index* idx) {
  if (!idx->pend_keys) return KMI_OK;
  idx->pend_keys = 0;                                     // (first: insert_overflow below goes through view())
  CU(cudaSetDevice(idx->cfg.device));
  CU(cudaStreamSynchronize(idx->stream));
  harvest_commit_times(idx);
  const int log_cap = __builtin_ctzll(idx->cap);
  const int log_seg = log_cap - idx->pend_log_segs;
  FineSink fs;
  fs.keys = idx->fine_keys.as<u64>(); fs.cursor = idx->fine_cur.as<u32>(); fs.capf = idx->pend_capf; fs.log_f = 0;
  BucketSink sp;                                          // only its spill list is used
  sp.keys = nullptr; sp.cursor = nullptr; sp.capb = 0; sp.hll = nullptr; sp.bases = nullptr;
  sp.ovf_keys = idx->pend_ovf.as<u64>(); sp.ovf_cursor = idx->pend_ovf_cur.as<u32>(); sp.ovf_cap = idx->pend_ovf_cap;
  CU(cudaMemsetAsync(sp.ovf_cursor, 0, sizeof(u32), idx->stream));
  CU(cudaEventRecord(idx->kev[0], idx->stream));
  CU(cudaEventRecord(idx->kev[1], idx->stream));
  const bool fresh_table = idx->virgin;
  const u64 n_regions = (u64)1 << idx->pend_log_segs;
  int e = idx->ops->insert_segments(idx->stream, idx->raw_view(), fs, sp, (u32)log_seg, (u32)n_regions,
                                    reinterpret_cast<u32*>(idx->ctr + CTR_PERR), fresh_table);
  idx->virgin = false;
  if (e) return fail(KMI_ERR_CUDA, "segment kernel configuration: %s", cudaGetErrorString((cudaError_t)e));
  CU(cudaEventRecord(idx->kev[2], idx->stream));
  idx->ops->insert_overflow(idx->stream, idx->raw_view(), sp);
  CU(cudaGetLastError());
  CU(cudaEventRecord(idx->kev[3], idx->stream));
  idx->kev_pending = true;
  idx->launches += 2;
  CU(cudaMemsetAsync(idx->fine_cur.p, 0, n_regions * sizeof(u32), idx->stream));
  idx->size_exact = ~0ULL;
  idx->flushes++;
  return KMI_OK;
}

namespace {
// One round of a bulk-load build: size the table for everything seen so far, fine-scatter the round's coarse buckets BEHIND
// what is staged, insert nothing. *staged = false: the segment path cannot run for this table / sink and the caller commits
// the round the regular way (nothing of it has been touched).
int bucket_stage(kmi_index* idx, const BucketSink& bs, u64 n_keys, u64 size_now, const std::vector<u32>& reg, bool* staged) {
  *staged = false;
  u64 need = need_after(idx, idx->pend_keys + n_keys, size_now, reg);
  if (!idx->slots || (double)(need + idx->tombstones) > idx->cap * MAX_LOAD) {
    if (idx->pend_keys) {                                 // what is staged fits the table as it is: insert it, then grow
      OK(flush_pending(idx));
      OK(read_ctr(idx, CTR_SIZE, &size_now));
      need = need_after(idx, n_keys, size_now, reg);
    }
    if (!idx->slots || (double)(need + idx->tombstones) > idx->cap * MAX_LOAD) {
      u64 want = pow2_at_least((u64)(need / MAX_LOAD) + 1);
      if (want > idx->cap || idx->tombstones) OK(table_grow(idx, std::max(want, idx->cap)));
    }
  }
  const SegPlan pl = seg_plan(idx, bs);
  if (!pl.ok || pl.log_f == 0) return KMI_OK;             // regular commit (it flushes what is staged first)
  // staging area: laid out for this geometry; its size is the budget, clamped by what the device has left
  if (idx->pend_log_segs != pl.log_segs || idx->pend_cap != idx->cap || !idx->fine_keys.p || idx->pend_req != idx->defer_budget) {
    OK(flush_pending(idx));
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    free_b += idx->fine_keys.cap;                         // (the old staging area is released by ensure)
    const u64 by_mem = (u64)((double)free_b * 0.6) / (sizeof(u64) * idx->nw);
    u64 budget = std::min<u64>(idx->defer_budget, by_mem);
    if (budget < 2 * n_keys) return KMI_OK;               // no room to hold more than this round: nothing to gain
    const double mean = (double)budget / (double)pl.n_regions;
    const u64 capf = (u64)(mean + 8.0 * std::sqrt(mean) + 32.0);
    if (capf >= 0xffffffffULL) return KMI_OK;
    OK(idx->fine_keys.ensure(pl.n_regions * capf * sizeof(u64) * idx->nw));
    OK(idx->fine_cur.ensure(pl.n_regions * sizeof(u32)));
    OK(idx->pend_cur_save.ensure(pl.n_regions * sizeof(u32)));
    const u64 spill = budget / 16 + 4096;
    OK(idx->pend_ovf.ensure(spill * sizeof(u64) * idx->nw));
    OK(idx->pend_ovf_cur.ensure(sizeof(u32) * 2));
    idx->pend_ovf_cap = (u32)std::min<u64>(spill, 0xffffffffULL);
    CU(cudaMemsetAsync(idx->fine_cur.p, 0, pl.n_regions * sizeof(u32), idx->stream));
    idx->pend_log_segs = pl.log_segs; idx->pend_cap = idx->cap; idx->pend_capf = (u32)capf; idx->pend_budget = budget;
    idx->pend_req = idx->defer_budget;
  }
  if (idx->pend_keys + n_keys > idx->pend_budget) OK(flush_pending(idx));   // staging area full: one sweep, then go on
  FineSink fs;
  fs.keys = idx->fine_keys.as<u64>(); fs.cursor = idx->fine_cur.as<u32>(); fs.capf = idx->pend_capf; fs.log_f = pl.log_f;
  if (idx->kev_pending) { CU(cudaStreamSynchronize(idx->stream)); harvest_commit_times(idx); }   // (a flush above left its events)
  // the cursors as they are now: a scatter that exhausts the overflow region is rolled back to them
  CU(cudaMemcpyAsync(idx->pend_cur_save.p, idx->fine_cur.p, pl.n_regions * sizeof(u32), cudaMemcpyDeviceToDevice, idx->stream));
  CU(cudaEventRecord(idx->kev[0], idx->stream));
  OK(fine_scatter_launch(idx, bs, pl, fs, n_keys));
  CU(cudaEventRecord(idx->kev[1], idx->stream));
  CU(cudaEventRecord(idx->kev[2], idx->stream));
  CU(cudaEventRecord(idx->kev[3], idx->stream));
  u64 perr = 0;
  u32 n_ovf = 0;
  CU(cudaMemcpyAsync(&n_ovf, ovf_cur(bs), sizeof(u32), cudaMemcpyDeviceToHost, idx->stream));
  OK(read_ctr(idx, CTR_PERR, &perr));                     // (synchronises)
  idx->kev_pending = true;
  harvest_commit_times(idx);
  if (perr & PERR_OVERFLOW) {
    // skewed fine buckets exhausted the round's overflow region: take this round's appends back, settle what was staged
    // before it, and let the caller commit the round on the regular path (which ends at the L2-atomic insert)
    OK(zero_ctr(idx, CTR_PERR));
    CU(cudaMemcpyAsync(ovf_cur(bs), &idx->h_ovf_cursor, sizeof(u32), cudaMemcpyHostToDevice, idx->stream));
    CU(cudaMemcpyAsync(idx->fine_cur.p, idx->pend_cur_save.p, pl.n_regions * sizeof(u32), cudaMemcpyDeviceToDevice, idx->stream));
    CU(cudaStreamSynchronize(idx->stream));
    return KMI_OK;
  }
  if (n_ovf) {
    // keys that did not fit their coarse / fine region sit in the round's own overflow region, which the next round reuses:
    // they go into the table now, with global atomics (the table then is no longer untouched: the flush will read it)
    if (idx->virgin && idx->slots) { idx->ops->table_clear(idx->stream, idx->slots, idx->cap); idx->launches++; idx->virgin = false; }
    idx->ops->insert_overflow(idx->stream, idx->raw_view(), bs);
    CU(cudaMemsetAsync(ovf_cur(bs), 0, sizeof(u32), idx->stream));
    idx->launches++;
  }
  idx->pend_keys += n_keys;
  idx->size_upper = need;
  idx->size_exact = ~0ULL;
  *staged = true;
  return KMI_OK;
}
}  // namespace

// Size the table for what the buckets hold, then insert them segment by segment. n_keys = keys
// scattered this round; size_now = live keys in the table before the insert.
int kmi::bucket_commit(kmi_index* idx, const BucketSink& bs, u64 n_keys, u64 size_now, const std::vector<u32>& reg, bool* not_inserted) {
  if (not_inserted) *not_inserted = false;
  if (n_keys == 0) return KMI_OK;
  if (idx->defer_budget) {
    bool staged = false;
    OK(bucket_stage(idx, bs, n_keys, size_now, reg, &staged));
    if (staged) return KMI_OK;
  }
  if (idx->pend_keys) {                                   // the regular path uses the staging buffers and may grow the table
    OK(flush_pending(idx));
    OK(read_ctr(idx, CTR_SIZE, &size_now));
  }
  const u64 need = need_after(idx, n_keys, size_now, reg);
  if (!idx->slots || (double)(need + idx->tombstones) > idx->cap * MAX_LOAD) {
    u64 want = pow2_at_least((u64)(need / MAX_LOAD) + 1);
    if (want > idx->cap || idx->tombstones) OK(table_grow(idx, std::max(want, idx->cap)));   // same capacity: purge tombstones
  }
  idx->size_upper = need;
  idx->size_exact = ~0ULL;  // stale until read back
  const SegPlan pl = seg_plan(idx, bs);
  if (pl.ok) {
    const u32 log_f = pl.log_f;
    const u64 n_regions = pl.n_regions;
    const int log_seg = pl.log_seg;
    FineSink fs;
    fs.log_f = log_f;
    bool fine_ok = true;
    CU(cudaEventRecord(idx->kev[0], idx->stream));
    if (log_f == 0) {
      fs.keys = bs.keys; fs.cursor = bs.cursor; fs.capf = bs.capb;
    } else {
      const double mean = (double)n_keys / (double)n_regions;
      const u64 capf = (u64)(mean + 8.0 * std::sqrt(mean) + 32.0);
      OK(idx->fine_keys.ensure(n_regions * capf * sizeof(u64) * idx->nw));
      OK(idx->fine_cur.ensure(n_regions * sizeof(u32)));
      idx->pend_log_segs = -1;                             // (the staging layout, if there was one, is gone)
      CU(cudaMemsetAsync(idx->fine_cur.p, 0, n_regions * sizeof(u32), idx->stream));
      fs.keys = idx->fine_keys.as<u64>(); fs.cursor = idx->fine_cur.as<u32>(); fs.capf = (u32)capf;
      OK(fine_scatter_launch(idx, bs, pl, fs, n_keys));
      u64 perr = 0;
      OK(read_ctr(idx, CTR_PERR, &perr));
      if (perr & PERR_OVERFLOW) {
        // skewed fine buckets exhausted the overflow region: undo its tail and take the L2 path below
        OK(zero_ctr(idx, CTR_PERR));
        CU(cudaMemcpyAsync(ovf_cur(bs), &idx->h_ovf_cursor, sizeof(u32), cudaMemcpyHostToDevice, idx->stream));
        fine_ok = false;
      }
    }
    if (fine_ok) {
      CU(cudaEventRecord(idx->kev[1], idx->stream));
      const bool fresh_table = idx->virgin;             // every segment is initialised in shared memory and written
      int e = idx->ops->insert_segments(idx->stream, idx->raw_view(), fs, bs, (u32)log_seg, (u32)n_regions,
                                        reinterpret_cast<u32*>(idx->ctr + CTR_PERR), fresh_table);
      idx->virgin = false;
      if (e) return fail(KMI_ERR_CUDA, "segment kernel configuration: %s", cudaGetErrorString((cudaError_t)e));
      CU(cudaEventRecord(idx->kev[2], idx->stream));
      idx->ops->insert_overflow(idx->stream, idx->view(), bs);
      CU(cudaGetLastError());
      CU(cudaEventRecord(idx->kev[3], idx->stream));
      idx->kev_pending = true;
      idx->launches += 2;
      return KMI_OK;
    }
  }
  if (bs.sub > 1) {   // the L2-atomic kernel walks one region per prefix: the caller inserts the regions as key streams
    if (!not_inserted) return fail(KMI_ERR_ARG, "receive buffer needs the segment path");
    *not_inserted = true;
    return KMI_OK;
  }
  CU(cudaEventRecord(idx->kev[0], idx->stream));
  CU(cudaEventRecord(idx->kev[1], idx->stream));
  idx->ops->insert_buckets(idx->stream, idx->view(), bs, n_keys);
  CU(cudaEventRecord(idx->kev[2], idx->stream));
  CU(cudaEventRecord(idx->kev[3], idx->stream));
  idx->kev_pending = true;
  idx->launches++;
  return KMI_OK;
}

static int insert_keys_chunked(kmi_index* idx, const u64* d_keys, size_t n, bool canon, const u64* d_counts = nullptr);

// several key streams -> the same buckets -> table (one table sweep for all of them); total <= BUCKET_MAX_KEYS
int kmi::insert_streams_bucketed(kmi_index* idx, const u64* const* d_keys, const u64* counts, int n_streams, bool canon) {
  u64 total = 0;
  for (int i = 0; i < n_streams; ++i) total += counts[i];
  if (total == 0) return KMI_OK;
  std::vector<u32> reg;
  BucketSink bs;
  OK(bucket_prepare(idx, total, &bs));
  OK(zero_ctr(idx, CTR_PERR));
  for (int i = 0; i < n_streams; ++i) {
    if (!counts[i]) continue;
    idx->ops->scatter_keys(idx->stream, d_keys[i], counts[i], bs, reinterpret_cast<u32*>(idx->ctr + CTR_PERR), canon);
    idx->launches++;
  }
  u64 host[CTR_N];
  OK(bucket_readback(idx, bs, host, reg));
  if (host[CTR_PERR] & PERR_OVERFLOW) {
    // extreme key skew: nothing of this round is in the table yet, redo it with direct atomics
    OK(zero_ctr(idx, CTR_PERR));
    for (int i = 0; i < n_streams; ++i) OK(insert_keys_chunked(idx, d_keys[i], counts[i], canon));
    return KMI_OK;
  }
  return bucket_commit(idx, bs, total, host[CTR_SIZE], reg);
}

// key stream -> buckets -> table, in rounds of at most BUCKET_MAX_KEYS keys
static int insert_keys_bucketed(kmi_index* idx, const u64* d_keys, size_t n, bool canon) {
  size_t done = 0;
  while (done < n) {
    const u64 m = std::min<u64>(n - done, BUCKET_MAX_KEYS);
    const u64* p = d_keys + done * idx->nw;
    OK(insert_streams_bucketed(idx, &p, &m, 1, canon));
    done += m;
  }
  return KMI_OK;
}

// ---- a2: where the first whole record of a byte range starts (host side, as in the reference: four lines) -------
// FASTQParser::find_first_record, src/io/fastq_loader.hpp:269-357 / :696-873. `bytes` holds file bytes
// [mem_offset, mem_offset + n); the first characters of four consecutive non-empty lines at or after search_begin
// are matched against @?+? / ?@?+ / +?@? / ?+?@. This is what cuts a file into the record-aligned blocks that
// kmi_build_fastq / kmi_parse_fastq_dev take (one block per GPU, or per pipelined chunk).
static inline bool host_is_eol(uint8_t c) { return c == '\n' || c == '\r'; }
// a6 (host): how far past a block's end its k-mers can reach — K-1 bytes that are not EOL, counted as
// KmerFileHelper::read_block does for a FASTA sequence that runs into the overlap (kmer_file_helper.hpp:153-173)
extern "C" int kmi_fasta_halo_end(const uint8_t* bytes, uint64_t mem_offset, size_t n, uint64_t valid_end, int k, uint64_t* halo_end) {
  if ((!bytes && n) || !halo_end) return fail(KMI_ERR_ARG, "null argument");
  if (k < 1 || valid_end < mem_offset || valid_end > mem_offset + n) return fail(KMI_ERR_ARG, "valid_end outside the bytes given");
  uint64_t p = valid_end - mem_offset;
  int need = k - 1;
  while (p < n && need > 0) {
    const uint8_t c = bytes[p++];
    if (c != '\n' && c != '\r') --need;
  }
  *halo_end = mem_offset + p;
  return KMI_OK;
}

extern "C" int kmi_fastq_find_first_record(const uint8_t* bytes, uint64_t mem_offset, size_t n, uint64_t search_begin,
                                           uint64_t search_end, uint64_t parent_begin, uint64_t parent_end, uint64_t* offset) {
  if ((!bytes && n) || !offset) return fail(KMI_ERR_ARG, "null argument");
  if (parent_begin > mem_offset || parent_end < mem_offset + n) return fail(KMI_ERR_ARG, "parent range does not contain the bytes in memory");
  const uint64_t ts = std::max(search_begin, mem_offset);
  const uint64_t te = std::max(ts, std::min<uint64_t>(search_end, mem_offset + n));
  auto at = [&](uint64_t i) { return bytes[i - mem_offset]; };
  uint64_t i = ts;
  *offset = te;
  if (i == te) return KMI_OK;
  if (ts > parent_begin && !host_is_eol(at(i))) while (i < te && !host_is_eol(at(i))) ++i;   // not a line start: drop the partial line
  while (i < te && host_is_eol(at(i))) ++i;
  if (i == te) return KMI_OK;
  uint8_t first[4] = {at(i), 0, 0, 0};
  uint64_t off[4] = {i, te, te, te};
  for (int j = 1; j < 4; ++j) {
    while (i < te && !host_is_eol(at(i))) ++i;
    while (i < te && host_is_eol(at(i))) ++i;
    off[j] = i;
    if (i != te) first[j] = at(i);
  }
  if (i == te) return KMI_OK;                             // fewer than four lines: no complete record in range
  if (first[0] == '@' && first[2] == '+') { *offset = off[0]; return KMI_OK; }
  if (first[1] == '@' && first[3] == '+') { *offset = off[1]; return KMI_OK; }
  if (first[0] == '+' && first[2] == '@') { *offset = off[2]; return KMI_OK; }
  if (first[1] == '+' && first[3] == '@') { *offset = off[3]; return KMI_OK; }
  if (te == parent_end) return KMI_OK;
  return fail(KMI_ERR_FORMAT, "bytes [%llu, %llu) do not contain valid FASTQ markers (first chars %c %c %c %c)",
              (unsigned long long)ts, (unsigned long long)te, first[0], first[1], first[2], first[3]);
}

// ---- record boundaries (pass A + scan), shared by parse and fused build ------------------------
int kmi::fastq_scan_alloc(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64* n_tiles_out) {
  if ((reinterpret_cast<uintptr_t>(d_bytes) & 15) != 0) return fail(KMI_ERR_ARG, "device text buffer must be 16-byte aligned");
  u64 n_tiles = (n + TILE_BYTES - 1) / TILE_BYTES;
  *n_tiles_out = n_tiles;
  if (n_tiles == 0) return KMI_OK;
  if (n_tiles > 0x7fffffffULL / TILE_CHUNKS) return fail(KMI_ERR_ARG, "block of %zu bytes too large for one call (max 1 TiB)", n);
  OK(idx->tile_lines.ensure(sizeof(u32) * n_tiles * TILE_CHUNKS));                       // line starts before every 512-byte chunk
  OK(idx->block_sums.ensure(sizeof(u32) * ((n_tiles * TILE_CHUNKS + 1023) / 1024)));
  OK(zero_ctr(idx, CTR_CARRY));
  return KMI_OK;
}
// line starts of tiles [t0, t1) (t0 a multiple of 1024) -> per-chunk prefixes; the running total of
// earlier ranges is carried in CTR_CARRY so that ranges can be scanned as their bytes arrive
int kmi::fastq_scan_range(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64 t0, u64 t1) {
  if (t1 <= t0) return KMI_OK;
  const u64 nt = t1 - t0, nc = nt * TILE_CHUNKS, c0 = t0 * TILE_CHUNKS, nb = (nc + 1023) / 1024;
  k_fastq_count_lines<<<(unsigned)std::min<u64>((nt + 7) / 8, 148 * 16), COUNT_THREADS, 0, idx->stream>>>(d_bytes, n, idx->tile_lines.as<u32>(), (u32)t0, (u32)nt);
  k_scan_l1<<<(unsigned)nb, 1024, 0, idx->stream>>>(idx->tile_lines.as<u32>() + c0, nc, idx->block_sums.as<u32>() + c0 / 1024);
  k_scan_tiles<<<1, 1024, 0, idx->stream>>>(idx->block_sums.as<u32>() + c0 / 1024, nb, reinterpret_cast<u32*>(idx->ctr + CTR_CARRY));
  idx->launches += 3;
  return KMI_OK;
}
int kmi::fastq_scan(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64* n_tiles_out) {
  OK(fastq_scan_alloc(idx, d_bytes, n, n_tiles_out));
  return fastq_scan_range(idx, d_bytes, n, 0, *n_tiles_out);
}
// record start that is open at the first byte of every tile (FASTQ position ids); after fastq_scan
static int fastq_rec_carry(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64 n_tiles) {
  if (n_tiles == 0) return KMI_OK;
  OK(idx->tile_rec.ensure(sizeof(u64) * n_tiles));
  k_fastq_rec_carry<<<(unsigned)std::min<u64>((n_tiles + 7) / 8, 148 * 16), COUNT_THREADS, 0, idx->stream>>>(
      d_bytes, n, idx->tile_lines.as<u32>(), idx->block_sums.as<u32>(), idx->tile_rec.as<u64>(), (u32)n_tiles);
  idx->launches++;
  return KMI_OK;
}
int kmi::parse_error(u64 perr) {
  if (perr & PERR_NO_AT) return fail(KMI_ERR_FORMAT, "malformed FASTQ: missing @ on first line of a record");
  if (perr & PERR_NO_PLUS) return fail(KMI_ERR_FORMAT, "malformed FASTQ: missing + on third line of a record");
  if (perr & PERR_ID_RANGE) return fail(KMI_ERR_ARG, "FASTQ record longer than 64 KiB: ShortSequenceKmerId cannot hold the k-mer's offset in its record");
  if (perr & PERR_OVERFLOW) return fail(KMI_ERR_CAPACITY, "k-mer output buffer too small");
  if (perr & PERR_HALO) return fail(KMI_ERR_FORMAT, "FASTA line structure too sparse for the tile halo");
  return KMI_OK;
}

extern "C" int kmi_parse_fastq_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset,
                                   uint64_t* d_keys, uint64_t* d_ids, size_t cap, uint64_t* n_kmers) {
  if (!idx || (!d_bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  u64 n_tiles = 0;
  OK(fastq_scan(idx, d_bytes, n, &n_tiles));
  if (d_ids) OK(fastq_rec_carry(idx, d_bytes, n, n_tiles));
  OK(zero_ctr(idx, CTR_NKMERS, 2));  // NKMERS, CURSOR
  OK(zero_ctr(idx, CTR_PERR));
  KeySink sink;
  sink.keys = reinterpret_cast<u64*>(d_keys); sink.ids = reinterpret_cast<u64*>(d_ids); sink.cap = cap; sink.cursor = idx->ctr + CTR_CURSOR;
  sink.file_offset = file_offset; sink.tile_rec = d_ids ? idx->tile_rec.as<u64>() : nullptr;
  TableView none; none.slots = nullptr; none.mask = 0; none.shift = 63; none.size = nullptr; none.error = nullptr; none.saturate = 0;
  idx->ops->fastq_kmers(idx->stream, 0, d_bytes, n, n_tiles, idx->tile_lines.as<u32>(), idx->block_sums.as<u32>(), sink, none, BucketSink(),
                        idx->ctr + CTR_NKMERS, reinterpret_cast<u32*>(idx->ctr + CTR_PERR), idx->cfg.canonical != 0,
                        idx->cfg.nsplit != 0, 0, 1);
  idx->launches++;
  u64 host[CTR_N];
  CU(cudaMemcpyAsync(host, idx->ctr, sizeof(u64) * CTR_N, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  if (n_kmers) *n_kmers = host[CTR_NKMERS];
  return parse_error(host[CTR_PERR]);
}

// S1-S5 with direct atomics: text -> table, no key stream in HBM (tiles [tile, tile_end))
static int build_fastq_direct(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64 tile, u64 tile_end) {
  KeySink nosink; nosink.keys = nullptr; nosink.ids = nullptr; nosink.cap = 0; nosink.cursor = nullptr;
  while (tile < tile_end) {
    // a FASTQ record spends at most half of its bytes on bases
    u64 bytes_left = std::min<u64>(n, tile_end * TILE_BYTES) - tile * TILE_BYTES;
    u64 grant = 0;
    OK(table_room(idx, bytes_left / 2 + 1, &grant));
    u64 tiles = std::max<u64>(1, (grant * 2) / TILE_BYTES);
    tiles = std::min(tiles, tile_end - tile);
    idx->ops->fastq_kmers(idx->stream, 1, d_bytes, n, tiles, idx->tile_lines.as<u32>(), idx->block_sums.as<u32>(), nosink,
                          idx->view(), BucketSink(), idx->ctr + CTR_NKMERS, reinterpret_cast<u32*>(idx->ctr + CTR_PERR),
                          idx->cfg.canonical != 0, idx->cfg.nsplit != 0, tile, 1);
    idx->launches++;
    idx->non_hll_upper += grant;
    tile += tiles;
  }
  return KMI_OK;
}

// S1-S5 through the hash-prefix buckets, in rounds: [H2D copy] -> line scan -> parse + scatter ->
// (HLL read-back, table sizing) -> fine scatter -> segment insert. With `host_src` the round's bytes
// are copied into d_bytes on the copy stream while the previous round computes, so the PCIe transfer
// hides the whole build except the last round. A round's last tile needs 64 bytes of the next round
// (k-mer halo), so it is parsed with the next round.
static int build_fastq_bucketed(kmi_index* idx, uint8_t* d_bytes, size_t n, u64 n_tiles, const HostSrc& host_src) {
  if (n_tiles == 0) return KMI_OK;
  const u64 MAX_TILES = (2 * BUCKET_MAX_KEYS) / TILE_BYTES;
  u64 per_round;
  if (host_src) {
    // 12 rounds: a round costs ~2.3 ms of fixed work (one sweep of the table) against 3.2 ms of PCIe time, and only the last
    // round is exposed (measured at 1 Gbp: 41.5 ms with 8 rounds, 41.0 with 12, 45.3 with 16; the copy alone is 38.1 ms)
    per_round = std::min<u64>(std::max<u64>(n_tiles / 12, (64u << 20) / TILE_BYTES), MAX_TILES);
  } else {
    const u64 n_rounds = (n_tiles + MAX_TILES - 1) / MAX_TILES;
    per_round = (n_tiles + n_rounds - 1) / std::max<u64>(n_rounds, 1);
  }
  per_round = (per_round + 1023) / 1024 * 1024;  // the two-level line scan works in blocks of 1024 tiles
  const u64 n_rounds = (n_tiles + per_round - 1) / per_round;
  HostFeed feed;   // (its destructor never returns while a copy from the caller's memory is in flight)
  if (host_src) {
    std::vector<u64> ends(n_rounds);
    for (u64 r = 0; r < n_rounds; ++r) ends[r] = std::min<u64>(n, (r + 1) * per_round * TILE_BYTES);
    OK(feed.start(idx, d_bytes, host_src, ends, idx->stream));
  }
  KeySink nosink; nosink.keys = nullptr; nosink.ids = nullptr; nosink.cap = 0; nosink.cursor = nullptr;
  std::vector<u32> reg;
  u64 kmers_before = 0;
  for (u64 r = 0; r < n_rounds; ++r) {
    const u64 t0 = r * per_round, t1 = std::min(n_tiles, t0 + per_round);
    const bool last = r + 1 == n_rounds;
    // (a tile's halo continues at the next tile's line count: resident text is counted whole before the first round)
    if (host_src) { OK(feed.wait(r, idx->stream)); OK(fastq_scan_range(idx, d_bytes, n, t0, t1)); }
    else if (r == 0) OK(fastq_scan_range(idx, d_bytes, n, 0, n_tiles));
    const u64 p0 = (host_src && r > 0) ? t0 - 1 : t0;
    const u64 p1 = (host_src && !last) ? t1 - 1 : t1;
    const u64 tiles = p1 - p0;
    BucketSink bs;
    OK(bucket_prepare(idx, tiles * TILE_BYTES / 2 + 1, &bs));
    CU(cudaEventRecord(idx->ev[4], idx->stream));
    idx->ops->fastq_kmers(idx->stream, 2, d_bytes, n, tiles, idx->tile_lines.as<u32>(), idx->block_sums.as<u32>(), nosink,
                          idx->raw_view() /* unused by sink 2: no lazy clear */, bs, idx->ctr + CTR_NKMERS, reinterpret_cast<u32*>(idx->ctr + CTR_PERR),
                          idx->cfg.canonical != 0, idx->cfg.nsplit != 0, p0, 1);
    idx->launches++;
    CU(cudaEventRecord(idx->ev[5], idx->stream));
    u64 host[CTR_N];
    OK(bucket_readback(idx, bs, host, reg));
    { float ms = 0; cudaEventElapsedTime(&ms, idx->ev[4], idx->ev[5]); idx->scatter_ms += ms; }
    if (host[CTR_PERR] & (PERR_NO_AT | PERR_NO_PLUS)) return KMI_OK;  // reported by the caller; this round is not inserted
    if (host[CTR_PERR] & PERR_OVERFLOW) {
      // extreme key skew: redo this round with direct atomics (nothing of it is in the table yet)
      OK(zero_ctr(idx, CTR_PERR));
      CU(cudaMemcpyAsync(idx->ctr + CTR_NKMERS, &kmers_before, sizeof(u64), cudaMemcpyHostToDevice, idx->stream));
      CU(cudaStreamSynchronize(idx->stream));
      OK(build_fastq_direct(idx, d_bytes, n, p0, p1));
      OK(read_ctr(idx, CTR_NKMERS, &kmers_before));
      continue;
    }
    OK(bucket_commit(idx, bs, host[CTR_NKMERS] - kmers_before, host[CTR_SIZE], reg));
    kmers_before = host[CTR_NKMERS];
  }
  return KMI_OK;
}


// PositionIndex from FASTQ: (k-mer, ShortSequenceKmerId) tuples, chunk by chunk through the tuple log + table
// (KmerPositionTupleParser, kmer_parser.hpp:465-567; densehash_multimap::insert)
static int build_fastq_positions(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64 file_offset, uint64_t* n_kmers) {
  if (n_kmers) *n_kmers = 0;
  u64 n_tiles = 0;
  OK(fastq_scan(idx, d_bytes, n, &n_tiles));
  if (n_tiles == 0) return KMI_OK;
  OK(fastq_rec_carry(idx, d_bytes, n, n_tiles));
  OK(zero_ctr(idx, CTR_NKMERS));
  OK(zero_ctr(idx, CTR_PERR));
  const u64 CHUNK_TILES = 65536;  // 256 MiB of text -> <= 128 Mi k-mers per chunk (every chunk costs one sweep of the table)
  const u64 cap = std::min<u64>(n_tiles, CHUNK_TILES) * TILE_BYTES / 2 + 1;
  DevBuf& keys = idx->parse_keys; DevBuf& ids = idx->parse_ids;
  int r = keys.ensure(cap * sizeof(u64) * idx->nw);
  if (r == KMI_OK) r = ids.ensure(cap * sizeof(u64));
  TableView none; none.slots = nullptr; none.mask = 0; none.shift = 63; none.size = nullptr; none.error = nullptr; none.saturate = 0;
  u64 total = 0, perr = 0;
  for (u64 tile = 0; r == KMI_OK && tile < n_tiles; tile += CHUNK_TILES) {
    const u64 tiles = std::min<u64>(CHUNK_TILES, n_tiles - tile);
    KeySink sink;
    sink.keys = keys.as<u64>(); sink.ids = ids.as<u64>(); sink.cap = cap; sink.cursor = idx->ctr + CTR_CURSOR;
    sink.file_offset = file_offset; sink.tile_rec = idx->tile_rec.as<u64>();
    r = zero_ctr(idx, CTR_CURSOR);
    if (r != KMI_OK) break;
    idx->ops->fastq_kmers(idx->stream, 0, d_bytes, n, tiles, idx->tile_lines.as<u32>(), idx->block_sums.as<u32>(), sink, none,
                          BucketSink(), idx->ctr + CTR_NKMERS, reinterpret_cast<u32*>(idx->ctr + CTR_PERR),
                          idx->cfg.canonical != 0, idx->cfg.nsplit != 0, tile, 1);
    idx->launches++;
    u64 produced = 0;
    r = read_ctr(idx, CTR_CURSOR, &produced);
    if (r == KMI_OK) r = read_ctr(idx, CTR_PERR, &perr);
    if (r != KMI_OK || perr) break;
    const int32_t canon = idx->cfg.canonical;   // keys are canonical already
    idx->cfg.canonical = 0;
    r = kmi_insert_dev(idx, keys.as<uint64_t>(), ids.as<uint64_t>(), produced);
    idx->cfg.canonical = canon;
    cudaStreamSynchronize(idx->stream);
    total += produced;
  }
  if (r != KMI_OK) return r;
  if (perr) return parse_error(perr);
  if (n_kmers) *n_kmers = total;
  return check_table_error(idx);
}

static int build_fastq_impl(kmi_index* idx, uint8_t* d_bytes, size_t n, HostSrc host_src, uint64_t* n_kmers,
                            u64 file_offset = 0) {
  if (idx->cfg.mode != KMI_MODE_COUNT) {
    if (host_src) OK(copy_to_device(idx, d_bytes, host_src, n));
    return build_fastq_positions(idx, d_bytes, n, file_offset, n_kmers);
  }
  u64 n_tiles = 0;
  const bool bucketed = idx->build_path == 2 || (idx->build_path == 0 && n >= AUTO_BUCKET_BYTES);
  CU(cudaEventRecord(idx->ev[0], idx->stream));
  if (host_src && !bucketed) {
    OK(copy_to_device(idx, d_bytes, host_src, n));
    host_src = HostSrc();
  }
  OK(fastq_scan_alloc(idx, d_bytes, n, &n_tiles));
  if (!bucketed) OK(fastq_scan_range(idx, d_bytes, n, 0, n_tiles));
  CU(cudaEventRecord(idx->ev[1], idx->stream));
  OK(zero_ctr(idx, CTR_NKMERS));
  OK(zero_ctr(idx, CTR_PERR));
  idx->scatter_ms = 0;
  for (int i = 0; i < 5; ++i) idx->kernel_ms[i] = 0;
  idx->kev_pending = false;
  if (bucketed) OK(build_fastq_bucketed(idx, d_bytes, n, n_tiles, host_src));
  else OK(build_fastq_direct(idx, d_bytes, n, 0, n_tiles));
  CU(cudaEventRecord(idx->ev[2], idx->stream));
  u64 host[CTR_N];
  CU(cudaMemcpyAsync(host, idx->ctr, sizeof(u64) * CTR_N, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  cudaEventElapsedTime(&idx->stage_ms[0], idx->ev[0], idx->ev[1]);
  cudaEventElapsedTime(&idx->stage_ms[1], idx->ev[1], idx->ev[2]);
  idx->stage_ms[2] = 0;
  harvest_commit_times(idx);
  idx->kernel_ms[0] = idx->stage_ms[0] + idx->stage_ms[1];
  idx->kernel_ms[1] = bucketed ? idx->scatter_ms : idx->stage_ms[1];
  if (bucketed) {  // [0] nothing (the scan is per round), [1] parse + scatter, [2] scan + table sizing + insert
    idx->stage_ms[2] = idx->stage_ms[1] - idx->scatter_ms;
    idx->stage_ms[1] = idx->scatter_ms;
  }
  if (!idx->pend_keys) idx->size_exact = idx->size_upper = host[CTR_SIZE];   // (staged keys are not in the counter yet)
  if (n_kmers) *n_kmers = host[CTR_NKMERS];
  OK(table_error(host[CTR_ERR]));
  return parse_error(host[CTR_PERR]);
}

extern "C" int kmi_build_fastq_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  if (!idx || (!d_bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  return build_fastq_impl(idx, const_cast<uint8_t*>(d_bytes), n, nullptr, n_kmers, file_offset);
}

// host bytes: the H2D copy is pipelined with the build (bucketed path), see build_fastq_bucketed
extern "C" int kmi_build_fastq(kmi_index* idx, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  if (!idx || (!bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  OK(idx->stage_in.ensure(n + 64));
  return build_fastq_impl(idx, idx->stage_in.as<uint8_t>(), n, bytes, n_kmers, file_offset);
}

// ---- insert --------------------------------------------------------------------------------------------
static int log_append(kmi_index* idx, const u64* d_keys, const u64* d_vals, size_t n) {
  size_t kb = sizeof(u64) * idx->nw;
  if ((idx->log_n + n) * kb > idx->log_keys.cap) {
    // grow geometrically, preserving contents
    DevBuf nk, nv;
    size_t want = std::max<size_t>((idx->log_n + n) * 2, 1024);
    OK(nk.ensure(want * kb));
    OK(nv.ensure(want * sizeof(u64)));
    if (idx->log_n) {
      CU(cudaMemcpyAsync(nk.p, idx->log_keys.p, idx->log_n * kb, cudaMemcpyDeviceToDevice, idx->stream));
      CU(cudaMemcpyAsync(nv.p, idx->log_ids.p, idx->log_n * sizeof(u64), cudaMemcpyDeviceToDevice, idx->stream));
      CU(cudaStreamSynchronize(idx->stream));
    }
    idx->log_keys.release(); idx->log_ids.release();
    idx->log_keys = nk; idx->log_ids = nv;
  }
  u64* dst = idx->log_keys.as<u64>() + idx->log_n * idx->nw;
  if (idx->cfg.canonical) { idx->ops->canonicalize(idx->stream, d_keys, dst, n); idx->launches++; }
  else CU(cudaMemcpyAsync(dst, d_keys, n * kb, cudaMemcpyDeviceToDevice, idx->stream));
  CU(cudaMemcpyAsync(idx->log_ids.as<u64>() + idx->log_n, d_vals, n * sizeof(u64), cudaMemcpyDeviceToDevice, idx->stream));
  idx->log_n += n;
  return KMI_OK;
}

static int insert_keys_chunked(kmi_index* idx, const u64* d_keys, size_t n, bool canon, const u64* d_counts) {
  size_t done = 0;
  while (done < n) {
    u64 grant = 0;
    OK(table_room(idx, n - done, &grant));
    idx->ops->insert_keys(idx->stream, idx->view(), d_keys + done * idx->nw, grant, canon, d_counts ? d_counts + done : nullptr);
    idx->launches++;
    idx->non_hll_upper += grant;
    done += grant;
  }
  return KMI_OK;
}

extern "C" int kmi_insert_dev(kmi_index* idx, const uint64_t* d_keys_, const uint64_t* d_vals_, size_t n) {
  if (!idx || (!d_keys_ && n)) return fail(KMI_ERR_ARG, "null argument");
  const u64* d_keys = reinterpret_cast<const u64*>(d_keys_);
  const u64* d_vals = reinterpret_cast<const u64*>(d_vals_);
  CU(cudaSetDevice(idx->cfg.device));
  if (n == 0) return KMI_OK;
  if (idx->nw == 2 && (reinterpret_cast<uintptr_t>(d_keys) & 15)) return fail(KMI_ERR_ARG, "2-word keys must be 16-byte aligned");
  if (idx->cfg.mode == KMI_MODE_POSITION) {
    if (!d_vals) return fail(KMI_ERR_ARG, "POSITION insert needs ids");
    OK(log_append(idx, d_keys, d_vals, n));
    // multiplicities go into the table now (keys in the log are already canonical), large batches through the buckets
    const u64* logged = idx->log_keys.as<u64>() + (idx->log_n - n) * idx->nw;
    if (idx->build_path == 2 || (idx->build_path == 0 && n >= AUTO_BUCKET_KEYS)) OK(insert_keys_bucketed(idx, logged, n, false));
    else OK(insert_keys_chunked(idx, logged, n, false));
    idx->csr_dirty = true;
  } else if (d_vals) {
    // (key, count) tuples: reduction insert, count added to the entry (direct atomics; the bucketed paths count keys)
    OK(insert_keys_chunked(idx, d_keys, n, idx->cfg.canonical != 0, d_vals));
  } else if (idx->build_path == 2 || (idx->build_path == 0 && n >= AUTO_BUCKET_KEYS)) {
    OK(insert_keys_bucketed(idx, d_keys, n, idx->cfg.canonical != 0));
  } else {
    OK(insert_keys_chunked(idx, d_keys, n, idx->cfg.canonical != 0));
  }
  return KMI_OK;
}

extern "C" int kmi_insert(kmi_index* idx, const uint64_t* keys, const uint64_t* vals, size_t n) {
  if (!idx || (!keys && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n == 0) return KMI_OK;
  size_t kb = sizeof(u64) * idx->nw;
  OK(idx->stage_keys.ensure(n * kb));
  OK(copy_to_device(idx, idx->stage_keys.p, keys, n * kb));
  const uint64_t* dv = nullptr;
  if (idx->cfg.mode == KMI_MODE_POSITION || vals) {
    if (!vals) return fail(KMI_ERR_ARG, "POSITION insert needs ids");
    OK(idx->stage_vals.ensure(n * sizeof(u64)));
    OK(copy_to_device(idx, idx->stage_vals.p, vals, n * sizeof(u64)));
    dv = idx->stage_vals.as<uint64_t>();
  }
  OK(kmi_insert_dev(idx, idx->stage_keys.as<uint64_t>(), dv, n));
  CU(cudaStreamSynchronize(idx->stream));
  return check_table_error(idx);
}

// ---- POSITION mode: (re)build the CSR from the tuple log ----------------------------------------------
static int csr_finalize(kmi_index* idx) {
  if (!idx->csr_dirty) return KMI_OK;
  if (!idx->slots) { idx->csr_dirty = false; return KMI_OK; }
  u64 n_blocks = (idx->cap + 1023) / 1024;
  OK(idx->offsets.ensure(sizeof(u64) * idx->cap));
  OK(idx->csr_sums.ensure(sizeof(u64) * n_blocks));
  OK(idx->values.ensure(sizeof(u64) * std::max<u64>(idx->log_n, 1)));
  idx->touch();
  idx->ops->zero_aux(idx->stream, idx->slots, idx->cap);
  idx->ops->csr_scan_l1(idx->stream, idx->slots, idx->cap, idx->offsets.as<u64>(), idx->csr_sums.as<u64>());
  k_scan64<<<1, 1024, 0, idx->stream>>>(idx->csr_sums.as<u64>(), n_blocks);
  k_csr_add_block<<<(unsigned)n_blocks, 1024, 0, idx->stream>>>(idx->offsets.as<u64>(), idx->cap, idx->csr_sums.as<u64>());
  idx->ops->csr_fill(idx->stream, idx->view(), idx->log_keys.as<u64>(), idx->log_ids.as<u64>(), idx->log_n,
                     idx->offsets.as<u64>(), idx->values.as<u64>());
  idx->launches += 5;
  idx->csr_dirty = false;
  return KMI_OK;
}

int kmi::scan64_dev(kmi_index* idx, u64* d_v, u64 n) {
  if (n == 0) return KMI_OK;
  if (n <= 8192) {
    k_scan64<<<1, 1024, 0, idx->stream>>>(d_v, n);
    idx->launches++;
    return KMI_OK;
  }
  const u64 nb = (n + 1023) / 1024;
  OK(idx->scan_tmp.ensure(nb * sizeof(u64)));
  k_scan64_l1<<<(unsigned)nb, 1024, 0, idx->stream>>>(d_v, n, idx->scan_tmp.as<u64>());
  k_scan64<<<1, 1024, 0, idx->stream>>>(idx->scan_tmp.as<u64>(), nb);
  k_csr_add_block<<<(unsigned)nb, 1024, 0, idx->stream>>>(d_v, n, idx->scan_tmp.as<u64>());
  idx->launches += 3;
  return KMI_OK;
}

// ---- batched query -----------------------------------------------------------------------------------
extern "C" int kmi_find_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_vals, uint8_t* d_found) {
  if (!idx || (!d_keys && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n == 0) return KMI_OK;
  if (!idx->slots) {
    if (d_vals) CU(cudaMemsetAsync(d_vals, 0, n * sizeof(u64), idx->stream));
    if (d_found) CU(cudaMemsetAsync(d_found, 0, n, idx->stream));
    return KMI_OK;
  }
  idx->ops->lookup(idx->stream, idx->view(), reinterpret_cast<const u64*>(d_keys), n, reinterpret_cast<u64*>(d_vals), d_found,
                   nullptr, idx->cfg.canonical != 0);
  idx->launches++;
  return KMI_OK;
}
// count(): presence (0/1) for a COUNT index, multiplicity for a POSITION index
extern "C" int kmi_count_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_counts) {
  if (!idx || (!d_keys && n) || (!d_counts && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n == 0) return KMI_OK;
  if (idx->cfg.mode == KMI_MODE_POSITION) return kmi_find_dev(idx, d_keys, n, d_counts, nullptr);
  OK(idx->stage_found.ensure(n));
  OK(kmi_find_dev(idx, d_keys, n, nullptr, idx->stage_found.as<uint8_t>()));
  k_widen_u8<<<grid_for(n, 256), 256, 0, idx->stream>>>(idx->stage_found.as<uint8_t>(), reinterpret_cast<u64*>(d_counts), n);
  idx->launches++;
  return KMI_OK;
}

static int stage_query(kmi_index* idx, const uint64_t* keys, size_t n) {
  size_t kb = sizeof(u64) * idx->nw;
  OK(idx->stage_keys.ensure(std::max<size_t>(n * kb, 16)));
  OK(idx->stage_out.ensure(std::max<size_t>((n + 1) * sizeof(u64), 16)));
  OK(idx->stage_found.ensure(std::max<size_t>(n, 16)));
  return copy_to_device(idx, idx->stage_keys.p, keys, n * kb);
}

extern "C" int kmi_count(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* counts) {
  if (!idx || (!keys && n) || (!counts && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n == 0) return KMI_OK;
  OK(stage_query(idx, keys, n));
  // counts land in stage_out (stage_query sized it): kmi_count_dev itself only uses stage_found
  OK(kmi_count_dev(idx, idx->stage_keys.as<uint64_t>(), n, idx->stage_out.as<uint64_t>()));
  OK(copy_to_host(idx, counts, idx->stage_out.p, n * sizeof(u64)));
  CU(cudaStreamSynchronize(idx->stream));
  return KMI_OK;
}

extern "C" int kmi_find(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* vals, uint8_t* found) {
  if (!idx || (!keys && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n == 0) return KMI_OK;
  OK(stage_query(idx, keys, n));
  OK(kmi_find_dev(idx, idx->stage_keys.as<uint64_t>(), n, idx->stage_out.as<uint64_t>(), idx->stage_found.as<uint8_t>()));
  if (vals) OK(copy_to_host(idx, vals, idx->stage_out.p, n * sizeof(u64)));
  if (found) OK(copy_to_host(idx, found, idx->stage_found.p, n));
  CU(cudaStreamSynchronize(idx->stream));
  return KMI_OK;
}

int kmi::find_positions_dev(kmi_index* idx, const u64* d_keys, u64 m, u64* d_off, DevBuf& slot_idx, DevBuf& vals, u64* total) {
  *total = 0;
  CU(cudaMemsetAsync(d_off, 0, (m + 1) * sizeof(u64), idx->stream));
  if (m == 0 || !idx->slots) return KMI_OK;
  OK(csr_finalize(idx));
  OK(slot_idx.ensure(m * sizeof(u64)));
  // multiplicities -> exclusive scan (m + 1 entries) -> gather
  idx->ops->lookup(idx->stream, idx->view(), d_keys, m, d_off, nullptr, slot_idx.as<u64>(), false);
  idx->launches++;
  OK(scan64_dev(idx, d_off, m + 1));
  CU(cudaMemcpyAsync(total, d_off + m, sizeof(u64), cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  if (*total) {
    OK(vals.ensure(*total * sizeof(u64)));
    k_csr_gather<<<grid_for(m, 256), 256, 0, idx->stream>>>(slot_idx.as<u64>(), d_off, m, idx->offsets.as<u64>(), idx->values.as<u64>(),
                                                          vals.as<u64>(), *total);
    idx->launches++;
  }
  CU(cudaGetLastError());
  return KMI_OK;
}

extern "C" int kmi_find_positions(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* offsets, uint64_t* vals, size_t cap) {
  if (!idx || (!keys && n) || !offsets) return fail(KMI_ERR_ARG, "null argument");
  if (idx->cfg.mode != KMI_MODE_POSITION) return fail(KMI_ERR_ARG, "find_positions needs a POSITION index");
  CU(cudaSetDevice(idx->cfg.device));
  offsets[0] = 0;
  if (n == 0) return KMI_OK;
  if (!idx->slots) { for (size_t i = 0; i <= n; ++i) offsets[i] = 0; return KMI_OK; }
  OK(csr_finalize(idx));
  OK(stage_query(idx, keys, n));
  DevBuf slot_idx, out_off, out_vals;
  int r = slot_idx.ensure(n * sizeof(u64));
  if (r == KMI_OK) r = out_off.ensure((n + 1) * sizeof(u64));
  if (r != KMI_OK) { slot_idx.release(); out_off.release(); return r; }
  // multiplicities -> exclusive scan (n+1 entries) -> gather
  cudaMemsetAsync(out_off.p, 0, (n + 1) * sizeof(u64), idx->stream);
  idx->ops->lookup(idx->stream, idx->view(), idx->stage_keys.as<u64>(), n, out_off.as<u64>(), nullptr, slot_idx.as<u64>(),
                   idx->cfg.canonical != 0);
  idx->launches++;
  r = scan64_dev(idx, out_off.as<u64>(), n + 1);
  if (r == KMI_OK) r = copy_to_host(idx, offsets, out_off.p, (n + 1) * sizeof(u64));
  cudaStreamSynchronize(idx->stream);
  if (r != KMI_OK) { slot_idx.release(); out_off.release(); return r; }
  u64 total = offsets[n];
  if (total > cap) { slot_idx.release(); out_off.release(); return fail(KMI_ERR_CAPACITY, "find_positions: %llu values, capacity %zu", (unsigned long long)total, cap); }
  if (total) {
    r = out_vals.ensure(total * sizeof(u64));
    if (r == KMI_OK) {
      k_csr_gather<<<grid_for(n, 256), 256, 0, idx->stream>>>(slot_idx.as<u64>(), out_off.as<u64>(), n, idx->offsets.as<u64>(),
                                                            idx->values.as<u64>(), out_vals.as<u64>(), total);
      idx->launches++;
      r = copy_to_host(idx, vals, out_vals.p, total * sizeof(u64));
      cudaStreamSynchronize(idx->stream);
    }
  }
  slot_idx.release(); out_off.release(); out_vals.release();
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(KMI_ERR_CUDA, "find_positions: %s", cudaGetErrorString(e));
  return r;
}

extern "C" int kmi_erase(kmi_index* idx, const uint64_t* keys, size_t n) {
  if (!idx || (!keys && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n == 0 || !idx->slots) return KMI_OK;
  OK(flush_pending(idx));
  OK(stage_query(idx, keys, n));
  OK(zero_ctr(idx, CTR_REMOVED));
  u64 live_before = 0;
  OK(read_ctr(idx, CTR_SIZE, &live_before));
  idx->ops->erase(idx->stream, idx->view(), idx->stage_keys.as<u64>(), n, idx->ctr + CTR_REMOVED, idx->cfg.canonical != 0);
  idx->launches++;
  if (idx->cfg.mode == KMI_MODE_POSITION && idx->log_n) {
    // drop the erased keys' tuples from the log
    size_t kb = sizeof(u64) * idx->nw;
    OK(idx->log_keys2.ensure(idx->log_keys.cap));
    OK(idx->log_ids2.ensure(idx->log_ids.cap));
    OK(zero_ctr(idx, CTR_CURSOR));
    idx->ops->log_filter(idx->stream, idx->view(), idx->log_keys.as<u64>(), idx->log_ids.as<u64>(), idx->log_n,
                         idx->log_keys2.as<u64>(), idx->log_ids2.as<u64>(), idx->ctr + CTR_CURSOR);
    idx->launches++;
    u64 kept = 0;
    OK(read_ctr(idx, CTR_CURSOR, &kept));
    std::swap(idx->log_keys, idx->log_keys2);
    std::swap(idx->log_ids, idx->log_ids2);
    idx->log_n = kept;
    idx->csr_dirty = true;
    (void)kb;
  }
  OK(read_ctr(idx, CTR_SIZE, &idx->size_exact));
  idx->size_upper = idx->size_exact;
  idx->tombstones += live_before - idx->size_exact;
  return KMI_OK;
}

// ---- accessors -----------------------------------------------------------------------------------------
extern "C" int kmi_num_keys(kmi_index* idx, uint64_t* n) {
  if (!idx || !n) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  OK(flush_pending(idx));
  u64 v = 0;
  OK(read_ctr(idx, CTR_SIZE, &v));
  idx->size_exact = idx->size_upper = v;
  *n = v;
  return check_table_error(idx);
}
extern "C" int kmi_local_size(kmi_index* idx, uint64_t* n) {
  if (!idx || !n) return fail(KMI_ERR_ARG, "null argument");
  if (idx->cfg.mode == KMI_MODE_POSITION) {
    CU(cudaSetDevice(idx->cfg.device));
    CU(cudaStreamSynchronize(idx->stream));
    *n = idx->log_n;
    return KMI_OK;
  }
  return kmi_num_keys(idx, n);
}

extern "C" int kmi_dump(kmi_index* idx, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out) {
  if (!idx || !n_out) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  OK(flush_pending(idx));
  *n_out = 0;
  size_t kb = sizeof(u64) * idx->nw;
  if (idx->cfg.mode == KMI_MODE_POSITION) {
    if (idx->log_n > cap) return fail(KMI_ERR_CAPACITY, "dump: %llu tuples, capacity %zu", (unsigned long long)idx->log_n, cap);
    if (idx->log_n) {
      if (keys) OK(copy_to_host(idx, keys, idx->log_keys.p, idx->log_n * kb));
      if (vals) OK(copy_to_host(idx, vals, idx->log_ids.p, idx->log_n * sizeof(u64)));
    }
    CU(cudaStreamSynchronize(idx->stream));
    *n_out = idx->log_n;
    return KMI_OK;
  }
  u64 size = 0;
  OK(read_ctr(idx, CTR_SIZE, &size));
  if (size > cap) return fail(KMI_ERR_CAPACITY, "dump: %llu entries, capacity %zu", (unsigned long long)size, cap);
  if (size == 0 || !idx->slots) return KMI_OK;
  DevBuf dk, dv;
  int r = dk.ensure(size * kb);
  if (r == KMI_OK) r = dv.ensure(size * sizeof(u64));
  if (r == KMI_OK) {
    zero_ctr(idx, CTR_CURSOR);
    idx->touch();
    idx->ops->dump(idx->stream, idx->slots, idx->cap, dk.as<u64>(), dv.as<u64>(), size, idx->ctr + CTR_CURSOR, -1, 0);
    idx->launches++;
    if (keys) r = copy_to_host(idx, keys, dk.p, size * kb);
    if (r == KMI_OK && vals) r = copy_to_host(idx, vals, dv.p, size * sizeof(u64));
    const cudaError_t e = cudaStreamSynchronize(idx->stream);   // also when a copy failed: dk / dv are freed below
    if (r == KMI_OK && e != cudaSuccess) r = fail(KMI_ERR_CUDA, "dump: %s", cudaGetErrorString(e));
    if (r == KMI_OK) *n_out = size;
  }
  dk.release(); dv.release();
  return r;
}

// ---- f2: predicates on the stored value, evaluated on the device (COUNT index: the frequency) -----------------------------
static int pred_args(kmi_index* idx, int op) {
  if (!idx) return fail(KMI_ERR_ARG, "null index");
  if (op < KMI_PRED_GE || op > KMI_PRED_EQ) return fail(KMI_ERR_ARG, "predicate must be KMI_PRED_GE, _LE or _EQ");
  if (idx->cfg.mode != KMI_MODE_COUNT) return fail(KMI_ERR_UNSUPPORTED, "value predicates are evaluated on the device for COUNT indices");
  return KMI_OK;
}
extern "C" int kmi_count_if(kmi_index* idx, int op, uint64_t value, uint64_t* n) {
  OK(pred_args(idx, op));
  if (!n) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  *n = 0;
  if (!idx->slots) return KMI_OK;
  OK(zero_ctr(idx, CTR_CURSOR));
  idx->touch();
  idx->ops->dump(idx->stream, idx->slots, idx->cap, nullptr, nullptr, 0, idx->ctr + CTR_CURSOR, op, value);
  idx->launches++;
  u64 v = 0;
  OK(read_ctr(idx, CTR_CURSOR, &v));
  *n = v;
  return check_table_error(idx);
}
extern "C" int kmi_dump_if(kmi_index* idx, int op, uint64_t value, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out) {
  OK(pred_args(idx, op));
  if (!n_out || (!keys && cap)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  *n_out = 0;
  if (!idx->slots || cap == 0) return KMI_OK;
  const size_t kb = sizeof(u64) * idx->nw;
  DevBuf dk, dv;
  int r = dk.ensure(cap * kb);
  if (r == KMI_OK) r = dv.ensure(cap * sizeof(u64));
  if (r == KMI_OK) {
    zero_ctr(idx, CTR_CURSOR);
    idx->touch();
    idx->ops->dump(idx->stream, idx->slots, idx->cap, dk.as<u64>(), dv.as<u64>(), cap, idx->ctr + CTR_CURSOR, op, value);
    idx->launches++;
    u64 m = 0;
    r = read_ctr(idx, CTR_CURSOR, &m);
    if (r == KMI_OK && m > cap) r = fail(KMI_ERR_CAPACITY, "dump_if: %llu entries satisfy the predicate, capacity %zu (kmi_count_if sizes it)", (unsigned long long)m, cap);
    if (r == KMI_OK && m) {
      r = copy_to_host(idx, keys, dk.p, m * kb);
      if (r == KMI_OK && vals) r = copy_to_host(idx, vals, dv.p, m * sizeof(u64));
      const cudaError_t e = cudaStreamSynchronize(idx->stream);
      if (r == KMI_OK && e != cudaSuccess) r = fail(KMI_ERR_CUDA, "dump_if: %s", cudaGetErrorString(e));
    }
    if (r == KMI_OK) *n_out = m;
  }
  dk.release(); dv.release();
  return r;
}
extern "C" int kmi_erase_if(kmi_index* idx, int op, uint64_t value, uint64_t* n_removed) {
  OK(pred_args(idx, op));
  CU(cudaSetDevice(idx->cfg.device));
  if (n_removed) *n_removed = 0;
  if (!idx->slots) return KMI_OK;
  OK(zero_ctr(idx, CTR_REMOVED));
  idx->touch();
  idx->ops->erase_if(idx->stream, idx->slots, idx->cap, idx->ctr + CTR_SIZE, idx->ctr + CTR_REMOVED, op, value);
  idx->launches++;
  u64 removed = 0;
  OK(read_ctr(idx, CTR_REMOVED, &removed));
  OK(read_ctr(idx, CTR_SIZE, &idx->size_exact));
  idx->size_upper = idx->size_exact;
  idx->tombstones += removed;
  if (n_removed) *n_removed = removed;
  return KMI_OK;
}

extern "C" int kmi_checksum(kmi_index* idx, uint64_t* key_value_sum, uint64_t* value_sum) {
  if (!idx) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  OK(zero_ctr(idx, CTR_CK0, 2));
  idx->touch();
  if (idx->slots) { idx->ops->checksum(idx->stream, idx->slots, idx->cap, idx->ctr + CTR_CK0); idx->launches++; }
  u64 host[2];
  CU(cudaMemcpyAsync(host, idx->ctr + CTR_CK0, sizeof(host), cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  if (key_value_sum) *key_value_sum = host[0];
  if (value_sum) *value_sum = host[1];
  return KMI_OK;
}

extern "C" int kmi_table_info(kmi_index* idx, uint64_t* slots, uint64_t* slot_bytes, uint64_t* device_bytes) {
  if (!idx) return fail(KMI_ERR_ARG, "null argument");
  if (slots) *slots = idx->cap;
  if (slot_bytes) *slot_bytes = idx->slot_bytes;
  if (device_bytes) {
    u64 b = idx->cap * idx->slot_bytes;
    const DevBuf* bufs[] = {&idx->tile_lines, &idx->block_sums, &idx->stage_in, &idx->stage_keys, &idx->stage_vals,
                            &idx->stage_out, &idx->stage_found, &idx->hist, &idx->hist_sums, &idx->log_keys, &idx->log_ids,
                            &idx->log_keys2, &idx->log_ids2, &idx->offsets, &idx->csr_sums, &idx->values, &idx->bkt_keys, &idx->bkt_meta, &idx->fine_keys, &idx->fine_cur, &idx->parse_keys, &idx->parse_ids, &idx->tile_rec, &idx->mid_keys, &idx->mid_cur,
                            &idx->pend_ovf, &idx->pend_ovf_cur, &idx->pend_cur_save, &idx->scan_tmp};
    for (const DevBuf* d : bufs) b += d->cap;
    *device_bytes = b;
  }
  return KMI_OK;
}
extern "C" int kmi_launch_count(kmi_index* idx, uint64_t* n) {
  if (!idx || !n) return fail(KMI_ERR_ARG, "null argument");
  *n = idx->launches;
  return KMI_OK;
}
extern "C" int kmi_last_kernel_ms(kmi_index* idx, float* ms5) {
  if (!idx || !ms5) return fail(KMI_ERR_ARG, "null argument");
  for (int i = 0; i < 5; ++i) ms5[i] = idx->kernel_ms[i];
  return KMI_OK;
}
extern "C" int kmi_last_stage_ms(kmi_index* idx, float* ms3) {
  if (!idx || !ms3) return fail(KMI_ERR_ARG, "null argument");
  for (int i = 0; i < 3; ++i) ms3[i] = idx->stage_ms[i];
  return KMI_OK;
}

// ---- stages ----------------------------------------------------------------------------------------------
extern "C" int kmi_canonicalize_dev(kmi_index* idx, const uint64_t* d_keys, uint64_t* d_out, size_t n) {
  if (!idx || ((!d_keys || !d_out) && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  idx->ops->canonicalize(idx->stream, reinterpret_cast<const u64*>(d_keys), reinterpret_cast<u64*>(d_out), n);
  idx->launches++;
  return KMI_OK;
}
extern "C" int kmi_hash_dev(kmi_index* idx, int which, const uint64_t* d_keys, uint64_t* d_hash, size_t n) {
  if (!idx || ((!d_keys || !d_hash) && n)) return fail(KMI_ERR_ARG, "null argument");
  if (which != 0 && which != 1) return fail(KMI_ERR_ARG, "which must be 0 (store) or 1 (dist)");
  CU(cudaSetDevice(idx->cfg.device));
  idx->ops->hash_keys(idx->stream, which, reinterpret_cast<const u64*>(d_keys), reinterpret_cast<u64*>(d_hash), n, (u32)idx->cfg.nranks);
  idx->launches++;
  return KMI_OK;
}

extern "C" int kmi_partition_dev(kmi_index* idx, const uint64_t* d_keys, const uint64_t* d_vals, size_t n,
                                 uint64_t* d_keys_out, uint64_t* d_vals_out, uint64_t* h_counts) {
  if (!idx || !h_counts || ((!d_keys || !d_keys_out) && n) || (d_vals && !d_vals_out)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  const u32 p = (u32)idx->cfg.nranks;
  for (u32 r = 0; r < p; ++r) h_counts[r] = 0;
  if (n == 0) return KMI_OK;
  if (n >= 0xffffffffULL) return fail(KMI_ERR_ARG, "partition: at most 2^32-2 keys per call");
  u64 n_chunks = (n + PART_ITEMS - 1) / PART_ITEMS;
  u64 n_hist = n_chunks * p;
  u64 n_blocks = (n_hist + 1023) / 1024;
  OK(idx->hist.ensure(sizeof(u32) * (n_hist + 1)));
  OK(idx->hist_sums.ensure(sizeof(u32) * (n_blocks + 1)));
  idx->ops->part_hist(idx->stream, reinterpret_cast<const u64*>(d_keys), n, p, idx->hist.as<u32>(), n_chunks, idx->dist_trans);
  k_scan_l1<<<(unsigned)n_blocks, 1024, 0, idx->stream>>>(idx->hist.as<u32>(), n_hist, idx->hist_sums.as<u32>());
  k_scan_tiles<<<1, 1024, 0, idx->stream>>>(idx->hist_sums.as<u32>(), n_blocks);
  idx->ops->part_scatter(idx->stream, reinterpret_cast<const u64*>(d_keys), reinterpret_cast<const u64*>(d_vals), n, p,
                         idx->hist.as<u32>(), idx->hist_sums.as<u32>(), n_chunks, reinterpret_cast<u64*>(d_keys_out),
                         reinterpret_cast<u64*>(d_vals_out), idx->dist_trans);
  idx->launches += 4;
  // owner r starts at scanned[r * n_chunks]: read the p starts back
  std::vector<u32> starts(p), sums(p);
  for (u32 r = 0; r < p; ++r) {
    u64 hi = (u64)r * n_chunks;
    CU(cudaMemcpyAsync(&starts[r], idx->hist.as<u32>() + hi, sizeof(u32), cudaMemcpyDeviceToHost, idx->stream));
    CU(cudaMemcpyAsync(&sums[r], idx->hist_sums.as<u32>() + (hi >> 10), sizeof(u32), cudaMemcpyDeviceToHost, idx->stream));
  }
  CU(cudaStreamSynchronize(idx->stream));
  for (u32 r = 0; r < p; ++r) {
    u64 b = (u64)starts[r] + sums[r];
    u64 e = (r + 1 < p) ? (u64)starts[r + 1] + sums[r + 1] : (u64)n;
    h_counts[r] = e - b;
  }
  return KMI_OK;
}

// ---- partition without allocation ----------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it). Same three kernels as kmi_partition_dev; their
// scratch — the [owner][chunk] histogram and its block sums — lives in ONE file-scope device array, held by every GPU that
// loads the library until exit. g_part_scratch is a host-side stand-in here: host code must not hand it to a kernel, only the
// address cudaGetSymbolAddress returns for the current device. Every call shares the array (calls that overlap on one
// GPU race); every word a call reads was written earlier by the same call (k_part_hist fills all n_chunks * p histogram
// words, k_scan_l1 all block sums), so nothing has to start at zero.
static constexpr u32 NOALLOC_HIST_WORDS = KMI_PARTITION_NOALLOC_MAX_HIST;            // histogram words (n_chunks * nranks)
static constexpr u32 NOALLOC_SUMS_AT = NOALLOC_HIST_WORDS + 64;                      // block sums start here (256-byte aligned)
static constexpr u32 NOALLOC_WORDS = NOALLOC_SUMS_AT + NOALLOC_HIST_WORDS / 1024 + 64;
static __device__ __align__(256) u32 g_part_scratch[NOALLOC_WORDS];                  // 4.2 MB

static constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static u32* g_noalloc_ptr[NOALLOC_MAX_DEVICES];   // device address of g_part_scratch, per device

// counts[r] = start of owner r+1 - start of owner r, from the scanned histogram (no host read-back)
__global__ void k_part_counts(const u32* __restrict__ hist_scanned, const u32* __restrict__ sums, u64 n_chunks, u32 p, u64 n,
                              u64* __restrict__ counts) {
  u32 r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= p) return;
  u64 hi = (u64)r * n_chunks, hj = hi + n_chunks;
  u64 b = (u64)hist_scanned[hi] + sums[hi >> 10];
  u64 e = (r + 1 < p) ? (u64)hist_scanned[hj] + sums[hj >> 10] : n;
  counts[r] = e - b;
}

extern "C" size_t kmi_partition_noalloc_max_keys(int nranks) {
  if (nranks < 1 || nranks > MAX_RANKS) return 0;
  return (size_t)(NOALLOC_HIST_WORDS / (u32)nranks) * PART_ITEMS;
}

extern "C" int kmi_partition_noalloc_dev(kmi_index* idx, const uint64_t* d_keys, const uint64_t* d_vals, size_t n,
                                         uint64_t* d_keys_out, uint64_t* d_vals_out, uint64_t* d_counts) {
  if (!idx || !d_counts || ((!d_keys || !d_keys_out) && n) || (d_vals && !d_vals_out)) return fail(KMI_ERR_ARG, "null argument");
  const int dev = idx->cfg.device;
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(KMI_ERR_ARG, "partition_noalloc: device ordinal out of range");
  CU(cudaSetDevice(dev));
  const u32 p = (u32)idx->cfg.nranks;
  const u64 n_chunks = (n + PART_ITEMS - 1) / PART_ITEMS;
  const u64 n_hist = n_chunks * p;
  if (n_hist > NOALLOC_HIST_WORDS)
    return fail(KMI_ERR_ARG, "partition_noalloc: %llu keys over %u ranks exceed the static scratch (at most %llu keys; use kmi_partition_dev)",
                (unsigned long long)n, p, (unsigned long long)kmi_partition_noalloc_max_keys((int)p));
  u32* scratch;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);     // set-up: once per device, with that device current
    if (!g_noalloc_ptr[dev]) {
      void* sym = nullptr;
      CU(cudaGetSymbolAddress(&sym, g_part_scratch));
      g_noalloc_ptr[dev] = static_cast<u32*>(sym);
    }
    scratch = g_noalloc_ptr[dev];
  }
  u32* hist = scratch;
  u32* sums = scratch + NOALLOC_SUMS_AT;
  if (n == 0) {
    CU(cudaMemsetAsync(d_counts, 0, sizeof(u64) * p, idx->stream));
    return KMI_OK;
  }
  const u64 n_blocks = (n_hist + 1023) / 1024;
  idx->ops->part_hist(idx->stream, reinterpret_cast<const u64*>(d_keys), n, p, hist, n_chunks, idx->dist_trans);
  k_scan_l1<<<(unsigned)n_blocks, 1024, 0, idx->stream>>>(hist, n_hist, sums);
  k_scan_tiles<<<1, 1024, 0, idx->stream>>>(sums, n_blocks);
  idx->ops->part_scatter(idx->stream, reinterpret_cast<const u64*>(d_keys), reinterpret_cast<const u64*>(d_vals), n, p, hist, sums, n_chunks,
                         reinterpret_cast<u64*>(d_keys_out), reinterpret_cast<u64*>(d_vals_out), idx->dist_trans);
  k_part_counts<<<(p + 255) / 256, 256, 0, idx->stream>>>(hist, sums, n_chunks, p, n, reinterpret_cast<u64*>(d_counts));
  CU(cudaGetLastError());
  idx->launches += 5;
  return KMI_OK;
}

// ---- synthetic workload ------------------------------------------------------------------------------------
extern "C" size_t kmi_synth_record_bytes(uint32_t read_len) { return 13 + 2 * (size_t)read_len + 4; }
extern "C" int kmi_synth_fastq_dev(int device, uint8_t* d_out, uint64_t first_read, uint64_t n_reads, uint32_t read_len,
                                   uint64_t genome_len, uint64_t seed, uint32_t err_ppm, uint32_t n_ppm) {
  if (!d_out && n_reads) return fail(KMI_ERR_ARG, "null argument");
  if (genome_len < read_len || read_len == 0) return fail(KMI_ERR_ARG, "genome shorter than a read");
  CU(cudaSetDevice(device));
  if (n_reads == 0) return KMI_OK;
  u64 total = n_reads * kmi_synth_record_bytes(read_len);
  k_synth_fastq<<<grid_for(total, 256, 148u * 32u), 256>>>(d_out, first_read, n_reads, read_len, genome_len, seed, err_ppm, n_ppm);
  CU(cudaGetLastError());
  return KMI_OK;
}

extern "C" size_t kmi_synth_fasta_bytes(uint32_t n_seq, uint64_t seq_len, uint32_t width) {
  if (width == 0) return 0;
  return (size_t)n_seq * (7 + seq_len + (seq_len + width - 1) / width);
}
extern "C" int kmi_synth_fasta_dev(int device, uint8_t* d_out, uint64_t file_offset, uint64_t n_bytes, uint32_t n_seq, uint64_t seq_len,
                                   uint32_t width, uint64_t seed) {
  if (!d_out && n_bytes) return fail(KMI_ERR_ARG, "null argument");
  if (width == 0 || n_seq == 0 || n_seq > 100) return fail(KMI_ERR_ARG, "1..100 sequences, line width > 0");
  CU(cudaSetDevice(device));
  if (n_bytes == 0) return KMI_OK;
  k_synth_fasta<<<grid_for(n_bytes, 256, 148u * 32u), 256>>>(d_out, file_offset, n_bytes, n_seq, seq_len, width, seed);
  CU(cudaGetLastError());
  return KMI_OK;
}
extern "C" int kmi_synth_queries_dev(int device, uint64_t* d_keys, uint64_t first, uint64_t n, int k, uint32_t n_seq, uint64_t seq_len,
                                     uint64_t genome_seed, uint64_t seed) {
  if (!d_keys && n) return fail(KMI_ERR_ARG, "null argument");
  if (k < 2 || k > 63 || n_seq == 0 || seq_len < (uint64_t)k) return fail(KMI_ERR_ARG, "bad query shape");
  CU(cudaSetDevice(device));
  if (n == 0) return KMI_OK;
  k_synth_queries<<<grid_for(n, 256, 148u * 32u), 256>>>(reinterpret_cast<u64*>(d_keys), first, n, k, n_seq, seq_len, genome_seed, seed);
  CU(cudaGetLastError());
  return KMI_OK;
}

// ---- FASTA ---------------------------------------------------------------------------------------------------
// per-tile side arrays of one block: enc[n_tiles + 1] (max-scanned in place) | first_t[n_tiles] | runs[n_tiles + 1]
// (sum-scanned in place); the extra entry is a zero whose exclusive scan value is the whole block's total.
struct FastaArrays {
  u64 n_tiles, n_blocks;
  u32 *enc, *first_t, *runs, *enc_max, *runs_sums;
};
static FastaArrays fasta_arrays(kmi_index* idx, u64 n_tiles) {
  FastaArrays a;
  a.n_tiles = n_tiles; a.n_blocks = (n_tiles + 1 + 1023) / 1024;
  a.enc = idx->tile_lines.as<u32>(); a.first_t = a.enc + n_tiles + 1; a.runs = a.first_t + n_tiles;
  a.enc_max = idx->block_sums.as<u32>(); a.runs_sums = a.enc_max + a.n_blocks;
  return a;
}
static FaBlock whole_file_block(size_t n) {
  FaBlock b;
  b.file_offset = 0; b.n_valid = n; b.runs0 = 0; b.kind0 = 0; b.bol = 1; b.first_is_header = 2;  // 2: look at bytes[0]
  return b;
}
static int fasta_block_from_abi(const kmi_fasta_block* in, size_t n, FaBlock* out) {
  if (!in) { *out = whole_file_block(n); return KMI_OK; }
  if (in->n_valid > n) return fail(KMI_ERR_ARG, "kmi_fasta_block: n_valid %llu exceeds the %zu bytes given", (unsigned long long)in->n_valid, n);
  out->file_offset = in->file_offset; out->n_valid = in->n_valid; out->runs0 = in->header_runs_before;
  out->kind0 = (in->flags & KMI_FASTA_PREV_HEADER) ? 1u : 0u;
  out->bol = (in->flags & KMI_FASTA_BOL) ? 1u : 0u;
  out->first_is_header = (in->flags & KMI_FASTA_FILE_STARTS_WITH_HEADER) ? 1u : 0u;
  return KMI_OK;
}
// pass A: line kinds entering every tile (max-scan) and header runs before every tile (sum-scan), over the block's own
// bytes [0, n_valid); the halo needs no scan (the parse kernel reads its line structure itself)
static int fasta_scan(kmi_index* idx, const uint8_t* d_bytes, size_t n, FaBlock* blk, u64* n_tiles_out) {
  if ((reinterpret_cast<uintptr_t>(d_bytes) & 15) != 0) return fail(KMI_ERR_ARG, "device text buffer must be 16-byte aligned");
  u64 n_tiles = (blk->n_valid + TILE_BYTES - 1) / TILE_BYTES;
  *n_tiles_out = n_tiles;
  if (n_tiles == 0) return KMI_OK;
  if (n_tiles > 0x3fffffffULL) return fail(KMI_ERR_ARG, "block of %zu bytes too large for one call", n);
  if (blk->first_is_header == 2) {   // whole file: its first byte says whether it opens with a header
    uint8_t c = 0;
    CU(cudaMemcpyAsync(&c, d_bytes, 1, cudaMemcpyDeviceToHost, idx->stream));
    CU(cudaStreamSynchronize(idx->stream));
    blk->first_is_header = (c == '>' || c == ';') ? 1u : 0u;
  }
  u64 n_blocks = (n_tiles + 1 + 1023) / 1024;
  OK(idx->tile_lines.ensure(sizeof(u32) * (n_tiles * 3 + 2)));
  OK(idx->block_sums.ensure(sizeof(u32) * n_blocks * 2));
  FastaArrays a = fasta_arrays(idx, n_tiles);
  CU(cudaMemsetAsync(idx->ctr + CTR_FIRST_LS, 0xFF, sizeof(u64), idx->stream));
  k_fasta_tiles<<<(unsigned)n_tiles, TILE_THREADS, 0, idx->stream>>>(d_bytes, n, blk->n_valid, blk->bol, a.enc, a.first_t, idx->ctr + CTR_FIRST_LS);
  k_maxscan_l1<<<(unsigned)n_blocks, 1024, 0, idx->stream>>>(a.enc, n_tiles + 1, a.enc_max);
  k_maxscan_l2<<<1, 32, 0, idx->stream>>>(a.enc_max, n_blocks);
  k_fasta_tile_starts<<<grid_for(n_tiles + 1, 256), 256, 0, idx->stream>>>(a.enc, a.enc_max, a.first_t, n_tiles, blk->kind0, a.runs);
  k_scan_l1<<<(unsigned)n_blocks, 1024, 0, idx->stream>>>(a.runs, n_tiles + 1, a.runs_sums);
  k_scan_tiles<<<1, 1024, 0, idx->stream>>>(a.runs_sums, n_blocks);
  idx->launches += 6;
  CU(cudaGetLastError());
  return KMI_OK;
}

// parse tiles [tile_begin, tile_begin + tiles) into the sink; retries with the wide halo when needed
static int fasta_parse_range(kmi_index* idx, const uint8_t* d_bytes, size_t n, const FaBlock& blk, u64 n_tiles_total, u64 tile_begin,
                             u64 tiles, KeySink sink, u64* produced, u64 cursor_start = 0) {
  FastaArrays a = fasta_arrays(idx, n_tiles_total);
  static_assert(CTR_CURSOR == CTR_NKMERS + 1, "the two counters are set together");
  for (int attempt = 0; attempt < 2; ++attempt) {
    const u64 init[2] = {0, cursor_start};   // output is appended at cursor_start
    CU(cudaMemcpyAsync(idx->ctr + CTR_NKMERS, init, sizeof(init), cudaMemcpyHostToDevice, idx->stream));
    OK(zero_ctr(idx, CTR_PERR));
    idx->ops->fasta_kmers(idx->stream, attempt == 1, d_bytes, n, tiles, a.enc, a.enc_max, a.runs, a.runs_sums, sink, idx->ctr + CTR_NKMERS,
                          reinterpret_cast<u32*>(idx->ctr + CTR_PERR), idx->cfg.canonical != 0, idx->cfg.nsplit != 0, tile_begin, blk);
    idx->launches++;
    u64 host[CTR_N];
    CU(cudaMemcpyAsync(host, idx->ctr, sizeof(u64) * CTR_N, cudaMemcpyDeviceToHost, idx->stream));
    CU(cudaStreamSynchronize(idx->stream));
    *produced = host[CTR_NKMERS];
    if ((host[CTR_PERR] & PERR_HALO) && attempt == 0) continue;
    return parse_error(host[CTR_PERR]);
  }
  return KMI_OK;
}

// read a scanned block's totals off the sentinel entries (valid after fasta_scan on the same stream)
static int fasta_read_summary(kmi_index* idx, u64 n_tiles, u32* out) {
  FastaArrays a = fasta_arrays(idx, n_tiles);
  u32 enc = 0, enc_b = 0, runs = 0, runs_b = 0;
  u64 first = 0;
  CU(cudaMemcpyAsync(&enc, a.enc + n_tiles, 4, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaMemcpyAsync(&enc_b, a.enc_max + (n_tiles >> 10), 4, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaMemcpyAsync(&runs, a.runs + n_tiles, 4, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaMemcpyAsync(&runs_b, a.runs_sums + (n_tiles >> 10), 4, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaMemcpyAsync(&first, idx->ctr + CTR_FIRST_LS, 8, cudaMemcpyDeviceToHost, idx->stream));
  CU(cudaStreamSynchronize(idx->stream));
  const u32 last = std::max(enc, enc_b);
  out[0] = last ? 1u : 0u;                                     // the block holds a line start
  out[1] = (last && first != ~0ULL) ? (u32)(first & 1u) : 0u;  // its first line start begins a header line
  out[2] = last & 1u;                                          // its last one does
  out[3] = runs + runs_b;                                      // header runs starting in the block, given the kind0 it was scanned with
  return KMI_OK;
}

extern "C" int kmi_fasta_block_summary_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n_valid, int bol, uint32_t* out) {
  if (!idx || (!d_bytes && n_valid) || !out) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  out[0] = out[1] = out[2] = out[3] = 0;
  FaBlock blk = whole_file_block(n_valid);
  blk.bol = bol ? 1u : 0u; blk.first_is_header = 0;
  u64 n_tiles = 0;
  OK(fasta_scan(idx, d_bytes, n_valid, &blk, &n_tiles));
  if (n_tiles == 0) return KMI_OK;
  OK(fasta_read_summary(idx, n_tiles, out));
  return KMI_OK;
}

extern "C" int kmi_parse_fasta_block_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, const kmi_fasta_block* block,
                                         uint64_t* d_keys, uint64_t* d_ids, size_t cap, uint64_t* n_kmers) {
  if (!idx || (!d_bytes && n) || (!d_keys && cap)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n_kmers) *n_kmers = 0;
  FaBlock blk;
  OK(fasta_block_from_abi(block, n, &blk));
  u64 n_tiles = 0;
  OK(fasta_scan(idx, d_bytes, n, &blk, &n_tiles));
  if (n_tiles == 0) return KMI_OK;
  KeySink sink;
  sink.keys = reinterpret_cast<u64*>(d_keys); sink.ids = reinterpret_cast<u64*>(d_ids); sink.cap = cap; sink.cursor = idx->ctr + CTR_CURSOR;
  u64 produced = 0;
  OK(fasta_parse_range(idx, d_bytes, n, blk, n_tiles, 0, n_tiles, sink, &produced));
  if (n_kmers) *n_kmers = produced;
  return KMI_OK;
}
extern "C" int kmi_parse_fasta_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t* d_keys, uint64_t* d_ids,
                                   size_t cap, uint64_t* n_kmers) {
  return kmi_parse_fasta_block_dev(idx, d_bytes, n, nullptr, d_keys, d_ids, cap, n_kmers);
}

// scan + parse + insert of one block whose bytes are resident; summary (optional) = the block's totals for chaining
static int fasta_build_block(kmi_index* idx, const uint8_t* d_bytes, size_t n, FaBlock blk, u64* n_kmers, u32* summary,
                             float* scan_ms, float* parse_ms_out) {
  u64 n_tiles = 0;
  CU(cudaEventRecord(idx->ev[0], idx->stream));
  OK(fasta_scan(idx, d_bytes, n, &blk, &n_tiles));
  CU(cudaEventRecord(idx->ev[1], idx->stream));
  if (summary) { summary[0] = summary[1] = summary[2] = summary[3] = 0; }
  if (n_tiles == 0) return KMI_OK;
  if (summary) OK(fasta_read_summary(idx, n_tiles, summary));
  // bounded staging: at most CHUNK_TILES tiles of text become keys (+ids) at a time
  const u64 CHUNK_TILES = 65536;  // 256 MiB of text -> <= 256 Mi k-mers (4 GiB of key + id staging): every chunk costs one sweep of the table
  const bool pos = idx->cfg.mode == KMI_MODE_POSITION;
  size_t kb = sizeof(u64) * idx->nw;
  u64 cap = std::min<u64>(n_tiles, CHUNK_TILES) * TILE_BYTES;
  DevBuf& keys = idx->parse_keys; DevBuf& ids = idx->parse_ids;
  int r = keys.ensure(cap * kb);
  if (r == KMI_OK && pos) r = ids.ensure(cap * sizeof(u64));
  u64 total = 0;
  float parse_ms = 0;
  for (u64 tile = 0; r == KMI_OK && tile < n_tiles; tile += CHUNK_TILES) {
    u64 tiles = std::min<u64>(CHUNK_TILES, n_tiles - tile);
    KeySink sink;
    sink.keys = keys.as<u64>(); sink.ids = pos ? ids.as<u64>() : nullptr; sink.cap = cap; sink.cursor = idx->ctr + CTR_CURSOR;
    u64 produced = 0;
    cudaEventRecord(idx->ev[2], idx->stream);
    r = fasta_parse_range(idx, d_bytes, n, blk, n_tiles, tile, tiles, sink, &produced);
    cudaEventRecord(idx->ev[3], idx->stream);
    if (r != KMI_OK) break;
    // keys are already canonical: insert them as they are
    if (pos) {
      int32_t canon = idx->cfg.canonical;
      idx->cfg.canonical = 0;
      r = kmi_insert_dev(idx, keys.as<uint64_t>(), ids.as<uint64_t>(), produced);
      idx->cfg.canonical = canon;
    } else {
      r = insert_keys_chunked(idx, keys.as<u64>(), produced, false);
    }
    cudaStreamSynchronize(idx->stream);
    float a = 0;
    cudaEventElapsedTime(&a, idx->ev[2], idx->ev[3]);
    parse_ms += a;
    total += produced;
  }
  if (r != KMI_OK) return r;
  float sm = 0;
  cudaEventElapsedTime(&sm, idx->ev[0], idx->ev[1]);
  *scan_ms += sm;
  *parse_ms_out += parse_ms;
  *n_kmers += total;
  return KMI_OK;
}

extern "C" int kmi_build_fasta_block_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, const kmi_fasta_block* block,
                                         uint64_t* n_kmers) {
  if (!idx || (!d_bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n_kmers) *n_kmers = 0;
  FaBlock blk;
  OK(fasta_block_from_abi(block, n, &blk));
  u64 total = 0;
  float scan_ms = 0, parse_ms = 0;
  OK(fasta_build_block(idx, d_bytes, n, blk, &total, nullptr, &scan_ms, &parse_ms));
  idx->stage_ms[0] = scan_ms; idx->stage_ms[1] = parse_ms; idx->stage_ms[2] = 0;
  if (n_kmers) *n_kmers = total;
  return check_table_error(idx);
}
extern "C" int kmi_build_fasta_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t* n_kmers) {
  return kmi_build_fasta_block_dev(idx, d_bytes, n, nullptr, n_kmers);
}

// Host-pointer FASTA build: the text goes to the device in chunks on a copy stream, and chunk c is built as a BLOCK of the
// file (its halo is chunk c+1, its inherited state the chained totals of the chunks before it) while chunks c+2.. are still
// in flight, so the PCIe transfer hides behind the build (or the other way round).
static int build_fasta_host(kmi_index* idx, const uint8_t* bytes, size_t n, FaBlock base, uint64_t* n_kmers, u32* runs_out,
                            u32* kind_out) {
  if (base.first_is_header == 2) base.first_is_header = (n && (bytes[0] == '>' || bytes[0] == ';')) ? 1u : 0u;
  OK(idx->stage_in.ensure(n + 64));
  uint8_t* d = idx->stage_in.as<uint8_t>();
  // Text is staged into keys (+ids) one 256 MiB ROUND at a time (every round costs one sweep of the table when it is
  // inserted); inside a round the copy and the parse go in CHUNKs of a quarter round, so only the first chunk's copy and the
  // inserts are exposed. KMI_FASTA_CHUNK_BYTES overrides the chunk size, a round is then 4 chunks (tests).
  const u64 ROUND_MAX = 256ull << 20;
  u64 CHUNK = std::min<u64>(std::max<u64>(base.n_valid / 4, 8ull << 20), ROUND_MAX / 4);
  if (const char* e = getenv("KMI_FASTA_CHUNK_BYTES")) CHUNK = std::max<u64>(strtoull(e, nullptr, 10), TILE_BYTES);
  CHUNK = (CHUNK + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES;
  const u64 ROUND = std::min<u64>(ROUND_MAX, 4 * CHUNK);
  const u64 n_chunks = std::max<u64>((base.n_valid + CHUNK - 1) / CHUNK, 1);
  HostFeed feed;
  {   // the last copy takes the caller's halo along
    std::vector<u64> ends(n_chunks);
    for (u64 c = 0; c < n_chunks; ++c) ends[c] = (c + 1 == n_chunks) ? n : std::min<u64>(n, (c + 1) * CHUNK);
    OK(feed.start(idx, d, bytes, ends, idx->stream));
  }
  const bool pos = idx->cfg.mode == KMI_MODE_POSITION;
  const u64 cap = std::min<u64>(ROUND, (base.n_valid + TILE_BYTES - 1) / TILE_BYTES * TILE_BYTES) + 1;
  OK(idx->parse_keys.ensure(cap * sizeof(u64) * idx->nw));
  if (pos) OK(idx->parse_ids.ensure(cap * sizeof(u64)));
  auto flush = [&](u64 count) -> int {   // keys are already canonical: insert them as they are
    if (count == 0) return KMI_OK;
    int r;
    if (pos) {
      int32_t canon = idx->cfg.canonical;
      idx->cfg.canonical = 0;
      r = kmi_insert_dev(idx, idx->parse_keys.as<uint64_t>(), idx->parse_ids.as<uint64_t>(), count);
      idx->cfg.canonical = canon;
    } else {
      r = insert_keys_chunked(idx, idx->parse_keys.as<u64>(), count, false);
    }
    return r;
  };
  u64 total = 0, staged = 0;
  u32 runs = base.runs0, kind = base.kind0;
  for (u64 c = 0; c < n_chunks; ++c) {
    const u64 b0 = c * CHUNK;
    const bool last = c + 1 == n_chunks;
    FaBlock blk = base;
    blk.file_offset = base.file_offset + b0;
    blk.n_valid = std::min<u64>(CHUNK, base.n_valid - std::min<u64>(base.n_valid, b0));
    blk.runs0 = runs; blk.kind0 = kind;
    blk.bol = (c == 0) ? base.bol : (bytes[b0 - 1] == '\n' ? 1u : 0u);
    const u64 avail = (last ? n : std::min<u64>(n, (c + 2) * CHUNK)) - b0;
    if (staged + blk.n_valid > cap) { OK(flush(staged)); staged = 0; }
    OK(feed.wait(last ? c : c + 1, idx->stream));   // own bytes + halo = the next chunk
    u64 n_tiles = 0;
    OK(fasta_scan(idx, d + b0, avail, &blk, &n_tiles));
    if (n_tiles == 0) continue;
    u32 sm[4];
    OK(fasta_read_summary(idx, n_tiles, sm));
    if (sm[0]) { runs += sm[3]; kind = sm[2]; }
    KeySink sink;
    sink.keys = idx->parse_keys.as<u64>(); sink.ids = pos ? idx->parse_ids.as<u64>() : nullptr; sink.cap = cap; sink.cursor = idx->ctr + CTR_CURSOR;
    u64 produced = 0;
    OK(fasta_parse_range(idx, d + b0, avail, blk, n_tiles, 0, n_tiles, sink, &produced, staged));
    staged += produced;
    total += produced;
  }
  OK(flush(staged));
  CU(cudaStreamSynchronize(idx->stream));
  idx->stage_ms[0] = 0; idx->stage_ms[1] = 0; idx->stage_ms[2] = 0;
  if (n_kmers) *n_kmers = total;
  if (runs_out) *runs_out = runs;
  if (kind_out) *kind_out = kind;
  return check_table_error(idx);
}
extern "C" int kmi_build_fasta_block(kmi_index* idx, const uint8_t* bytes, size_t n, const kmi_fasta_block* block,
                                    uint64_t* n_kmers) {
  if (!idx || (!bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  if (n_kmers) *n_kmers = 0;
  FaBlock base;
  OK(fasta_block_from_abi(block, n, &base));
  return build_fasta_host(idx, bytes, n, base, n_kmers, nullptr, nullptr);
}
extern "C" int kmi_build_fasta(kmi_index* idx, const uint8_t* bytes, size_t n, uint64_t* n_kmers) {
  return kmi_build_fasta_block(idx, bytes, n, nullptr, n_kmers);
}

// ---- f1: build straight from a file (Index::build_posix<FASTQParser|FASTAParser>, kmer_index.hpp:231-260,332-372; the
// posix_file reader of src/io/file.hpp:552-900 replaced by pread into two pinned buffers) -------------------------------
// A reader thread fills buffer c+1 from the file while the GPU builds from buffer c (whose copy to the device is itself
// pipelined with the parse). FASTQ chunks end at a record boundary (kmi_fastq_find_first_record near the chunk's end, the
// remainder is carried to the front of the next buffer); FASTA chunks are blocks of the file (the last K-1 bases stay in the
// buffer as halo and are carried too, header runs / line kind chained from chunk to chunk).
#include <fcntl.h>
#include <future>
#include <sys/stat.h>
#include <unistd.h>
namespace {
int pinned_ensure(kmi_index* idx, int i, size_t n) {
  if (n <= idx->file_buf_cap[i]) return KMI_OK;
  if (idx->file_buf[i]) cudaFreeHost(idx->file_buf[i]);
  idx->file_buf[i] = nullptr; idx->file_buf_cap[i] = 0;
  if (cudaHostAlloc(reinterpret_cast<void**>(&idx->file_buf[i]), n, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    return fail(KMI_ERR_NOMEM, "cannot pin %zu bytes of host memory", n);
  }
  idx->file_buf_cap[i] = n;
  return KMI_OK;
}
struct FdGuard { int fd; ~FdGuard() { if (fd >= 0) close(fd); } };
// bytes read (short only at the end of the file), or -1
long long read_slice(int fd, uint8_t* dst, size_t want, unsigned long long at) {
  size_t got = 0;
  while (got < want) {
    ssize_t r = pread(fd, dst + got, want - got, (off_t)(at + got));
    if (r < 0) return -1;
    if (r == 0) break;
    got += (size_t)r;
  }
  return (long long)got;
}
// large reads are cut into slices read by parallel threads (one thread copying out of the page cache is ~5 GB/s, far below
// what the PCIe copy and the build take)
long long read_at(int fd, uint8_t* dst, size_t want, unsigned long long at) {
  const size_t SLICE = 32u << 20;
  const size_t n_slices = (want + SLICE - 1) / SLICE;
  if (n_slices <= 1) return read_slice(fd, dst, want, at);
  const size_t n_thr = std::min<size_t>(n_slices, 8);
  std::vector<std::future<long long>> jobs;
  std::vector<long long> got(n_slices, 0);
  for (size_t t = 0; t < n_thr; ++t)
    jobs.push_back(std::async(std::launch::async, [=, &got]() -> long long {
      for (size_t i = t; i < n_slices; i += n_thr) {
        const size_t off = i * SLICE, len = std::min(SLICE, want - off);
        got[i] = read_slice(fd, dst + off, len, at + off);
        if (got[i] < (long long)len) break;   // error or end of file: later slices of this thread stay 0
      }
      return 0;
    }));
  for (auto& j : jobs) j.get();
  long long total = 0;
  for (size_t i = 0; i < n_slices; ++i) {   // contiguous prefix that was read
    if (got[i] < 0) return -1;
    total += got[i];
    if ((size_t)got[i] < std::min(SLICE, want - i * SLICE)) break;
  }
  return total;
}
bool has_suffix(const char* s, const char* e) {
  size_t a = strlen(s), b = strlen(e);
  return a >= b && strcmp(s + a - b, e) == 0;
}
}  // namespace

int kmi::file_format(const char* path, int* format) {
  if (*format == 0) {
    if (has_suffix(path, ".fastq") || has_suffix(path, ".fq")) *format = 1;
    else if (has_suffix(path, ".fasta") || has_suffix(path, ".fa")) *format = 2;
    else return fail(KMI_ERR_ARG, "input filename extension is not supported.");   // kmer_index.hpp:245-254
  }
  if (*format != 1 && *format != 2) return fail(KMI_ERR_ARG, "format must be 0 (by extension), 1 (FASTQ) or 2 (FASTA)");
  return KMI_OK;
}
long long kmi::file_read(int fd, uint8_t* dst, size_t want, u64 at) { return read_at(fd, dst, want, at); }

// first whole FASTQ record at or after `start` (FASTQParser::find_first_record on a growing window of the file): what
// partitioned_file::read_file does at every block boundary (file.hpp:1342-1431). `size` when no record starts after it.
static int fastq_cut_at(int fd, u64 size, u64 start, u64* cut) {
  if (start == 0 || start >= size) { *cut = std::min(start, size); return KMI_OK; }
  std::vector<uint8_t> win;
  for (u64 w = std::min<u64>(size - start, 64u << 10);; w = std::min<u64>(size - start, w * 4)) {
    win.resize(w);
    if (pread_full(fd, win.data(), (size_t)w, start) != (long long)w) return fail(KMI_ERR_ARG, "read error at offset %llu", (unsigned long long)start);
    uint64_t off = 0;
    const int r = kmi_fastq_find_first_record(win.data(), start, (size_t)w, start, start + w, 0, size, &off);
    if (r == KMI_OK && off < start + w) { *cut = off; return KMI_OK; }
    if (start + w >= size) {                               // the window reached the end of the file: no record starts after `start`
      if (r != KMI_OK) return r;
      *cut = size;
      return KMI_OK;
    }
  }
}
int kmi::file_block_fd(int fd, u64 size, int format, int rank, int nranks, u64* begin, u64* end) {
  if (nranks < 1 || rank < 0 || rank >= nranks) return fail(KMI_ERR_ARG, "rank %d of %d", rank, nranks);
  // BlockPartitioner: equal byte ranges (128-bit product: size * rank does not overflow for any file)
  const u64 b0 = (u64)((unsigned __int128)size * (unsigned)rank / (unsigned)nranks);
  const u64 e0 = (u64)((unsigned __int128)size * (unsigned)(rank + 1) / (unsigned)nranks);
  if (format == 2) { *begin = b0; *end = e0; return KMI_OK; }
  OK(fastq_cut_at(fd, size, b0, begin));
  if (rank + 1 == nranks) *end = size;
  else OK(fastq_cut_at(fd, size, e0, end));
  if (*end < *begin) *end = *begin;
  return KMI_OK;
}
int kmi::file_fasta_halo_end(int fd, u64 size, u64 valid_end, int k, u64* halo_end) {
  if (valid_end >= size) { *halo_end = size; return KMI_OK; }
  std::vector<uint8_t> win;
  for (u64 w = std::min<u64>(size - valid_end, 4096 + 2 * (u64)k);; w = std::min<u64>(size - valid_end, w * 4)) {
    win.resize(w);
    if (pread_full(fd, win.data(), (size_t)w, valid_end) != (long long)w) return fail(KMI_ERR_ARG, "read error at offset %llu", (unsigned long long)valid_end);
    uint64_t he = 0;
    OK(kmi_fasta_halo_end(win.data(), valid_end, (size_t)w, valid_end, k, &he));
    if (he < valid_end + w || valid_end + w >= size) { *halo_end = he; return KMI_OK; }   // K-1 bases found, or the file ends
  }
}
extern "C" int kmi_file_block(const char* path, int format, int rank, int nranks, uint64_t* begin, uint64_t* end, uint64_t* file_size) {
  if (!path || !begin || !end) return fail(KMI_ERR_ARG, "null argument");
  OK(file_format(path, &format));
  FdGuard f{open(path, O_RDONLY)};
  if (f.fd < 0) return fail(KMI_ERR_ARG, "cannot open %s", path);
  struct stat st;
  if (fstat(f.fd, &st) != 0) return fail(KMI_ERR_ARG, "cannot stat %s", path);
  if (file_size) *file_size = (uint64_t)st.st_size;
  u64 b = 0, e = 0;
  OK(file_block_fd(f.fd, (u64)st.st_size, format, rank, nranks, &b, &e));
  *begin = b; *end = e;
  return KMI_OK;
}

// FASTQ files when the pinned slice ring is available: no copy of the file in host memory at all. The file is cut into
// chunks at record boundaries found by reading only a chunk's last bytes (FASTQParser::find_first_record on a tail window),
// and every chunk is ONE pipelined build whose bytes the ring's threads pread straight into pinned slices (HostSrc::file):
// the reads, the H2D copies and the rounds of the build overlap from the first slice on, where the chunked reader below
// reads a whole chunk before its build starts.
static int build_file_fastq_stream(kmi_index* idx, int fd, u64 size, u64 CH, const char* path, uint64_t* n_kmers) {
  std::vector<uint8_t> tail;
  u64 pos = 0, total = 0;
  while (pos < size) {
    const u64 end = std::min<u64>(size, pos + CH), len = end - pos;
    u64 cut = end;
    if (end < size) {
      cut = 0;
      for (u64 w = std::min<u64>(len, std::max<u64>(std::min<u64>(len / 8, 1u << 20), 2048));; w = std::min<u64>(len, w * 4)) {
        tail.resize(w);
        if (pread_full(fd, tail.data(), (size_t)w, end - w) != (long long)w) return fail(KMI_ERR_ARG, "read error on %s", path);
        uint64_t off = 0;
        const int r = kmi_fastq_find_first_record(tail.data(), end - w, (size_t)w, end - w, end, 0, size, &off);
        if (r == KMI_OK && off < end && off > pos) { cut = off; break; }
        if (w == len) break;
      }
      if (cut == 0) return fail(KMI_ERR_FORMAT, "no FASTQ record boundary inside a chunk of %llu bytes at offset %llu", (unsigned long long)len, (unsigned long long)pos);
    }
    OK(idx->stage_in.ensure(cut - pos + 64));
    uint64_t produced = 0;
    OK(build_fastq_impl(idx, idx->stage_in.as<uint8_t>(), (size_t)(cut - pos), HostSrc::file(fd, pos), &produced, pos));
    total += produced;
    pos = cut;
  }
  if (n_kmers) *n_kmers = total;
  return KMI_OK;
}

extern "C" int kmi_build_file(kmi_index* idx, const char* path, int format, uint64_t* n_kmers) {
  if (!idx || !path) return fail(KMI_ERR_ARG, "null argument");
  if (n_kmers) *n_kmers = 0;
  OK(file_format(path, &format));
  CU(cudaSetDevice(idx->cfg.device));
  FdGuard f{open(path, O_RDONLY)};
  if (f.fd < 0) return fail(KMI_ERR_ARG, "cannot open %s", path);
  struct stat st;
  if (fstat(f.fd, &st) != 0) return fail(KMI_ERR_ARG, "cannot stat %s", path);
  const u64 size = (u64)st.st_size;
  u64 CH = 1ull << 30;   // measured on the 2.1 GB bench file (page cache): 134 ms with 256 MiB chunks, 120 with 1 GiB
  const bool stream = format == 1 && size > 0 && host_ring_ready(idx);
  if (stream) CH = 4ull << 30;   // (no pinned buffer of the chunk's size: only the device staging buffer grows with it)
  if (const char* e = getenv("KMI_FILE_CHUNK_BYTES")) CH = std::max<u64>(strtoull(e, nullptr, 10), 256);
  if (stream) return build_file_fastq_stream(idx, f.fd, size, CH, path, n_kmers);
  CH = std::min<u64>(CH, std::max<u64>(size, 1));
  const u64 CARRY_MAX = std::min<u64>(CH, 16ull << 20);   // a record / the K-1 tail never comes near this
  OK(pinned_ensure(idx, 0, CH + CARRY_MAX + 64));
  if (size > CH) OK(pinned_ensure(idx, 1, CH + CARRY_MAX + 64));
  struct { uint8_t* p; } buf[2] = {{idx->file_buf[0]}, {idx->file_buf[1]}};
  const int K = idx->cfg.k;
  u64 file_pos = 0;      // next file byte to read
  u64 base_off = 0;      // file offset of buf[cur][0]
  u64 carry = 0;         // bytes at the front of buf[cur] carried over from the previous chunk
  u64 total = 0;
  u32 runs = 0, kind = 0, first_is_header = 0;
  uint8_t prev_byte = '\n';
  int cur = 0;
  long long got = read_at(f.fd, buf[0].p, (size_t)std::min<u64>(CH, size), 0);
  if (got < 0) return fail(KMI_ERR_ARG, "read error on %s", path);
  file_pos = (u64)got;
  if (size) first_is_header = (buf[0].p[0] == '>' || buf[0].p[0] == ';') ? 1u : 0u;
  for (;;) {
    uint8_t* b = buf[cur].p;
    const u64 len = carry + (file_pos - (base_off + carry));   // bytes in the buffer: [base_off, file_pos)
    const bool eof = file_pos >= size;
    // -- where this chunk ends
    u64 cut = len;
    if (!eof) {
      if (format == 1) {
        cut = 0;
        for (u64 tail = std::min<u64>(len, std::max<u64>(std::min<u64>(len / 8, 1u << 20), 2048));; tail = std::min<u64>(len, tail * 4)) {
          uint64_t off = 0;
          int r = kmi_fastq_find_first_record(b, base_off, len, base_off + len - tail, base_off + len, 0, size, &off);
          if (r == KMI_OK && off < base_off + len) { cut = off - base_off; break; }
          if (tail == len) break;
        }
        if (cut == 0) return fail(KMI_ERR_FORMAT, "no FASTQ record boundary inside a chunk of %llu bytes at offset %llu", (unsigned long long)len, (unsigned long long)base_off);
      } else {
        int need = K - 1;
        while (cut > 0 && need > 0) { --cut; if (b[cut] != '\n' && b[cut] != '\r') --need; }
        if (need > 0) return fail(KMI_ERR_FORMAT, "chunk of %llu bytes holds fewer than K-1 bases", (unsigned long long)len);
      }
    }
    // -- start reading the next chunk behind the carried bytes of the other buffer
    std::future<long long> next;
    const u64 next_carry = len - cut;
    if (!eof) {
      if (next_carry > CARRY_MAX) return fail(KMI_ERR_FORMAT, "a record longer than %llu bytes", (unsigned long long)CARRY_MAX);
      uint8_t* nb = buf[cur ^ 1].p;
      memcpy(nb, b + cut, next_carry);
      const u64 want = std::min<u64>(CH, size - file_pos), at = file_pos;
      const int fd = f.fd;
      next = std::async(std::launch::async, [fd, nb, next_carry, want, at]() { return read_at(fd, nb + next_carry, (size_t)want, at); });
    }
    // -- build from this buffer
    uint64_t produced = 0;
    int r;
    if (format == 1) {
      r = kmi_build_fastq(idx, b, cut, base_off, &produced);
    } else {
      FaBlock blk;
      blk.file_offset = base_off; blk.n_valid = cut; blk.runs0 = runs; blk.kind0 = kind;
      blk.bol = (base_off == 0 || prev_byte == '\n') ? 1u : 0u;
      blk.first_is_header = first_is_header;
      r = build_fasta_host(idx, b, len, blk, &produced, &runs, &kind);
    }
    total += produced;
    if (cut > 0) prev_byte = b[cut - 1];
    long long g = 0;
    if (!eof) g = next.get();   // always collect the reader before leaving
    if (r != KMI_OK) return r;
    if (eof) break;
    if (g < 0) return fail(KMI_ERR_ARG, "read error on %s", path);
    base_off += cut;
    carry = next_carry;
    file_pos += (u64)g;
    if (g == 0 && file_pos < size) return fail(KMI_ERR_ARG, "%s is shorter than its size says", path);
    cur ^= 1;
  }
  if (n_kmers) *n_kmers = total;
  return KMI_OK;
}

// ---- f2: persistence (the reference has no on-disk index format; this defines one: what to_vector() returns,
// distributed_densehash_map.hpp:1103-1113, sorted by (key, value), one file per shard) -------------------------------------
namespace {
struct KmiFileHeader {
  char magic[8];        // "KMIDX001"
  uint32_t k, mode, nw, canonical;
  uint64_t n_entries;
  uint32_t hash, rank, nranks, flags;   // hash kind and shard of the index that wrote the file (flags bit 0: fields valid)
  uint64_t reserved[2];
};
static_assert(sizeof(KmiFileHeader) == 64, "header is 64 bytes");
struct FileGuard { FILE* f; ~FileGuard() { if (f) fclose(f); } };
}  // namespace

extern "C" int kmi_save(kmi_index* idx, const char* path) {
  if (!idx || !path) return fail(KMI_ERR_ARG, "null argument");
  uint64_t n = 0, got = 0;
  OK(kmi_local_size(idx, &n));
  const size_t nw = (size_t)idx->nw;
  std::vector<uint64_t> keys(n * nw + 1), vals(n + 1);
  OK(kmi_dump_sorted(idx, keys.data(), vals.data(), n, &got));   // (key, value) order from the device radix sort (kmi_sorted.cu)
  FileGuard g{fopen(path, "wb")};
  if (!g.f) return fail(KMI_ERR_ARG, "cannot create %s", path);
  KmiFileHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, "KMIDX001", 8);
  h.k = (uint32_t)idx->cfg.k; h.mode = (uint32_t)idx->cfg.mode; h.nw = (uint32_t)nw; h.canonical = idx->cfg.canonical ? 1u : 0u;
  h.n_entries = got;
  h.hash = (uint32_t)idx->cfg.hash; h.rank = (uint32_t)idx->cfg.rank; h.nranks = (uint32_t)idx->cfg.nranks; h.flags = 1u;
  bool ok = fwrite(&h, sizeof(h), 1, g.f) == 1;
  if (ok && got) ok = fwrite(keys.data(), 8, got * nw, g.f) == got * nw;   // keys, then values, in sorted order
  if (ok && got) ok = fwrite(vals.data(), 8, got, g.f) == got;
  if (!ok) return fail(KMI_ERR_ARG, "write error on %s", path);
  return KMI_OK;
}

extern "C" int kmi_load(kmi_index* idx, const char* path, uint64_t* n_entries) {
  if (!idx || !path) return fail(KMI_ERR_ARG, "null argument");
  if (n_entries) *n_entries = 0;
  FileGuard g{fopen(path, "rb")};
  if (!g.f) return fail(KMI_ERR_ARG, "cannot open %s", path);
  KmiFileHeader h;
  if (fread(&h, sizeof(h), 1, g.f) != 1 || memcmp(h.magic, "KMIDX001", 8) != 0) return fail(KMI_ERR_FORMAT, "%s is not a kmi index file", path);
  if ((int)h.k != idx->cfg.k || (int)h.mode != idx->cfg.mode || (int)h.nw != idx->nw)
    return fail(KMI_ERR_ARG, "%s holds a K=%u mode-%u index, this one is K=%d mode-%d", path, h.k, h.mode, idx->cfg.k, idx->cfg.mode);
  // A canonical index stores lex_less keys: loading non-canonical entries into it would merge strands silently, and
  // entries of a canonical file are not what a non-canonical index would hold. A shard file belongs to owner `rank` of
  // `nranks` under `hash`: it may be loaded into a single-shard index (merge) or into the same shard of the same layout.
  if ((h.canonical != 0) != (idx->cfg.canonical != 0))
    return fail(KMI_ERR_ARG, "%s holds %s keys, this index is %s", path, h.canonical ? "canonical" : "non-canonical", idx->cfg.canonical ? "canonical" : "non-canonical");
  if ((h.flags & 1u) && idx->cfg.nranks > 1 &&
      ((int)h.nranks != idx->cfg.nranks || (int)h.rank != idx->cfg.rank || (int)h.hash != idx->cfg.hash))
    return fail(KMI_ERR_ARG, "%s is shard %u of %u (hash %u); this index is shard %d of %d (hash %d)", path, h.rank, h.nranks, h.hash,
                idx->cfg.rank, idx->cfg.nranks, idx->cfg.hash);
  const size_t nw = (size_t)idx->nw;
  const uint64_t CH = 1u << 24;                      // entries per insert
  std::vector<uint64_t> keys, vals;
  for (uint64_t done = 0; done < h.n_entries; done += CH) {
    const uint64_t m = std::min<uint64_t>(CH, h.n_entries - done);
    keys.resize(m * nw); vals.resize(m);
    if (fseeko(g.f, (off_t)(sizeof(h) + done * nw * 8), SEEK_SET) != 0 || fread(keys.data(), 8, m * nw, g.f) != m * nw ||
        fseeko(g.f, (off_t)(sizeof(h) + h.n_entries * nw * 8 + done * 8), SEEK_SET) != 0 || fread(vals.data(), 8, m, g.f) != m)
      return fail(KMI_ERR_FORMAT, "%s is truncated", path);
    OK(kmi_insert(idx, keys.data(), vals.data(), m));   // COUNT: counts are added; POSITION: ids appended
  }
  if (n_entries) *n_entries = h.n_entries;
  return KMI_OK;
}
