// kmi_internal.cuh — what the translation units of libkmi.so share: the index object, device scratch buffers, the device
// counter block and the handful of host-side helpers that the single-shard API (kmi_api.cu) and the sharded layer
// (kmi_dist.cu) both use. Not part of the ABI (include/kmi.h is).
#pragma once
#include "../../include/kmi.h"
#include "kmi_ops.cuh"

#include <cstdarg>
#include <cstdio>
#include <vector>

namespace kmi {
int fail(int code, const char* fmt, ...);   // sets the thread-local message of kmi_last_error(), returns `code`
}
#define CU(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) return fail(KMI_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)
#define OK(call)            \
  do {                      \
    int r_ = (call);        \
    if (r_ != KMI_OK) return r_; \
  } while (0)

namespace kmi {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return KMI_OK;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { p = nullptr; return fail(KMI_ERR_NOMEM, "cudaMalloc(%zu): %s", want, cudaGetErrorString(e)); }
    cap = want;
    return KMI_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

// device counters block
enum { CTR_SIZE = 0, CTR_NKMERS = 1, CTR_CURSOR = 2, CTR_CK0 = 3, CTR_CK1 = 4, CTR_REMOVED = 5, CTR_ERR = 6, CTR_PERR = 7, CTR_CARRY = 8, CTR_FIRST_LS = 9, CTR_N = 10 };

// bkt_meta layout (u32 words): cursors [0, N_BUCKETS], HLL registers from META_HLL
enum { META_HLL = 512, META_WORDS = META_HLL + HLL_M };

}  // namespace kmi

struct kmi_dist_state;   // kmi_dist.cu
struct kmi_index;
namespace kmi { int flush_pending(kmi_index* idx); }   // kmi_api.cu: the deferred segment insert of bulk-load mode
namespace kmi { struct HostRing; }                     // kmi_api.cu: pinned slice ring for pageable host memory (kmi_hoststage.hpp)

struct kmi_index {
  kmi_config cfg;
  kmi_dist_state* dist = nullptr;   // sharded layer (kmi_dist_attach), owned by the index
  const kmi::Ops* ops = nullptr;
  cudaStream_t stream = 0;
  int nw = 1;
  size_t slot_bytes = 16;
  // table
  void* slots = nullptr;
  kmi::u64 cap = 0;             // number of slots (power of two)
  kmi::u64* ctr = nullptr;      // device counters (CTR_N x kmi::u64)
  kmi::u64 size_exact = 0;      // last size read back
  kmi::u64 size_upper = 0;      // upper bound on the current size (exact + keys issued since)
  kmi::u64 tombstones = 0;      // DELETED slots in the table: they occupy probe sequences like live keys (dense_hash_map counts
                           // them in its resize decision too); dropped by the next rehash
  // scratch
  kmi::DevBuf mid_keys, mid_cur;        // first of two fine passes (very large tables)
  kmi::DevBuf parse_keys, parse_ids;   // staging of the parse -> insert builds (FASTA, FASTQ positions): kept between calls
  kmi::DevBuf tile_rec;   // FASTQ position ids: record start open at every tile (k_fastq_rec_carry)
  kmi::DevBuf tile_lines, block_sums, stage_in, stage_keys, stage_vals, stage_out, stage_found, hist, hist_sums, scan_tmp;
  // POSITION mode: tuple log + CSR
  kmi::DevBuf log_keys, log_ids, log_keys2, log_ids2, offsets, csr_sums, values;
  kmi::u64 log_n = 0;
  bool csr_dirty = false;
  // hash-prefix buckets (kmi_bucket.cuh): key regions, cursors + HyperLogLog registers
  kmi::DevBuf bkt_keys, bkt_meta, fine_keys, fine_cur;
  kmi::u32 h_ovf_cursor = 0;       // last read-back of the overflow region's cursor (bucket_readback)
  int seg_path = 1;        // KMI_PARAM_SEGMENT_INSERT: 1 = shared-memory segment insert where possible, 0 = L2 atomics
  bool hll_ready = false;
  kmi::u64 non_hll_upper = 0;   // keys inserted by paths that do not feed the HLL (upper bound on their distinct keys)
  // KMI_PARAM_DEFER_INSERT (bulk load): fine buckets of several rounds / calls staged in fine_keys, inserted with ONE table sweep
  kmi::u64 defer_budget = 0;    // 0 = off; else the number of keys the staging area is laid out for (clamped by free memory)
  kmi::u64 pend_keys = 0;       // keys staged, not yet in the table
  kmi::u64 pend_budget = 0;     // keys the current staging layout holds (the budget, clamped by free memory)
  kmi::u64 pend_req = 0;        // the budget that layout was made for
  int pend_log_segs = -1;       // the layout: 2^pend_log_segs regions, cut for a table of pend_cap slots
  kmi::u64 pend_cap = 0;
  kmi::u32 pend_capf = 0, pend_ovf_cap = 0;
  kmi::DevBuf pend_ovf, pend_ovf_cur, pend_cur_save;   // spill list of the deferred insert; cursor snapshot for roll-back
  kmi::u64 flushes = 0;         // deferred segment inserts run so far
  double max_load = 0.7;   // KMI_PARAM_MAX_LOAD: the table is sized so that live + deleted keys stay below this fraction of its slots
  kmi::u32 dist_trans = 0; // KMI_PARAM_DIST_TRANS: transform under the dist hash of a non-canonical index (0 iden, 1 lex_less, 2 xor)
  bool saturate = false;   // KMI_PARAM_SATURATE: counts stick at 2^32 - 1; such an index builds with direct atomics (table_add)
  int build_path = 0;      // kmi_param KMI_PARAM_BUILD_PATH: 0 auto, 1 direct atomics, 2 bucketed
  // accounting
  kmi::u64 launches = 0;
  float stage_ms[3] = {0, 0, 0};
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  float scatter_ms = 0;
  cudaEvent_t kev[4] = {nullptr, nullptr, nullptr, nullptr};  // around the kernels of bucket_commit
  bool kev_pending = false;
  float kernel_ms[5] = {0, 0, 0, 0, 0};  // last build: total, parse+scatter, fine scatter, segment/L2 insert, overflow insert
  uint8_t* file_buf[2] = {nullptr, nullptr};   // pinned buffers of kmi_build_file (kept between calls: pinning is slow)
  size_t file_buf_cap[2] = {0, 0};
  cudaStream_t copy_stream = nullptr;        // H2D copies of the pipelined host-pointer build
  kmi::HostRing* hring = nullptr;            // pageable host memory goes through this ring of pinned slices (created on first use)
  bool hring_tried = false;
  std::vector<cudaEvent_t> copy_events;    // parse+scatter share of the last bucketed build

  // Lazy clear: `virgin` = the table is logically empty but its memory has not been written. The segment insert of a
  // bucketed build initialises every segment in shared memory and writes all of them, so a build into an empty table
  // never pays the 16 B/slot clear nor the read of empty segments; everything else goes through view(), which clears first.
  bool virgin = false;
  void touch() {
    if (pend_keys) kmi::flush_pending(this);   // bulk-load mode: whatever looks at the table sees every staged key
    if (virgin && slots) { ops->table_clear(stream, slots, cap); launches++; }
    virgin = false;
  }
  kmi::TableView view() { touch(); return raw_view(); }
  kmi::TableView raw_view() const {
    kmi::TableView t;
    t.slots = slots; t.mask = cap - 1; t.shift = 64u - (kmi::u32)__builtin_ctzll(cap ? cap : 1); t.size = ctr + kmi::CTR_SIZE; t.saturate = saturate ? 1u : 0u; t.error = reinterpret_cast<kmi::u32*>(ctr + kmi::CTR_ERR);
    return t;
  }
};


// ---- helpers defined in kmi_api.cu ---------------------------------------------------------------------------------
namespace kmi {
static const u64 BUCKET_MAX_KEYS = 1ULL << 31;   // per bucket round; keeps every cursor and their sum below 2^32
static const size_t AUTO_BUCKET_KEYS = 1u << 21; // below this the direct path wins (bucket set-up dominates)
static const size_t AUTO_BUCKET_BYTES = 8u << 20;
int read_ctr(kmi_index* idx, int which, u64* out);
int zero_ctr(kmi_index* idx, int which, int count = 1);
int table_error(u64 e);
int check_table_error(kmi_index* idx);
int parse_error(u64 perr);
int bucket_meta_prepare(kmi_index* idx);                         // cursors zeroed, HLL registers kept (zeroed on first use)
int bucket_readback(kmi_index* idx, const BucketSink& bs, u64* host_ctr, std::vector<u32>& reg);
// not_inserted (needed when bs.sub > 1): set when the table is too small for the segment path — nothing was inserted and the
// caller must insert the regions as key streams (insert_streams_bucketed)
int bucket_commit(kmi_index* idx, const BucketSink& bs, u64 n_keys, u64 size_now, const std::vector<u32>& reg, bool* not_inserted = nullptr);
int insert_streams_bucketed(kmi_index* idx, const u64* const* d_keys, const u64* counts, int n_streams, bool canon);
int fastq_scan_alloc(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64* n_tiles_out);
int fastq_scan_range(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64 t0, u64 t1);
int bucket_prepare(kmi_index* idx, u64 max_keys, BucketSink* bs);
int fastq_scan(kmi_index* idx, const uint8_t* d_bytes, size_t n, u64* n_tiles_out);
void harvest_commit_times(kmi_index* idx);
// in-place exclusive scan of n u64 values on the index's stream (two levels above 8192 values)
int scan64_dev(kmi_index* idx, u64* d_v, u64 n);
// POSITION index, everything left on the device: for m keys AS GIVEN (no transform) d_off[0..m] = exclusive scan of their
// multiplicities, vals = the ids of every hit in query order (grown as needed), *total = d_off[m] (read back: one sync)
int find_positions_dev(kmi_index* idx, const u64* d_keys, u64 m, u64* d_off, DevBuf& slot_idx, DevBuf& vals, u64* total);
// ---- host <-> device copies (kmi_api.cu). Large copies whose host side is PAGEABLE go through the index's ring of pinned
// slices, filled / drained by host threads (kmi_hoststage.hpp); pinned or small ones are plain cudaMemcpyAsync.
// copy_to_device: on return the caller's memory has been read (or is pinned and the copy enqueued) and idx->stream is ordered
// behind the copy. copy_to_host: like cudaMemcpyAsync(.., idx->stream) for the caller — the data is valid after the next
// cudaStreamSynchronize(idx->stream) (a staged copy has in fact completed on return).
int copy_to_device(kmi_index* idx, void* d_dst, const void* src, size_t bytes);
int copy_to_host(kmi_index* idx, void* dst, const void* d_src, size_t bytes);
void host_ring_release(kmi_index* idx);
// where the bytes of a host-side build come from: the caller's memory, or a file read with pread straight into the pinned
// slices (kmi_build_file: no intermediate copy of the file in host memory; needs the ring — host_ring_ready)
struct HostSrc {
  const uint8_t* mem = nullptr;
  int fd = -1;
  u64 fd_off = 0;                                   // file offset of byte 0 of the source
  HostSrc() {}
  HostSrc(const uint8_t* m) : mem(m) {}             // (implicit: every host-pointer entry point passes its pointer)
  static HostSrc file(int fd_, u64 off) { HostSrc s; s.fd = fd_; s.fd_off = off; return s; }
  explicit operator bool() const { return mem != nullptr || fd >= 0; }
};
// file helpers of kmi_build_file, shared with the sharded file build (kmi_dist.cu)
int file_format(const char* path, int* format);                          // 0 = by extension -> 1 FASTQ / 2 FASTA (KMI_ERR_ARG)
long long file_read(int fd, uint8_t* dst, size_t want, u64 at);          // parallel pread; bytes read (short at EOF) or -1
int file_block_fd(int fd, u64 size, int format, int rank, int nranks, u64* begin, u64* end);
int file_fasta_halo_end(int fd, u64 size, u64 valid_end, int k, u64* halo_end);
bool host_ring_ready(kmi_index* idx);               // the ring exists (or can be created) and no feed is running on it
int copy_to_device(kmi_index* idx, void* d_dst, const HostSrc& src, size_t bytes);
// HostFeed: the host -> device transfer of a pipelined build. Range i = bytes [ends[i-1], ends[i]) of `src` goes to the same
// offsets of d_dst on idx->copy_stream (created on demand, ordered behind the work already queued on `after`), and
// idx->copy_events[i] is recorded behind it. wait(i, s) makes stream s wait for range i (blocking the host until the range
// has been enqueued when the source is pageable and therefore staged by threads). The destructor never returns while a copy
// from the caller's memory is in flight.
class HostFeed {
 public:
  HostFeed() = default;
  HostFeed(const HostFeed&) = delete;
  HostFeed& operator=(const HostFeed&) = delete;
  ~HostFeed();
  int start(kmi_index* idx, uint8_t* d_dst, const HostSrc& src, const std::vector<u64>& ends, cudaStream_t after);
  int wait(size_t i, cudaStream_t consumer);
 private:
  struct Impl;
  Impl* impl_ = nullptr;
  kmi_index* idx_ = nullptr;
};
// kmi_sorted.cu: stable LSD radix sort of (key, value) pairs; keys/vals ping-pong with their *_tmp twins (see there)
int device_sort_pairs(cudaStream_t st, DevBuf& hist, DevBuf& sums, u64*& keys, u64*& vals, u64*& keys_tmp, u64*& vals_tmp, u64 n,
                      int nw, int key_bits, int val_bits, u64* launches);
}  // namespace kmi
void kmi_dist_release(kmi_index* idx);   // kmi_dist.cu: frees idx->dist
