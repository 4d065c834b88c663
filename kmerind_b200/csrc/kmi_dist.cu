// kmi_dist.cu — the sharded index: one process per GPU, keys owned by dist_hash(key) % nranks.
//
// C++ counterpart of the reference's distributed maps: every batch operation is COLLECTIVE — all ranks call it, also with
// empty input (distributed_densehash_map.hpp:2567-2573) — and consists of a local CUDA stage of libkmi.so plus one
// exchange step (imxx::distribute, src/io/incremental_mxx.hpp:1039-1109):
//   insert   reduction_densehash_map::insert   distributed_densehash_map.hpp:2565-2628, multimap :2329-2384
//   count    densehash_map_base::count          :1122-1216      find  :327-448      erase  :1602-1663
//   size     map_base::size (allreduce)         distributed_map_base.hpp:243
// Plumbing: NCCL (loaded with dlopen at the first kmi_comm_* call, so single-GPU users need no libnccl) for the small
// collectives and the generic all-to-all; the count-index BUILD moves its keys with k_owner_push instead — plain stores into
// the owners' receive buffers (CUDA IPC mappings, NVLink / NVSwitch), overlapped with the parsing of the next round.
#define KMI_DIST_KERNELS
#include "kmi_internal.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>

using namespace kmi;

// kmi::u64 is unsigned long long (CUDA's 64-bit atomics), the ABI speaks uint64_t: same size, different type
static inline const uint64_t* A(const u64* p) { return reinterpret_cast<const uint64_t*>(p); }
static inline uint64_t* A(u64* p) { return reinterpret_cast<uint64_t*>(p); }
static inline const u64* B(const uint64_t* p) { return reinterpret_cast<const u64*>(p); }

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

// ---- NCCL, bound at run time ---------------------------------------------------------------------------------------
namespace {
struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
} g_nccl;

int nccl_load() {
  if (g_nccl.lib) return KMI_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
  if (!h) return fail(KMI_ERR_UNSUPPORTED, "cannot load libnccl.so.2 (%s): the sharded index needs NCCL", dlerror());
#define KMI_SYM(field, name)                                                         \
  *reinterpret_cast<void**>(&g_nccl.field) = dlsym(h, name);                          \
  if (!g_nccl.field) return fail(KMI_ERR_UNSUPPORTED, "libnccl lacks %s", name);
  KMI_SYM(GetUniqueId, "ncclGetUniqueId") KMI_SYM(CommInitRank, "ncclCommInitRank") KMI_SYM(CommDestroy, "ncclCommDestroy")
  KMI_SYM(AllReduce, "ncclAllReduce") KMI_SYM(AllGather, "ncclAllGather") KMI_SYM(Send, "ncclSend") KMI_SYM(Recv, "ncclRecv")
  KMI_SYM(GroupStart, "ncclGroupStart") KMI_SYM(GroupEnd, "ncclGroupEnd") KMI_SYM(GetErrorString, "ncclGetErrorString")
#undef KMI_SYM
  g_nccl.lib = h;
  return KMI_OK;
}
}  // namespace
#define NC(call)                                                                                                \
  do {                                                                                                          \
    ncclResult_t r_ = (call);                                                                                   \
    if (r_ != ncclSuccess) return fail(KMI_ERR_CUDA, "%s: %s (%s:%d)", #call, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
  } while (0)

struct kmi_comm {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1, device = 0;
  u64* d_tmp = nullptr;                 // device scratch for the small collectives (TMP_WORDS u64)
  static constexpr size_t TMP_WORDS = 1 << 17;
};

// ---- state of one sharded index ----------------------------------------------------------------------------------------
struct kmi_dist_state {
  kmi_comm* comm = nullptr;
  // NVLink push path. Receive block of this rank (one cudaMalloc, exported with CUDA IPC):
  //   [counts: N_BUCKETS * p + 1 u32 (the last one is the overflow cursor), padded to HDR_BYTES]
  //   [N_BUCKETS * p regions of cap_sub keys: region prefix * p + source][overflow region of ovf_cap keys]
  static constexpr size_t HDR_BYTES = 16384;
  bool peer_ok = false;                 // every rank could map every receive block
  void* recv = nullptr;
  u64 cap_sub = 0, ovf_cap = 0;
  std::vector<void*> peer_block;        // base of rank r's receive block in this process
  DevBuf peer_keys, peer_cnt;           // device arrays of p pointers (regions / count words of every rank's block)
  DevBuf push_cursor;                   // N_BUCKETS * p u32, local
  DevBuf status;                        // 2 x u32 in, 2 x u32 out: [0] a region overflowed somewhere, [1] malformed input somewhere
  DevBuf stage_keys[2], stage_meta[2];  // double-buffered staging buckets of the parser (local)
  cudaStream_t push_stream = nullptr;
  cudaEvent_t ev_parsed[2] = {nullptr, nullptr}, ev_pushed[2] = {nullptr, nullptr}, ev_t[4] = {nullptr, nullptr, nullptr, nullptr};
  // generic exchange scratch
  DevBuf part_keys, part_vals, recv_keys, recv_vals, q_keys, q_tags, q_vals, q_found, q_back, q_backf, text, keys, ids;
  DevBuf fp_cnt, fp_slot, fp_part_off, fp_out_off, fp_out_vals;   // sharded find_positions: multiplicities, slots, the two offset arrays, the CSR values
  u64 n_received = 0;
  float ms[4] = {0, 0, 0, 0};           // last build: [0] whole, [1] parse + push + exchange, [2] insert of received keys, [3] unused
  int last_path = 0;                    // 1 = NVLink push, 2 = generic NCCL exchange
};

namespace {

int comm_allreduce(kmi_comm* c, u64* host_vals, int n, ncclRedOp_t op, cudaStream_t st) {
  if ((size_t)n > kmi_comm::TMP_WORDS / 2) return fail(KMI_ERR_ARG, "allreduce of %d words", n);
  CU(cudaMemcpyAsync(c->d_tmp, host_vals, sizeof(u64) * n, cudaMemcpyHostToDevice, st));
  NC(g_nccl.AllReduce(c->d_tmp, c->d_tmp + n, (size_t)n, ncclUint64, op, c->comm, st));
  CU(cudaMemcpyAsync(host_vals, c->d_tmp + n, sizeof(u64) * n, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return KMI_OK;
}
// every rank contributes n words; out holds nranks * n words in rank order
int comm_allgather(kmi_comm* c, const u64* mine, int n, u64* out, cudaStream_t st) {
  if ((size_t)n * (c->nranks + 1) > kmi_comm::TMP_WORDS) return fail(KMI_ERR_ARG, "allgather of %d words", n);
  CU(cudaMemcpyAsync(c->d_tmp, mine, sizeof(u64) * n, cudaMemcpyHostToDevice, st));
  NC(g_nccl.AllGather(c->d_tmp, c->d_tmp + n, (size_t)n, ncclUint64, c->comm, st));
  CU(cudaMemcpyAsync(out, c->d_tmp + n, sizeof(u64) * n * c->nranks, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return KMI_OK;
}

// mxx::all2allv: rows of `row_bytes` bytes, grouped by destination rank in `d_send` (send_counts[r] rows for rank r);
// recv_counts[s] rows arrive from rank s, grouped by source in rank order
int exchange_rows(kmi_dist_state* d, cudaStream_t st, const void* d_send, const u64* send_counts, const u64* recv_counts, size_t row_bytes,
                  void* d_recv) {
  kmi_comm* c = d->comm;
  const char* s = reinterpret_cast<const char*>(d_send);
  char* r = reinterpret_cast<char*>(d_recv);
  NC(g_nccl.GroupStart());
  for (int i = 0; i < c->nranks; ++i) {
    if (send_counts[i]) NC(g_nccl.Send(s, send_counts[i] * row_bytes, ncclUint8, i, c->comm, st));
    if (recv_counts[i]) NC(g_nccl.Recv(r, recv_counts[i] * row_bytes, ncclUint8, i, c->comm, st));
    s += send_counts[i] * row_bytes;
    r += recv_counts[i] * row_bytes;
  }
  NC(g_nccl.GroupEnd());
  return KMI_OK;
}
// mxx::all2all of the counts (incremental_mxx.hpp:1087)
int exchange_counts(kmi_dist_state* d, cudaStream_t st, const u64* send_counts, u64* recv_counts) {
  kmi_comm* c = d->comm;
  const int p = c->nranks;
  std::vector<u64> all((size_t)p * p);
  OK(comm_allgather(c, send_counts, p, all.data(), st));
  for (int s = 0; s < p; ++s) recv_counts[s] = all[(size_t)s * p + c->rank];
  return KMI_OK;
}

kmi_dist_state* dist_of(kmi_index* idx) { return idx ? idx->dist : nullptr; }

void peers_close(kmi_index* idx) {
  kmi_dist_state* d = idx->dist;
  for (size_t r = 0; r < d->peer_block.size(); ++r)
    if (d->peer_block[r] && (int)r != d->comm->rank) cudaIpcCloseMemHandle(d->peer_block[r]);
  d->peer_block.clear();
  if (d->recv) cudaFree(d->recv);
  d->recv = nullptr;
  d->cap_sub = d->ovf_cap = 0;
}

// (re)allocate the receive blocks for cap_sub keys per (prefix, source) region and map every peer's. Collective; the
// arguments must be the same on every rank.
int peers_ensure(kmi_index* idx, u64 cap_sub, u64 ovf_cap) {
  kmi_dist_state* d = idx->dist;
  kmi_comm* c = d->comm;
  const int p = c->nranks, me = c->rank;
  if (d->recv && d->cap_sub >= cap_sub && d->ovf_cap >= ovf_cap) return KMI_OK;
  CU(cudaStreamSynchronize(idx->stream));
  CU(cudaStreamSynchronize(d->push_stream));
  peers_close(idx);
  cap_sub = (cap_sub + cap_sub / 8 + 1) & ~1ULL;          // some room to grow without another collective re-mapping; even
  const size_t nw = (size_t)idx->nw;
  const size_t bytes = kmi_dist_state::HDR_BYTES + ((size_t)N_BUCKETS * p * cap_sub + ovf_cap) * sizeof(u64) * nw;
  cudaError_t e = cudaMalloc(&d->recv, bytes);
  u64 ok = e == cudaSuccess ? 1 : 0;
  if (!ok) { d->recv = nullptr; cudaGetLastError(); }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  u64 mine[9] = {0, 0, 0, 0, 0, 0, 0, 0, ok};
  if (ok && p > 1) {
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, d->recv) != cudaSuccess) { mine[8] = 0; cudaGetLastError(); }
    else memcpy(mine, &h, sizeof(h));
  }
  std::vector<u64> all((size_t)9 * p);
  OK(comm_allgather(c, mine, 9, all.data(), idx->stream));
  bool every = true;
  for (int r = 0; r < p; ++r) every = every && all[(size_t)r * 9 + 8] == 1;
  u64 mapped = every ? 1 : 0;
  d->peer_block.assign(p, nullptr);
  if (every) {
    for (int r = 0; r < p && mapped; ++r) {
      void* base = d->recv;
      if (r != me) {
        cudaIpcMemHandle_t h;
        memcpy(&h, &all[(size_t)r * 9], sizeof(h));
        if (cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); mapped = 0; base = nullptr; }
      }
      d->peer_block[r] = base;
    }
  }
  OK(comm_allreduce(c, &mapped, 1, ncclMin, idx->stream));   // every rank leaves the push path together
  if (!mapped) {
    peers_close(idx);
    d->peer_ok = false;
    return ok ? KMI_OK : fail(KMI_ERR_NOMEM, "cudaMalloc receive block %zu bytes", bytes);
  }
  d->cap_sub = cap_sub; d->ovf_cap = ovf_cap;
  std::vector<u64*> keys(p);
  std::vector<u32*> cnts(p);
  for (int r = 0; r < p; ++r) {
    cnts[r] = reinterpret_cast<u32*>(d->peer_block[r]);
    keys[r] = reinterpret_cast<u64*>(reinterpret_cast<char*>(d->peer_block[r]) + kmi_dist_state::HDR_BYTES);
  }
  OK(d->peer_keys.ensure(sizeof(u64*) * p));
  OK(d->peer_cnt.ensure(sizeof(u32*) * p));
  CU(cudaMemcpy(d->peer_keys.p, keys.data(), sizeof(u64*) * p, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(d->peer_cnt.p, cnts.data(), sizeof(u32*) * p, cudaMemcpyHostToDevice));
  OK(d->push_cursor.ensure(sizeof(u32) * N_BUCKETS * p));
  return KMI_OK;
}

// staging sink of one parse round (local coarse buckets, same layout as the single-GPU build's)
int stage_sink(kmi_index* idx, int which, u64 max_keys, BucketSink* bs) {
  kmi_dist_state* d = idx->dist;
  const u64 capb = (max_keys / N_BUCKETS + max_keys / (N_BUCKETS * 16) + 1024 + 7) & ~7ULL;
  const u64 ovf = max_keys / 8 + 4096;
  OK(d->stage_keys[which].ensure((capb * N_BUCKETS + ovf) * sizeof(u64) * idx->nw));
  OK(d->stage_meta[which].ensure(sizeof(u32) * (N_BUCKETS + 1)));
  *bs = BucketSink();
  bs->keys = d->stage_keys[which].as<u64>();
  bs->cursor = d->stage_meta[which].as<u32>();
  bs->capb = (u32)capb; bs->ovf_cap = (u32)ovf; bs->hll = nullptr; bs->bases = nullptr;
  return KMI_OK;
}

// keys (+ values) on the device, any owner -> canonical -> stable partition by owner -> exchange -> what this rank owns.
// Outputs live in d->recv_keys / d->recv_vals (n_recv rows); send order bookkeeping (counts) is returned for the way back.
int distribute(kmi_index* idx, const u64* d_keys, const u64* d_vals, size_t n, bool canonicalize, std::vector<u64>& send_counts,
               std::vector<u64>& recv_counts, u64* n_recv_out) {
  kmi_dist_state* d = idx->dist;
  const int p = d->comm->nranks;
  const size_t kb = sizeof(u64) * idx->nw;
  send_counts.assign(p, 0); recv_counts.assign(p, 0);
  OK(d->part_keys.ensure(std::max<size_t>(n, 1) * kb));
  if (d_vals) OK(d->part_vals.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
  const u64* src = d_keys;
  if (canonicalize && n) {
    OK(d->keys.ensure(n * kb));
    OK(kmi_canonicalize_dev(idx, A(d_keys), d->keys.as<uint64_t>(), n));
    src = d->keys.as<u64>();
  }
  OK(kmi_partition_dev(idx, A(src), A(d_vals), n, d->part_keys.as<uint64_t>(), d_vals ? d->part_vals.as<uint64_t>() : nullptr, A(send_counts.data())));
  OK(exchange_counts(d, idx->stream, send_counts.data(), recv_counts.data()));
  u64 n_recv = 0;
  for (int s = 0; s < p; ++s) n_recv += recv_counts[s];
  OK(d->recv_keys.ensure(std::max<u64>(n_recv, 1) * kb));
  OK(exchange_rows(d, idx->stream, d->part_keys.p, send_counts.data(), recv_counts.data(), kb, d->recv_keys.p));
  if (d_vals) {
    OK(d->recv_vals.ensure(std::max<u64>(n_recv, 1) * sizeof(u64)));
    OK(exchange_rows(d, idx->stream, d->part_vals.p, send_counts.data(), recv_counts.data(), sizeof(u64), d->recv_vals.p));
  }
  *n_recv_out = n_recv;
  return KMI_OK;
}

// insert rows this rank owns (already canonical) without running the input transform again
int insert_owned(kmi_index* idx, const u64* d_keys, const u64* d_vals, u64 n) {
  const int32_t canon = idx->cfg.canonical;
  idx->cfg.canonical = 0;
  const int r = kmi_insert_dev(idx, A(d_keys), A(d_vals), n);
  idx->cfg.canonical = canon;
  return r;
}

// cnt[i] = off[i + 1] - off[i]
__global__ void k_adj_diff(const u64* __restrict__ off, u64* __restrict__ cnt, u64 n) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) cnt[i] = off[i + 1] - off[i];
}
// sharded find_positions: query i (partition order, input position tags[i]) got cnt[i] ids at src_off[i] of the stream that
// came back; they belong at dst_off[tags[i]] of the caller's CSR
__global__ void k_place_ids(const u64* __restrict__ tags, const u64* __restrict__ cnt, const u64* __restrict__ src_off,
                            const u64* __restrict__ dst_off, const u64* __restrict__ got, u64* __restrict__ out, u64 n) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    const u64 c = cnt[i], s0 = src_off[i], d0 = dst_off[tags[i]];
    for (u64 j = 0; j < c; ++j) out[d0 + j] = got[s0 + j];
  }
}
__global__ void k_iota(u64* out, u64 n) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) out[i] = i;
}
// answers come back in the order the (partitioned) queries were sent: out[tag[i]] = back[i]
__global__ void k_unpermute(const u64* __restrict__ tags, const u64* __restrict__ back, const uint8_t* __restrict__ backf,
                            u64* __restrict__ out, uint8_t* __restrict__ outf, u64 n) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    const u64 t = tags[i];
    if (out) out[t] = back[i];
    if (outf) outf[t] = backf[i];
  }
}

}  // namespace

void kmi_dist_release(kmi_index* idx) {
  kmi_dist_state* d = idx->dist;
  if (!d) return;
  cudaSetDevice(idx->cfg.device);
  cudaStreamSynchronize(idx->stream);
  if (d->push_stream) cudaStreamSynchronize(d->push_stream);
  peers_close(idx);
  DevBuf* bufs[] = {&d->peer_keys, &d->peer_cnt, &d->push_cursor, &d->status, &d->stage_keys[0], &d->stage_keys[1], &d->stage_meta[0],
                    &d->stage_meta[1], &d->part_keys, &d->part_vals, &d->recv_keys, &d->recv_vals, &d->q_keys, &d->q_tags, &d->q_vals,
                    &d->q_found, &d->q_back, &d->q_backf, &d->text, &d->keys, &d->ids, &d->fp_cnt, &d->fp_slot, &d->fp_part_off,
                    &d->fp_out_off, &d->fp_out_vals};
  for (DevBuf* b : bufs) b->release();
  for (int i = 0; i < 2; ++i) { if (d->ev_parsed[i]) cudaEventDestroy(d->ev_parsed[i]); if (d->ev_pushed[i]) cudaEventDestroy(d->ev_pushed[i]); }
  for (int i = 0; i < 4; ++i) if (d->ev_t[i]) cudaEventDestroy(d->ev_t[i]);
  if (d->push_stream) cudaStreamDestroy(d->push_stream);
  delete d;
  idx->dist = nullptr;
}

// ---- communicator ------------------------------------------------------------------------------------------------------
extern "C" int kmi_comm_unique_id(uint8_t* id_out) {
  if (!id_out) return fail(KMI_ERR_ARG, "null argument");
  OK(nccl_load());
  static_assert(sizeof(ncclUniqueId) == KMI_COMM_ID_BYTES, "NCCL unique id size");
  ncclUniqueId id;
  NC(g_nccl.GetUniqueId(&id));
  memcpy(id_out, &id, sizeof(id));
  return KMI_OK;
}

extern "C" int kmi_comm_create(const uint8_t* id_bytes, int nranks, int rank, int device, kmi_comm** out) {
  if (!id_bytes || !out) return fail(KMI_ERR_ARG, "null argument");
  *out = nullptr;
  if (nranks < 1 || rank < 0 || rank >= nranks) return fail(KMI_ERR_ARG, "bad rank %d / nranks %d", rank, nranks);
  OK(nccl_load());
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(KMI_ERR_CUDA, "no CUDA device: this library has no CPU path");
  if (device < 0 || device >= ndev) return fail(KMI_ERR_ARG, "device %d of %d", device, ndev);
  CU(cudaSetDevice(device));
  kmi_comm* c = new (std::nothrow) kmi_comm();
  if (!c) return fail(KMI_ERR_NOMEM, "host allocation");
  c->rank = rank; c->nranks = nranks; c->device = device;
  ncclUniqueId id;
  memcpy(&id, id_bytes, sizeof(id));
  ncclResult_t r = g_nccl.CommInitRank(&c->comm, nranks, id, rank);
  if (r != ncclSuccess) { delete c; return fail(KMI_ERR_CUDA, "ncclCommInitRank: %s", g_nccl.GetErrorString(r)); }
  if (cudaMalloc(&c->d_tmp, sizeof(u64) * kmi_comm::TMP_WORDS) != cudaSuccess) { g_nccl.CommDestroy(c->comm); delete c; return fail(KMI_ERR_NOMEM, "cudaMalloc"); }
  *out = c;
  return KMI_OK;
}

extern "C" void kmi_comm_destroy(kmi_comm* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  if (c->comm) g_nccl.CommDestroy(c->comm);
  if (c->d_tmp) cudaFree(c->d_tmp);
  delete c;
}
extern "C" int kmi_comm_rank(const kmi_comm* c) { return c ? c->rank : -1; }
extern "C" int kmi_comm_size(const kmi_comm* c) { return c ? c->nranks : -1; }
extern "C" int kmi_comm_allreduce_u64(kmi_comm* c, uint64_t* vals, int n, int op) {
  if (!c || (!vals && n)) return fail(KMI_ERR_ARG, "null argument");
  if (op < 0 || op > 2) return fail(KMI_ERR_ARG, "op must be 0 (sum), 1 (max) or 2 (min)");
  CU(cudaSetDevice(c->device));
  const int CH = (int)(kmi_comm::TMP_WORDS / 2);
  for (int at = 0; at < n; at += CH)
    OK(comm_allreduce(c, reinterpret_cast<u64*>(vals) + at, std::min(CH, n - at), op == 0 ? ncclSum : op == 1 ? ncclMax : ncclMin, 0));
  return KMI_OK;
}
extern "C" int kmi_comm_allgather_u64(kmi_comm* c, const uint64_t* mine, int n, uint64_t* all) {
  if (!c || ((!mine || !all) && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(c->device));
  return comm_allgather(c, reinterpret_cast<const u64*>(mine), n, reinterpret_cast<u64*>(all), 0);
}
extern "C" int kmi_comm_barrier(kmi_comm* c) {
  uint64_t v = 0;
  return kmi_comm_allreduce_u64(c, &v, 1, 0);
}

// ---- attach ------------------------------------------------------------------------------------------------------------
extern "C" int kmi_dist_attach(kmi_index* idx, kmi_comm* comm) {
  if (!idx || !comm) return fail(KMI_ERR_ARG, "null argument");
  if (idx->cfg.rank != comm->rank || idx->cfg.nranks != comm->nranks)
    return fail(KMI_ERR_ARG, "index is shard %d of %d, communicator is rank %d of %d", idx->cfg.rank, idx->cfg.nranks, comm->rank, comm->nranks);
  if (idx->cfg.device != comm->device) return fail(KMI_ERR_ARG, "index and communicator are on different devices");
  CU(cudaSetDevice(idx->cfg.device));
  if (idx->dist) kmi_dist_release(idx);
  kmi_dist_state* d = new (std::nothrow) kmi_dist_state();
  if (!d) return fail(KMI_ERR_NOMEM, "host allocation");
  d->comm = comm;
  idx->dist = d;
  int lo = 0, hi = 0;
  CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CU(cudaStreamCreateWithPriority(&d->push_stream, cudaStreamNonBlocking, hi));   // the pushes are NVLink-bound: let their CTAs in first
  for (int i = 0; i < 2; ++i) {
    CU(cudaEventCreateWithFlags(&d->ev_parsed[i], cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&d->ev_pushed[i], cudaEventDisableTiming));
  }
  for (int i = 0; i < 4; ++i) CU(cudaEventCreate(&d->ev_t[i]));
  OK(d->status.ensure(sizeof(u32) * 8));
  // the NVLink push path needs <= MAX_SUB ranks and peer mappings; both are probed now with a small receive block so that
  // every rank knows the answer (KMI_DIST_PUSH=0 forces the generic NCCL exchange)
  const char* env = getenv("KMI_DIST_PUSH");
  u64 want = (comm->nranks <= MAX_SUB && !(env && env[0] == '0')) ? 1 : 0;
  OK(comm_allreduce(comm, &want, 1, ncclMin, idx->stream));
  d->peer_ok = want != 0;
  if (d->peer_ok) OK(peers_ensure(idx, 1024, 4096));      // clears peer_ok when the mapping fails anywhere
  return KMI_OK;
}

extern "C" int kmi_dist_size(kmi_index* idx, uint64_t* n) {
  kmi_dist_state* d = dist_of(idx);
  if (!d || !n) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  uint64_t local = 0;
  const int r = kmi_local_size(idx, &local);                // errors must not leave the peers alone in the collective
  u64 v[2] = {local, (u64)(r != KMI_OK)};
  OK(comm_allreduce(d->comm, v, 2, ncclSum, idx->stream));
  *n = v[0];
  if (r != KMI_OK) return r;
  if (v[1]) return fail(KMI_ERR_CAPACITY, "another rank reported a table error");
  return KMI_OK;
}

extern "C" int kmi_dist_last_stats(kmi_index* idx, float* ms4, uint64_t* n_received, int32_t* path) {
  kmi_dist_state* d = dist_of(idx);
  if (!d) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  if (ms4) for (int i = 0; i < 4; ++i) ms4[i] = d->ms[i];
  if (n_received) *n_received = d->n_received;
  if (path) *path = d->last_path;
  return KMI_OK;
}

// ---- insert ------------------------------------------------------------------------------------------------------------
static int dist_insert_dev(kmi_index* idx, const u64* d_keys, const u64* d_vals, size_t n, bool canonicalize) {
  kmi_dist_state* d = idx->dist;
  std::vector<u64> sc, rc;
  u64 n_recv = 0;
  OK(distribute(idx, d_keys, d_vals, n, canonicalize, sc, rc, &n_recv));
  d->n_received = n_recv;
  return insert_owned(idx, d->recv_keys.as<u64>(), d_vals ? d->recv_vals.as<u64>() : nullptr, n_recv);
}

extern "C" int kmi_dist_insert_dev(kmi_index* idx, const uint64_t* d_keys, const uint64_t* d_vals, size_t n) {
  if (!dist_of(idx) || (!d_keys && n)) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  if (idx->cfg.mode == KMI_MODE_POSITION && !d_vals && n) return fail(KMI_ERR_ARG, "POSITION insert needs ids");
  CU(cudaSetDevice(idx->cfg.device));
  // an index without ids/counts on SOME ranks would desynchronise the value exchange: POSITION always has them, COUNT
  // ranks agree through a flag
  u64 has_vals = (d_vals || idx->cfg.mode == KMI_MODE_POSITION) ? 1 : 0;
  OK(comm_allreduce(idx->dist->comm, &has_vals, 1, ncclMax, idx->stream));
  const u64* vals = B(d_vals);
  if (has_vals && !d_vals) {                              // this rank has plain keys (or none): they count 1 each
    OK(idx->dist->q_vals.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
    std::vector<u64> ones(n, 1);
    if (n) CU(cudaMemcpyAsync(idx->dist->q_vals.p, ones.data(), n * sizeof(u64), cudaMemcpyHostToDevice, idx->stream));
    CU(cudaStreamSynchronize(idx->stream));
    vals = idx->dist->q_vals.as<u64>();
  }
  OK(dist_insert_dev(idx, B(d_keys), vals, n, idx->cfg.canonical != 0));
  CU(cudaStreamSynchronize(idx->stream));
  return check_table_error(idx);
}

extern "C" int kmi_dist_insert(kmi_index* idx, const uint64_t* keys, const uint64_t* vals, size_t n) {
  kmi_dist_state* d = dist_of(idx);
  if (!d || (!keys && n)) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  CU(cudaSetDevice(idx->cfg.device));
  const size_t kb = sizeof(u64) * idx->nw;
  OK(d->q_keys.ensure(std::max<size_t>(n, 1) * kb));
  OK(copy_to_device(idx, d->q_keys.p, keys, n * kb));
  const uint64_t* dv = nullptr;
  if (vals) {
    OK(d->q_back.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
    OK(copy_to_device(idx, d->q_back.p, vals, n * sizeof(u64)));
    dv = d->q_back.as<uint64_t>();
  }
  return kmi_dist_insert_dev(idx, d->q_keys.as<uint64_t>(), dv, n);
}

// ---- batched query: canonicalise -> distribute -> local probe -> answers back along the transposed counts
// (distributed_densehash_map.hpp:1122-1216); answers are delivered in INPUT order
// `dev`: keys / vals / found are DEVICE pointers (queries already resident in HBM, answers left there)
static int dist_query(kmi_index* idx, const uint64_t* keys, size_t n, int what /*0 count, 1 find, 2 erase*/, uint64_t* vals, uint8_t* found,
                      bool dev = false) {
  kmi_dist_state* d = dist_of(idx);
  if (!d || (!keys && n)) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  CU(cudaSetDevice(idx->cfg.device));
  const size_t kb = sizeof(u64) * idx->nw;
  cudaStream_t st = idx->stream;
  OK(d->q_keys.ensure(std::max<size_t>(n, 1) * kb));
  OK(d->q_tags.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
  if (n) {
    if (dev) CU(cudaMemcpyAsync(d->q_keys.p, keys, n * kb, cudaMemcpyDeviceToDevice, st));
    else OK(copy_to_device(idx, d->q_keys.p, keys, n * kb));
    k_iota<<<grid_for(n, 256), 256, 0, st>>>(d->q_tags.as<u64>(), n);
  }
  std::vector<u64> sc, rc;
  u64 m = 0;
  // the tags travel through the partition only (they stay here): exchange keys alone
  {
    const int p = d->comm->nranks;
    sc.assign(p, 0); rc.assign(p, 0);
    OK(d->part_keys.ensure(std::max<size_t>(n, 1) * kb));
    OK(d->part_vals.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
    const u64* src = d->q_keys.as<u64>();
    if (idx->cfg.canonical && n) {
      OK(kmi_canonicalize_dev(idx, d->q_keys.as<uint64_t>(), d->q_keys.as<uint64_t>(), n));
    }
    OK(kmi_partition_dev(idx, A(src), d->q_tags.as<uint64_t>(), n, d->part_keys.as<uint64_t>(), d->part_vals.as<uint64_t>(), A(sc.data())));
    OK(exchange_counts(d, st, sc.data(), rc.data()));
    for (int s = 0; s < p; ++s) m += rc[s];
    OK(d->recv_keys.ensure(std::max<u64>(m, 1) * kb));
    OK(exchange_rows(d, st, d->part_keys.p, sc.data(), rc.data(), kb, d->recv_keys.p));
  }
  if (what == 2) {                                        // erase: nothing comes back
    OK(d->keys.ensure(std::max<u64>(m, 1) * kb));
    int r = KMI_OK;
    if (m) {
      std::vector<u64> hk(m * idx->nw);                     // kmi_erase takes host keys (it also compacts the tuple log)
      CU(cudaMemcpyAsync(hk.data(), d->recv_keys.p, m * kb, cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      const int32_t canon = idx->cfg.canonical;
      idx->cfg.canonical = 0;
      r = kmi_erase(idx, A(hk.data()), m);
      idx->cfg.canonical = canon;
    }
    return r;
  }
  OK(d->q_vals.ensure(std::max<u64>(m, 1) * sizeof(u64)));
  OK(d->q_found.ensure(std::max<u64>(m, 1)));
  {
    const int32_t canon = idx->cfg.canonical;               // received keys are canonical already
    idx->cfg.canonical = 0;
    int r = KMI_OK;
    if (m) r = what == 0 ? kmi_count_dev(idx, d->recv_keys.as<uint64_t>(), m, d->q_vals.as<uint64_t>())
                         : kmi_find_dev(idx, d->recv_keys.as<uint64_t>(), m, d->q_vals.as<uint64_t>(), d->q_found.as<uint8_t>());
    idx->cfg.canonical = canon;
    OK(r);
  }
  OK(d->q_back.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
  OK(exchange_rows(d, st, d->q_vals.p, rc.data(), sc.data(), sizeof(u64), d->q_back.p));
  if (what == 1 && found) {
    OK(d->q_backf.ensure(std::max<size_t>(n, 1)));
    OK(exchange_rows(d, st, d->q_found.p, rc.data(), sc.data(), 1, d->q_backf.p));
  }
  if (n) {
    // out[tag] = back: reuse recv_keys (n * 8 <= its size is not guaranteed) -> dedicated buffers
    OK(d->recv_vals.ensure(n * sizeof(u64)));
    OK(d->ids.ensure(n));
    k_unpermute<<<grid_for(n, 256), 256, 0, st>>>(d->part_vals.as<u64>(), d->q_back.as<u64>(), (what == 1 && found) ? d->q_backf.as<uint8_t>() : nullptr,
                                                 d->recv_vals.as<u64>(), (what == 1 && found) ? d->ids.as<uint8_t>() : nullptr, n);
    if (vals) {
      if (dev) CU(cudaMemcpyAsync(vals, d->recv_vals.p, n * sizeof(u64), cudaMemcpyDeviceToDevice, st));
      else OK(copy_to_host(idx, vals, d->recv_vals.p, n * sizeof(u64)));
    }
    if (what == 1 && found) {
      if (dev) CU(cudaMemcpyAsync(found, d->ids.p, n, cudaMemcpyDeviceToDevice, st));
      else OK(copy_to_host(idx, found, d->ids.p, n));
    }
  }
  CU(cudaStreamSynchronize(st));
  return KMI_OK;
}
extern "C" int kmi_dist_count(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* counts) {
  if (!counts && n) return fail(KMI_ERR_ARG, "null argument");
  return dist_query(idx, keys, n, 0, counts, nullptr);
}
extern "C" int kmi_dist_find(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* vals, uint8_t* found) {
  return dist_query(idx, keys, n, 1, vals, found);
}
extern "C" int kmi_dist_erase(kmi_index* idx, const uint64_t* keys, size_t n) { return dist_query(idx, keys, n, 2, nullptr, nullptr); }
extern "C" int kmi_dist_count_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_counts) {
  if (!d_counts && n) return fail(KMI_ERR_ARG, "null argument");
  return dist_query(idx, d_keys, n, 0, d_counts, nullptr, true);
}
extern "C" int kmi_dist_find_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_vals, uint8_t* d_found) {
  return dist_query(idx, d_keys, n, 1, d_vals, d_found, true);
}

// find on a sharded PositionIndex: every id of every query key (LocalFind of the multimap, :2094-2133), CSR in input order.
// Two round trips: the multiplicities come back first (they size the second answer), then the ids.
extern "C" int kmi_dist_find_positions(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* offsets, uint64_t* vals, size_t cap) {
  kmi_dist_state* d = dist_of(idx);
  if (!d || (!keys && n) || !offsets) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  if (idx->cfg.mode != KMI_MODE_POSITION) return fail(KMI_ERR_ARG, "find_positions needs a POSITION index");
  CU(cudaSetDevice(idx->cfg.device));
  const int p = d->comm->nranks;
  const size_t kb = sizeof(u64) * idx->nw;
  cudaStream_t st = idx->stream;
  // 1. partition the queries (tags stay here), send the keys
  OK(d->q_keys.ensure(std::max<size_t>(n, 1) * kb));
  OK(d->q_tags.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
  OK(d->part_keys.ensure(std::max<size_t>(n, 1) * kb));
  OK(d->part_vals.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
  if (n) {
    OK(copy_to_device(idx, d->q_keys.p, keys, n * kb));
    k_iota<<<grid_for(n, 256), 256, 0, st>>>(d->q_tags.as<u64>(), n);
    if (idx->cfg.canonical) OK(kmi_canonicalize_dev(idx, d->q_keys.as<uint64_t>(), d->q_keys.as<uint64_t>(), n));
  }
  std::vector<u64> sc(p, 0), rc(p, 0);
  OK(kmi_partition_dev(idx, d->q_keys.as<uint64_t>(), d->q_tags.as<uint64_t>(), n, d->part_keys.as<uint64_t>(), d->part_vals.as<uint64_t>(), A(sc.data())));
  OK(exchange_counts(d, st, sc.data(), rc.data()));
  u64 m = 0;
  for (int s = 0; s < p; ++s) m += rc[s];
  OK(d->recv_keys.ensure(std::max<u64>(m, 1) * kb));
  OK(exchange_rows(d, st, d->part_keys.p, sc.data(), rc.data(), kb, d->recv_keys.p));
  // 2. local answers for the m received queries, all on the device: offsets (exclusive scan of the multiplicities) and the ids
  // of every hit in query order; the multiplicities go back to the asking ranks
  OK(d->q_vals.ensure((m + 1) * sizeof(u64)));
  OK(d->fp_cnt.ensure(std::max<u64>(m, 1) * sizeof(u64)));
  u64 total_send = 0;
  const int lr = find_positions_dev(idx, d->recv_keys.as<u64>(), m, d->q_vals.as<u64>(), d->fp_slot, d->recv_vals, &total_send);
  u64 bad = lr != KMI_OK;
  OK(comm_allreduce(d->comm, &bad, 1, ncclMax, st));
  if (bad) return lr != KMI_OK ? lr : fail(KMI_ERR_CUDA, "another rank failed in find_positions");
  if (m) k_adj_diff<<<grid_for(m, 256), 256, 0, st>>>(d->q_vals.as<u64>(), d->fp_cnt.as<u64>(), m);
  OK(d->q_back.ensure(std::max<size_t>(n, 1) * sizeof(u64)));
  OK(exchange_rows(d, st, d->fp_cnt.p, rc.data(), sc.data(), sizeof(u64), d->q_back.p));      // back_cnt, in partition order
  // where every query's ids will sit in the stream of ids that comes back (groups arrive in rank order, each in the order the
  // queries were sent: the partition order), and where they belong in the caller's order
  OK(d->fp_part_off.ensure((n + 1) * sizeof(u64)));
  OK(d->fp_out_off.ensure((n + 1) * sizeof(u64)));
  if (n) CU(cudaMemcpyAsync(d->fp_part_off.p, d->q_back.p, n * sizeof(u64), cudaMemcpyDeviceToDevice, st));
  CU(cudaMemsetAsync(d->fp_part_off.as<u64>() + n, 0, sizeof(u64), st));
  OK(scan64_dev(idx, d->fp_part_off.as<u64>(), n + 1));
  CU(cudaMemsetAsync(d->fp_out_off.p, 0, (n + 1) * sizeof(u64), st));
  if (n) k_unpermute<<<grid_for(n, 256), 256, 0, st>>>(d->part_vals.as<u64>(), d->q_back.as<u64>(), nullptr, d->fp_out_off.as<u64>(), nullptr, n);
  OK(scan64_dev(idx, d->fp_out_off.as<u64>(), n + 1));
  // 3. ids back: rank s answers my queries for it with sum(back_cnt over that group) ids. The per-rank sums are differences
  // of the two offset arrays at the group boundaries: p + 1 words each way
  std::vector<u64> ids_send(p, 0), ids_recv(p, 0), bs(p + 1, 0), br(p + 1, 0);
  {
    u64 qs = 0, qr = 0;
    for (int r = 0; r <= p; ++r) {
      CU(cudaMemcpyAsync(&bs[r], d->q_vals.as<u64>() + qs, sizeof(u64), cudaMemcpyDeviceToHost, st));
      CU(cudaMemcpyAsync(&br[r], d->fp_part_off.as<u64>() + qr, sizeof(u64), cudaMemcpyDeviceToHost, st));
      if (r < p) { qs += rc[r]; qr += sc[r]; }
    }
    CU(cudaStreamSynchronize(st));
    for (int r = 0; r < p; ++r) { ids_send[r] = bs[r + 1] - bs[r]; ids_recv[r] = br[r + 1] - br[r]; }
  }
  const u64 total_recv = br[p];
  OK(d->keys.ensure(std::max<u64>(total_recv, 1) * sizeof(u64)));
  OK(d->recv_vals.ensure(std::max<u64>(total_send, 1) * sizeof(u64)));                        // (untouched when there were no hits)
  OK(exchange_rows(d, st, d->recv_vals.p, ids_send.data(), ids_recv.data(), sizeof(u64), d->keys.p));
  // 4. CSR in input order
  if (total_recv > cap) { CU(cudaStreamSynchronize(st)); return fail(KMI_ERR_CAPACITY, "find_positions: %llu values, capacity %zu", (unsigned long long)total_recv, cap); }
  OK(d->fp_out_vals.ensure(std::max<u64>(total_recv, 1) * sizeof(u64)));
  if (n && total_recv) k_place_ids<<<grid_for(n, 256), 256, 0, st>>>(d->part_vals.as<u64>(), d->q_back.as<u64>(), d->fp_part_off.as<u64>(), d->fp_out_off.as<u64>(),
                                                                     d->keys.as<u64>(), d->fp_out_vals.as<u64>(), n);
  OK(copy_to_host(idx, offsets, d->fp_out_off.p, (n + 1) * sizeof(u64)));
  if (total_recv) OK(copy_to_host(idx, vals, d->fp_out_vals.p, total_recv * sizeof(u64)));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  return KMI_OK;
}

// ---- build ---------------------------------------------------------------------------------------------------------------
namespace {

// generic path for one block of text resident on the device: parse -> (keys, ids) -> distribute -> insert
int build_fastq_generic(kmi_index* idx, const uint8_t* d_text, size_t n, u64 file_offset, u64* n_kmers) {
  kmi_dist_state* d = idx->dist;
  const bool pos = idx->cfg.mode == KMI_MODE_POSITION;
  const size_t kb = sizeof(u64) * idx->nw;
  const size_t cap = n / 2 + 1;
  OK(d->keys.ensure(cap * kb));
  if (pos) OK(d->ids.ensure(cap * sizeof(u64)));
  uint64_t produced = 0;
  int r = n ? kmi_parse_fastq_dev(idx, d_text, n, file_offset, d->keys.as<uint64_t>(), pos ? d->ids.as<uint64_t>() : nullptr, cap, &produced) : KMI_OK;
  // a malformed block on one rank must not strand the others in the exchange: agree first
  u64 bad = r != KMI_OK;
  OK(comm_allreduce(d->comm, &bad, 1, ncclMax, idx->stream));
  if (bad) return r != KMI_OK ? r : fail(KMI_ERR_FORMAT, "another rank reported malformed FASTQ");
  // parse output is canonical already; kmi_parse keeps its keys in d->keys, distribute() must not canonicalise into it
  std::vector<u64> sc, rc;
  u64 n_recv = 0;
  OK(distribute(idx, d->keys.as<u64>(), pos ? d->ids.as<u64>() : nullptr, produced, false, sc, rc, &n_recv));
  d->n_received += n_recv;
  OK(insert_owned(idx, d->recv_keys.as<u64>(), pos ? d->recv_vals.as<u64>() : nullptr, n_recv));
  *n_kmers += produced;
  return KMI_OK;
}

// The count-index build over NVLink. Text: `n` bytes at d_text (device). With host_src the bytes are copied round by
// round on the copy stream while earlier rounds are parsed and pushed.
int build_fastq_push(kmi_index* idx, uint8_t* d_text, size_t n, const HostSrc& host_src, u64 n_max, u64* n_kmers, bool* redo) {
  kmi_dist_state* d = idx->dist;
  kmi_comm* c = d->comm;
  const int p = c->nranks, me = c->rank;
  cudaStream_t st = idx->stream;
  *redo = false;
  // ---- geometry every rank derives from the agreed maximum text size
  const u64 TB = TILE_BYTES;
  const u64 tiles_max = (n_max + TB - 1) / TB;
  u64 per_round = std::max<u64>((tiles_max + 7) / 8, (64u << 20) / TB);                    // <= 8 rounds of >= 64 MiB
  if (const char* e = getenv("KMI_DIST_ROUND_BYTES")) per_round = std::max<u64>(strtoull(e, nullptr, 10) / TB, 1);
  per_round = std::min<u64>(per_round, std::max<u64>(tiles_max, 1));
  per_round = (per_round + 1023) / 1024 * 1024;                                             // the line scan works in blocks of 1024 tiles
  const u64 n_rounds = std::max<u64>((tiles_max + per_round - 1) / per_round, 1);
  const u64 upper = n_max / 2 + 1;                                                          // k-mers one rank can generate
  const double mean = (double)upper / ((double)N_BUCKETS * p);
  const u64 cap_sub = (u64)(mean * 1.06 + 8.0 * std::sqrt(mean) + 1024.0);
  const u64 ovf_cap = upper / 8 + 4096;
  if ((u64)N_BUCKETS * p * cap_sub >= (1ULL << 32) || (double)upper * 1.2 > (double)BUCKET_MAX_KEYS)
    return fail(KMI_ERR_ARG, "block of %llu bytes per rank is too large for one sharded build call (split it)", (unsigned long long)n_max);
  OK(peers_ensure(idx, cap_sub, ovf_cap));
  if (!d->peer_ok) { *redo = true; return KMI_OK; }
  u32* my_counts = reinterpret_cast<u32*>(d->recv);
  const u32 n_regions = (u32)N_BUCKETS * p;
  u32* st_in = d->status.as<u32>();
  u32* st_out = st_in + 2;
  CU(cudaMemsetAsync(my_counts, 0, sizeof(u32) * (n_regions + 1), st));
  CU(cudaMemsetAsync(d->push_cursor.p, 0, sizeof(u32) * n_regions, st));
  CU(cudaMemsetAsync(st_in, 0, sizeof(u32) * 4, st));
  OK(zero_ctr(idx, CTR_NKMERS));
  OK(zero_ctr(idx, CTR_PERR));
  OK(bucket_meta_prepare(idx));                                                             // HLL registers (cursors unused here)
  // nobody pushes into a receive block before its owner has reset it (and finished reading the previous build's keys)
  NC(g_nccl.AllReduce(st_in, st_out, 2, ncclUint32, ncclMax, c->comm, st));
  CU(cudaEventRecord(d->ev_t[0], st));
  // ---- rounds
  const u64 n_tiles = (n + TB - 1) / TB;
  u64 scan_tiles = 0;
  OK(fastq_scan_alloc(idx, d_text, n, &scan_tiles));
  HostFeed feed;   // declared before the guard below: the streams are drained first, then the feed's workers stop
  if (host_src) {
    std::vector<u64> ends(n_rounds);
    for (u64 r = 0; r < n_rounds; ++r) ends[r] = std::min<u64>(n, (r + 1) * per_round * TB);
    OK(feed.start(idx, d_text, host_src, ends, st));
  }
  struct Guard { cudaStream_t a, b; ~Guard() { if (a) cudaStreamSynchronize(a); if (b) cudaStreamSynchronize(b); } } guard{host_src ? idx->copy_stream : nullptr, d->push_stream};
  KeySink nosink; nosink.keys = nullptr; nosink.ids = nullptr; nosink.cap = 0; nosink.cursor = nullptr;
  PushSink ps;
  ps.peer_keys = d->peer_keys.as<u64*>(); ps.cursor = d->push_cursor.as<u32>(); ps.cap_sub = (u32)d->cap_sub; ps.p = (u32)p; ps.me = (u32)me;
  ps.status = st_in;
  { const char* e = getenv("KMI_PUSH_BULK"); ps.bulk = (e && e[0] == '0') ? 0u : 1u; }   // TMA bulk copies of the runs (0: plain stores, A/B)
  u64 next_tile = 0;                                        // first tile not parsed yet
  for (u64 r = 0; r < n_rounds; ++r) {
    const int w = (int)(r & 1);
    const u64 t0 = std::min(n_tiles, r * per_round), t1 = std::min(n_tiles, t0 + per_round);
    // line counts: a tile's halo continues at the NEXT tile's count, so resident text is counted whole before the first
    // round; arriving text is counted round by round and the last tile that has arrived waits for the next round
    if (host_src) { OK(feed.wait(r, st)); OK(fastq_scan_range(idx, d_text, n, t0, t1)); }
    else if (r == 0) OK(fastq_scan_range(idx, d_text, n, 0, n_tiles));
    // with a host source the last tile that has arrived needs the first bytes of the next round (k-mer halo), unless the
    // text ends there: it is parsed with the next round
    const u64 p0 = next_tile;
    const u64 p1 = std::max<u64>(p0, (host_src && t1 < n_tiles && t1 > 0) ? t1 - 1 : t1);
    const u64 tiles = p1 - p0;
    next_tile = p1;
    BucketSink bs;
    OK(stage_sink(idx, w, std::max<u64>(per_round + 1, 1) * TB / 2 + 1, &bs));
    if (r >= 2) CU(cudaStreamWaitEvent(st, d->ev_pushed[w], 0));     // staging buffer w was read by the push of round r - 2
    else CU(cudaMemsetAsync(bs.cursor, 0, sizeof(u32) * (N_BUCKETS + 1), st));
    if (tiles) {
      idx->ops->fastq_kmers(st, 2, d_text, n, tiles, idx->tile_lines.as<u32>(), idx->block_sums.as<u32>(), nosink, idx->raw_view(), bs,
                            idx->ctr + CTR_NKMERS, reinterpret_cast<u32*>(idx->ctr + CTR_PERR), idx->cfg.canonical != 0, idx->cfg.nsplit != 0, p0, 1);
      idx->launches++;
    }
    CU(cudaEventRecord(d->ev_parsed[w], st));
    CU(cudaStreamWaitEvent(d->push_stream, d->ev_parsed[w], 0));
    idx->ops->owner_push(d->push_stream, bs, ps, tiles * TB / 2 + 1);
    k_push_publish<<<8, 256, 0, d->push_stream>>>(ps, d->peer_cnt.as<u32*>(), bs.cursor, reinterpret_cast<const u32*>(idx->ctr + CTR_PERR));
    idx->launches += 2;
    CU(cudaEventRecord(d->ev_pushed[w], d->push_stream));
  }
  CU(cudaStreamWaitEvent(st, d->ev_pushed[0], 0));
  CU(cudaStreamWaitEvent(st, d->ev_pushed[1], 0));
  // completion + agreement: after this all-reduce every rank's pushes into every receive block have finished, and every
  // rank knows whether any region overflowed or any block was malformed
  NC(g_nccl.AllReduce(st_in, st_out, 2, ncclUint32, ncclMax, c->comm, st));
  CU(cudaEventRecord(d->ev_t[1], st));
  u32 h_status[2] = {0, 0};
  std::vector<u32> h_counts(n_regions + 1);
  u64 host_ctr[CTR_N];
  CU(cudaMemcpyAsync(h_status, st_out, sizeof(h_status), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(h_counts.data(), my_counts, sizeof(u32) * (n_regions + 1), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(host_ctr, idx->ctr, sizeof(host_ctr), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (h_status[1]) {
    const int r = parse_error(host_ctr[CTR_PERR] & (PERR_NO_AT | PERR_NO_PLUS));
    return r != KMI_OK ? r : fail(KMI_ERR_FORMAT, "another rank reported malformed FASTQ");
  }
  if (h_status[0]) { *redo = true; return KMI_OK; }         // skewed keys: nothing is in any table yet
  *n_kmers += host_ctr[CTR_NKMERS];
  u64 n_recv = 0;
  for (u32 j = 0; j < n_regions; ++j) n_recv += h_counts[j];
  d->n_received += n_recv;
  // ---- received regions -> table: region prefix * p + source, `p` inputs per prefix for the fine scatter
  BucketSink rs;
  rs.keys = reinterpret_cast<u64*>(reinterpret_cast<char*>(d->recv) + kmi_dist_state::HDR_BYTES);
  rs.cursor = my_counts;
  rs.capb = (u32)d->cap_sub; rs.ovf_cap = (u32)std::min<u64>(d->ovf_cap, 0xffffffffULL);
  rs.hll = idx->bkt_meta.as<u32>() + META_HLL;
  rs.sub = (u32)p;
  std::vector<u32> reg;
  OK(bucket_readback(idx, rs, host_ctr, reg));
  bool not_inserted = false;
  OK(bucket_commit(idx, rs, n_recv, host_ctr[CTR_SIZE], reg, &not_inserted));
  if (not_inserted) {                                       // small table: the regions go in as key streams
    std::vector<const u64*> ptrs;
    std::vector<u64> cnts;
    for (u32 j = 0; j < n_regions; ++j)
      if (h_counts[j]) { ptrs.push_back(rs.keys + (u64)j * rs.capb * idx->nw); cnts.push_back(h_counts[j]); }
    if (!ptrs.empty()) OK(insert_streams_bucketed(idx, ptrs.data(), cnts.data(), (int)ptrs.size(), false));
  }
  CU(cudaEventRecord(d->ev_t[2], st));
  return KMI_OK;
}

int dist_build_fastq(kmi_index* idx, uint8_t* d_text, size_t n, const HostSrc& host_src, u64 file_offset, uint64_t* n_kmers) {
  kmi_dist_state* d = idx->dist;
  cudaStream_t st = idx->stream;
  if (n_kmers) *n_kmers = 0;
  d->n_received = 0;
  for (int i = 0; i < 4; ++i) d->ms[i] = 0;
  for (int i = 0; i < 5; ++i) idx->kernel_ms[i] = 0;
  idx->kev_pending = false;
  // what every rank must agree on before any collective with data-dependent shape: the largest block, and the path
  u64 v[1] = {n};
  OK(comm_allreduce(d->comm, v, 1, ncclMax, st));
  const u64 n_max = v[0];
  u64 total = 0;
  const bool count_mode = idx->cfg.mode == KMI_MODE_COUNT;
  const char* force = getenv("KMI_DIST_PUSH");
  const bool big = n_max >= AUTO_BUCKET_BYTES || (force && force[0] == '1');
  bool use_push = count_mode && d->peer_ok && big && !idx->saturate && idx->ops->owner_push != nullptr && idx->dist_trans == 0;   // (saturating counts add with direct atomics)
  if (use_push) {
    bool redo = false;
    int r = build_fastq_push(idx, d_text, n, host_src, n_max, &total, &redo);
    if (r != KMI_OK) return r;
    if (!redo) {
      d->last_path = 1;
      CU(cudaStreamSynchronize(st));
      cudaEventElapsedTime(&d->ms[1], d->ev_t[0], d->ev_t[1]);
      cudaEventElapsedTime(&d->ms[2], d->ev_t[1], d->ev_t[2]);
      cudaEventElapsedTime(&d->ms[0], d->ev_t[0], d->ev_t[2]);
      harvest_commit_times(idx);
      if (n_kmers) *n_kmers = total;
      return check_table_error(idx);
    }
    total = 0; d->n_received = 0;                           // (a host text is simply copied again below: rare path)
  }
  d->last_path = 2;
  if (host_src && n) OK(copy_to_device(idx, d_text, host_src, n));
  // generic path in slices (bounded key staging); every rank runs the same number of slices
  const u64 SLICE = 1ULL << 30;
  const u64 n_slices = std::max<u64>((n_max + SLICE - 1) / SLICE, 1);
  CU(cudaEventRecord(d->ev_t[0], st));
  if (n_slices == 1) {
    OK(build_fastq_generic(idx, d_text, n, file_offset, &total));
  } else {
    // slices must hold whole records: cut on the host copy of the slice boundaries' neighbourhood
    std::vector<u64> cuts(n_slices + 1, n);
    cuts[0] = 0;
    for (u64 s = 1; s < n_slices; ++s) {
      const u64 at = std::min<u64>(n, s * SLICE);
      const u64 win = std::min<u64>(n - at, 16u << 20);
      std::vector<uint8_t> hb(win);
      if (win) CU(cudaMemcpy(hb.data(), d_text + at, win, cudaMemcpyDeviceToHost));
      uint64_t off = at;
      // (a window that holds no record boundary makes this slice empty and leaves its bytes to the next one)
      if (win && kmi_fastq_find_first_record(hb.data(), at, win, at, at + win, 0, n, &off) != KMI_OK) off = at + win;
      cuts[s] = std::max<u64>(cuts[s - 1], std::min<u64>(off, n));
    }
    for (u64 s = 0; s < n_slices; ++s) {
      // parse needs 16-byte aligned text: slices that start unaligned are moved to the front of a scratch buffer
      const u64 b0 = cuts[s], b1 = cuts[s + 1];
      const uint8_t* src = d_text + b0;
      if ((b0 & 15) && b1 > b0) {
        OK(d->text.ensure(b1 - b0 + 64));
        CU(cudaMemcpyAsync(d->text.p, d_text + b0, b1 - b0, cudaMemcpyDeviceToDevice, st));
        src = d->text.as<uint8_t>();
      }
      OK(build_fastq_generic(idx, src, b1 - b0, file_offset + b0, &total));
    }
  }
  CU(cudaEventRecord(d->ev_t[2], st));
  CU(cudaStreamSynchronize(st));
  cudaEventElapsedTime(&d->ms[0], d->ev_t[0], d->ev_t[2]);
  if (n_kmers) *n_kmers = total;
  return check_table_error(idx);
}

}  // namespace

extern "C" int kmi_dist_build_fastq_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  if (!dist_of(idx) || (!d_bytes && n)) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  CU(cudaSetDevice(idx->cfg.device));
  return dist_build_fastq(idx, const_cast<uint8_t*>(d_bytes), n, nullptr, file_offset, n_kmers);
}
extern "C" int kmi_dist_build_fastq(kmi_index* idx, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  if (!dist_of(idx) || (!bytes && n)) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  CU(cudaSetDevice(idx->cfg.device));
  OK(idx->stage_in.ensure(n + 64));
  return dist_build_fastq(idx, idx->stage_in.as<uint8_t>(), n, bytes, file_offset, n_kmers);
}

// Index::build_posix(filename, comm) at p ranks: every rank takes its block of the ONE file (kmi_file_block) and builds from it
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
extern "C" int kmi_dist_build_file(kmi_index* idx, const char* path, int format, uint64_t* n_kmers) {
  kmi_dist_state* d = dist_of(idx);
  if (!d || !path) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  CU(cudaSetDevice(idx->cfg.device));
  if (n_kmers) *n_kmers = 0;
  const int p = d->comm->nranks, me = d->comm->rank;
  struct Fd { int fd; ~Fd() { if (fd >= 0) close(fd); } } f{-1};
  u64 size = 0, b = 0, e = 0, he = 0;
  std::vector<uint8_t> buf;                                // FASTA block + halo; FASTQ block when the slice ring is off
  int r = file_format(path, &format);
  if (r == KMI_OK) {
    f.fd = open(path, O_RDONLY);
    struct stat st;
    if (f.fd < 0 || fstat(f.fd, &st) != 0) r = fail(KMI_ERR_ARG, "cannot open %s", path);
    else size = (u64)st.st_size;
  }
  if (r == KMI_OK) r = file_block_fd(f.fd, size, format, me, p, &b, &e);
  const bool stream = r == KMI_OK && format == 1 && e > b && host_ring_ready(idx);
  if (r == KMI_OK && format == 2) r = file_fasta_halo_end(f.fd, size, e, idx->cfg.k, &he);
  if (r == KMI_OK && !stream) {
    const u64 hi = format == 2 ? he : e;
    buf.resize((size_t)(hi - b));
    if (file_read(f.fd, buf.data(), buf.size(), b) != (long long)buf.size()) r = fail(KMI_ERR_ARG, "read error on %s", path);
  }
  // nobody enters the build's collectives unless every rank has its block
  char msg[400];
  snprintf(msg, sizeof(msg), "%s", kmi_last_error());
  u64 bad = r == KMI_OK ? 0 : 1;
  OK(comm_allreduce(d->comm, &bad, 1, ncclMax, idx->stream));
  if (r != KMI_OK) return fail(r, "%s", msg);
  if (bad) return fail(KMI_ERR_ARG, "another rank could not read its block of %s", path);
  if (format == 1) {
    const size_t n = (size_t)(e - b);
    OK(idx->stage_in.ensure(n + 64));
    const HostSrc src = stream ? HostSrc::file(f.fd, b) : HostSrc(buf.data());
    return dist_build_fastq(idx, idx->stage_in.as<uint8_t>(), n, src, b, n_kmers);
  }
  uint8_t before = '\n';
  if (b > 0 && pread(f.fd, &before, 1, (off_t)(b - 1)) != 1) before = 0;   // (cannot fail after the reads above; 0 = not a line start)
  kmi_fasta_part part;
  part.file_offset = b; part.n_valid = e - b; part.bol = (b == 0 || before == '\n') ? 1 : 0; part.reserved = 0;
  return kmi_dist_build_fasta(idx, buf.data(), buf.size(), &part, n_kmers);
}

// FASTA: rank r holds block r of ONE file (blocks in rank order, cut at any byte) or — part == NULL — a whole file of its own.
extern "C" int kmi_dist_build_fasta_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, const kmi_fasta_part* part, uint64_t* n_kmers) {
  kmi_dist_state* d = dist_of(idx);
  if (!d || (!d_bytes && n)) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  CU(cudaSetDevice(idx->cfg.device));
  if (n_kmers) *n_kmers = 0;
  const int p = d->comm->nranks, me = d->comm->rank;
  const bool pos = idx->cfg.mode == KMI_MODE_POSITION;
  const size_t kb = sizeof(u64) * idx->nw;
  kmi_fasta_block blk;
  const kmi_fasta_block* bp = nullptr;
  int r = KMI_OK;
  if (part) {
    if (part->n_valid > n) r = fail(KMI_ERR_ARG, "kmi_fasta_part: n_valid exceeds the bytes given");
    // the cross-rank part of FASTAParser::init_parser (fasta_loader.hpp:202-483): every block's own line structure, then an
    // exclusive scan over the blocks (kmi::fasta_chain)
    uint32_t sum[4] = {0, 0, 0, 0};
    if (r == KMI_OK) r = kmi_fasta_block_summary_dev(idx, d_bytes, part->n_valid, part->bol, sum);
    uint8_t first = 0;
    if (r == KMI_OK && part->file_offset == 0 && part->n_valid) CU(cudaMemcpy(&first, d_bytes, 1, cudaMemcpyDeviceToHost));
    u64 mine[8] = {sum[0], sum[1], sum[2], sum[3], (u64)(part->bol != 0), part->file_offset, (u64)(first == '>' || first == ';'), (u64)(r != KMI_OK)};
    std::vector<u64> all((size_t)8 * p);
    OK(comm_allgather(d->comm, mine, 8, all.data(), idx->stream));
    for (int q = 0; q < p; ++q) if (all[(size_t)q * 8 + 7]) return r != KMI_OK ? r : fail(KMI_ERR_ARG, "another rank rejected its FASTA block");
    bool starts_with_header = false;
    for (int q = 0; q < p; ++q) if (all[(size_t)q * 8] && all[(size_t)q * 8 + 5] == 0) { starts_with_header = all[(size_t)q * 8 + 6] != 0; break; }
    u32 runs = 0, kind = 0;
    for (int q = 0; q < me; ++q) {
      const u64* s = &all[(size_t)q * 8];
      if (s[0]) { runs += (u32)s[3] - ((s[1] && kind) ? 1u : 0u); kind = (u32)s[2]; }
    }
    blk.file_offset = part->file_offset; blk.n_valid = part->n_valid; blk.header_runs_before = runs;
    blk.flags = (part->bol ? KMI_FASTA_BOL : 0u) | (kind ? KMI_FASTA_PREV_HEADER : 0u) | (starts_with_header ? KMI_FASTA_FILE_STARTS_WITH_HEADER : 0u);
    bp = &blk;
  }
  const size_t cap = std::max<size_t>(part ? part->n_valid : n, 1);
  OK(d->keys.ensure(cap * kb));
  if (pos) OK(d->ids.ensure(cap * sizeof(u64)));
  uint64_t produced = 0;
  r = n ? kmi_parse_fasta_block_dev(idx, d_bytes, n, bp, d->keys.as<uint64_t>(), pos ? d->ids.as<uint64_t>() : nullptr, cap, &produced) : KMI_OK;
  u64 bad = r != KMI_OK;
  OK(comm_allreduce(d->comm, &bad, 1, ncclMax, idx->stream));
  if (bad) return r != KMI_OK ? r : fail(KMI_ERR_FORMAT, "another rank failed to parse its FASTA block");
  std::vector<u64> sc, rc;
  u64 n_recv = 0;
  OK(distribute(idx, d->keys.as<u64>(), pos ? d->ids.as<u64>() : nullptr, produced, false, sc, rc, &n_recv));
  d->n_received = n_recv;
  d->last_path = 2;
  OK(insert_owned(idx, d->recv_keys.as<u64>(), pos ? d->recv_vals.as<u64>() : nullptr, n_recv));
  CU(cudaStreamSynchronize(idx->stream));
  if (n_kmers) *n_kmers = produced;
  return check_table_error(idx);
}
extern "C" int kmi_dist_build_fasta(kmi_index* idx, const uint8_t* bytes, size_t n, const kmi_fasta_part* part, uint64_t* n_kmers) {
  kmi_dist_state* d = dist_of(idx);
  if (!d || (!bytes && n)) return fail(KMI_ERR_ARG, "kmi_dist_attach first");
  CU(cudaSetDevice(idx->cfg.device));
  OK(d->text.ensure(n + 64));
  OK(copy_to_device(idx, d->text.p, bytes, n));
  return kmi_dist_build_fasta_dev(idx, d->text.as<uint8_t>(), n, part, n_kmers);
}
