// kmi_sorted.cu — the sort-based back end: a device LSD radix sort of (key, value) pairs and, on top of it,
// the single-shard form of the reference's sorted maps:
//   dsc::counting_sorted_map   src/containers/distributed_sorted_map.hpp:2386-2700   (COUNT: key -> frequency)
//   dsc::sorted_multimap       src/containers/distributed_sorted_map.hpp:1700-2000   (POSITION: key -> ids, duplicates kept)
// whose local container is a std::vector kept sorted: insert() appends, rehash() sorts and reduces (local_reduction,
// :1521-1540, :2560-2600), count()/find() are binary searches (sorted_map_base::Intersect, :228-480).
// Here: insert / build_* append canonical keys (+ ids) in HBM, kmi_sorted_rehash runs one radix sort over old + new
// entries and a run-length reduction (COUNT: counts summed as uint32_t, wrapping like the reference's mapped type), the
// queries are one binary search per key. The sort also serves kmi_save (entries sorted by (key, value)).
// Nothing here is a hash table: it is the alternative strategy the reference offers for the same Index<> facade.
// The reference's cross-rank half (sample sort with splitters, :1544-1700) is NOT built: one shard only.
#include "kmi_internal.cuh"

#include <algorithm>
#include <new>

using namespace kmi;

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_ROUNDS = 8;
constexpr int RS_ITEMS = RS_THREADS * RS_ROUNDS;   // keys per CTA and pass
constexpr int RS_BINS = 256;                       // 8-bit digits

// digit of entry i: words [0, nw) are the key (word 0 least significant), word nw is the value
__device__ __forceinline__ u32 rs_digit(const u64* __restrict__ keys, const u64* __restrict__ vals, u64 i, int nw, int word, int shift) {
  const u64 w = (word < nw) ? keys[i * nw + word] : vals[i];
  return (u32)(w >> shift) & (RS_BINS - 1);
}

// per-chunk digit histogram, bin-major (hist[bin * n_chunks + chunk]) so that ONE exclusive scan gives every (bin, chunk) base
__global__ void __launch_bounds__(RS_THREADS)
k_rs_hist(const u64* __restrict__ keys, const u64* __restrict__ vals, u64 n, int nw, int word, int shift, u32* __restrict__ hist, u64 n_chunks) {
  __shared__ u32 h[RS_BINS];
  h[threadIdx.x] = 0;
  __syncthreads();
  const u64 base = (u64)blockIdx.x * RS_ITEMS;
#pragma unroll
  for (int rd = 0; rd < RS_ROUNDS; ++rd) {
    const u64 i = base + (u64)rd * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[rs_digit(keys, vals, i, nw, word, shift)], 1u);
  }
  __syncthreads();
  hist[(u64)threadIdx.x * n_chunks + blockIdx.x] = h[threadIdx.x];
}

// stable scatter of one chunk: ranks inside a round by __match_any_sync (lanes with the same digit, lower lanes first) and
// per-warp counts, rounds in order — so equal digits keep their input order, which is what makes LSD passes compose
__global__ void __launch_bounds__(RS_THREADS)
k_rs_scatter(const u64* __restrict__ keys, const u64* __restrict__ vals, u64 n, int nw, int word, int shift,
             const u32* __restrict__ hist_scanned, const u32* __restrict__ block_sums, u64 n_chunks,
             u64* __restrict__ keys_out, u64* __restrict__ vals_out) {
  constexpr int NWARPS = RS_THREADS / 32;
  __shared__ u32 run[RS_BINS];
  __shared__ u32 wcnt[NWARPS][RS_BINS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  {
    const u64 hi = (u64)threadIdx.x * n_chunks + blockIdx.x;
    run[threadIdx.x] = hist_scanned[hi] + block_sums[hi >> 10];
  }
  const u64 base = (u64)blockIdx.x * RS_ITEMS;
  for (int rd = 0; rd < RS_ROUNDS; ++rd) {
#pragma unroll
    for (int w = 0; w < NWARPS; ++w) wcnt[w][threadIdx.x] = 0;
    __syncthreads();
    const u64 i = base + (u64)rd * RS_THREADS + threadIdx.x;
    const bool ok = i < n;
    u64 k0 = 0, k1 = 0, v = 0;
    u32 d = 0xFFFFFFFFu;
    if (ok) {
      k0 = keys[i * nw];
      if (nw == 2) k1 = keys[i * nw + 1];
      if (vals) v = vals[i];
      const u64 w = (word == 0) ? k0 : (word < nw) ? k1 : v;
      d = (u32)(w >> shift) & (RS_BINS - 1);
    }
    const u32 peers = __match_any_sync(0xffffffffu, d);
    const u32 rank_in_warp = __popc(peers & ((1u << lane) - 1u));
    if (ok && rank_in_warp == 0) wcnt[warp][d] = __popc(peers);
    __syncthreads();
    if (ok) {
      u32 o = run[d] + rank_in_warp;
      for (int w = 0; w < warp; ++w) o += wcnt[w][d];
      keys_out[(u64)o * nw] = k0;
      if (nw == 2) keys_out[(u64)o * nw + 1] = k1;
      if (vals) vals_out[o] = v;
    }
    __syncthreads();
    {
      u32 s = 0;
#pragma unroll
      for (int w = 0; w < NWARPS; ++w) s += wcnt[w][threadIdx.x];
      run[threadIdx.x] += s;
    }
    __syncthreads();
  }
}

// two-level exclusive scan of u32 (1024 per block, then the block sums by one block; consumers add the two)
__global__ void __launch_bounds__(1024) k_rs_scan_l1(u32* __restrict__ v, u64 n, u32* __restrict__ block_sums) {
  __shared__ u32 tmp[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const u64 i = (u64)blockIdx.x * 1024 + threadIdx.x;
  const u32 x = (i < n) ? v[i] : 0;
  u32 inc = x;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const u32 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const u32 w = tmp[lane];
    u32 winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const u32 o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += o;
    }
    tmp[lane] = winc - w;
    if (lane == 31) tmp[32] = winc;
  }
  __syncthreads();
  if (i < n) v[i] = tmp[warp] + inc - x;
  if (threadIdx.x == 0) block_sums[blockIdx.x] = tmp[32];
}
__global__ void __launch_bounds__(1024) k_rs_scan_top(u32* __restrict__ v, u64 n, u32* __restrict__ total) {
  __shared__ u32 tmp[33];
  __shared__ u32 carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (u64 base = 0; base < n; base += 1024) {
    const u64 i = base + threadIdx.x;
    const u32 x = (i < n) ? v[i] : 0;
    u32 inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const u32 o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) tmp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      const u32 w = tmp[lane];
      u32 winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const u32 o = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += o;
      }
      tmp[lane] = winc - w;
      if (lane == 31) tmp[32] = winc;
    }
    __syncthreads();
    const u32 carry = carry_s;
    if (i < n) v[i] = carry + tmp[warp] + inc - x;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + tmp[32];
    __syncthreads();
  }
  if (threadIdx.x == 0 && total) *total = carry_s;
}

// OR of all values (how many value digits a sort by (key, value) has to look at)
__global__ void __launch_bounds__(256) k_rs_or(const u64* __restrict__ v, u64 n, u64* __restrict__ out) {
  u64 a = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) a |= v[i];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) a |= __shfl_xor_sync(0xffffffffu, a, d);
  if ((threadIdx.x & 31) == 0 && a) atomicOr(reinterpret_cast<unsigned long long*>(out), (unsigned long long)a);
}
__global__ void __launch_bounds__(256) k_fill_u64(u64* __restrict__ v, u64 n, u64 x) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) v[i] = x;
}

__device__ __forceinline__ bool rs_key_eq(const u64* __restrict__ keys, u64 a, u64 b, int nw) {
  if (keys[a * nw] != keys[b * nw]) return false;
  return nw == 1 || keys[a * nw + 1] == keys[b * nw + 1];
}
// run heads of a sorted key array: flags[i] = key[i] != key[i-1]
__global__ void __launch_bounds__(256) k_run_heads(const u64* __restrict__ keys, u64 n, int nw, u32* __restrict__ flags) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
    flags[i] = (i == 0 || !rs_key_eq(keys, i, i - 1, nw)) ? 1u : 0u;
}
// local_reduction of the counting map (distributed_sorted_map.hpp:2560-2600): one entry per run, values summed as uint32_t.
// heads_before = exclusive scan of the run-head flags (+ block sums). Lanes of a warp that fall into the same run add their
// values with one warp reduction and one atomic by the run's first lane of that warp. vals == nullptr: every entry counts 1.
__global__ void __launch_bounds__(256)
k_reduce_runs(const u64* __restrict__ keys, const u64* __restrict__ vals, u64 n, int nw, const u32* __restrict__ heads_before,
              const u32* __restrict__ block_sums, u64* __restrict__ out_keys, u64* __restrict__ out_vals) {
  const int lane = threadIdx.x & 31;
  for (u64 base = (u64)blockIdx.x * blockDim.x; base < n; base += (u64)gridDim.x * blockDim.x) {
    const u64 i = base + threadIdx.x;
    u32 u = 0xFFFFFFFFu, v = 0;
    if (i < n) {
      const bool head = (i == 0 || !rs_key_eq(keys, i, i - 1, nw));
      u = heads_before[i] + block_sums[i >> 10] + (head ? 1u : 0u) - 1u;
      v = vals ? (u32)vals[i] : 1u;
      if (head) {
        out_keys[(u64)u * nw] = keys[i * nw];
        if (nw == 2) out_keys[(u64)u * nw + 1] = keys[i * nw + 1];
      }
    }
    const u32 peers = __match_any_sync(0xffffffffu, u);
    const u32 sum = __reduce_add_sync(peers, v);
    if (i < n && (peers & ((1u << lane) - 1u)) == 0) atomicAdd(reinterpret_cast<u32*>(out_vals + u), sum);   // low half of a zeroed u64
  }
}

// Kmer::operator< on (hi, lo)
__device__ __forceinline__ int rs_cmp(const u64* __restrict__ keys, u64 i, int nw, u64 q0, u64 q1) {
  if (nw == 2) {
    const u64 h = keys[i * 2 + 1];
    if (h != q1) return h < q1 ? -1 : 1;
  }
  const u64 l = keys[i * nw];
  return l < q0 ? -1 : (l > q0 ? 1 : 0);
}
// binary search of every query: lo = first entry >= q, cnt = number of entries == q (map: 0/1; multimap: multiplicity)
__global__ void __launch_bounds__(256)
k_sorted_range(const u64* __restrict__ keys, u64 n, int nw, const u64* __restrict__ q, u64 nq, u64* __restrict__ lo_out, u64* __restrict__ cnt_out) {
  for (u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x; t < nq; t += (u64)gridDim.x * blockDim.x) {
    const u64 q0 = q[t * nw], q1 = nw == 2 ? q[t * nw + 1] : 0;
    u64 a = 0, b = n;                       // lower bound
    while (a < b) { const u64 m = (a + b) >> 1; if (rs_cmp(keys, m, nw, q0, q1) < 0) a = m + 1; else b = m; }
    const u64 lo = a;
    b = n;                                  // upper bound
    while (a < b) { const u64 m = (a + b) >> 1; if (rs_cmp(keys, m, nw, q0, q1) <= 0) a = m + 1; else b = m; }
    lo_out[t] = lo;
    cnt_out[t] = a - lo;
  }
}
__global__ void __launch_bounds__(256)
k_sorted_pick(const u64* __restrict__ vals, const u64* __restrict__ lo, const u64* __restrict__ cnt, u64 nq, u64* __restrict__ out_vals,
              uint8_t* __restrict__ out_found) {
  for (u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x; t < nq; t += (u64)gridDim.x * blockDim.x) {
    const bool hit = cnt[t] != 0;
    if (out_vals) out_vals[t] = hit ? vals[lo[t]] : 0;
    if (out_found) out_found[t] = hit ? 1 : 0;
  }
}
__global__ void __launch_bounds__(256)
k_sorted_gather(const u64* __restrict__ vals, const u64* __restrict__ lo, const u64* __restrict__ off, u64 nq, u64* __restrict__ out, u64 cap) {
  for (u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x; t < nq; t += (u64)gridDim.x * blockDim.x) {
    const u64 b = off[t], e = off[t + 1], src = lo[t];
    for (u64 j = b; j < e && j < cap; ++j) out[j] = vals[src + (j - b)];
  }
}

unsigned grid1(u64 n, unsigned block) { return grid_for(n, block, 148u * 16u); }

}  // namespace

namespace kmi {

// Stable LSD radix sort of n entries by (key, value): value digits first when val_bits > 0, then the key words from the least
// significant. keys/vals and their *_tmp twins ping-pong; on return keys/vals point at the sorted data. vals may be null
// (keys only). hist / sums are scratch. Everything is queued on `st`; *launches is advanced.
int device_sort_pairs(cudaStream_t st, DevBuf& hist, DevBuf& sums, u64*& keys, u64*& vals, u64*& keys_tmp, u64*& vals_tmp, u64 n,
                      int nw, int key_bits, int val_bits, u64* launches) {
  if (n < 2) return KMI_OK;
  if (n >= 0xffffffffULL) return fail(KMI_ERR_ARG, "sort: at most 2^32-2 entries per call");
  const u64 n_chunks = (n + RS_ITEMS - 1) / RS_ITEMS;
  const u64 n_hist = n_chunks * RS_BINS;
  const u64 n_blocks = (n_hist + 1023) / 1024;
  OK(hist.ensure(sizeof(u32) * (n_hist + 1)));
  OK(sums.ensure(sizeof(u32) * (n_blocks + 1)));
  auto pass = [&](int word, int shift) {
    k_rs_hist<<<(unsigned)n_chunks, RS_THREADS, 0, st>>>(keys, vals, n, nw, word, shift, hist.as<u32>(), n_chunks);
    k_rs_scan_l1<<<(unsigned)n_blocks, 1024, 0, st>>>(hist.as<u32>(), n_hist, sums.as<u32>());
    k_rs_scan_top<<<1, 1024, 0, st>>>(sums.as<u32>(), n_blocks, nullptr);
    k_rs_scatter<<<(unsigned)n_chunks, RS_THREADS, 0, st>>>(keys, vals, n, nw, word, shift, hist.as<u32>(), sums.as<u32>(), n_chunks,
                                                           keys_tmp, vals_tmp);
    std::swap(keys, keys_tmp);
    std::swap(vals, vals_tmp);
    if (launches) *launches += 4;
  };
  if (vals && val_bits > 0)
    for (int s = 0; s < val_bits; s += 8) pass(nw, s);
  for (int w = 0; w < nw; ++w) {
    const int bits = std::min(64, key_bits - 64 * w);
    for (int s = 0; s < bits; s += 8) pass(w, s);
  }
  CU(cudaGetLastError());
  return KMI_OK;
}

}  // namespace kmi

// ---- the sorted map ---------------------------------------------------------------------------------------------------------
struct kmi_sorted {
  kmi_index* base = nullptr;      // parser / canonicaliser / stream of the same configuration (its hash table is never built)
  int nw = 1, mode = 0;
  DevBuf keys[2], vals[2];        // [cur]: n_sorted reduced entries followed by n_pend appended ones; [cur ^ 1]: sort scratch
  int cur = 0;
  u64 n_sorted = 0, n_pend = 0;
  bool pend_all_ones = true;      // every pending COUNT entry has value 1 (keys-only sort of a fresh build)
  DevBuf hist, sums, flags, fsums, q_keys, q_lo, q_cnt, q_out, q_found, scratch;
  u64 launches = 0;
  float ms[3] = {0, 0, 0};        // last rehash: sort, reduce; last build: parse
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};

namespace {

cudaStream_t sstream(kmi_sorted* s) { return s->base->stream; }

// room for `extra` more entries behind what is held (grows both ping-pong halves, keeps the contents)
int sorted_reserve(kmi_sorted* s, u64 extra) {
  const u64 have = s->n_sorted + s->n_pend, need = have + extra;
  const size_t kb = sizeof(u64) * (size_t)s->nw;
  if (need * kb > s->keys[s->cur].cap || need * sizeof(u64) > s->vals[s->cur].cap) {
    DevBuf nk, nv;
    const u64 want = need + need / 4 + 1024;
    OK(nk.ensure(want * kb));
    OK(nv.ensure(want * sizeof(u64)));
    if (have) {
      CU(cudaMemcpyAsync(nk.p, s->keys[s->cur].p, have * kb, cudaMemcpyDeviceToDevice, sstream(s)));
      CU(cudaMemcpyAsync(nv.p, s->vals[s->cur].p, have * sizeof(u64), cudaMemcpyDeviceToDevice, sstream(s)));
      CU(cudaStreamSynchronize(sstream(s)));
    }
    s->keys[s->cur].release(); s->vals[s->cur].release();
    s->keys[s->cur] = nk; s->vals[s->cur] = nv;
  }
  return KMI_OK;
}

// entries [n_sorted + n_pend, +n) were just written (keys always; vals when given): account for them
int sorted_appended(kmi_sorted* s, u64 n, bool vals_written) {
  if (!n) return KMI_OK;
  u64* v = s->vals[s->cur].as<u64>() + s->n_sorted + s->n_pend;
  if (s->mode == KMI_MODE_COUNT) {
    if (!vals_written) { k_fill_u64<<<grid1(n, 256), 256, 0, sstream(s)>>>(v, n, 1ull); s->launches++; }
    else s->pend_all_ones = false;
  }
  s->n_pend += n;
  return KMI_OK;
}

int sorted_rehash(kmi_sorted* s) {
  if (!s->n_pend) return KMI_OK;
  CU(cudaSetDevice(s->base->cfg.device));
  cudaStream_t st = sstream(s);
  const u64 n = s->n_sorted + s->n_pend;
  const size_t kb = sizeof(u64) * (size_t)s->nw;
  const int o = s->cur ^ 1;
  OK(s->keys[o].ensure(n * kb));
  OK(s->vals[o].ensure(n * sizeof(u64)));
  u64 *k = s->keys[s->cur].as<u64>(), *v = s->vals[s->cur].as<u64>(), *kt = s->keys[o].as<u64>(), *vt = s->vals[o].as<u64>();
  const bool count = s->mode == KMI_MODE_COUNT;
  const bool keys_only = count && s->n_sorted == 0 && s->pend_all_ones;
  int val_bits = 0;
  if (!count) {                       // multimap entries are ordered by (key, id): how many id digits are in use
    OK(s->scratch.ensure(sizeof(u64)));
    CU(cudaMemsetAsync(s->scratch.p, 0, sizeof(u64), st));
    k_rs_or<<<grid1(n, 256), 256, 0, st>>>(v, n, s->scratch.as<u64>());
    u64 h = 0;
    CU(cudaMemcpyAsync(&h, s->scratch.p, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    while (h) { ++val_bits; h >>= 1; }
    s->launches++;
  }
  cudaEventRecord(s->ev[0], st);
  u64* vs = keys_only ? nullptr : v;
  u64* vts = keys_only ? nullptr : vt;
  OK(device_sort_pairs(st, s->hist, s->sums, k, vs, kt, vts, n, s->nw, 2 * s->base->cfg.k, val_bits, &s->launches));
  cudaEventRecord(s->ev[1], st);
  // k (and vs) now point at the sorted data, which lives in one of the two halves
  const int where = (k == s->keys[s->cur].as<u64>()) ? s->cur : o;
  u64 n_out = n;
  if (count) {
    const u64 n_blocks = (n + 1023) / 1024;
    OK(s->flags.ensure(sizeof(u32) * (n + 1)));
    OK(s->fsums.ensure(sizeof(u32) * (n_blocks + 2)));
    k_run_heads<<<grid1(n, 256), 256, 0, st>>>(k, n, s->nw, s->flags.as<u32>());
    k_rs_scan_l1<<<(unsigned)n_blocks, 1024, 0, st>>>(s->flags.as<u32>(), n, s->fsums.as<u32>());
    k_rs_scan_top<<<1, 1024, 0, st>>>(s->fsums.as<u32>(), n_blocks, s->fsums.as<u32>() + n_blocks + 1);
    u32 n_runs = 0;
    CU(cudaMemcpyAsync(&n_runs, s->fsums.as<u32>() + n_blocks + 1, sizeof(u32), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    // reduce into the other half
    const int dst = where ^ 1;
    OK(s->keys[dst].ensure((u64)n_runs * kb + kb));
    OK(s->vals[dst].ensure((u64)n_runs * sizeof(u64) + 8));
    CU(cudaMemsetAsync(s->vals[dst].p, 0, (u64)n_runs * sizeof(u64), st));
    const u64* sv = keys_only ? nullptr : s->vals[where].as<u64>();
    k_reduce_runs<<<grid1(n, 256), 256, 0, st>>>(s->keys[where].as<u64>(), sv, n, s->nw, s->flags.as<u32>(), s->fsums.as<u32>(),
                                                s->keys[dst].as<u64>(), s->vals[dst].as<u64>());
    s->launches += 4;
    s->cur = dst;
    n_out = n_runs;
  } else {
    s->cur = where;
  }
  cudaEventRecord(s->ev[2], st);
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  cudaEventElapsedTime(&s->ms[0], s->ev[0], s->ev[1]);
  cudaEventElapsedTime(&s->ms[1], s->ev[1], s->ev[2]);
  s->n_sorted = n_out;
  s->n_pend = 0;
  s->pend_all_ones = true;
  return KMI_OK;
}

// queries to the device, canonicalised like the reference's transform_input; lo / cnt per query in q_lo / q_cnt
int sorted_ranges(kmi_sorted* s, const uint64_t* keys, u64 n, bool host) {
  OK(sorted_rehash(s));
  CU(cudaSetDevice(s->base->cfg.device));
  cudaStream_t st = sstream(s);
  const size_t kb = sizeof(u64) * (size_t)s->nw;
  OK(s->q_keys.ensure(n * kb + 8));
  OK(s->q_lo.ensure(n * sizeof(u64) + 8));
  OK(s->q_cnt.ensure(n * sizeof(u64) + 8));
  if (!n) return KMI_OK;
  if (host) CU(cudaMemcpyAsync(s->q_keys.p, keys, n * kb, cudaMemcpyHostToDevice, st));
  else CU(cudaMemcpyAsync(s->q_keys.p, keys, n * kb, cudaMemcpyDeviceToDevice, st));
  if (s->base->cfg.canonical) OK(kmi_canonicalize_dev(s->base, s->q_keys.as<uint64_t>(), s->q_keys.as<uint64_t>(), n));
  k_sorted_range<<<grid_for(n, 256, 148u * 32u), 256, 0, st>>>(s->keys[s->cur].as<u64>(), s->n_sorted, s->nw, s->q_keys.as<u64>(), n,
                                                              s->q_lo.as<u64>(), s->q_cnt.as<u64>());
  s->launches++;
  CU(cudaGetLastError());
  return KMI_OK;
}

}  // namespace

extern "C" int kmi_sorted_create(const kmi_config* cfg, kmi_sorted** out) {
  if (!cfg || !out) return fail(KMI_ERR_ARG, "null argument");
  *out = nullptr;
  if (cfg->nranks != 1) return fail(KMI_ERR_UNSUPPORTED, "the sorted back end is one shard (the reference's sample-sort redistribution is not built)");
  kmi_index* base = nullptr;
  OK(kmi_create(cfg, &base));
  kmi_sorted* s = new (std::nothrow) kmi_sorted();
  if (!s) { kmi_destroy(base); return fail(KMI_ERR_NOMEM, "host allocation"); }
  s->base = base; s->nw = base->nw; s->mode = cfg->mode;
  for (int i = 0; i < 4; ++i) cudaEventCreate(&s->ev[i]);
  *out = s;
  return KMI_OK;
}

extern "C" void kmi_sorted_destroy(kmi_sorted* s) {
  if (!s) return;
  cudaSetDevice(s->base->cfg.device);
  for (int i = 0; i < 2; ++i) { s->keys[i].release(); s->vals[i].release(); }
  DevBuf* bufs[] = {&s->hist, &s->sums, &s->flags, &s->fsums, &s->q_keys, &s->q_lo, &s->q_cnt, &s->q_out, &s->q_found, &s->scratch};
  for (DevBuf* b : bufs) b->release();
  for (int i = 0; i < 4; ++i) if (s->ev[i]) cudaEventDestroy(s->ev[i]);
  kmi_destroy(s->base);
  delete s;
}

// parse straight behind the held entries; fasta: 0 FASTQ, 1 FASTA
static int sorted_build_dev(kmi_sorted* s, const uint8_t* d_bytes, size_t n, uint64_t file_offset, int fasta, uint64_t* n_kmers) {
  if (!s || (!d_bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  if (n_kmers) *n_kmers = 0;
  if (!n) return KMI_OK;
  CU(cudaSetDevice(s->base->cfg.device));
  const u64 cap = fasta ? (u64)n : (u64)n / 2 + 64;       // a FASTQ record spends at least as many bytes on qualities as on bases
  OK(sorted_reserve(s, cap));
  const u64 at = s->n_sorted + s->n_pend;
  u64* dk = s->keys[s->cur].as<u64>() + at * s->nw;
  u64* dv = s->mode == KMI_MODE_POSITION ? s->vals[s->cur].as<u64>() + at : nullptr;
  uint64_t got = 0;
  cudaEventRecord(s->ev[3], sstream(s));
  if (fasta) OK(kmi_parse_fasta_dev(s->base, d_bytes, n, reinterpret_cast<uint64_t*>(dk), reinterpret_cast<uint64_t*>(dv), cap, &got));
  else OK(kmi_parse_fastq_dev(s->base, d_bytes, n, file_offset, reinterpret_cast<uint64_t*>(dk), reinterpret_cast<uint64_t*>(dv), cap, &got));
  cudaEventRecord(s->ev[0], sstream(s));
  CU(cudaStreamSynchronize(sstream(s)));
  cudaEventElapsedTime(&s->ms[2], s->ev[3], s->ev[0]);
  OK(sorted_appended(s, got, dv != nullptr));
  if (n_kmers) *n_kmers = got;
  return KMI_OK;
}
static int sorted_build_host(kmi_sorted* s, const uint8_t* bytes, size_t n, uint64_t file_offset, int fasta, uint64_t* n_kmers) {
  if (!s || (!bytes && n)) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(s->base->cfg.device));
  DevBuf text;
  OK(text.ensure(n + 64));
  int r = copy_to_device(s->base, text.p, bytes, n);
  if (r == KMI_OK) r = sorted_build_dev(s, text.as<uint8_t>(), n, file_offset, fasta, n_kmers);
  cudaStreamSynchronize(sstream(s));
  text.release();
  return r;
}
extern "C" int kmi_sorted_build_fastq_dev(kmi_sorted* s, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  return sorted_build_dev(s, d_bytes, n, file_offset, 0, n_kmers);
}
extern "C" int kmi_sorted_build_fasta_dev(kmi_sorted* s, const uint8_t* d_bytes, size_t n, uint64_t* n_kmers) {
  return sorted_build_dev(s, d_bytes, n, 0, 1, n_kmers);
}
extern "C" int kmi_sorted_build_fastq(kmi_sorted* s, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers) {
  return sorted_build_host(s, bytes, n, file_offset, 0, n_kmers);
}
extern "C" int kmi_sorted_build_fasta(kmi_sorted* s, const uint8_t* bytes, size_t n, uint64_t* n_kmers) {
  return sorted_build_host(s, bytes, n, 0, 1, n_kmers);
}

static int sorted_insert_any(kmi_sorted* s, const uint64_t* keys, const uint64_t* vals, size_t n, bool host) {
  if (!s || (!keys && n)) return fail(KMI_ERR_ARG, "null argument");
  if (s->mode == KMI_MODE_POSITION && !vals && n) return fail(KMI_ERR_ARG, "a POSITION map inserts (key, id) tuples");
  if (!n) return KMI_OK;
  CU(cudaSetDevice(s->base->cfg.device));
  OK(sorted_reserve(s, n));
  const u64 at = s->n_sorted + s->n_pend;
  const size_t kb = sizeof(u64) * (size_t)s->nw;
  u64* dk = s->keys[s->cur].as<u64>() + at * s->nw;
  u64* dv = s->vals[s->cur].as<u64>() + at;
  const cudaMemcpyKind kind = host ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
  CU(cudaMemcpyAsync(dk, keys, n * kb, kind, sstream(s)));
  if (vals) CU(cudaMemcpyAsync(dv, vals, n * sizeof(u64), kind, sstream(s)));
  if (s->base->cfg.canonical) OK(kmi_canonicalize_dev(s->base, reinterpret_cast<uint64_t*>(dk), reinterpret_cast<uint64_t*>(dk), n));
  if (host) CU(cudaStreamSynchronize(sstream(s)));
  return sorted_appended(s, n, vals != nullptr);
}
extern "C" int kmi_sorted_insert(kmi_sorted* s, const uint64_t* keys, const uint64_t* vals, size_t n) { return sorted_insert_any(s, keys, vals, n, true); }
extern "C" int kmi_sorted_insert_dev(kmi_sorted* s, const uint64_t* d_keys, const uint64_t* d_vals, size_t n) { return sorted_insert_any(s, d_keys, d_vals, n, false); }

extern "C" int kmi_sorted_rehash(kmi_sorted* s) {
  if (!s) return fail(KMI_ERR_ARG, "null map");
  return sorted_rehash(s);
}

extern "C" int kmi_sorted_local_size(kmi_sorted* s, uint64_t* n) {
  if (!s || !n) return fail(KMI_ERR_ARG, "null argument");
  OK(sorted_rehash(s));
  *n = s->n_sorted;
  return KMI_OK;
}

extern "C" int kmi_sorted_count(kmi_sorted* s, const uint64_t* keys, size_t n, uint64_t* counts) {
  if (!s || ((!keys || !counts) && n)) return fail(KMI_ERR_ARG, "null argument");
  OK(sorted_ranges(s, keys, n, true));
  if (n) CU(cudaMemcpyAsync(counts, s->q_cnt.p, n * sizeof(u64), cudaMemcpyDeviceToHost, sstream(s)));
  CU(cudaStreamSynchronize(sstream(s)));
  return KMI_OK;
}

extern "C" int kmi_sorted_find(kmi_sorted* s, const uint64_t* keys, size_t n, uint64_t* vals, uint8_t* found) {
  if (!s || (!keys && n)) return fail(KMI_ERR_ARG, "null argument");
  if (s->mode != KMI_MODE_COUNT) return fail(KMI_ERR_ARG, "find with one value per key belongs to a COUNT map; use kmi_sorted_find_positions");
  OK(sorted_ranges(s, keys, n, true));
  if (!n) return KMI_OK;
  OK(s->q_out.ensure(n * sizeof(u64)));
  OK(s->q_found.ensure(n));
  k_sorted_pick<<<grid_for(n, 256), 256, 0, sstream(s)>>>(s->vals[s->cur].as<u64>(), s->q_lo.as<u64>(), s->q_cnt.as<u64>(), n, s->q_out.as<u64>(),
                                                        s->q_found.as<uint8_t>());
  s->launches++;
  if (vals) CU(cudaMemcpyAsync(vals, s->q_out.p, n * sizeof(u64), cudaMemcpyDeviceToHost, sstream(s)));
  if (found) CU(cudaMemcpyAsync(found, s->q_found.p, n, cudaMemcpyDeviceToHost, sstream(s)));
  CU(cudaStreamSynchronize(sstream(s)));
  CU(cudaGetLastError());
  return KMI_OK;
}

extern "C" int kmi_sorted_find_positions(kmi_sorted* s, const uint64_t* keys, size_t n, uint64_t* offsets, uint64_t* vals, size_t cap) {
  if (!s || (!keys && n) || !offsets) return fail(KMI_ERR_ARG, "null argument");
  if (s->mode != KMI_MODE_POSITION) return fail(KMI_ERR_ARG, "find_positions belongs to a POSITION map");
  OK(sorted_ranges(s, keys, n, true));
  offsets[0] = 0;
  if (!n) return KMI_OK;
  std::vector<u64> cnt(n);
  CU(cudaMemcpyAsync(cnt.data(), s->q_cnt.p, n * sizeof(u64), cudaMemcpyDeviceToHost, sstream(s)));
  CU(cudaStreamSynchronize(sstream(s)));
  for (size_t i = 0; i < n; ++i) offsets[i + 1] = offsets[i] + cnt[i];
  if (offsets[n] > cap) return fail(KMI_ERR_CAPACITY, "find_positions: %llu values, capacity %zu", (unsigned long long)offsets[n], cap);
  if (!offsets[n] || !vals) return KMI_OK;
  OK(s->q_out.ensure(offsets[n] * sizeof(u64)));
  OK(s->q_found.ensure((n + 1) * sizeof(u64)));
  CU(cudaMemcpyAsync(s->q_found.p, offsets, (n + 1) * sizeof(u64), cudaMemcpyHostToDevice, sstream(s)));
  k_sorted_gather<<<grid_for(n, 256), 256, 0, sstream(s)>>>(s->vals[s->cur].as<u64>(), s->q_lo.as<u64>(), s->q_found.as<u64>(), n, s->q_out.as<u64>(),
                                                          offsets[n]);
  s->launches++;
  CU(cudaMemcpyAsync(vals, s->q_out.p, offsets[n] * sizeof(u64), cudaMemcpyDeviceToHost, sstream(s)));
  CU(cudaStreamSynchronize(sstream(s)));
  CU(cudaGetLastError());
  return KMI_OK;
}

extern "C" int kmi_sorted_dump(kmi_sorted* s, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out) {
  if (!s || !n_out) return fail(KMI_ERR_ARG, "null argument");
  OK(sorted_rehash(s));
  const u64 m = std::min<u64>(s->n_sorted, cap);
  if (m && keys) CU(cudaMemcpyAsync(keys, s->keys[s->cur].p, m * sizeof(u64) * s->nw, cudaMemcpyDeviceToHost, sstream(s)));
  if (m && vals) CU(cudaMemcpyAsync(vals, s->vals[s->cur].p, m * sizeof(u64), cudaMemcpyDeviceToHost, sstream(s)));
  CU(cudaStreamSynchronize(sstream(s)));
  *n_out = m;
  if (s->n_sorted > cap) return fail(KMI_ERR_CAPACITY, "dump: %llu entries, capacity %zu", (unsigned long long)s->n_sorted, cap);
  return KMI_OK;
}

extern "C" int kmi_sorted_stats(kmi_sorted* s, float* ms3, uint64_t* launches) {
  if (!s) return fail(KMI_ERR_ARG, "null map");
  if (ms3) { ms3[0] = s->ms[2]; ms3[1] = s->ms[0]; ms3[2] = s->ms[1]; }
  if (launches) *launches = s->launches + s->base->launches;
  return KMI_OK;
}

// the sort by itself, on caller-owned device arrays (d_keys_tmp / d_vals_tmp: scratch of the same sizes). On return
// *sorted_in_tmp says which pair holds the result.
extern "C" int kmi_sort_pairs_dev(kmi_index* idx, uint64_t* d_keys, uint64_t* d_vals, uint64_t* d_keys_tmp, uint64_t* d_vals_tmp, size_t n,
                                  int val_bits, int* sorted_in_tmp) {
  if (!idx || !d_keys || !d_keys_tmp || (d_vals && !d_vals_tmp) || !sorted_in_tmp) return fail(KMI_ERR_ARG, "null argument");
  if (val_bits < 0 || val_bits > 64) return fail(KMI_ERR_ARG, "val_bits %d", val_bits);
  CU(cudaSetDevice(idx->cfg.device));
  u64 *k = reinterpret_cast<u64*>(d_keys), *v = reinterpret_cast<u64*>(d_vals), *kt = reinterpret_cast<u64*>(d_keys_tmp), *vt = reinterpret_cast<u64*>(d_vals_tmp);
  OK(device_sort_pairs(idx->stream, idx->hist, idx->hist_sums, k, v, kt, vt, n, idx->nw, 2 * idx->cfg.k, d_vals ? val_bits : 0, &idx->launches));
  *sorted_in_tmp = (k == reinterpret_cast<u64*>(d_keys)) ? 0 : 1;
  return KMI_OK;
}

// to_vector() in key order (COUNT: by key; POSITION: by (key, id)): the table's entries / the tuple log go through the device
// sort instead of a host std::sort — what kmi_save writes.
extern "C" int kmi_dump_sorted(kmi_index* idx, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out) {
  if (!idx || !n_out) return fail(KMI_ERR_ARG, "null argument");
  CU(cudaSetDevice(idx->cfg.device));
  *n_out = 0;
  uint64_t n = 0;
  OK(kmi_local_size(idx, &n));
  if (n > cap) return fail(KMI_ERR_CAPACITY, "dump: %llu entries, capacity %zu", (unsigned long long)n, cap);
  if (n == 0) return KMI_OK;
  const size_t kb = sizeof(u64) * (size_t)idx->nw;
  cudaStream_t st = idx->stream;
  DevBuf dk[2], dv[2], orb;
  struct Release { DevBuf* a; DevBuf* b; DevBuf* c; ~Release() { a[0].release(); a[1].release(); b[0].release(); b[1].release(); c->release(); } } rel{dk, dv, &orb};
  for (int i = 0; i < 2; ++i) { OK(dk[i].ensure(n * kb)); OK(dv[i].ensure(n * sizeof(u64))); }
  int val_bits = 0;
  if (idx->cfg.mode == KMI_MODE_POSITION) {
    CU(cudaMemcpyAsync(dk[0].p, idx->log_keys.p, n * kb, cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(dv[0].p, idx->log_ids.p, n * sizeof(u64), cudaMemcpyDeviceToDevice, st));
    OK(orb.ensure(sizeof(u64)));
    CU(cudaMemsetAsync(orb.p, 0, sizeof(u64), st));
    k_rs_or<<<grid1(n, 256), 256, 0, st>>>(dv[0].as<u64>(), n, orb.as<u64>());
    u64 h = 0;
    CU(cudaMemcpyAsync(&h, orb.p, sizeof(u64), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    while (h) { ++val_bits; h >>= 1; }
    idx->launches++;
  } else {
    OK(zero_ctr(idx, CTR_CURSOR));
    idx->touch();
    idx->ops->dump(st, idx->slots, idx->cap, dk[0].as<u64>(), dv[0].as<u64>(), n, idx->ctr + CTR_CURSOR, -1, 0);
    idx->launches++;
  }
  u64 *k = dk[0].as<u64>(), *v = dv[0].as<u64>(), *kt = dk[1].as<u64>(), *vt = dv[1].as<u64>();
  OK(device_sort_pairs(st, idx->hist, idx->hist_sums, k, v, kt, vt, n, idx->nw, 2 * idx->cfg.k, val_bits, &idx->launches));
  if (keys) OK(copy_to_host(idx, keys, k, n * kb));
  if (vals) OK(copy_to_host(idx, vals, v, n * sizeof(u64)));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  *n_out = n;
  return KMI_OK;
}
