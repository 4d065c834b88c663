// kmi_kernels.cuh — table, query, partition and utility kernels (sm_100a).
#pragma once
#include "kmi_block.cuh"

namespace kmi {

// ---- table maintenance ---------------------------------------------------------------------
template <int NW> __global__ void k_table_clear(Slot<NW>* __restrict__ slots, u64 n) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    if (NW == 1) {
      reinterpret_cast<ulonglong2*>(slots)[i] = make_ulonglong2(KEY_EMPTY_TOP, 0ULL);
    } else {
      ulonglong2* p = reinterpret_cast<ulonglong2*>(slots) + 2 * i;
      p[0] = make_ulonglong2(0ULL, KEY_EMPTY_TOP);
      p[1] = make_ulonglong2(0ULL, 0ULL);
    }
  }
}

// a15: counting insert, distributed_densehash_map.hpp:2463-2482 / :2714-2899 (val += 1 per key); with `vals` the
// reduction insert of (key, count) tuples: val += count, std::plus<uint32_t> (wraps at 2^32), :2565-2628
template <int K, int HASH>
__global__ void __launch_bounds__(256) k_insert_keys(TableView t, const u64* __restrict__ keys, u64 n, const bool CANON,
                                                     const u64* __restrict__ vals) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  u32 n_new = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    Key<NW> key = load_key<NW>(keys, i);
    if (CANON) key = canonical<K>(key);
    bool was_new;
    u64 s = table_upsert<NW>(t, key, slot_hash<K>(key), was_new);
    if (s != ~0ULL) {
      table_add<NW>(t, s, vals ? (u32)vals[i] : 1u);
      n_new += was_new ? 1u : 0u;
    }
  }
  n_new = warp_sum(n_new);
  if ((threadIdx.x & 31) == 0 && n_new) atomicAdd(t.size, (u64)n_new);
}

// growth: re-insert every live entry of the old table (dense_hash_map's rehash on resize)
template <int K, int HASH>
__global__ void __launch_bounds__(256) k_rehash(const Slot<(K <= 32 ? 1 : 2)>* __restrict__ old, u64 n_old, TableView t) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  u32 n_new = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n_old; i += (u64)gridDim.x * blockDim.x) {
    Key<NW> key = slot_key<NW>(old + i);
    if (!is_live<NW>(key)) continue;
    bool was_new;
    u64 s = table_upsert<NW>(t, key, slot_hash<K>(key), was_new);
    if (s != ~0ULL) {
      table_add<NW>(t, s, old[i].val);
      n_new += was_new ? 1u : 0u;
    }
  }
  n_new = warp_sum(n_new);
  if ((threadIdx.x & 31) == 0 && n_new) atomicAdd(t.size, (u64)n_new);
}

// a18: batched probe. vals[i] = stored value (0 when absent), found[i] = 0/1; slot_idx optional.
template <int K, int HASH>
__global__ void __launch_bounds__(256)
k_lookup(TableView t, const u64* __restrict__ keys, u64 n, u64* __restrict__ vals, uint8_t* __restrict__ found,
         u64* __restrict__ slot_idx, const bool CANON) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    Key<NW> key = load_key<NW>(keys, i);
    if (CANON) key = canonical<K>(key);
    u32 v;
    u64 s = table_find_val<NW>(t, key, slot_hash<K>(key), v);
    if (vals) vals[i] = (u64)v;
    if (found) found[i] = (s == ~0ULL) ? 0 : 1;
    if (slot_idx) slot_idx[i] = s;
  }
}

// erase: tombstone the key (dense_hash_map deleted key), distributed_densehash_map.hpp:1602-1663
template <int K, int HASH>
__global__ void __launch_bounds__(256) k_erase(TableView t, const u64* __restrict__ keys, u64 n, u64* __restrict__ vals_removed, const bool CANON) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    Key<NW> key = load_key<NW>(keys, i);
    if (CANON) key = canonical<K>(key);
    u64 s = table_find<NW>(t, key, slot_hash<K>(key));
    if (s == ~0ULL) continue;
    Slot<NW>* sl = reinterpret_cast<Slot<NW>*>(t.slots) + s;
    bool won;
    if (NW == 1) {
      won = atomicCAS(&reinterpret_cast<Slot<1>*>(sl)->key, key.w[0], KEY_DELETED_TOP) == key.w[0];
    } else {
      u64 olo, ohi;
      cas128(sl, key.w[0], key.w[NW - 1], 0ULL, KEY_DELETED_TOP, olo, ohi);
      won = (olo == key.w[0] && ohi == key.w[NW - 1]);
    }
    if (won) {
      atomicAdd(t.size, ~0ULL);  // -1
      if (vals_removed) atomicAdd(vals_removed, (u64)sl->val);
    }
  }
}

// value predicate evaluated on the device (find(pred) / count(pred) / erase(pred) of the reference with the predicates its
// drivers use: frequency thresholds; kmer_index.hpp:156-194, distributed_densehash_map.hpp:1666-1686,1835-1903).
// op: -1 always, 0 val >= x, 1 val <= x, 2 val == x
__device__ __forceinline__ bool value_pred(int op, u64 x, u32 val) {
  return op < 0 || (op == 0 ? (u64)val >= x : op == 1 ? (u64)val <= x : (u64)val == x);
}
// to_vector() / find(pred): compact the live entries that satisfy the predicate (unordered). keys == nullptr: count only.
template <int NW>
__global__ void __launch_bounds__(256)
k_dump(const Slot<NW>* __restrict__ slots, u64 n_slots, u64* __restrict__ keys, u64* __restrict__ vals, u64 cap,
       u64* __restrict__ cursor, const int op, const u64 x) {
  for (u64 base = (u64)blockIdx.x * blockDim.x; base < n_slots; base += (u64)gridDim.x * blockDim.x) {
    u64 i = base + threadIdx.x;
    Key<NW> key; key.w[NW - 1] = KEY_EMPTY_TOP;
    if (i < n_slots) key = slot_key<NW>(slots + i);
    bool live = is_live<NW>(key);
    if (live && op >= 0) live = value_pred(op, x, slots[i].val);
    u32 m = __ballot_sync(0xffffffffu, live);
    u64 wbase = 0;
    if ((threadIdx.x & 31) == 0 && m) wbase = atomicAdd(cursor, (u64)__popc(m));
    wbase = __shfl_sync(0xffffffffu, wbase, 0);
    if (live && keys) {
      u64 at = wbase + __popc(m & ((1u << (threadIdx.x & 31)) - 1));
      if (at < cap) {
        store_key<NW>(keys, at, key);
        if (vals) vals[at] = slots[i].val;
      }
    }
  }
}

// erase(pred): tombstone every live entry whose value satisfies the predicate; *removed += entries erased
template <int NW>
__global__ void __launch_bounds__(256)
k_erase_if(Slot<NW>* __restrict__ slots, u64 n_slots, u64* __restrict__ size, u64* __restrict__ removed, const int op, const u64 x) {
  u32 n = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (u64)gridDim.x * blockDim.x) {
    const Key<NW> key = slot_key<NW>(slots + i);
    if (!is_live<NW>(key) || !value_pred(op, x, slots[i].val)) continue;
    if (NW == 1) reinterpret_cast<Slot<1>*>(slots + i)->key = KEY_DELETED_TOP;   // one thread per slot: no race inside this kernel
    else { Slot<2>* s2 = reinterpret_cast<Slot<2>*>(slots + i); s2->lo = 0; s2->hi = KEY_DELETED_TOP; }
    ++n;
  }
  n = warp_sum(n);
  if ((threadIdx.x & 31) == 0 && n) { atomicAdd(size, ~(u64)n + 1); atomicAdd(removed, (u64)n); }
}

// order-independent checksums: sum mix(key)*val, sum val
template <int NW>
__global__ void __launch_bounds__(256) k_checksum(const Slot<NW>* __restrict__ slots, u64 n_slots, u64* __restrict__ out2) {
  u64 a = 0, b = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots; i += (u64)gridDim.x * blockDim.x) {
    Key<NW> key = slot_key<NW>(slots + i);
    if (!is_live<NW>(key)) continue;
    u64 v = slots[i].val;
    u64 m = mix64(key.w[0]);
    if (NW == 2) m = mix64(m ^ key.w[NW - 1]);
    a += m * v; b += v;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, d);
    b += __shfl_xor_sync(0xffffffffu, b, d);
  }
  if ((threadIdx.x & 31) == 0) { atomicAdd(out2, a); atomicAdd(out2 + 1, b); }
}

// ---- value transforms --------------------------------------------------------------------------
template <int K> __global__ void __launch_bounds__(256) k_canonicalize(const u64* in, u64* out, u64 n) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x)
    store_key<NW>(out, i, canonical<K>(load_key<NW>(in, i)));
}
template <int K, int HASH> __global__ void __launch_bounds__(256) k_hash(int which, const u64* __restrict__ in, u64* __restrict__ out, u64 n, const u32 pbits) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    u64 s, d;
    kmer_hashes<K, HASH>(load_key<NW>(in, i), s, d, pbits);
    out[i] = which ? d : s;
  }
}

// ---- a13/a14: owner = dist_hash % p, stable counting-sort by owner -----------------------------------
// Replaces imxx::local::assign_to_buckets / bucket_to_permutation / permute
// (reference src/io/incremental_mxx.hpp:273-364,596-640). Chunk = PART_ITEMS consecutive keys per
// block; histogram is [owner][chunk] so that one exclusive scan yields every (owner, chunk) base.
static constexpr int PART_THREADS = 256;
static constexpr int PART_ROUNDS = 8;
static constexpr int PART_ITEMS = PART_THREADS * PART_ROUNDS;
static constexpr int MAX_RANKS = 256;

// `trans` = the DistTrans of the reference's map parameters (SingleStrandHashMapParams<Key, DistHash, StoreHash, DistTrans>,
// kmer_index.hpp; BenchmarkKmerIndex.cpp:152-161): 0 identity, 1 lex_less, 2 xor_rev_comp (x ^ rc(x),
// kmer_transform.hpp:91-107) — applied to the key for the OWNER computation only; the key itself travels unchanged.
// (A canonical index has lex_less keys already, and lex_less of a canonical key is itself.)
template <int K, int HASH>
__device__ __forceinline__ u32 owner_of(const Key<(K <= 32 ? 1 : 2)>& key, u32 p, u32 trans = 0) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  Key<NW> x = key;
  if (trans) {
    const Key<NW> rc = revcomp<K>(key);
    if (trans == 1) { if (!key_less<NW>(key, rc)) x = rc; }
    else {
#pragma unroll
      for (int w = 0; w < NW; ++w) x.w[w] = key.w[w] ^ rc.w[w];
    }
  }
  u64 s, d;
  kmer_hashes<K, HASH>(x, s, d, HASH >= 3 ? ceil_log2_u32(p) : 0u);
  return mod_ranks(d, p);
}

template <int K, int HASH>
__global__ void __launch_bounds__(PART_THREADS)
k_part_hist(const u64* __restrict__ keys, u64 n, u32 p, u32* __restrict__ hist, u64 n_chunks, const u32 trans) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  __shared__ u32 h[MAX_RANKS];
  for (u32 r = threadIdx.x; r < p; r += PART_THREADS) h[r] = 0;
  __syncthreads();
  const u64 base = (u64)blockIdx.x * PART_ITEMS;
  for (int rd = 0; rd < PART_ROUNDS; ++rd) {
    u64 i = base + (u64)rd * PART_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&h[owner_of<K, HASH>(load_key<NW>(keys, i), p, trans)], 1u);
  }
  __syncthreads();
  for (u32 r = threadIdx.x; r < p; r += PART_THREADS) hist[(u64)r * n_chunks + blockIdx.x] = h[r];
}

template <int K, int HASH>
__global__ void __launch_bounds__(PART_THREADS)
k_part_scatter(const u64* __restrict__ keys, const u64* __restrict__ vals, u64 n, u32 p,
               const u32* __restrict__ hist_scanned, const u32* __restrict__ hist_block_sums, u64 n_chunks,
               u64* __restrict__ keys_out, u64* __restrict__ vals_out, const u32 trans) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  constexpr int NWARPS = PART_THREADS / 32;
  __shared__ u32 run[MAX_RANKS];              // next free offset per owner for this chunk
  __shared__ u32 wcnt[NWARPS][MAX_RANKS];     // per-warp counts of the current round
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (u32 r = threadIdx.x; r < p; r += PART_THREADS) {
    u64 hi = (u64)r * n_chunks + blockIdx.x;
    run[r] = hist_scanned[hi] + hist_block_sums[hi >> 10];
  }
  const u64 base = (u64)blockIdx.x * PART_ITEMS;
  for (int rd = 0; rd < PART_ROUNDS; ++rd) {
    for (u32 r = threadIdx.x; r < p * NWARPS; r += PART_THREADS) wcnt[r / p][r % p] = 0;
    __syncthreads();
    u64 i = base + (u64)rd * PART_THREADS + threadIdx.x;
    bool ok = i < n;
    Key<NW> key;
    u32 own = 0xFFFFFFFFu;
    if (ok) { key = load_key<NW>(keys, i); own = owner_of<K, HASH>(key, p, trans); }
    u32 peers = __match_any_sync(0xffffffffu, own);
    u32 rank_in_warp = __popc(peers & ((1u << lane) - 1));
    if (ok && rank_in_warp == 0) wcnt[warp][own] = __popc(peers);
    __syncthreads();
    if (ok) {
      u32 o = run[own] + rank_in_warp;
      for (int w = 0; w < warp; ++w) o += wcnt[w][own];
      store_key<NW>(keys_out, o, key);
      if (vals) vals_out[o] = vals[i];
    }
    __syncthreads();
    for (u32 r = threadIdx.x; r < p; r += PART_THREADS) {
      u32 s = 0;
      for (int w = 0; w < NWARPS; ++w) s += wcnt[w][r];
      run[r] += s;
    }
    __syncthreads();
  }
}

// ---- a17: multimap as CSR (key -> [offset, offset+multiplicity) into a value array) --------------
// Replaces fsc::densehash_multimap's singleton/vector split (reference
// src/containers/densehash_map.hpp:2022-2044,2190-2249): count -> exclusive scan -> fill.
// scan level 1 over slot multiplicities (dead slots count 0); 1024 slots per block
template <int NW>
__global__ void __launch_bounds__(1024)
k_csr_scan_l1(const Slot<NW>* __restrict__ slots, u64 n_slots, u64* __restrict__ offsets, u64* __restrict__ block_sums) {
  __shared__ u64 tmp[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u64 i = (u64)blockIdx.x * 1024 + threadIdx.x;
  u64 x = 0;
  if (i < n_slots) {
    Key<NW> key = slot_key<NW>(slots + i);
    if (is_live<NW>(key)) x = slots[i].val;
  }
  u64 inc = x;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    u64 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    u64 w = tmp[lane], winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u64 o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += o;
    }
    tmp[lane] = winc - w;
    if (lane == 31) tmp[32] = winc;
  }
  __syncthreads();
  if (i < n_slots) offsets[i] = tmp[warp] + inc - x;
  if (threadIdx.x == 0) block_sums[blockIdx.x] = tmp[32];
}
#ifdef KMI_SHARED_KERNELS
__global__ void __launch_bounds__(1024) k_scan64(u64* __restrict__ v, u64 n) {
  __shared__ u64 tmp[33];
  __shared__ u64 carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (u64 base = 0; base < n; base += 1024) {
    u64 i = base + threadIdx.x;
    u64 x = (i < n) ? v[i] : 0, inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u64 o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) tmp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      u64 w = tmp[lane], winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        u64 o = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += o;
      }
      tmp[lane] = winc - w;
      if (lane == 31) tmp[32] = winc;
    }
    __syncthreads();
    u64 carry = carry_s;
    if (i < n) v[i] = carry + tmp[warp] + inc - x;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + tmp[32];
    __syncthreads();
  }
}
// level 1 of a multi-block exclusive scan of u64 values: every block scans its 1024 values in place and writes its total
// (k_scan64 over the totals and k_csr_add_block finish it: kmi::scan64_dev)
__global__ void __launch_bounds__(1024) k_scan64_l1(u64* __restrict__ v, u64 n, u64* __restrict__ block_sums) {
  __shared__ u64 tmp[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const u64 i = (u64)blockIdx.x * 1024 + threadIdx.x;
  const u64 x = (i < n) ? v[i] : 0;
  u64 inc = x;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const u64 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const u64 w = tmp[lane];
    u64 winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const u64 o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += o;
    }
    tmp[lane] = winc - w;
    if (lane == 31) tmp[32] = winc;
  }
  __syncthreads();
  if (i < n) v[i] = tmp[warp] + inc - x;
  if (threadIdx.x == 0) block_sums[blockIdx.x] = tmp[32];
}
__global__ void __launch_bounds__(1024) k_csr_add_block(u64* __restrict__ offsets, u64 n, const u64* __restrict__ block_sums) {
  u64 i = (u64)blockIdx.x * 1024 + threadIdx.x;
  if (i < n) offsets[i] += block_sums[blockIdx.x];
}
#endif  // KMI_SHARED_KERNELS
// fill: every logged (key, id) goes to values[offset[slot] + cursor++]; aux is the cursor
template <int K, int HASH>
__global__ void __launch_bounds__(256)
k_csr_fill(TableView t, const u64* __restrict__ keys, const u64* __restrict__ ids, u64 n,
           const u64* __restrict__ offsets, u64* __restrict__ values) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    Key<NW> key = load_key<NW>(keys, i);
    u64 s = table_find<NW>(t, key, slot_hash<K>(key));
    if (s == ~0ULL) continue;
    u32 c = atomicAdd(&reinterpret_cast<Slot<NW>*>(t.slots)[s].aux, 1u);
    values[offsets[s] + c] = ids[i];
  }
}
#ifdef KMI_SHARED_KERNELS
__global__ void __launch_bounds__(256) k_widen_u8(const uint8_t* __restrict__ in, u64* __restrict__ out, u64 n) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) out[i] = in[i];
}
// find (multimap): per query multiplicity, then gather
__global__ void __launch_bounds__(256)
k_csr_gather(const u64* __restrict__ slot_idx, const u64* __restrict__ out_off, u64 n, const u64* __restrict__ offsets,
             const u64* __restrict__ values, u64* __restrict__ out_vals, u64 cap) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    u64 s = slot_idx[i];
    if (s == ~0ULL) continue;
    u64 b = out_off[i], e = out_off[i + 1], src = offsets[s];
    for (u64 j = b; j < e && j < cap; ++j) out_vals[j] = values[src + (j - b)];
  }
}
#endif  // KMI_SHARED_KERNELS
// drop logged tuples whose key is no longer live (after erase); stable not required
template <int K, int HASH>
__global__ void __launch_bounds__(256)
k_log_filter(TableView t, const u64* __restrict__ keys, const u64* __restrict__ ids, u64 n, u64* __restrict__ keys_out,
             u64* __restrict__ ids_out, u64* __restrict__ cursor) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  for (u64 base = (u64)blockIdx.x * blockDim.x; base < n; base += (u64)gridDim.x * blockDim.x) {
    u64 i = base + threadIdx.x;
    bool keep = false;
    Key<NW> key;
    if (i < n) { key = load_key<NW>(keys, i); keep = table_find<NW>(t, key, slot_hash<K>(key)) != ~0ULL; }
    u32 m = __ballot_sync(0xffffffffu, keep);
    u64 wbase = 0;
    if ((threadIdx.x & 31) == 0 && m) wbase = atomicAdd(cursor, (u64)__popc(m));
    wbase = __shfl_sync(0xffffffffu, wbase, 0);
    if (keep) {
      u64 at = wbase + __popc(m & ((1u << (threadIdx.x & 31)) - 1));
      store_key<NW>(keys_out, at, key);
      ids_out[at] = ids[i];
    }
  }
}
template <int NW> __global__ void k_zero_aux(Slot<NW>* __restrict__ slots, u64 n) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) slots[i].aux = 0;
}

// ---- synthetic FASTQ (same bytes as oracle/kmi_oracle.c: kmio_synth_fastq) ---------------------
__device__ __forceinline__ u64 sm64(u64 x) { return mix64(x); }
__device__ __forceinline__ u32 syn_genome_base(u64 seed, u64 i) {
  return (u32)((sm64(seed * 0x100000001b3ULL + (i >> 5)) >> (2 * (i & 31))) & 3);
}
#ifdef KMI_SHARED_KERNELS
__global__ void __launch_bounds__(256)
k_synth_fastq(uint8_t* __restrict__ out, u64 first_read, u64 n_reads, u32 L, u64 G, u64 seed, u32 err_ppm, u32 n_ppm) {
  const u32 HDR = 13;
  const u64 rec = HDR + 2ULL * L + 4, total = n_reads * rec;
  for (u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (u64)gridDim.x * blockDim.x) {
    u64 r = first_read + idx / rec;
    u32 o = (u32)(idx % rec);
    uint8_t ch;
    if (o < HDR) {
      if (o == 0) ch = '@';
      else if (o == 1) ch = 'S';
      else if (o == HDR - 1) ch = '\n';
      else { u64 d = r; for (u32 k = HDR - 2; k > o; --k) d /= 10; ch = (uint8_t)('0' + d % 10); }
    } else if (o - HDR < L) {
      u32 q = o - HDR;
      u64 h = sm64((seed + 1) * 0x9e3779b97f4a7c15ULL + r);
      u64 start = (h >> 1) % (G - L + 1);
      u32 b = (h & 1) ? 3u - syn_genome_base(seed, start + (L - 1 - q)) : syn_genome_base(seed, start + q);
      u64 e = sm64((seed + 2) * 0xd1b54a32d192ed03ULL + r * L + q);
      if ((u32)(e % 1000000u) < err_ppm) b = (b + 1 + (u32)((e >> 40) % 3)) & 3;
      ch = ((u32)((e >> 20) % 1000000u) < n_ppm) ? (uint8_t)'N' : (uint8_t)"ACGT"[b];
    } else {
      u32 q = o - HDR - L;
      ch = (q == 0) ? '\n' : (q == 1) ? '+' : (q == 2) ? '\n' : (q - 3 < L) ? 'I' : '\n';
    }
    out[idx] = ch;
  }
}

// bytes [file_offset, file_offset + n_bytes) of the synthetic FASTA genome (same bytes as oracle/kmi_oracle.c kmio_synth_fasta):
// n_seq sequences of seq_len bases, `width` bases per line, header ">chrNN\n"
__global__ void __launch_bounds__(256)
k_synth_fasta(uint8_t* __restrict__ out, u64 file_offset, u64 n_bytes, u32 n_seq, u64 seq_len, u32 width, u64 seed) {
  const u64 lines = (seq_len + width - 1) / width, seq_bytes = 7 + seq_len + lines;
  for (u64 idx = (u64)blockIdx.x * blockDim.x + threadIdx.x; idx < n_bytes; idx += (u64)gridDim.x * blockDim.x) {
    const u64 f = file_offset + idx;
    const u64 s = f / seq_bytes;
    u64 o = f % seq_bytes;
    uint8_t ch;
    if (s >= n_seq) ch = '\n';
    else if (o < 7) ch = (o == 0) ? '>' : (o == 1) ? 'c' : (o == 2) ? 'h' : (o == 3) ? 'r' : (o == 4) ? (uint8_t)('0' + (s / 10) % 10) : (o == 5) ? (uint8_t)('0' + s % 10) : '\n';
    else {
      o -= 7;
      const u64 line = o / (width + 1), col = o % (width + 1), i = line * width + col;
      ch = (col == width || i >= seq_len) ? (uint8_t)'\n' : (uint8_t)"ACGT"[syn_genome_base(seed + 977 * (s + 1), i)];
    }
    out[idx] = ch;
  }
}
// query k-mers of config 5: even j = K bases of the synthetic genome at a random place (random strand),
// odd j = uniform random k-mer; nw words per key
__global__ void __launch_bounds__(256)
k_synth_queries(u64* __restrict__ out, u64 first, u64 n, int k, u32 n_seq, u64 seq_len, u64 genome_seed, u64 seed) {
  const int nw = k <= 32 ? 1 : 2;
  for (u64 j = (u64)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (u64)gridDim.x * blockDim.x) {
    const u64 h = sm64((seed + 5) * 0x9e3779b97f4a7c15ULL + first + j);
    u64 lo = 0, hi = 0;
    if (((first + j) & 1) == 0) {
      const u64 s = (h >> 1) % n_seq, pos = sm64(h) % (seq_len - (u64)k + 1);
      const bool rc = (h & 1) != 0;
      for (int i = 0; i < k; ++i) {
        const u32 b = rc ? 3u - syn_genome_base(genome_seed + 977 * (s + 1), pos + (u64)(k - 1 - i)) : syn_genome_base(genome_seed + 977 * (s + 1), pos + (u64)i);
        hi = (hi << 2) | (lo >> 62);
        lo = (lo << 2) | b;
      }
    } else {
      lo = sm64(h); hi = sm64(h ^ 0x5555555555555555ULL);
    }
    if (nw == 1) out[j] = k == 32 ? lo : (lo & ((1ULL << (2 * k)) - 1));
    else { out[2 * j] = lo; out[2 * j + 1] = hi & ((1ULL << (2 * (k - 32))) - 1); }
  }
}

#endif  // KMI_SHARED_KERNELS

}  // namespace kmi
