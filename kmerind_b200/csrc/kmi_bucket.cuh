// kmi_bucket.cuh — hash-prefix bucketing of the key stream in front of the table (sm_100a).
//
// Why: a count insert is a random read-modify-write of a 16/32-byte slot in a multi-GB table; done
// straight from the parser it drags ~165 B of DRAM traffic per k-mer (profiles/r01_fused_direct_
// atomics.md). Because the home slot is the TOP bits of the store hash (kmi_table.cuh), the top
// LOG_BUCKETS bits name a contiguous 1/256 segment of the table whatever its capacity. Keys are
// therefore first scattered into 256 hash-prefix buckets (streaming, coalesced runs assembled in
// shared memory) and then inserted bucket after bucket, so that the table segment being updated is
// L2-resident and every slot sector crosses the HBM interface once per sweep instead of once per k-mer.
//
// This replaces nothing in the reference one-to-one: it is the GPU counterpart of the cache
// blocking the reference gets for free from per-rank tables that are 1/p of the data
// (distributed_densehash_map.hpp:2565-2628), and it reuses the histogram -> scan -> scatter shape of
// imxx::local::assign_to_buckets / permute (src/io/incremental_mxx.hpp:273-364,596-640), unordered.
#pragma once
#include "kmi_block.cuh"

namespace kmi {

static constexpr int LOG_BUCKETS = 8;
static constexpr int N_BUCKETS = 1 << LOG_BUCKETS;        // == TILE_THREADS (one scan slot per thread)
static constexpr int HLL_P = 14;
static constexpr int HLL_M = 1 << HLL_P;                  // HyperLogLog registers (sigma = 0.81 %)
static constexpr u32 BKT_NONE = 0xFFFFFFFFu;

template <int NW> struct BktRound { static constexpr int R = (NW == 1 ? 8 : 4); };  // keys / thread / round

struct BucketSink {
  u64* keys;     // N_BUCKETS regions of `capb` keys, then one overflow region of `ovf_cap` keys
  u32* cursor;   // N_BUCKETS + 1 fill counters; a counter may run past its capacity (clamp when reading)
  u32 capb;
  u32 ovf_cap;
  u32* hll;      // HLL_M registers: distinct-key estimate that sizes the table before the insert
  u64* const* bases;  // optional: explicit base pointer per region (peer-GPU receive regions of the
                      // fused exchange, kmi_exchange_*); null = regions are keys + b * capb
  // What the keys of region r have in common: the top `log_pre` bits of their slot hash equal r & pre_mask.
  // Local build: log_pre = LOG_BUCKETS, pre_mask = 255 (region = hash prefix). Receive buffer of the fused exchange:
  // region = source rank * CB + prefix with CB = N_BUCKETS / nranks sub-bins, log_pre = log2(CB), pre_mask = CB - 1.
  u32 log_pre = LOG_BUCKETS;
  u32 pre_mask = N_BUCKETS - 1;
  // A sink that views 256 consecutive regions of a LARGER set (third scatter level of very large tables: the output
  // regions of a first fine pass, taken 256 at a time): prefix of region r = pre_base + (r & pre_mask); the overflow
  // region then is not "behind region 255" but wherever the owner of the whole set keeps it.
  u32 pre_base = 0;
  // Receive buffer of the sharded build (kmi_dist.cu): every hash prefix has `sub` regions, one per SOURCE rank, filled by
  // that rank's k_owner_push over NVLink: region index = prefix * sub + source, N_BUCKETS * sub regions and cursors in all.
  // The scatter kernels of the local build always write sub == 1 sinks; the fine scatter and the HLL kernel read any sub.
  u32 sub = 1;
  u64* ovf_keys = nullptr;        // null: keys + N_BUCKETS * sub * capb (* words per key)
  u32* ovf_cursor = nullptr;      // null: cursor + N_BUCKETS * sub
};
static constexpr int MAX_SUB = 8;                         // ranks of one NVLink domain that push into one receive buffer
template <int NW> __host__ __device__ __forceinline__ u64* ovf_base(const BucketSink& bs) {
  return bs.ovf_keys ? bs.ovf_keys : bs.keys + (u64)N_BUCKETS * bs.sub * bs.capb * NW;
}
__host__ __device__ __forceinline__ u32* ovf_cur(const BucketSink& bs) { return bs.ovf_cursor ? bs.ovf_cursor : bs.cursor + N_BUCKETS * bs.sub; }

template <int NW> struct BucketSmem;
template <> struct BucketSmem<0> { u32 unused; };   // kernels instantiated without the bucket sink
template <int NW> struct BucketSmem {
  static constexpr int R = BktRound<NW>::R;
  static constexpr int ROUND_KEYS = TILE_THREADS * R;
  u64 sorted[ROUND_KEYS * NW];
  u32 hist[N_BUCKETS];    // counts while ranking, exclusive prefix afterwards
  u32 delta[N_BUCKETS];
  u64 dst[N_BUCKETS];     // where sorted[0] would land if the bucket's run started there (bytes); bit 63: run does not fit
  u32 tmp[TILE_THREADS / 32 + 1];
  uint8_t sbid[ROUND_KEYS];
};

__device__ __forceinline__ u32 bucket_of(u64 store_hash) { return (u32)(store_hash >> (64 - LOG_BUCKETS)); }

// HyperLogLog over a 1/HLL_SAMPLE hash-sampled subset of the keys (distinct sampling: the keys whose
// hash prefix names one of HLL_SAMPLE_BUCKETS evenly spaced coarse buckets hold 1/HLL_SAMPLE of the distinct
// keys; the estimate is scaled back on the host). Because the sample IS a set of buckets, the registers are
// filled by a small kernel over those buckets after the scatter (k_hll_buckets, ~0.1 GB read) and the
// per-k-mer path carries no HLL work at all (it used to cost 13 % of the parser: profiles/r01_*).
// Register = low HLL_P bits of the folded hash, rank from the 42 bits above them (the top 8 bits pick the
// bucket and stay out of it).
static constexpr int HLL_SAMPLE_BUCKETS = 4;
static constexpr int HLL_SAMPLE = N_BUCKETS / HLL_SAMPLE_BUCKETS;
__device__ __forceinline__ bool hll_sampled(u32 bucket) { return (bucket & (HLL_SAMPLE - 1)) == 0; }
__device__ __forceinline__ void hll_update(u32* __restrict__ hll, u64 h) {
  h ^= h >> 29;                                          // slot_hash mixes upwards: fold the high half into the low bits
  const u32 reg = (u32)h & (HLL_M - 1);
  const u64 w = (h >> HLL_P) << (HLL_P + LOG_BUCKETS);   // 42 significant bits, left aligned
  const u32 rank = (u32)__clzll(w | (1ULL << (HLL_P + LOG_BUCKETS - 1))) + 1u;
  if (hll[reg] < rank) atomicMax(hll + reg, rank);       // the plain read may be stale: then the max is redundant, never wrong
}

// block scan without the trailing barrier of block_excl_scan: the caller guarantees a barrier between
// this call's reads of tmp and the next write to it
__device__ __forceinline__ u32 block_excl_scan_nb(u32 v, u32* tmp, u32& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u32 inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    u32 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  u32 w = (lane < TILE_THREADS / 32) ? tmp[lane] : 0;    // every warp scans the 8 warp totals itself
  u32 winc = w;
#pragma unroll
  for (int d = 1; d < TILE_THREADS / 32; d <<= 1) {
    u32 o = __shfl_up_sync(0xffffffffu, winc, d);
    if (lane >= d) winc += o;
  }
  total = __shfl_sync(0xffffffffu, winc, TILE_THREADS / 32 - 1);
  const u32 wex = __shfl_sync(0xffffffffu, winc - w, warp);
  return wex + inc - v;
}

// zero the histogram; call once per kernel before the first bucket_round
template <int NW> __device__ __forceinline__ void bucket_init(BucketSmem<NW>& sm) {
  sm.hist[threadIdx.x] = 0;
  __syncthreads();
}

// One round: every thread brings up to R keys (bkt[q] == BKT_NONE: no key). The block ranks them by
// bucket in shared memory, reserves one run per non-empty bucket with a single global atomicAdd (in
// flight while the keys are sorted), and writes the runs coalesced. Keys that do not fit their bucket
// go to the overflow region; if that is full too, PERR_OVERFLOW is raised and the host redoes the chunk
// on the direct path. (Ranking by ballots over the digit bits with warp-private counters instead of
// shared-memory atomics was measured slower: parse+scatter 11.8 vs 9.4 ms, fine scatter 7.5 vs 5.6 ms.)
template <int NW>
__device__ __forceinline__ void bucket_round(const BucketSink& bs, BucketSmem<NW>& sm, const Key<NW> (&key)[BktRound<NW>::R],
                                             const u32 (&bkt)[BktRound<NW>::R], u32* __restrict__ perr) {
  constexpr int R = BktRound<NW>::R;
  constexpr u64 DST_SLOW = 1ULL << 63;
  const int t = threadIdx.x;
  u32 rank[R];
#pragma unroll
  for (int q = 0; q < R; ++q) {
    rank[q] = 0u;
    if (bkt[q] != BKT_NONE) rank[q] = atomicAdd(&sm.hist[bkt[q]], 1u);
  }
  __syncthreads();
  const u32 cnt = sm.hist[t];
  const u32 gbase = cnt ? atomicAdd(bs.cursor + t, cnt) : 0u;    // in flight during the scan and the sort
  u32 total;
  const u32 ex = block_excl_scan_nb(cnt, sm.tmp, total);
  sm.hist[t] = ex;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < R; ++q) {
    if (bkt[q] != BKT_NONE) {
      const u32 pos = sm.hist[bkt[q]] + rank[q];
      if (NW == 1) sm.sorted[pos] = key[q].w[0];
      else reinterpret_cast<ulonglong2*>(sm.sorted)[pos] = make_ulonglong2(key[q].w[0], key[q].w[NW - 1]);
      sm.sbid[pos] = (uint8_t)bkt[q];
    }
  }
  {
    // run of bucket t: sorted[ex .. ex+cnt) -> region[gbase .. gbase+cnt). One 64-bit destination per bucket, so
    // that the copy-out below is load / add / store per key; runs that cross the region's capacity take the
    // slow branch (overflow region).
    u64* const region = bs.bases ? bs.bases[t] : bs.keys + (u64)t * bs.capb * NW;
    u64 d = reinterpret_cast<u64>(region) + ((u64)gbase - (u64)ex) * (8u * NW);
    if ((u64)gbase + cnt > (u64)bs.capb) d = DST_SLOW;
    sm.dst[t] = d;
    sm.delta[t] = gbase - ex;
  }
  __syncthreads();
  sm.hist[t] = 0;                                         // dead now: cleared for the next round
  for (u32 r = t; r < total; r += TILE_THREADS) {
    const u32 b = sm.sbid[r];
    const u64 d = sm.dst[b];
    if (!(d & DST_SLOW)) {
      if (NW == 1) *reinterpret_cast<u64*>(d + (u64)r * 8u) = sm.sorted[r];
      else *reinterpret_cast<ulonglong2*>(d + (u64)r * 16u) = reinterpret_cast<const ulonglong2*>(sm.sorted)[r];
      continue;
    }
    u32 at = sm.delta[b] + r;
    u64* base = bs.bases ? bs.bases[b] : bs.keys + (u64)b * bs.capb * NW;
    if (at >= bs.capb) {
      at = atomicAdd(ovf_cur(bs), 1u);
      base = ovf_base<NW>(bs);
      if (at >= bs.ovf_cap) { atomicOr(perr, PERR_OVERFLOW); continue; }
    }
    if (NW == 1) base[at] = sm.sorted[r];
    else reinterpret_cast<ulonglong2*>(base)[at] = reinterpret_cast<const ulonglong2*>(sm.sorted)[r];
  }
  __syncthreads();
}

// key stream -> buckets (used by kmi_insert_dev and the FASTA / sharded builds)
template <int K, int HASH>
__global__ void __launch_bounds__(TILE_THREADS)
k_scatter_keys(const u64* __restrict__ keys, u64 n, BucketSink bs, u32* __restrict__ perr, const bool CANON) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  constexpr int R = BktRound<NW>::R;
  __shared__ BucketSmem<NW> sm;
  const u64 per_round = (u64)TILE_THREADS * R;
  bucket_init<NW>(sm);
  Key<NW> nkey[R];                                       // next round's keys, requested one round ahead
  {
    const u64 base = (u64)blockIdx.x * per_round;
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const u64 i = base + (u64)q * TILE_THREADS + threadIdx.x;
      if (i < n) nkey[q] = load_key<NW>(keys, i);
    }
  }
  for (u64 base = (u64)blockIdx.x * per_round; base < n; base += (u64)gridDim.x * per_round) {
    Key<NW> key[R];
    u32 bkt[R];
#pragma unroll
    for (int q = 0; q < R; ++q) key[q] = nkey[q];
    {
      const u64 nbase = base + (u64)gridDim.x * per_round;
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const u64 i = nbase + (u64)q * TILE_THREADS + threadIdx.x;
        if (i < n) nkey[q] = load_key<NW>(keys, i);
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const u64 i = base + (u64)q * TILE_THREADS + threadIdx.x;
      bkt[q] = BKT_NONE;
      if (i < n) {
        if (CANON) key[q] = canonical<K>(key[q]);
        const u64 h = slot_hash<K>(key[q]);
        bkt[q] = bucket_of(h);
      }
    }
    bucket_round<NW>(bs, sm, key, bkt, perr);
  }
}

// HLL registers from the sampled buckets (and the sampled keys of the overflow region); grid.y = sampled
// bucket index, the last y-slice walks the overflow region
template <int K>
__global__ void __launch_bounds__(256) k_hll_buckets(BucketSink bs) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  // y-slice = (group, j, s): the region of group `grp` whose prefix covers sampled bucket j * HLL_SAMPLE, sub-region s; a
  // local build has one group and one sub-region (region = bucket); the last slice is the overflow region
  const u32 n_groups = N_BUCKETS >> bs.log_pre;          // regions per prefix value (two-pass layouts)
  const u32 n_slices = HLL_SAMPLE_BUCKETS * n_groups * bs.sub;
  const bool ovf = blockIdx.y == n_slices;
  const u32 sidx = blockIdx.y % bs.sub, y = blockIdx.y / bs.sub;
  const u32 grp = y / HLL_SAMPLE_BUCKETS, j = y % HLL_SAMPLE_BUCKETS;
  const u32 b = ((grp << bs.log_pre) + ((j * HLL_SAMPLE) >> (LOG_BUCKETS - bs.log_pre))) * bs.sub + sidx;
  const u32 n = ovf ? min(*ovf_cur(bs), bs.ovf_cap) : min(bs.cursor[b], bs.capb);
  const u64* src = ovf ? ovf_base<NW>(bs) : bs.keys + (u64)b * bs.capb * NW;
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const u64 h = slot_hash<K>(load_key<NW>(src, i));
    if (hll_sampled(bucket_of(h))) hll_update(bs.hll, h);
  }
}

// buckets -> table. The grid walks the concatenation of the buckets front to back (grid-stride), so
// at any moment all CTAs work on the same one or two table segments: those stay in L2.
template <int K, int HASH>
__global__ void __launch_bounds__(256) k_insert_buckets(TableView t, BucketSink bs) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  __shared__ u64 pre[N_BUCKETS + 2];
  __shared__ u32 tmp[TILE_THREADS / 32 + 1];
  {
    const u32 cnt = min(bs.cursor[threadIdx.x], bs.capb);
    u32 total;
    const u32 ex = block_excl_scan(cnt, tmp, total);   // chunk sizes keep the total below 2^32
    pre[threadIdx.x] = ex;
    if (threadIdx.x == 0) {
      pre[N_BUCKETS] = total;
      pre[N_BUCKETS + 1] = (u64)total + min(*ovf_cur(bs), bs.ovf_cap);
    }
  }
  __syncthreads();
  const u64 T = pre[N_BUCKETS + 1];
  u32 n_new = 0;
  for (u64 g = (u64)blockIdx.x * blockDim.x + threadIdx.x; g < T; g += (u64)gridDim.x * blockDim.x) {
    int lo = 0, hi = N_BUCKETS;                         // largest b with pre[b] <= g
#pragma unroll
    for (int s = 0; s < LOG_BUCKETS + 1; ++s) {
      const int mid = (lo + hi + 1) >> 1;
      if (pre[mid] <= g) lo = mid; else hi = mid - 1;
    }
    const Key<NW> key = lo < N_BUCKETS ? load_key<NW>(bs.keys, (u64)lo * bs.capb + (g - pre[lo])) : load_key<NW>(ovf_base<NW>(bs), g - pre[lo]);
    bool was_new;
    const u64 s = table_upsert<NW>(t, key, slot_hash<K>(key), was_new);
    if (s != ~0ULL) {
      table_add<NW>(t, s, 1u);
      n_new += was_new ? 1u : 0u;
    }
  }
  n_new = warp_sum(n_new);
  if ((threadIdx.x & 31) == 0 && n_new) atomicAdd(t.size, (u64)n_new);
}


// ======================================================================================================
// Second level: table segments small enough for shared memory.
//
// L2-resident atomics top out near 25 G k-mers/s (profiles/r01_bucketed_l2.md): the table update itself
// has to leave the L2 atomic units. A second 2^log_f-way scatter (same shape as the first, on the next
// hash bits) cuts every coarse bucket into fine buckets whose table segment (<= 8192 slots) fits in
// shared memory; one CTA then loads the segment, applies its fine bucket with shared-memory atomics,
// and writes the segment back with coalesced 128-bit stores: the table crosses HBM once per sweep.
// ======================================================================================================
#ifndef KMI_MAX_LOG_F
#define KMI_MAX_LOG_F 11
#endif
static constexpr int MAX_LOG_F = KMI_MAX_LOG_F;
#ifndef KMI_FINE_ROUND
#define KMI_FINE_ROUND 2048
#endif
static constexpr int FINE_ROUND = KMI_FINE_ROUND;         // keys per CTA round of the second scatter

struct FineSink {
  u64* keys;       // (N_BUCKETS << log_f) regions of capf keys
  u32* cursor;     // one fill counter per region (may run past capf)
  u32 capf;
  u32 log_f;
};

// dynamic shared memory of k_scatter_fine for 2^log_f fine buckets
__host__ __device__ inline size_t scatter_fine_smem(u32 log_f) {
  const size_t F = (size_t)1 << log_f;
  return FINE_ROUND * sizeof(u64) + F * sizeof(u64) + 2 * F * sizeof(u32) + FINE_ROUND * sizeof(unsigned short);
}

#ifndef KMI_FINE_MINB
#define KMI_FINE_MINB 4       // CTAs per SM of the (persistent) fine scatter
#endif
template <int K, int HASH>
__global__ void __launch_bounds__(TILE_THREADS, KMI_FINE_MINB)
k_scatter_fine(BucketSink bs, FineSink fs, u32* __restrict__ perr) {
  static_assert(K <= 32, "segment path is built for one-word keys");
  extern __shared__ __align__(16) unsigned char fine_raw[];
  __shared__ u32 cpre[N_BUCKETS * MAX_SUB + 1];
  __shared__ u32 tmp[TILE_THREADS / 32 + 1];
  constexpr int R = FINE_ROUND / TILE_THREADS;
  const int t = threadIdx.x;
  const u32 F = 1u << fs.log_f;
  u64* sorted = reinterpret_cast<u64*>(fine_raw);                       // FINE_ROUND keys
  u64* dst = sorted + FINE_ROUND;                                       // F: byte address sorted[0] would land at (bit 63: run does not fit)
  u32* hist = reinterpret_cast<u32*>(dst + F);                          // F: counts while ranking, exclusive prefix afterwards
  u32* delta = hist + F;                                                // F
  unsigned short* sbid = reinterpret_cast<unsigned short*>(delta + F);  // FINE_ROUND
  const u32 EPT = F >= TILE_THREADS ? F / TILE_THREADS : 1u;   // histogram entries per thread
  const u32 sub = bs.sub, NR = N_BUCKETS * sub;                // input regions: prefix * sub + source
  u32 n_chunks;
  {
    u32 mine = 0;
    for (u32 q = 0; q < sub; ++q) mine += (min(bs.cursor[t * sub + q], bs.capb) + FINE_ROUND - 1) / FINE_ROUND;
    u32 ex = block_excl_scan(mine, tmp, n_chunks);
    for (u32 q = 0; q < sub; ++q) { cpre[t * sub + q] = ex; ex += (min(bs.cursor[t * sub + q], bs.capb) + FINE_ROUND - 1) / FINE_ROUND; }
    if (t == 0) cpre[NR] = n_chunks;
  }
  for (u32 f = t; f < F; f += TILE_THREADS) hist[f] = 0;
  __syncthreads();
  // software pipeline: the keys of the NEXT chunk are requested before this chunk is ranked, so the
  // global-load latency hides behind a whole round
  auto locate = [&](u32 c, u32& b, u32& m, const u64*& src) {
    int lo = 0, hi = (int)NR - 1;                         // largest region with cpre[region] <= c: the non-empty one that holds chunk c
#pragma unroll
    for (int s = 0; s < LOG_BUCKETS; ++s) {
      const int mid = (lo + hi + 1) >> 1;
      if (cpre[mid] <= c) lo = mid; else hi = mid - 1;
    }
    if (sub > 1) {                                        // up to MAX_SUB * 256 regions: three more halvings
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int mid = (lo + hi + 1) >> 1;
        if (cpre[mid] <= c) lo = mid; else hi = mid - 1;
      }
    }
    const u32 first = (c - cpre[lo]) * FINE_ROUND;
    const u32 cnt_b = min(bs.cursor[lo], bs.capb);
    m = min((u32)FINE_ROUND, cnt_b - first);
    src = bs.keys + (u64)lo * bs.capb + first;
    b = sub == 1 ? (u32)lo : (u32)lo / sub;               // hash-prefix index of the region
  };
  u64 nkey[R];
  u32 nb = 0, nm = 0;
  if (blockIdx.x < n_chunks) {
    const u64* src;
    locate(blockIdx.x, nb, nm, src);
#pragma unroll
    for (int q = 0; q < R; ++q) { const u32 i = q * TILE_THREADS + t; nkey[q] = (i < nm) ? __ldg(src + i) : 0ULL; }
  }
  for (u32 c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    const u32 b = nb, m = nm;
    u64 key[R];
#pragma unroll
    for (int q = 0; q < R; ++q) key[q] = nkey[q];
    if (c + gridDim.x < n_chunks) {
      const u64* src;
      locate(c + gridDim.x, nb, nm, src);
#pragma unroll
      for (int q = 0; q < R; ++q) { const u32 i = q * TILE_THREADS + t; nkey[q] = (i < nm) ? __ldg(src + i) : 0ULL; }
    }
    u32 fid[R], rank[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const u32 i = q * TILE_THREADS + t;
      fid[q] = BKT_NONE;
      if (i < m) {
        Key<1> kk; kk.w[0] = key[q];
        const u64 h = slot_hash<K>(kk);
        fid[q] = (u32)(h >> (64 - bs.log_pre - fs.log_f)) & (F - 1);
        rank[q] = atomicAdd(&hist[fid[q]], 1u);
      }
    }
    __syncthreads();
    const u64 region0 = (u64)(bs.pre_base + (b & bs.pre_mask)) << fs.log_f;   // first fine region (= table segment) of this hash prefix
    u32 gb1 = 0, ex1 = 0, cnt1 = 0;                       // EPT == 1: the cursor atomic stays in flight during the sort
    {
      u32 local = 0;
      for (u32 e = 0; e < EPT; ++e) { const u32 f = t * EPT + e; if (f < F) local += hist[f]; }
      u32 total;
      u32 ex = block_excl_scan_nb(local, tmp, total);
      if (EPT == 1) {
        if ((u32)t < F) {
          if (local) gb1 = atomicAdd(fs.cursor + region0 + t, local);
          hist[t] = ex; ex1 = ex; cnt1 = local;
        }
      } else {
        for (u32 e = 0; e < EPT; ++e) {
          const u32 f = t * EPT + e;
          if (f < F) {
            const u32 cnt = hist[f];
            hist[f] = ex;
            if (cnt) delta[f] = atomicAdd(fs.cursor + region0 + f, cnt) - ex;
            ex += cnt;
          }
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (fid[q] != BKT_NONE) {
        const u32 pos = hist[fid[q]] + rank[q];
        sorted[pos] = key[q];
        sbid[pos] = (unsigned short)fid[q];
      }
    }
    if (EPT == 1 && (u32)t < F) {
      delta[t] = gb1 - ex1;
      const bool fits = (u64)gb1 + cnt1 <= (u64)fs.capf;
      dst[t] = fits ? reinterpret_cast<u64>(fs.keys + (region0 + t) * fs.capf) + ((u64)gb1 - (u64)ex1) * 8u : (1ULL << 63);
    }
    __syncthreads();
    for (u32 f = t; f < F; f += TILE_THREADS) hist[f] = 0;
    for (u32 r = t; r < m; r += TILE_THREADS) {
      const u32 f = sbid[r];
      if (EPT == 1) {
        const u64 d = dst[f];
        if (!(d >> 63)) { *reinterpret_cast<u64*>(d + (u64)r * 8u) = sorted[r]; continue; }
      }
      const u32 at = delta[f] + r;
      if (at < fs.capf) {
        fs.keys[(region0 + f) * fs.capf + at] = sorted[r];
      } else {
        const u32 o = atomicAdd(ovf_cur(bs), 1u);
        if (o < bs.ovf_cap) ovf_base<1>(bs)[o] = sorted[r];
        else atomicOr(perr, PERR_OVERFLOW);
      }
    }
    __syncthreads();
  }
}

// fine buckets -> table, one CTA per table segment resident in shared memory (K <= 31).
// A probe that runs off the end of the segment "spills": the key is appended to the overflow region
// and inserted afterwards by k_insert_overflow with global atomics, which continues the linear probe
// into the next segment exactly as table_upsert would have.
#ifndef KMI_SEG_THREADS
#define KMI_SEG_THREADS 384   // 4 CTAs/SM with the 48 KB keys[]+vals[] segment (measured: 5.01 ms vs 5.13 at 512 x 3)
#endif
#ifndef KMI_LOG_SEG_SLOTS
#define KMI_LOG_SEG_SLOTS 12  // log2 of the slots of one shared-memory table segment
#endif
#ifndef KMI_SEG_MINB
#define KMI_SEG_MINB 4
#endif
#ifndef KMI_SEG_SOA
#define KMI_SEG_SOA 1         // shared-memory segment as keys[] + vals[] (12 B/slot, fewer bank conflicts) instead of 16-byte slots
#endif
static constexpr int SEG_THREADS = KMI_SEG_THREADS;
__host__ __device__ inline size_t segment_smem(u32 log_seg) { return (size_t)(KMI_SEG_SOA ? 12 : 16) << log_seg; }
#ifndef KMI_SEG_PF
#define KMI_SEG_PF -1         // <0: the whole fine bucket is prefetched into L2 at segment start (-2: and the next table segment); >0: in-loop distance in chunks; 0 off
#endif
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }
// Probing is organised so that no lane waits for its neighbours' longer probe sequences: every warp owns a
// contiguous slice of the fine bucket, keeps the next 64 keys of it in two registers per lane (coalesced
// loads, one ahead), and a lane that has finished its key takes the next unclaimed one from that window
// (ballot + shuffle) in the same loop iteration in which the other lanes advance their probes. A loop
// iteration is therefore one probe for every lane, instead of max-over-lanes probes per key.
// EMPTY sentinel test on the top word alone: one-word keys are < 2^62, so no real key shares it
__device__ __forceinline__ bool idle(u64 k) { return (u32)(k >> 32) == (u32)(KEY_EMPTY_TOP >> 32); }
template <int K, int HASH>
__global__ void __launch_bounds__(SEG_THREADS, KMI_SEG_MINB)
k_insert_segments(TableView t, FineSink fs, BucketSink bs, const u32 log_seg, const u32 n_segs, u32* __restrict__ perr,
                  const u32 fresh) {
  static_assert(K <= 32, "segment path is built for one-word keys");
  extern __shared__ __align__(16) unsigned char seg_raw[];
  const u32 SEG = 1u << log_seg;
#if KMI_SEG_SOA
  u64* skey = reinterpret_cast<u64*>(seg_raw);
  u32* sval = reinterpret_cast<u32*>(skey + SEG);
#else
  Slot<1>* seg = reinterpret_cast<Slot<1>*>(seg_raw);
#endif
  constexpr int NWARP = SEG_THREADS / 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const u32 lt = (1u << lane) - 1u;
  u32 n_new = 0;
  for (u32 s = blockIdx.x; s < n_segs; s += gridDim.x) {
    const u32 cnt = min(fs.cursor[s], fs.capf);
    if (cnt == 0 && !fresh) continue;                     // block-uniform (a fresh table still needs its EMPTY pattern written)
    const u64* src = fs.keys + (u64)s * fs.capf;
    const u32 per = ((cnt + NWARP * 32 - 1) / (NWARP * 32)) * 32;   // keys per warp, whole 32-key chunks
    const u32 wb = min(cnt, (u32)warp * per), we = min(cnt, wb + per);
    u32 nxt = wb + lane;
    u64 buf0 = (nxt < we) ? __ldg(src + nxt) : KEY_EMPTY_TOP; nxt += 32;   // requested before the segment is loaded
    u64 buf1 = (nxt < we) ? __ldg(src + nxt) : KEY_EMPTY_TOP; nxt += 32;
    u64 buf2 = (nxt < we) ? __ldg(src + nxt) : KEY_EMPTY_TOP; nxt += 32;
    u32 ph = 0;
#if KMI_SEG_PF < 0
    for (u32 i = tid * 16u; i < cnt; i += SEG_THREADS * 16u) prefetch_l2(src + i);   // the whole fine bucket -> L2
#if KMI_SEG_PF < -1
    if (s + gridDim.x < n_segs) {                         // the CTA's next table segment -> L2
      const char* nslots = reinterpret_cast<const char*>(reinterpret_cast<const Slot<1>*>(t.slots) + ((u64)(s + gridDim.x) << log_seg));
      for (u32 i = tid * 128u; i < SEG * (u32)sizeof(Slot<1>); i += SEG_THREADS * 128u) prefetch_l2(nslots + i);
    }
#endif
#else
    if (KMI_SEG_PF) {                                     // chunks 2 .. PF+1 of the slice -> L2, one 128-byte line per lane
      const u32 i = wb + 64 + lane * 16;
      if (lane < 2 * KMI_SEG_PF && i < we) prefetch_l2(src + i);
    }
#endif
    uint4* gseg = reinterpret_cast<uint4*>(reinterpret_cast<Slot<1>*>(t.slots) + ((u64)s << log_seg));
    // fresh: the table is logically empty and its memory was never written (lazy clear, kmi_clear): nothing to load
#if KMI_SEG_SOA
    if (fresh) {
      for (u32 i = tid; i < SEG; i += SEG_THREADS) { skey[i] = KEY_EMPTY_TOP; sval[i] = 0u; }
    } else {
      for (u32 i = tid; i < SEG; i += SEG_THREADS) {
        const uint4 v = gseg[i];
        skey[i] = (u64)v.x | ((u64)v.y << 32); sval[i] = v.z;
      }
    }
#else
    if (fresh) {
      for (u32 i = tid; i < SEG; i += SEG_THREADS) reinterpret_cast<uint4*>(seg)[i] = make_uint4(0u, 0x80000000u, 0u, 0u);
    } else {
      for (u32 i = tid; i < SEG; i += SEG_THREADS) reinterpret_cast<uint4*>(seg)[i] = gseg[i];
    }
#endif
    __syncthreads();
    int remaining = (int)(we - wb);                       // keys of the slice not yet handed to a lane (warp-uniform)
    u32 pos = 0;                                          // consumed keys of buf0
    u64 key = KEY_EMPTY_TOP;
    u32 li = 0;
    for (;;) {
      const u32 need = __ballot_sync(0xffffffffu, idle(key));
      if (need && remaining > 0) {                        // warp-uniform
        const u32 idx = pos + __popc(need & lt);
        const u32 n = __popc(need);
        // C = chunk being handed out, N = the one after it; the exhausted chunk's register takes the new load, and
        // the roles rotate by phase instead of by register moves (a move would wait for the load in flight)
#define KMI_REFILL(C, N)                                                                            \
        {                                                                                          \
          u64 kx = __shfl_sync(0xffffffffu, C, idx & 31);                                          \
          if (pos + n > 32) {                                                                      \
            const u64 k1 = __shfl_sync(0xffffffffu, N, idx & 31);                                  \
            if (idx >= 32) kx = k1;                                                                \
          }                                                                                        \
          if (idle(key)) {                                                                         \
            key = kx;                                                                              \
            Key<1> kk; kk.w[0] = key;                                                              \
            li = (u32)(slot_hash<K>(kk) >> t.shift) & (SEG - 1);                                   \
          }                                                                                        \
          pos += n; remaining -= (int)n;                                                           \
          if (pos >= 32) {                                                                         \
            pos -= 32;                                                                             \
            C = (nxt < we) ? __ldg(src + nxt) : KEY_EMPTY_TOP;                                     \
            nxt += 32;                                                                             \
            ph = (ph == 2) ? 0 : ph + 1;                                                           \
          }                                                                                        \
        }
        if (ph == 0) KMI_REFILL(buf0, buf1)
        else if (ph == 1) KMI_REFILL(buf1, buf2)
        else KMI_REFILL(buf2, buf0)
#undef KMI_REFILL
      }
      // every lane holds a key while the slice still has some; only at its end does the warp have to vote
      if (remaining <= 0 && __ballot_sync(0xffffffffu, !idle(key)) == 0) break;
      if (!idle(key)) {
#if KMI_SEG_SOA
        u64* const kp = skey + li; u32* const vp = sval + li;
#else
        u64* const kp = &seg[li].key; u32* const vp = &seg[li].val;
#endif
        u64 cur = *kp;
        if (idle(cur)) {
          cur = atomicCAS(kp, KEY_EMPTY_TOP, key);
          if (idle(cur)) { ++n_new; cur = key; }
        }
        if (cur == key) {
          atomicAdd(vp, 1u);
          key = KEY_EMPTY_TOP;
        } else if (++li == SEG) {
          const u32 o = atomicAdd(ovf_cur(bs), 1u);
          if (o < bs.ovf_cap) ovf_base<1>(bs)[o] = key;
          else atomicOr(t.error, TERR_SPILL);   // dropped: sticky table error, every caller's check_table_error sees it
          key = KEY_EMPTY_TOP;
        }
      }
    }
    __syncthreads();
#if KMI_SEG_SOA
    // aux (the CSR fill cursor of the position index) is scratch that csr_finalize zeroes before use: written as 0
    for (u32 i = tid; i < SEG; i += SEG_THREADS) { const u64 k = skey[i]; gseg[i] = make_uint4((u32)k, (u32)(k >> 32), sval[i], 0u); }
#else
    for (u32 i = tid; i < SEG; i += SEG_THREADS) gseg[i] = reinterpret_cast<const uint4*>(seg)[i];
#endif
    __syncthreads();
  }
  n_new = warp_sum(n_new);
  if ((tid & 31) == 0 && n_new) atomicAdd(t.size, (u64)n_new);
}

// ======================================================================================================
// Two-word keys (33 <= K <= 63): the same second level with 16-byte keys and 32-byte slots.
// Shared memory has no 128-bit CAS, so a slot is claimed in two steps: CAS of its top word from EMPTY to LOCK,
// low word stored, top word published; a prober that sees LOCK looks at the same slot again (the claiming lane
// never waits in between, so this cannot dead-lock inside a warp).
// ======================================================================================================
static constexpr int FINE_ROUND2 = 2048;
static constexpr int LOG_SEG_SLOTS2 = 11;                 // 2048 slots x 20 B (hi[] + lo[] + val[]) = 40 KB per CTA
static constexpr u64 KEY_LOCK_TOP = 0xA000000000000000ULL; // top word of a slot whose low word is being written
__host__ __device__ inline size_t scatter_fine2_smem(u32 log_f) {
  const size_t F = (size_t)1 << log_f;
  return FINE_ROUND2 * sizeof(ulonglong2) + F * sizeof(u64) + 2 * F * sizeof(u32) + FINE_ROUND2 * sizeof(unsigned short);
}
__host__ __device__ inline size_t segment2_smem(u32 log_seg) { return (size_t)20 << log_seg; }

template <int K, int HASH>
__global__ void __launch_bounds__(TILE_THREADS, 4)
k_scatter_fine2(BucketSink bs, FineSink fs, u32* __restrict__ perr) {
  static_assert(K > 32, "two-word keys");
  extern __shared__ __align__(16) unsigned char fine_raw[];
  __shared__ u32 cpre[N_BUCKETS * MAX_SUB + 1];
  __shared__ u32 tmp[TILE_THREADS / 32 + 1];
  constexpr int R = FINE_ROUND2 / TILE_THREADS;
  const int t = threadIdx.x;
  const u32 F = 1u << fs.log_f;
  ulonglong2* sorted = reinterpret_cast<ulonglong2*>(fine_raw);
  u64* dst = reinterpret_cast<u64*>(sorted + FINE_ROUND2);
  u32* hist = reinterpret_cast<u32*>(dst + F);
  u32* delta = hist + F;
  unsigned short* sbid = reinterpret_cast<unsigned short*>(delta + F);
  const u32 sub = bs.sub, NR = N_BUCKETS * sub;
  u32 n_chunks;
  {
    u32 mine = 0;
    for (u32 q = 0; q < sub; ++q) mine += (min(bs.cursor[t * sub + q], bs.capb) + FINE_ROUND2 - 1) / FINE_ROUND2;
    u32 ex = block_excl_scan(mine, tmp, n_chunks);
    for (u32 q = 0; q < sub; ++q) { cpre[t * sub + q] = ex; ex += (min(bs.cursor[t * sub + q], bs.capb) + FINE_ROUND2 - 1) / FINE_ROUND2; }
    if (t == 0) cpre[NR] = n_chunks;
  }
  for (u32 f = t; f < F; f += TILE_THREADS) hist[f] = 0;
  __syncthreads();
  for (u32 c = blockIdx.x; c < n_chunks; c += gridDim.x) {
    int lo = 0, hi = (int)NR - 1;                         // largest region with cpre[region] <= c
#pragma unroll
    for (int s = 0; s < LOG_BUCKETS; ++s) {
      const int mid = (lo + hi + 1) >> 1;
      if (cpre[mid] <= c) lo = mid; else hi = mid - 1;
    }
    if (sub > 1) {                                        // up to MAX_SUB * 256 regions: three more halvings
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int mid = (lo + hi + 1) >> 1;
        if (cpre[mid] <= c) lo = mid; else hi = mid - 1;
      }
    }
    const u32 first = (c - cpre[lo]) * FINE_ROUND2;
    const u32 m = min((u32)FINE_ROUND2, min(bs.cursor[lo], bs.capb) - first);
    const ulonglong2* src = reinterpret_cast<const ulonglong2*>(bs.keys) + (u64)lo * bs.capb + first;
    const u32 b = sub == 1 ? (u32)lo : (u32)lo / sub;
    const u64 region0 = (u64)(bs.pre_base + (b & bs.pre_mask)) << fs.log_f;
    ulonglong2 key[R];
    u32 fid[R], rank[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const u32 i = q * TILE_THREADS + t;
      fid[q] = BKT_NONE;
      if (i < m) {
        key[q] = __ldg(src + i);
        Key<2> kk; kk.w[0] = key[q].x; kk.w[1] = key[q].y;
        const u64 h = slot_hash<K>(kk);
        fid[q] = (u32)(h >> (64 - bs.log_pre - fs.log_f)) & (F - 1);
        rank[q] = atomicAdd(&hist[fid[q]], 1u);
      }
    }
    __syncthreads();
    for (u32 f = t; f < F; f += TILE_THREADS) delta[f] = hist[f];          // counts
    __syncthreads();
    {                                                     // exclusive prefix of the counts into hist (F <= 2048: 8 per thread)
      const u32 EPT = (F + TILE_THREADS - 1) / TILE_THREADS;
      u32 local = 0;
      for (u32 e = 0; e < EPT; ++e) { const u32 f = t * EPT + e; if (f < F) local += delta[f]; }
      u32 total;
      u32 ex = block_excl_scan_nb(local, tmp, total);
      for (u32 e = 0; e < EPT; ++e) {
        const u32 f = t * EPT + e;
        if (f < F) {
          const u32 cnt = delta[f];
          hist[f] = ex;
          u64 d = 1ULL << 63;
          u32 gb = 0;
          if (cnt) {
            gb = atomicAdd(fs.cursor + region0 + f, cnt);
            if ((u64)gb + cnt <= (u64)fs.capf)
              d = reinterpret_cast<u64>(reinterpret_cast<ulonglong2*>(fs.keys) + (region0 + f) * fs.capf) + ((u64)gb - (u64)ex) * 16u;
          }
          dst[f] = d;
          delta[f] = gb - ex;                              // slow path: position in the region = delta + r
          ex += cnt;
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (fid[q] != BKT_NONE) {
        const u32 pos = hist[fid[q]] + rank[q];
        sorted[pos] = key[q];
        sbid[pos] = (unsigned short)fid[q];
      }
    }
    __syncthreads();
    for (u32 f = t; f < F; f += TILE_THREADS) hist[f] = 0;
    for (u32 r = t; r < m; r += TILE_THREADS) {
      const u32 f = sbid[r];
      const u64 d = dst[f];
      if (!(d >> 63)) { *reinterpret_cast<ulonglong2*>(d + (u64)r * 16u) = sorted[r]; continue; }
      const u32 at = delta[f] + r;                          // the run crosses the region's capacity: key by key
      if (at < fs.capf) { reinterpret_cast<ulonglong2*>(fs.keys)[(region0 + f) * fs.capf + at] = sorted[r]; continue; }
      const u32 o = atomicAdd(ovf_cur(bs), 1u);             // the fine region is full: overflow region
      if (o < bs.ovf_cap) reinterpret_cast<ulonglong2*>(ovf_base<2>(bs))[o] = sorted[r];
      else atomicOr(perr, PERR_OVERFLOW);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ u64 lds_volatile(const u64* p) { return *reinterpret_cast<const volatile u64*>(p); }

template <int K, int HASH>
__global__ void __launch_bounds__(384, 4)
k_insert_segments2(TableView t, FineSink fs, BucketSink bs, const u32 log_seg, const u32 n_segs, u32* __restrict__ perr,
                   const u32 fresh) {
  static_assert(K > 32, "two-word keys");
  extern __shared__ __align__(16) unsigned char seg_raw[];
  const u32 SEG = 1u << log_seg;
  u64* shi = reinterpret_cast<u64*>(seg_raw);
  u64* slo = shi + SEG;
  u32* sval = reinterpret_cast<u32*>(slo + SEG);
  const int tid = threadIdx.x;
  u32 n_new = 0;
  for (u32 s = blockIdx.x; s < n_segs; s += gridDim.x) {
    const u32 cnt = min(fs.cursor[s], fs.capf);
    if (cnt == 0 && !fresh) continue;
    const ulonglong2* src = reinterpret_cast<const ulonglong2*>(fs.keys) + (u64)s * fs.capf;
    for (u32 i = tid * 8u; i < cnt; i += 384u * 8u) prefetch_l2(src + i);       // the whole fine bucket -> L2
    Slot<2>* gseg = reinterpret_cast<Slot<2>*>(t.slots) + ((u64)s << log_seg);
    if (fresh) {
      for (u32 i = tid; i < SEG; i += 384) { shi[i] = KEY_EMPTY_TOP; slo[i] = 0; sval[i] = 0; }
    } else {
      for (u32 i = tid; i < SEG; i += 384) {
        const ulonglong2 k = *reinterpret_cast<const ulonglong2*>(&gseg[i]);
        slo[i] = k.x; shi[i] = k.y; sval[i] = gseg[i].val;
      }
    }
    __syncthreads();
    ulonglong2 nk = (tid < (int)cnt) ? __ldg(src + tid) : make_ulonglong2(0, 0);
    for (u32 i = tid; i < cnt; i += 384) {
      const ulonglong2 key = nk;
      if (i + 384 < cnt) nk = __ldg(src + i + 384);
      Key<2> kk; kk.w[0] = key.x; kk.w[1] = key.y;
      u32 li = (u32)(slot_hash<K>(kk) >> t.shift) & (SEG - 1);
      for (;;) {
        u64 h = lds_volatile(shi + li);
        if (h == KEY_EMPTY_TOP) {
          h = atomicCAS(shi + li, KEY_EMPTY_TOP, KEY_LOCK_TOP);
          if (h == KEY_EMPTY_TOP) {                        // claimed: low word, then publish the top word
            *reinterpret_cast<volatile u64*>(slo + li) = key.x;
            __threadfence_block();
            *reinterpret_cast<volatile u64*>(shi + li) = key.y;
            atomicAdd(sval + li, 1u);
            ++n_new;
            break;
          }
        }
        if (h == KEY_LOCK_TOP) continue;                   // being claimed: look again
        if (h == key.y && lds_volatile(slo + li) == key.x) { atomicAdd(sval + li, 1u); break; }
        if (++li == SEG) {                                 // ran off the segment: global linear probe continues there
          const u32 o = atomicAdd(ovf_cur(bs), 1u);
          if (o < bs.ovf_cap) reinterpret_cast<ulonglong2*>(ovf_base<2>(bs))[o] = key;
          else atomicOr(t.error, TERR_SPILL);   // dropped: sticky table error, every caller's check_table_error sees it
          break;
        }
      }
    }
    __syncthreads();
    for (u32 i = tid; i < SEG; i += 384) {                 // aux is scratch (zeroed before the CSR fill), pad unused
      uint4* g = reinterpret_cast<uint4*>(&gseg[i]);
      const u64 l = slo[i], h = shi[i];
      g[0] = make_uint4((u32)l, (u32)(l >> 32), (u32)h, (u32)(h >> 32));
      g[1] = make_uint4(sval[i], 0u, 0u, 0u);
    }
    __syncthreads();
  }
  n_new = warp_sum(n_new);
  if ((tid & 31) == 0 && n_new) atomicAdd(t.size, (u64)n_new);
}

// ======================================================================================================
// Sharded build: coarse buckets -> the owners' receive buffers over NVLink (replaces mxx::all2allv of
// imxx::distribute, reference src/io/incremental_mxx.hpp:1039-1109, for the count-index build).
//
// The parser has already cut this rank's k-mers into the 256 hash-prefix buckets of its LOCAL staging buffer (the same
// kernel as the single-GPU build). This kernel reads a bucket chunk by chunk, computes each key's owner with the
// reference's dist hash (KeyToRank = farm/murmur % p, distributed_densehash_map.hpp:148-170, bit-exact), ranks the chunk
// by owner in shared memory, and stores each owner's run — ~FINE_ROUND / p consecutive keys — straight into region
// (prefix, this rank) of that owner's receive buffer (peer memory mapped with CUDA IPC; plain stores that travel over
// NVLink / NVSwitch). Because every (prefix, source) region has exactly one writer, its fill cursor is a LOCAL counter:
// no remote atomics, and the receiver's table build starts at its fine scatter (the regions of one prefix are the
// `sub` inputs of k_scatter_fine) — no coarse pass over received keys. Runs while the parser produces the next round.
// TMA bulk copy shared -> global (cp.async.bulk, SASS UBLKCP): one thread hands a whole run to the copy engine of the SM,
// which writes it as full-width bursts while the CTA goes on hashing the next chunk. src / dst 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_store(void* gdst, const void* ssrc, u32 bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(__cvta_generic_to_global(gdst)), "r"((u32)__cvta_generic_to_shared(ssrc)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

struct PushSink {
  u64* const* peer_keys;  // [p]: base of owner r's receive regions as seen from THIS process (own buffer for r == me)
  u32* cursor;            // [N_BUCKETS * p], local: keys sent so far to owner r for prefix b at b * p + r (accumulates over rounds)
  u32 cap_sub;            // keys per (prefix, source) region of a receive buffer
  u32 p, me;              // ranks; this rank is sub-region `me` of every prefix in every receive buffer
  u32* status;            // |= 1: a receive region would overflow (skew) -> the round is redone through the generic exchange
  u32 bulk;               // 1: runs leave as TMA bulk copies (one thread per owner), 0: 8-byte stores by all threads
};
static constexpr int PUSH_ROUND = 2048;
template <int NW> struct PushSmem {
  u64 sorted[(PUSH_ROUND + 2 * MAX_SUB + 2) * NW];   // every owner's run may start one key late (16-byte alignment of the bulk copies)
  u64 dst[MAX_SUB];       // byte address sorted[0] would land at in owner r's region (0: the run does not fit)
  u32 hist[MAX_SUB];      // keys per owner while ranking, start of the owner's run in sorted[] afterwards
  u32 cnt[MAX_SUB];       // keys per owner of this round
  uint8_t sbid[PUSH_ROUND];
};
template <int K, int HASH>
__global__ void __launch_bounds__(TILE_THREADS, 3)
k_owner_push(BucketSink bs, PushSink ps) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  constexpr int R = PUSH_ROUND / TILE_THREADS;
  __shared__ PushSmem<NW> sm;
  __shared__ u32 cpre[N_BUCKETS + 2];
  __shared__ u32 tmp[TILE_THREADS / 32 + 1];
  const int t = threadIdx.x, lane = t & 31;
  const u32 lt = (1u << lane) - 1u;
  const u32 p = ps.p;
  // chunks of the 256 buckets, then of the overflow region (keys of any prefix: placed one by one)
  u32 n_chunks;
  {
    const u32 mine = (min(bs.cursor[t], bs.capb) + PUSH_ROUND - 1) / PUSH_ROUND;
    const u32 ex = block_excl_scan(mine, tmp, n_chunks);
    cpre[t] = ex;
    if (t == 0) {
      cpre[N_BUCKETS] = n_chunks;
      cpre[N_BUCKETS + 1] = n_chunks + (min(*ovf_cur(bs), bs.ovf_cap) + PUSH_ROUND - 1) / PUSH_ROUND;
    }
  }
  if (t < MAX_SUB) sm.hist[t] = 0;
  __syncthreads();
  const u32 n_all = cpre[N_BUCKETS + 1];
  // chunk c -> (region, its keys); the keys of the NEXT chunk are requested before this one is hashed and ranked, so the
  // global-load latency hides behind a whole round (the kernel was waiting on its key loads half of the time)
  auto locate = [&](u32 c, u32& b, u32& m, const u64*& src) {
    int lo = 0, hi = N_BUCKETS;                           // largest region with cpre[region] <= c (256 = overflow region)
#pragma unroll
    for (int s = 0; s < LOG_BUCKETS + 1; ++s) {
      const int mid = (lo + hi + 1) >> 1;
      if (cpre[mid] <= c) lo = mid; else hi = mid - 1;
    }
    b = (u32)lo;
    const u32 first = (c - cpre[b]) * PUSH_ROUND;
    const u32 cnt_b = b == N_BUCKETS ? min(*ovf_cur(bs), bs.ovf_cap) : min(bs.cursor[b], bs.capb);
    m = min((u32)PUSH_ROUND, cnt_b - first);
    src = (b == N_BUCKETS ? ovf_base<NW>(bs) : bs.keys + (u64)b * bs.capb * NW) + (u64)first * NW;
  };
  Key<NW> nkey[R];
  u32 nb = 0, nm = 0;
  if (blockIdx.x < n_all) {
    const u64* src;
    locate(blockIdx.x, nb, nm, src);
#pragma unroll
    for (int q = 0; q < R; ++q) { const u32 i = q * TILE_THREADS + t; if (i < nm) nkey[q] = load_key<NW>(src, i); }
  }
  for (u32 c = blockIdx.x; c < n_all; c += gridDim.x) {
    const u32 b = nb, m = nm;
    const bool is_ovf = b == N_BUCKETS;
    Key<NW> key[R];
#pragma unroll
    for (int q = 0; q < R; ++q) key[q] = nkey[q];
    if (c + gridDim.x < n_all) {
      const u64* src;
      locate(c + gridDim.x, nb, nm, src);
#pragma unroll
      for (int q = 0; q < R; ++q) { const u32 i = q * TILE_THREADS + t; if (i < nm) nkey[q] = load_key<NW>(src, i); }
    }
    u32 own[R], rank[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const u32 i = q * TILE_THREADS + t;
      own[q] = BKT_NONE;
      if (i < m) {
        u64 hs, hd;
        kmer_hashes<K, HASH>(key[q], hs, hd);
        own[q] = mod_ranks(hd, p);
      }
    }
    if (is_ovf) {                                          // rare: every key finds its own prefix and slot
#pragma unroll
      for (int q = 0; q < R; ++q) {
        if (own[q] == BKT_NONE) continue;
        const u32 pb = bucket_of(slot_hash<K>(key[q]));
        const u32 at = atomicAdd(ps.cursor + pb * p + own[q], 1u);
        if (at < ps.cap_sub) store_key<NW>(ps.peer_keys[own[q]] + ((u64)(pb * p + ps.me) * ps.cap_sub) * NW, at, key[q]);
        else atomicOr(ps.status, 1u);
      }
      continue;                                            // block-uniform
    }
    // rank by owner: one shared-memory atomic per distinct owner and warp instruction (p <= 8 owners)
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const u32 peers = __match_any_sync(0xffffffffu, own[q]);
      u32 base = 0;
      if (own[q] != BKT_NONE && (peers & lt) == 0) base = atomicAdd(&sm.hist[own[q]], (u32)__popc(peers));
      base = __shfl_sync(0xffffffffu, base, __ffs((int)peers) - 1);
      rank[q] = base + (u32)__popc(peers & lt);
    }
    __syncthreads();
    if (t < 32) {                                          // owners' counts -> exclusive prefix, global run reservation
      const u32 cnt = (u32)t < p ? sm.hist[t] : 0u;
      u32 inc = cnt;
#pragma unroll
      for (int d = 1; d < MAX_SUB; d <<= 1) {
        const u32 o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
      }
      u32 ex = inc - cnt;
      if ((u32)t < p) {
        u64 d = 0;
        u32 gb = 0;
        if (cnt) {
          gb = atomicAdd(ps.cursor + b * p + t, cnt);
          if ((u64)gb + cnt > (u64)ps.cap_sub) { atomicOr(ps.status, 1u); gb = ~0u; }
        }
        if (ps.bulk) {
          // the run starts where its first key is 16-byte aligned in shared memory exactly when it is in the region
          if (NW == 1) ex += 2u * (u32)t + ((ex ^ gb) & 1u);
          bulk_wait_read();                                // the copies of the last round have read their runs
        }
        if (cnt && gb != ~0u) d = reinterpret_cast<u64>(ps.peer_keys[t] + ((u64)(b * p + ps.me) * ps.cap_sub + gb) * NW) - (u64)ex * (8u * NW);
        sm.dst[t] = d;
        sm.hist[t] = ex;
        sm.cnt[t] = cnt;
      }
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (own[q] != BKT_NONE) {
        const u32 pos = sm.hist[own[q]] + rank[q];
        if (NW == 1) sm.sorted[pos] = key[q].w[0];
        else reinterpret_cast<ulonglong2*>(sm.sorted)[pos] = make_ulonglong2(key[q].w[0], key[q].w[NW - 1]);
        if (!ps.bulk) sm.sbid[pos] = (uint8_t)own[q];
      }
    }
    if (ps.bulk) fence_async_smem();                       // the copy engine reads what the threads just wrote
    __syncthreads();
    if (ps.bulk) {
      if ((u32)t < p) {                                    // one thread per owner: head key, bulk body, tail key
        const u64 d = sm.dst[t];
        u32 cnt = sm.cnt[t], at = sm.hist[t];
        if (d && cnt) {
          if (NW == 1) {
            if ((d + (u64)at * 8u) & 8u) { *reinterpret_cast<u64*>(d + (u64)at * 8u) = sm.sorted[at]; ++at; --cnt; }
            const u32 body = cnt & ~1u;
            if (body) bulk_store(reinterpret_cast<void*>(d + (u64)at * 8u), sm.sorted + at, body * 8u);
            if (cnt & 1u) *reinterpret_cast<u64*>(d + (u64)(at + body) * 8u) = sm.sorted[at + body];
          } else {
            bulk_store(reinterpret_cast<void*>(d + (u64)at * 16u), sm.sorted + 2 * at, cnt * 16u);
          }
        }
        bulk_commit();
      }
      if (t < MAX_SUB) sm.hist[t] = 0;
    } else {
      if (t < MAX_SUB) sm.hist[t] = 0;
      for (u32 r = t; r < m; r += TILE_THREADS) {
        const u64 d = sm.dst[sm.sbid[r]];
        if (!d) continue;
        if (NW == 1) *reinterpret_cast<u64*>(d + (u64)r * 8u) = sm.sorted[r];
        else *reinterpret_cast<ulonglong2*>(d + (u64)r * 16u) = reinterpret_cast<const ulonglong2*>(sm.sorted)[r];
      }
    }
    __syncthreads();
  }
  if (ps.bulk && (u32)t < p) bulk_wait_all();
}
#ifdef KMI_DIST_KERNELS  // non-template: compiled once, in kmi_dist.cu
// after k_owner_push of a round: publish this rank's running totals in the owners' receive buffers (count word of region
// (prefix, me) in owner r's count array) and clear the staging cursors for the next round
__global__ void __launch_bounds__(256) k_push_publish(PushSink ps, u32* const* __restrict__ peer_counts, u32* __restrict__ stage_cursor,
                                                      const u32* __restrict__ perr) {
  // what the parser reported for this round joins the status words the ranks all-reduce: [0] something overflowed (the
  // round must be redone through the generic exchange), [1] malformed input
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const u32 e = *perr;
    if (e & PERR_OVERFLOW) atomicOr(ps.status, 1u);
    if (e & (PERR_NO_AT | PERR_NO_PLUS)) ps.status[1] = 1u;
  }
  const u32 n = N_BUCKETS * ps.p;
  for (u32 j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
    const u32 b = j / ps.p, r = j % ps.p;
    peer_counts[r][b * ps.p + ps.me] = min(ps.cursor[j], ps.cap_sub);
  }
  if (blockIdx.x == 0) for (u32 j = threadIdx.x; j <= N_BUCKETS; j += blockDim.x) stage_cursor[j] = 0;
  __threadfence_system();
}
#endif  // KMI_DIST_KERNELS

// the overflow region (bucket overflow + segment spills) -> table with global atomics
template <int K, int HASH>
__global__ void __launch_bounds__(256) k_insert_overflow(TableView t, BucketSink bs) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  const u64 n = min(*ovf_cur(bs), bs.ovf_cap);
  const u64* ovf = ovf_base<NW>(bs);
  u32 n_new = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64)gridDim.x * blockDim.x) {
    Key<NW> key = load_key<NW>(ovf, i);
    bool was_new;
    const u64 s = table_upsert<NW>(t, key, slot_hash<K>(key), was_new);
    if (s != ~0ULL) {
      table_add<NW>(t, s, 1u);
      n_new += was_new ? 1u : 0u;
    }
  }
  n_new = warp_sum(n_new);
  if ((threadIdx.x & 31) == 0 && n_new) atomicAdd(t.size, (u64)n_new);
}

}  // namespace kmi
