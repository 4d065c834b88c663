// kmi_device.cuh — device-side value arithmetic of the k-mer index path (sm_100a).
// Keys are NW 64-bit words, word 0 least significant, laid out exactly as the reference's
// bliss::common::Kmer<K, DNA, uint64_t>::data (reference src/common/kmer.hpp:76-100,177).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace kmi {

typedef unsigned long long u64;
typedef unsigned int u32;

// Empty / deleted sentinels of the reference's dense-hash maps for k-mers with >= 2 pad bits:
// top word 0x8000.. / 0xC000.., other words 0 (reference src/index/kmer_hash.hpp:408-414).
static constexpr u64 KEY_EMPTY_TOP = 0x8000000000000000ULL;
static constexpr u64 KEY_DELETED_TOP = 0xC000000000000000ULL;

template <int NW> struct Key { u64 w[NW]; };

template <int NW> __host__ __device__ __forceinline__ bool key_eq(const Key<NW>& a, const Key<NW>& b) {
  bool e = true;
#pragma unroll
  for (int i = 0; i < NW; ++i) e = e && (a.w[i] == b.w[i]);
  return e;
}
// Kmer::operator< (kmer.hpp:820-823): unsigned multi-word compare, most significant word first.
template <int NW> __host__ __device__ __forceinline__ bool key_less(const Key<NW>& a, const Key<NW>& b) {
  if (NW == 2) {
    if (a.w[NW - 1] != b.w[NW - 1]) return a.w[NW - 1] < b.w[NW - 1];
  }
  return a.w[0] < b.w[0];
}

// reverse the order of the 32 2-bit groups of a word
__device__ __forceinline__ u64 rev2(u64 x) {
  u64 r = __brevll(x);
  return ((r >> 1) & 0x5555555555555555ULL) | ((r & 0x5555555555555555ULL) << 1);
}

// Kmer::reverse_complement for the 2-bit DNA alphabet (kmer.hpp:1118-1123,1726-1742):
// reverse the 2-bit groups over the whole word array, complement, shift out the pad bits.
template <int K> __device__ __forceinline__ Key<(K <= 32 ? 1 : 2)> revcomp(const Key<(K <= 32 ? 1 : 2)>& x) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  Key<NW> r;
  if (NW == 1) {
    r.w[0] = (~rev2(x.w[0])) >> (64 - 2 * K);
  } else {
    constexpr int S = 128 - 2 * K;  // pad bits, 2..62 for 33 <= K <= 63
    u64 hi = ~rev2(x.w[0]);         // reversed low word becomes the high word
    u64 lo = ~rev2(x.w[NW - 1]);
    r.w[0] = (lo >> S) | (hi << (64 - S));
    r.w[NW - 1] = hi >> S;
  }
  return r;
}
// lex_less (reference src/common/kmer_transform.hpp:109-126): (x < rc) ? x : rc
template <int K> __device__ __forceinline__ Key<(K <= 32 ? 1 : 2)> canonical(const Key<(K <= 32 ? 1 : 2)>& x) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  Key<NW> rc = revcomp<K>(x);
  return key_less<NW>(x, rc) ? x : rc;
}

// ---- MurmurHash3_x64_128 over the first NB = (2K+7)/8 bytes of the key, seed 42
// (reference ext/smhasher/MurmurHash3.cpp:255-332; src/index/kmer_hash.hpp:243-276).
__device__ __forceinline__ u64 rotl(u64 x, int r) { return (x << r) | (x >> (64 - r)); }
__device__ __forceinline__ u64 rotr(u64 x, int r) { return (x >> r) | (x << (64 - r)); }
__device__ __forceinline__ u64 fmix64(u64 k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
  return k;
}
template <int NB> __device__ __forceinline__ void murmur3_128(u64 lo, u64 hi, u64& h1o, u64& h2o) {
  const u64 c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
  u64 h1 = 42, h2 = 42;
  if (NB == 16) {
    u64 k1 = lo, k2 = hi;
    k1 *= c1; k1 = rotl(k1, 31); k1 *= c2; h1 ^= k1;
    h1 = rotl(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
    k2 *= c2; k2 = rotl(k2, 33); k2 *= c1; h2 ^= k2;
    h2 = rotl(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
  } else {
    if (NB > 8) {
      u64 k2 = hi & ((NB >= 16) ? ~0ULL : ((1ULL << (8 * (NB > 8 ? NB - 8 : 0) % 64)) - 1));
      k2 *= c2; k2 = rotl(k2, 33); k2 *= c1; h2 ^= k2;
    }
    u64 k1 = (NB >= 8) ? lo : (lo & ((1ULL << (8 * (NB % 8))) - 1));
    k1 *= c1; k1 = rotl(k1, 31); k1 *= c2; h1 ^= k1;
  }
  h1 ^= (u64)NB; h2 ^= (u64)NB;
  h1 += h2; h2 += h1;
  h1 = fmix64(h1); h2 = fmix64(h2);
  h1 += h2; h2 += h1;
  h1o = h1; h2o = h2;
}

// ---- farmhashna::Hash64WithSeed for len <= 16 (reference ext/farmhash/src/farmhash.cc:523-529,
// HashLen0to16 :388-410, HashLen16 :378-386, Hash128to64 farmhash.h:129-138, DebugTweak :334-343)
__device__ __forceinline__ u64 farm_hl16(u64 u, u64 v, u64 mul) {
  u64 a = (u ^ v) * mul; a ^= a >> 47;
  u64 b = (v ^ a) * mul; b ^= b >> 47; b *= mul;
  return b;
}
template <int NB> __device__ __forceinline__ u64 farm_len0to16(u64 lo, u64 hi) {
  const u64 k0 = 0xc3a5c85c97cb3127ULL, k2 = 0x9ae16a3b2f90404fULL;
  if (NB >= 8) {
    u64 mul = k2 + NB * 2;
    u64 a = lo + k2;
    u64 b = (NB == 8) ? lo : (NB == 16) ? hi : ((lo >> (8 * (NB - 8) % 64)) | (hi << ((64 - 8 * (NB - 8)) % 64)));
    u64 c = rotr(b, 37) * mul + a;
    u64 d = (rotr(a, 25) + b) * mul;
    return farm_hl16(c, d, mul);
  }
  if (NB >= 4) {
    u64 mul = k2 + NB * 2;
    u64 a = lo & 0xffffffffULL;
    u64 b = (lo >> (8 * (NB - 4) % 64)) & 0xffffffffULL;
    return farm_hl16(NB + (a << 3), b, mul);
  }
  if (NB > 0) {
    u32 a = (u32)(lo & 0xff), b = (u32)((lo >> (8 * (NB >> 1))) & 0xff), c = (u32)((lo >> (8 * (NB - 1))) & 0xff);
    u32 y = a + (b << 8);
    u32 z = NB + (c << 2);
    u64 v = (y * k2) ^ (z * k0);
    v ^= v >> 47;
    return v * k2;
  }
  return k2;
}
__device__ __forceinline__ u64 farm_finish(u64 h0to16, u64 seed, bool tweak) {
  const u64 k1 = 0xb492b66fbe98f273ULL, k2 = 0x9ae16a3b2f90404fULL;
  u64 h = farm_hl16(h0to16 - k2, seed, 0x9ddfea08eb382d69ULL);
  if (tweak) {
    u64 x = h * k1;
    u32 xl = (u32)x, xh = (u32)(x >> 32);
    h = ~(((u64)__byte_perm(xl, 0, 0x0123) << 32) | (u64)__byte_perm(xh, 0, 0x0123));
  }
  return h;
}

// HASH: 0 murmur, 1 farm (NDEBUG), 2 farm with DebugTweak, 3 identity, 4 cpp_std. Returns store and dist hash together
// (murmur: one evaluation gives both halves; farm: the seed only enters the last mixing step).
// identity / cpp_std (reference src/index/kmer_hash.hpp:157-232) take the number of prefix bits the reference's KeyToRank
// constructs its DistHash with, ceilLog2(comm_size) (distributed_densehash_map.hpp:153-156): `pbits`.
//   identity: store = Kmer::getSuffix(min(nBits, 64)) (kmer.hpp:1245-1249), dist = Kmer::getPrefix(min(nBits, pbits)):
//             the top bits of the most significant word, below the pad (kmer.hpp:1203-1210).
//   cpp_std:  h = XOR over the words of (std::hash<size_t>(word) << 1) — libstdc++'s std::hash<size_t> is the identity —
//             store = h, dist = h >> (min(nBits, 64) - min(pbits, nBits)).
__host__ __device__ __forceinline__ u32 ceil_log2_u32(u32 p) {
  u32 b = 0;
  while ((1ull << b) < (u64)p) ++b;
  return b;
}
template <int K, int HASH> __device__ __forceinline__ void kmer_hashes(const Key<(K <= 32 ? 1 : 2)>& x, u64& store, u64& dist,
                                                                       const u32 pbits = 0) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  constexpr int NB = (2 * K + 7) / 8;
  constexpr u32 NBITS = 2u * K, PAD = 64u * NW - NBITS;
  u64 lo = x.w[0], hi = (NW == 2) ? x.w[NW - 1] : 0;
  if (HASH == 0) {
    murmur3_128<NB>(lo, hi, store, dist);
  } else if (HASH == 3) {
    const u32 bits = pbits < NBITS ? pbits : NBITS;
    store = (NW == 1 && NBITS < 64) ? (lo & ((1ull << (NBITS & 63)) - 1ull)) : lo;
    if (!bits) dist = 0ull;
    else if (PAD + bits <= 64u) dist = x.w[NW - 1] >> (64u - bits - PAD);
    else {                                                    // prefix longer than the top word's used bits (two-word keys)
      const u32 sh = NBITS - bits;                            // 0 < sh < 64
      dist = ((lo >> sh) | (hi << (64u - sh))) & (bits >= 64u ? ~0ull : ((1ull << bits) - 1ull));
    }
  } else if (HASH == 4) {
    u64 h = lo << 1;
    if (NW == 2) h ^= hi << 1;
    const u32 nb64 = NBITS < 64u ? NBITS : 64u, bits = pbits < NBITS ? pbits : NBITS;
    store = h;
    dist = (nb64 - bits) >= 64u ? 0ull : (h >> (nb64 - bits));
  } else {
    u64 h = farm_len0to16<NB>(lo, hi);
    store = farm_finish(h, 42, HASH == 2);
    dist = farm_finish(h, 83, HASH == 2);
  }
}
template <int K, int HASH> __device__ __forceinline__ u64 store_hash(const Key<(K <= 32 ? 1 : 2)>& x) {
  u64 s, d;
  kmer_hashes<K, HASH>(x, s, d);
  return s;
}

// 16-byte asynchronous global -> shared copy (LDGSTS), bypassing L1
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// dist_hash % p (KeyToRank, distributed_densehash_map.hpp:148-170): p is uniform over the grid, and 2/4/8 ranks are
// the common case, so the ~40-instruction 64-bit division is kept for the other values of p only
__device__ __forceinline__ u32 mod_ranks(u64 d, u32 p) {
  return ((p & (p - 1u)) == 0u) ? ((u32)d & (p - 1u)) : (u32)(d % (u64)p);
}

// Placement hash of the local table: picks the home slot (top bits), the hash-prefix bucket and the
// HyperLogLog register. Where an entry sits inside one GPU's table is unobservable — the reference's
// StoreHash is a free template parameter for the same reason (kmer_index.hpp:437-466) — so this is a
// 2-multiply mixer (~15 instructions) instead of the ~90-instruction farm/murmur evaluation, which
// stays bit-exact where it IS observable: the dist hash that picks the owning rank, and kmi_hash_dev.
// Measured on 31-mers of the bench workload: bucket balance, probe lengths and HLL error equal farm's.
template <int K> __device__ __forceinline__ u64 slot_hash(const Key<(K <= 32 ? 1 : 2)>& x) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  u64 h = x.w[0] * 0x9E3779B97F4A7C15ULL;
  if (NW == 2) h += x.w[NW - 1] * 0xC2B2AE3D27D4EB4FULL;
  h ^= h >> 32;
  h *= 0xD6E8FEB86659FD93ULL;
  return h;
}

// order-independent content checksum mixer (splitmix64 finaliser)
__host__ __device__ __forceinline__ u64 mix64(u64 x) {
  u64 z = x + 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

}  // namespace kmi
