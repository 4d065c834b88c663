// TEST INFRASTRUCTURE — not product code.
// Thin extern "C" shim over the REFERENCE's own headers, compiled from where they lie under
// /root/reference (never copied) into oracle/_ref/libkmi_ref_{ndebug,debug}.so by oracle/Makefile.
// It executes the reference's real code for SURVEY.md §8 rows a7 (DNA::FROM_ASCII), a8
// (Kmer::nextFromChar), a10 (lex_less / reverse_complement), a11 (murmur), a12 (farm), a16
// (special_keys sentinels, google::dense_hash_map). The record walkers above them need mpi.h
// (absent here) and are restated in oracle/kmi_oracle.c instead.
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#include <utility>
#include <thread>

#include "common/kmer.hpp"
#include "common/kmer_transform.hpp"
#include "common/alphabets.hpp"
#include "index/kmer_hash.hpp"
#include <sparsehash/dense_hash_map>

using bliss::common::DNA;
template <unsigned K> using KM = bliss::common::Kmer<K, DNA, uint64_t>;

namespace {

template <unsigned K>
struct Ops {
  using Kmer = KM<K>;
  static constexpr int NW = Kmer::nWords;

  // chars -> forward k-mers (every window of one EOL-free sequence), reference sliding window.
  static size_t windows(const uint8_t* s, size_t n, uint64_t* out) {
    if (n < K) return 0;
    Kmer km;
    size_t m = 0;
    for (size_t i = 0; i < n; ++i) {
      km.nextFromChar(DNA::FROM_ASCII[s[i]]);
      if (i + 1 >= K) { std::memcpy(out + m * NW, km.getData(), NW * 8); ++m; }
    }
    return m;
  }
  static void revcomp(const uint64_t* in, uint64_t* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
      Kmer km; std::memcpy(km.getDataRef(), in + i * NW, NW * 8);
      Kmer rc = km.reverse_complement();
      std::memcpy(out + i * NW, rc.getData(), NW * 8);
    }
  }
  static void canonical(const uint64_t* in, uint64_t* out, size_t n) {
    bliss::kmer::transform::lex_less<Kmer> op;
    for (size_t i = 0; i < n; ++i) {
      Kmer km; std::memcpy(km.getDataRef(), in + i * NW, NW * 8);
      Kmer c = op(km);
      std::memcpy(out + i * NW, c.getData(), NW * 8);
    }
  }
  // which: 0 murmur store (h[0]), 1 murmur dist (h[1]), 2 farm store (seed 42), 3 farm dist (seed 83)
  static void hash(int which, const uint64_t* in, uint64_t* out, size_t n) {
    bliss::kmer::hash::murmur<Kmer, false> ms; bliss::kmer::hash::murmur<Kmer, true> md;
    bliss::kmer::hash::farm<Kmer, false> fs;   bliss::kmer::hash::farm<Kmer, true> fd;
    for (size_t i = 0; i < n; ++i) {
      Kmer km; std::memcpy(km.getDataRef(), in + i * NW, NW * 8);
      out[i] = which == 0 ? ms(km) : which == 1 ? md(km) : which == 2 ? fs(km) : fd(km);
    }
  }
  static void sentinels(uint64_t* out) {
    bliss::kmer::hash::sparsehash::special_keys<Kmer, true> sk;
    Kmer e = sk.generate(0), d = sk.generate(1);
    std::memcpy(out, e.getData(), NW * 8);
    std::memcpy(out + NW, d.getData(), NW * 8);
  }

  struct StoreHash {
    int which;
    bliss::kmer::hash::murmur<Kmer, false> ms; bliss::kmer::hash::farm<Kmer, false> fs;
    size_t operator()(const Kmer& k) const { return which == 0 ? ms(k) : fs(k); }
  };
  using Map = google::dense_hash_map<Kmer, uint32_t, StoreHash>;

  // The reference's p-rank build restated over threads (no MPI here): every "rank" walks its
  // block of sequences with the reference Kmer window, canonicalises with lex_less, buckets by
  // KeyToRank = dist_hash % p (distributed_densehash_map.hpp:148-170), the buckets are handed
  // over (the all2allv), and every rank inserts what it received into its own
  // google::dense_hash_map with the reference's load factors and sentinels
  // (densehash_map.hpp:960-974), count += 1 on a hit (distributed_densehash_map.hpp:2463-2482).
  // seqs: concatenated EOL-free sequences, off[i]..off[i+1]. Returns distinct keys; if out_keys
  // is non-null the (key,count) pairs are written unsorted, capacity cap.
  static size_t count_build(const uint8_t* seqs, const uint64_t* off, size_t nseq, int p,
                            int hash_kind /*0 murmur,1 farm*/, uint64_t* out_keys,
                            uint32_t* out_counts, size_t cap, uint64_t* per_rank_sizes) {
    if (p < 1) p = 1;
    std::vector<std::vector<std::vector<Kmer>>> send(p, std::vector<std::vector<Kmer>>(p));
    auto produce = [&](int r) {
      size_t b = nseq * (size_t)r / p, e = nseq * (size_t)(r + 1) / p;
      bliss::kmer::transform::lex_less<Kmer> canon;
      bliss::kmer::hash::murmur<Kmer, true> md; bliss::kmer::hash::farm<Kmer, true> fd;
      for (size_t i = b; i < e; ++i) {
        const uint8_t* s = seqs + off[i]; size_t n = off[i + 1] - off[i];
        if (n < K) continue;
        Kmer km;
        for (size_t j = 0; j < n; ++j) {
          km.nextFromChar(DNA::FROM_ASCII[s[j]]);
          if (j + 1 >= K) {
            Kmer c = canon(km);
            uint64_t h = hash_kind == 0 ? md(c) : fd(c);
            send[r][h % (uint64_t)p].push_back(c);
          }
        }
      }
    };
    std::vector<Map*> maps(p, nullptr);
    auto consume = [&](int r) {
      bliss::kmer::hash::sparsehash::special_keys<Kmer, true> sk;
      Map* m = new Map(128, StoreHash{hash_kind, {}, {}});
      m->set_empty_key(sk.generate(0)); m->set_deleted_key(sk.generate(1));
      m->max_load_factor(0.7); m->min_load_factor(0.3);
      for (int src = 0; src < p; ++src)
        for (const Kmer& k : send[src][r]) {
          auto res = m->insert(std::make_pair(k, (uint32_t)1));
          if (!res.second) res.first->second += 1;
        }
      maps[r] = m;
    };
    {
      std::vector<std::thread> th;
      for (int r = 1; r < p; ++r) th.emplace_back(produce, r);
      produce(0);
      for (auto& t : th) t.join();
    }
    {
      std::vector<std::thread> th;
      for (int r = 1; r < p; ++r) th.emplace_back(consume, r);
      consume(0);
      for (auto& t : th) t.join();
    }
    size_t total = 0, w = 0;
    for (int r = 0; r < p; ++r) {
      if (per_rank_sizes) per_rank_sizes[r] = maps[r]->size();
      total += maps[r]->size();
      if (out_keys)
        for (auto it = maps[r]->begin(); it != maps[r]->end() && w < cap; ++it, ++w) {
          std::memcpy(out_keys + w * NW, it->first.getData(), NW * 8);
          out_counts[w] = it->second;
        }
      delete maps[r];
    }
    return total;
  }
};

}  // namespace

// k-mer lengths instantiated (the ones libkmi.so builds)
#define DISPATCH(K, CALL, FALLBACK)                                                        \
  switch (K) {                                                                              \
    case 15: return Ops<15>::CALL; case 21: return Ops<21>::CALL; case 31: return Ops<31>::CALL; \
    case 35: return Ops<35>::CALL; case 47: return Ops<47>::CALL; case 63: return Ops<63>::CALL; \
    default: return FALLBACK;                                                               \
  }

extern "C" {

int kmiref_nwords(int K) { return K <= 32 ? 1 : 2; }
int kmiref_debug_mode(void) {
#ifdef NDEBUG
  return 0;
#else
  return 1;
#endif
}
uint8_t kmiref_from_ascii(uint8_t c) { return DNA::FROM_ASCII[c]; }

size_t kmiref_windows(int K, const uint8_t* s, size_t n, uint64_t* out) { DISPATCH(K, windows(s, n, out), 0) }
void kmiref_revcomp(int K, const uint64_t* in, uint64_t* out, size_t n) { DISPATCH(K, revcomp(in, out, n), (void)0) }
void kmiref_canonical(int K, const uint64_t* in, uint64_t* out, size_t n) { DISPATCH(K, canonical(in, out, n), (void)0) }
void kmiref_hash(int K, int which, const uint64_t* in, uint64_t* out, size_t n) { DISPATCH(K, hash(which, in, out, n), (void)0) }
void kmiref_sentinels(int K, uint64_t* out) { DISPATCH(K, sentinels(out), (void)0) }
size_t kmiref_count_build(int K, const uint8_t* seqs, const uint64_t* off, size_t nseq, int p, int hash_kind,
                          uint64_t* out_keys, uint32_t* out_counts, size_t cap, uint64_t* per_rank_sizes) {
  DISPATCH(K, count_build(seqs, off, nseq, p, hash_kind, out_keys, out_counts, cap, per_rank_sizes), 0)
}

}  // extern "C"
