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
#include <functional>

#include "common/kmer.hpp"
#include "common/kmer_transform.hpp"
#include "common/alphabets.hpp"
#include "index/kmer_hash.hpp"
#include "index/quality_score_iterator.hpp"
#include <sparsehash/dense_hash_map>

using bliss::common::DNA;
template <unsigned K> using KM = bliss::common::Kmer<K, DNA, uint64_t>;


namespace {

// ---- FASTQ record walk, restated (the MPI-dependent parser headers do not build here) -------------------------
// SequentialFASTQParser::get_next_record (reference src/io/fastq_loader.hpp:389-470): a line is a maximal run of
// non-EOL characters (findNonEOL / findEOL skip every '\n' and '\r', so blank lines vanish); a record is four
// consecutive lines, the first starting with '@', the third with '+'; the sequence is the second line.
inline bool is_eol(uint8_t c) { return c == '\n' || c == '\r'; }
// FASTQParser::find_first_record (fastq_loader.hpp:269-357): first record start at or after `from` (a partial first line is
// dropped unless from == 0), by the first characters of four consecutive lines: @?+? / ?@?+ / +?@? / ?+?@.
inline size_t first_record(const uint8_t* t, size_t n, size_t from) {
  size_t i = from;
  if (i >= n) return n;
  if (i > 0 && !is_eol(t[i - 1])) while (i < n && !is_eol(t[i])) ++i;
  while (i < n && is_eol(t[i])) ++i;
  size_t off[4]; uint8_t c[4] = {0, 0, 0, 0};
  for (int j = 0; j < 4; ++j) {
    if (i >= n) return n;
    off[j] = i; c[j] = t[i];
    while (i < n && !is_eol(t[i])) ++i;
    while (i < n && is_eol(t[i])) ++i;
  }
  if (c[0] == '@' && c[2] == '+') return off[0];
  if (c[1] == '@' && c[3] == '+') return off[1];
  if (c[0] == '+' && c[2] == '@') return off[2];
  if (c[1] == '+' && c[3] == '@') return off[3];
  return n;
}
// calls f(seq_begin, seq_len) for every record of the whole-record block [b, e); returns false on a malformed record
template <typename F> inline bool walk_fastq(const uint8_t* t, size_t b, size_t e, F f) {
  size_t i = b;
  for (;;) {
    while (i < e && is_eol(t[i])) ++i;
    if (i >= e) return true;
    if (t[i] != '@') return false;
    while (i < e && !is_eol(t[i])) ++i;
    while (i < e && is_eol(t[i])) ++i;
    const size_t s = i;
    while (i < e && !is_eol(t[i])) ++i;
    const size_t len = i - s;
    while (i < e && is_eol(t[i])) ++i;
    if (i >= e || t[i] != '+') return false;
    while (i < e && !is_eol(t[i])) ++i;
    while (i < e && is_eol(t[i])) ++i;
    while (i < e && !is_eol(t[i])) ++i;   // quality line
    f(s, len);
  }
}
// the order-independent content checksum of kmi_checksum (include/kmi.h): splitmix64 finaliser over the key words
inline uint64_t mix64(uint64_t x) {
  uint64_t z = x + 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
inline uint64_t key_mix(const uint64_t* w, int nw) {
  uint64_t m = mix64(w[0]);
  if (nw == 2) m = mix64(m ^ w[1]);
  return m;
}

}  // namespace

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
  // the two remaining hash flavours of the benchmark matrix (kmer_hash.hpp:157-232), constructed the way KeyToRank does
  // (prefix bits = ceilLog2(comm_size)): kind 0 identity store, 1 identity dist, 2 cpp_std store, 3 cpp_std dist
  static void hash_prefix(int kind, unsigned pbits, const uint64_t* in, uint64_t* out, size_t n) {
    bliss::kmer::hash::identity<Kmer, false> is;  bliss::kmer::hash::identity<Kmer, true> id(pbits);
    bliss::kmer::hash::cpp_std<Kmer, false> ss;   bliss::kmer::hash::cpp_std<Kmer, true> sd(pbits);
    for (size_t i = 0; i < n; ++i) {
      Kmer km; std::memcpy(km.getDataRef(), in + i * NW, NW * 8);
      out[i] = kind == 0 ? is(km) : kind == 1 ? id(km) : kind == 2 ? ss(km) : sd(km);
    }
  }
  // f4: the reference's k-mer quality score (QualityScoreGenerationIterator over Illumina18QualityScoreCodec<float>,
  // src/index/quality_score_iterator.hpp, quality_scores.hpp:529) of every K-window of one quality line
  static size_t quality(const uint8_t* q, size_t n, float* out) {
    if (n < K) return 0;
    typedef const unsigned char* It;
    bliss::index::QualityScoreGenerationIterator<It, K, bliss::index::Illumina18QualityScoreCodec<float> > it(q);
    const size_t m = n - K + 1;
    for (size_t i = 0; i < m; ++i, ++it) out[i] = *it;
    return m;
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
  // Source::operator()(r, f) calls f(chars, n) for every EOL-free sequence of rank r's block.
  // Returns distinct keys; if out_keys is non-null the (key,count) pairs are written unsorted, capacity cap;
  // per_rank_ck (optional) receives per rank [sum of mix(key) * count, sum of counts] as kmi_checksum defines them.
  template <typename Source>
  static size_t count_build_core(const Source& source, int p, int hash_kind /*0 murmur,1 farm*/, uint64_t* out_keys,
                                 uint32_t* out_counts, size_t cap, uint64_t* per_rank_sizes, uint64_t* per_rank_ck) {
    if (p < 1) p = 1;
    std::vector<std::vector<std::vector<Kmer>>> send(p, std::vector<std::vector<Kmer>>(p));
    auto produce = [&](int r) {
      bliss::kmer::transform::lex_less<Kmer> canon;
      bliss::kmer::hash::murmur<Kmer, true> md; bliss::kmer::hash::farm<Kmer, true> fd;
      source(r, [&](const uint8_t* s, size_t n) {
        if (n < K) return;
        Kmer km;
        for (size_t j = 0; j < n; ++j) {
          km.nextFromChar(DNA::FROM_ASCII[s[j]]);
          if (j + 1 >= K) {
            Kmer c = canon(km);
            uint64_t h = hash_kind == 0 ? md(c) : fd(c);
            send[r][h % (uint64_t)p].push_back(c);
          }
        }
      });
    };
    std::vector<Map*> maps(p, nullptr);
    auto consume = [&](int r) {
      bliss::kmer::hash::sparsehash::special_keys<Kmer, true> sk;
      Map* m = new Map(128, StoreHash{hash_kind, {}, {}});
      m->set_empty_key(sk.generate(0)); m->set_deleted_key(sk.generate(1));
      m->max_load_factor(0.7); m->min_load_factor(0.3);
      for (int src = 0; src < p; ++src) {
        for (const Kmer& k : send[src][r]) {
          auto res = m->insert(std::make_pair(k, (uint32_t)1));
          if (!res.second) res.first->second += 1;
        }
        std::vector<Kmer>().swap(send[src][r]);   // handed over: release
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
      if (per_rank_ck) {
        uint64_t a = 0, b = 0;
        for (auto it = maps[r]->begin(); it != maps[r]->end(); ++it) {
          a += key_mix(it->first.getData(), NW) * (uint64_t)it->second;
          b += it->second;
        }
        per_rank_ck[2 * r] = a; per_rank_ck[2 * r + 1] = b;
      }
      if (out_keys)
        for (auto it = maps[r]->begin(); it != maps[r]->end() && w < cap; ++it, ++w) {
          std::memcpy(out_keys + w * NW, it->first.getData(), NW * 8);
          out_counts[w] = it->second;
        }
      delete maps[r];
    }
    return total;
  }
  // seqs: concatenated EOL-free sequences, off[i]..off[i+1]; rank r takes sequences [nseq*r/p, nseq*(r+1)/p)
  static size_t count_build(const uint8_t* seqs, const uint64_t* off, size_t nseq, int p, int hash_kind, uint64_t* out_keys,
                            uint32_t* out_counts, size_t cap, uint64_t* per_rank_sizes) {
    if (p < 1) p = 1;
    auto source = [&](int r, const std::function<void(const uint8_t*, size_t)>& f) {
      size_t b = nseq * (size_t)r / p, e = nseq * (size_t)(r + 1) / p;
      for (size_t i = b; i < e; ++i) f(seqs + off[i], off[i + 1] - off[i]);
    };
    return count_build_core(source, p, hash_kind, out_keys, out_counts, cap, per_rank_sizes, nullptr);
  }
  // FASTQ text in, as the reference's ranks see it: rank r reads the byte range [n*r/p, n*(r+1)/p) moved forward to its
  // first whole record (partitioned_file + FASTQParser::find_first_record, file.hpp:1342-1431) and walks its records.
  // `cuts` (optional, p + 1 entries) overrides the equal split (rank r's block = [cuts[r], cuts[r+1]), record-aligned).
  // Returns distinct keys, or (size_t)-1 on a malformed record.
  static size_t count_build_fastq(const uint8_t* text, size_t n, int p, int hash_kind, const uint64_t* cuts, uint64_t* out_keys,
                                  uint32_t* out_counts, size_t cap, uint64_t* per_rank_sizes, uint64_t* per_rank_ck) {
    if (p < 1) p = 1;
    std::vector<size_t> cut(p + 1);
    for (int r = 0; r <= p; ++r) cut[r] = cuts ? (size_t)cuts[r] : (r == 0 ? 0 : r == p ? n : first_record(text, n, n * (size_t)r / p));
    for (int r = 1; r <= p; ++r) cut[r] = std::max(cut[r], cut[r - 1]);
    std::vector<int> bad(p, 0);
    auto source = [&](int r, const std::function<void(const uint8_t*, size_t)>& f) {
      if (!walk_fastq(text, cut[r], cut[r + 1], [&](size_t s, size_t len) { f(text + s, len); })) bad[r] = 1;
    };
    size_t d = count_build_core(source, p, hash_kind, out_keys, out_counts, cap, per_rank_sizes, per_rank_ck);
    for (int r = 0; r < p; ++r) if (bad[r]) return (size_t)-1;
    return d;
  }
  // What every OWNER rank (of p_owners) must end up holding from this text, without building tables: per owner
  // [sum of mix(key), number of k-mers] over all k-mer occurrences whose KeyToRank is that owner. Both are additive over
  // texts, so ranks can each walk their own block and sum; sum of mix(key) over occurrences == sum of mix(key) * count
  // over distinct keys (kmi_checksum). `threads` host threads split the text at record boundaries.
  static int owner_checksums_fastq(const uint8_t* text, size_t n, int threads, int p_owners, int hash_kind, uint64_t* out /* 2 * p_owners */) {
    if (threads < 1) threads = 1;
    std::vector<size_t> cut(threads + 1);
    for (int r = 0; r <= threads; ++r) cut[r] = r == 0 ? 0 : r == threads ? n : first_record(text, n, n * (size_t)r / threads);
    for (int r = 1; r <= threads; ++r) cut[r] = std::max(cut[r], cut[r - 1]);
    std::vector<std::vector<uint64_t>> part(threads, std::vector<uint64_t>(2 * (size_t)p_owners, 0));
    std::vector<int> bad(threads, 0);
    auto work = [&](int r) {
      bliss::kmer::transform::lex_less<Kmer> canon;
      bliss::kmer::hash::murmur<Kmer, true> md; bliss::kmer::hash::farm<Kmer, true> fd;
      uint64_t* acc = part[r].data();
      bool ok = walk_fastq(text, cut[r], cut[r + 1], [&](size_t s0, size_t len) {
        if (len < K) return;
        const uint8_t* s = text + s0;
        Kmer km;
        for (size_t j = 0; j < len; ++j) {
          km.nextFromChar(DNA::FROM_ASCII[s[j]]);
          if (j + 1 >= K) {
            Kmer c = canon(km);
            const uint64_t h = hash_kind == 0 ? md(c) : fd(c);
            const size_t o = (size_t)(h % (uint64_t)p_owners);
            acc[2 * o] += key_mix(c.getData(), NW);
            acc[2 * o + 1] += 1;
          }
        }
      });
      if (!ok) bad[r] = 1;
    };
    std::vector<std::thread> th;
    for (int r = 1; r < threads; ++r) th.emplace_back(work, r);
    work(0);
    for (auto& t : th) t.join();
    for (int o = 0; o < 2 * p_owners; ++o) { out[o] = 0; for (int r = 0; r < threads; ++r) out[o] += part[r][o]; }
    for (int r = 0; r < threads; ++r) if (bad[r]) return -1;
    return 0;
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
void kmiref_hash_prefix(int K, int kind, unsigned pbits, const uint64_t* in, uint64_t* out, size_t n) { DISPATCH(K, hash_prefix(kind, pbits, in, out, n), (void)0) }
size_t kmiref_quality(int K, const uint8_t* q, size_t n, float* out) { DISPATCH(K, quality(q, n, out), 0) }
void kmiref_quality_lut(float* out96) {
  for (int i = 0; i < 96; ++i) out96[i] = bliss::index::Illumina18QualityScoreCodec<float>::decode((unsigned char)(33 + i));
}
void kmiref_sentinels(int K, uint64_t* out) { DISPATCH(K, sentinels(out), (void)0) }
size_t kmiref_count_build(int K, const uint8_t* seqs, const uint64_t* off, size_t nseq, int p, int hash_kind,
                          uint64_t* out_keys, uint32_t* out_counts, size_t cap, uint64_t* per_rank_sizes) {
  DISPATCH(K, count_build(seqs, off, nseq, p, hash_kind, out_keys, out_counts, cap, per_rank_sizes), 0)
}
size_t kmiref_count_build_fastq(int K, const uint8_t* text, size_t n, int p, int hash_kind, const uint64_t* cuts, uint64_t* out_keys,
                                uint32_t* out_counts, size_t cap, uint64_t* per_rank_sizes, uint64_t* per_rank_ck) {
  DISPATCH(K, count_build_fastq(text, n, p, hash_kind, cuts, out_keys, out_counts, cap, per_rank_sizes, per_rank_ck), 0)
}
int kmiref_owner_checksums_fastq(int K, const uint8_t* text, size_t n, int threads, int p_owners, int hash_kind, uint64_t* out) {
  DISPATCH(K, owner_checksums_fastq(text, n, threads, p_owners, hash_kind, out), -2)
}

}  // extern "C"
