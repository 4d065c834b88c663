// kmi_index.hpp — header-only C++11 adapter over the C ABI (kmi.h).
//
// Two faces, both thin forwarding (no computation on the host):
//  1. kmi::b200_map<Key, T, Mode>: a MapType with the member set bliss::index::kmer::Index touches
//     (reference src/index/kmer_index.hpp:107-110,115,132-194,201-225,377-392), so that
//       using MapType   = kmi::b200_map<KmerType, uint32_t, KMI_MODE_COUNT>;
//       using IndexType = bliss::index::kmer::CountIndex<MapType>;
//     compiles in the reference's driver (test/benchmark/BenchmarkKmerIndex.cpp:262-306) with Key =
//     bliss::common::Kmer<K, DNA, uint64_t> passed through by reinterpret_cast (same layout:
//     nWords little-endian uint64_t words, src/common/kmer.hpp:177). See INTEGRATION.md.
//  2. kmi::CountIndex<K> / kmi::PositionIndex<K>: the same thing with a self-contained kmi::Kmer<K>,
//     for callers that do not include the reference's headers.
//
// Differences from the reference that a caller can observe (documented, deliberate):
//  - inputs are NOT mutated (the reference canonicalises and permutes them in place,
//    src/containers/distributed_map_base.hpp:286-301);
//  - find()/count() results are in input order (the reference returns them grouped by answering
//    rank, src/containers/distributed_densehash_map.hpp:1179-1194);
//  - predicates (find / count / erase with a functor) run on the host over the entries the device returns; the
//    common count-threshold predicates have device-side forms (find_if_count / count_if_count / erase_if_count).
// Errors: KMI_ERR_ARG -> std::invalid_argument, KMI_ERR_FORMAT -> std::logic_error (as
// fastq_loader.hpp:393,419,435 throws), anything else -> std::runtime_error.
#ifndef KMI_INDEX_HPP_
#define KMI_INDEX_HPP_

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "kmi.h"

namespace kmi {

/// Record-aligned block boundaries {0, c1, ..., n} of a FASTQ file held in memory, for `parts` ranks / GPUs: equal byte
/// ranges, each moved forward to its first whole record — what partitioned_file<...,FASTQParser>::read_file does with
/// BlockPartitioner + FASTQParser::find_first_record (reference src/io/file.hpp:1342-1431, fastq_loader.hpp:269-357).
inline std::vector<size_t> fastq_cuts(const uint8_t* bytes, size_t n, int parts) {
  std::vector<size_t> cuts(1, 0);
  for (int r = 1; r < parts; ++r) {
    uint64_t at = 0;
    if (kmi_fastq_find_first_record(bytes, 0, n, (uint64_t)n * r / parts, n, 0, n, &at) != KMI_OK)
      throw std::logic_error(kmi_last_error());
    cuts.push_back(std::max<size_t>(cuts.back(), (size_t)at));
  }
  cuts.push_back(n);
  return cuts;
}

/// What every block of a partitioned FASTA file inherits from the bytes before it, from the per-block summaries of
/// kmi_fasta_block_summary_dev (blocks in file order; summaries[r] = 4 words, bol[r] = block r starts a line): an exclusive
/// scan carrying (header runs so far, kind of the last line start) — the cross-rank part of FASTAParser::init_parser
/// (reference src/io/fasta_loader.hpp:202-483). Fills header_runs_before and flags of blocks[r]; file_offset and n_valid are
/// the caller's.
inline void fasta_chain(const uint32_t (*summaries)[4], const bool* bol, int parts, bool file_starts_with_header,
                        kmi_fasta_block* blocks) {
  uint32_t runs = 0, kind = 0;
  for (int r = 0; r < parts; ++r) {
    blocks[r].header_runs_before = runs;
    blocks[r].flags = (bol[r] ? KMI_FASTA_BOL : 0u) | (kind ? KMI_FASTA_PREV_HEADER : 0u) |
                      (file_starts_with_header ? KMI_FASTA_FILE_STARTS_WITH_HEADER : 0u);
    if (summaries[r][0]) {
      runs += summaries[r][3] - ((summaries[r][1] && kind) ? 1u : 0u);
      kind = summaries[r][2];
    }
  }
}

/// bliss::common::Kmer<K, DNA, uint64_t> look-alike: same size, same word order.
template <unsigned K_>
struct Kmer {
  static constexpr unsigned size = K_;
  static constexpr unsigned nWords = (K_ <= 32) ? 1 : 2;
  static constexpr unsigned nBits = 2 * K_;
  uint64_t data[nWords];
  Kmer() { for (unsigned i = 0; i < nWords; ++i) data[i] = 0; }
  const uint64_t* getData() const { return data; }
  uint64_t (&getDataRef())[nWords] { return data; }
  /// Kmer::nextFromChar (kmer.hpp:731-741) for a 2-bit code
  void nextFromChar(uint8_t code) {
    if (nWords == 1) {
      data[0] = ((data[0] << 2) | (code & 3)) & (K_ == 32 ? ~0ULL : ((1ULL << (2 * (K_ % 32))) - 1));
    } else {
      data[nWords - 1] = ((data[nWords - 1] << 2) | (data[0] >> 62)) & ((1ULL << (2 * (K_ - 32))) - 1);
      data[0] = (data[0] << 2) | (code & 3);
    }
  }
  /// 2-bit code of base i (0 = first base = most significant used bits)
  uint8_t base(unsigned i) const {
    const unsigned bit = 2 * (K_ - 1 - i);
    return (uint8_t)((data[bit >> 6] >> (bit & 63)) & 3);
  }
  /// Kmer::reverse_complement (kmer.hpp:1118-1123): DNA complement = 3 - code
  Kmer reverse_complement() const {
    Kmer r;
    for (unsigned i = 0; i < K_; ++i) r.nextFromChar((uint8_t)(3 - base(K_ - 1 - i)));
    return r;
  }
  /// lex_less (kmer_transform.hpp:109-126): the smaller of the k-mer and its reverse complement
  Kmer canonical() const { Kmer r = reverse_complement(); return r < *this ? r : *this; }
  bool operator==(const Kmer& o) const { return std::memcmp(data, o.data, sizeof(data)) == 0; }
  bool operator<(const Kmer& o) const {
    for (int i = (int)nWords - 1; i >= 0; --i) if (data[i] != o.data[i]) return data[i] < o.data[i];
    return false;
  }
};

/// lex_less on the word array of a 2-bit DNA k-mer (kmer_transform.hpp:109-126; reverse_complement kmer.hpp:1118-1123),
/// for ANY key type with the reference's layout — this is what lets b200_map take bliss::common::Kmer<K, DNA, uint64_t>
/// (whose `data` is private) as well as kmi::Kmer<K>. w: nw words, word 0 least significant.
inline void canonical_words(uint64_t* w, unsigned k) {
  const unsigned nw = k <= 32 ? 1 : 2;
  auto rev2 = [](uint64_t x) {   // reverse the 32 two-bit groups of a word
    x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
    x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
    x = ((x >> 8) & 0x00FF00FF00FF00FFULL) | ((x & 0x00FF00FF00FF00FFULL) << 8);
    x = ((x >> 16) & 0x0000FFFF0000FFFFULL) | ((x & 0x0000FFFF0000FFFFULL) << 16);
    return (x >> 32) | (x << 32);
  };
  uint64_t rc[2] = {0, 0};
  if (nw == 1) {
    rc[0] = (~rev2(w[0])) >> (64 - 2 * k);
  } else {
    const unsigned s = 128 - 2 * k;   // pad bits, 2..62
    const uint64_t hi = ~rev2(w[0]), lo = ~rev2(w[1]);
    rc[0] = (lo >> s) | (hi << (64 - s));
    rc[1] = hi >> s;
  }
  bool rc_less = false;
  for (int i = (int)nw - 1; i >= 0; --i)
    if (rc[i] != w[i]) { rc_less = rc[i] < w[i]; break; }
  if (rc_less) for (unsigned i = 0; i < nw; ++i) w[i] = rc[i];
}

/// what the adapter needs from mxx::comm (rank / size); one process drives one GPU shard
struct comm {
  int rank_, size_;
  void (*bcast_)(void* buf, int n, void* arg);   ///< rank 0's bytes to every rank (size() > 1 only)
  void* bcast_arg_;
  explicit comm(int rank = 0, int size = 1, void (*bcast)(void*, int, void*) = nullptr, void* arg = nullptr)
      : rank_(rank), size_(size), bcast_(bcast), bcast_arg_(arg) {}
  int rank() const { return rank_; }
  int size() const { return size_; }
};

/// How a communicator of the caller's world carries NCCL's unique id from rank 0 to every rank when a map is created over
/// more than one rank (kmi_comm_unique_id / kmi_comm_create). Specialise it for your communicator type; for the reference's
/// mxx::comm that is one line:
///   template <> struct kmi::comm_bootstrap<mxx::comm> {
///     static void bcast(const mxx::comm& c, void* buf, int n) { MPI_Bcast(buf, n, MPI_BYTE, 0, c); } };
/// kmi::comm carries a std::function for it (tests/cpp/dist_check.cpp hands the bytes through shared memory).
template <typename Comm> struct comm_bootstrap {
  static void bcast(const Comm&, void*, int) {
    throw std::logic_error("kmi::b200_map over more than one rank: specialise kmi::comm_bootstrap<Comm>::bcast (see INTEGRATION.md)");
  }
};

template <> struct comm_bootstrap<comm> {
  static void bcast(const comm& c, void* buf, int n) {
    if (!c.bcast_) throw std::logic_error("kmi::comm of more than one rank needs a bcast function");
    c.bcast_(buf, n, c.bcast_arg_);
  }
};

inline void check(int status) {
  if (status == KMI_OK) return;
  std::string msg = kmi_last_error();
  if (status == KMI_ERR_ARG) throw std::invalid_argument(msg);
  if (status == KMI_ERR_FORMAT) throw std::logic_error(msg);
  throw std::runtime_error(msg);
}

/// Predicates on the stored value that the library evaluates on the device (kmi_count_if / kmi_dump_if / kmi_erase_if): pass
/// one to find / erase / count_if of a COUNT map and only the matching entries cross PCIe. They are ordinary functors too,
/// so the generic (host-side) predicate overloads accept them as well.
struct value_pred {
  int op;
  uint64_t x;
  template <typename V> bool operator()(const V& v) const {
    const uint64_t c = (uint64_t)v.second;
    return op == KMI_PRED_GE ? c >= x : op == KMI_PRED_LE ? c <= x : c == x;
  }
};
inline value_pred count_at_least(uint64_t c) { value_pred p = {KMI_PRED_GE, c}; return p; }
inline value_pred count_at_most(uint64_t c) { value_pred p = {KMI_PRED_LE, c}; return p; }
inline value_pred count_equal(uint64_t c) { value_pred p = {KMI_PRED_EQ, c}; return p; }

struct map_options {
  int hash = KMI_HASH_MURMUR;   // reference default MurmurHash (kmer_index.hpp:437-438); benchmarks use FARM
  bool canonical = true;        // CanonicalHashMapParams (kmer_index.hpp:453-466)
  bool nsplit = false;          // NSplitSequencesIterator instead of SequencesIterator
  int device = 0;
  uint64_t reserve = 0;
  int dist_trans = 0;           // non-canonical maps: DistTrans under the dist hash, 0 identity / 1 lex_less / 2 xor_rev_comp
                                // (SingleStrandHashMapParams<Key, DistHash, StoreHash, DistTrans>, BenchmarkKmerIndex.cpp:152-161)
  uint64_t bulk_load_keys = 0;  // COUNT maps: KMI_PARAM_DEFER_INSERT staging budget (0 = every build inserts at once)
};

template <typename Key, typename T, int Mode>
class b200_map {
  static_assert(sizeof(Key) == 8 || sizeof(Key) == 16, "Key must be 1 or 2 uint64_t words");

 public:
  using key_type = Key;
  using mapped_type = T;
  using value_type = std::pair<Key, T>;
  using const_iterator = typename std::vector<value_type>::const_iterator;
  static constexpr unsigned K = Key::size;

  /// Comm: anything with rank() and size() — mxx::comm (Index::Index(const mxx::comm&), kmer_index.hpp:115) or kmi::comm
  template <typename Comm>
  explicit b200_map(const Comm& c, const map_options& o = map_options()) : comm_(c.rank(), c.size()), canonical_(o.canonical) {
    uint8_t id[KMI_COMM_ID_BYTES];
    if (c.size() > 1) {                                   // collective: every rank constructs its map
      std::memset(id, 0, sizeof(id));
      if (c.rank() == 0) kmi_comm_unique_id(id);           // (a failure leaves zeros: kmi_comm_create then fails on every rank alike)
      comm_bootstrap<Comm>::bcast(c, id, (int)sizeof(id));
    }
    create(o);
    if (c.size() > 1) {
      int r = kmi_comm_create(id, c.size(), c.rank(), o.device, &kc_);
      if (r == KMI_OK) r = kmi_dist_attach(h_, kc_);
      if (r != KMI_OK) { kmi_destroy(h_); h_ = nullptr; if (kc_) kmi_comm_destroy(kc_); kc_ = nullptr; check(r); }
    }
  }
  explicit b200_map(const map_options& o = map_options()) : comm_(), canonical_(o.canonical) { create(o); }
  ~b200_map() { kmi_destroy(h_); if (kc_) kmi_comm_destroy(kc_); }
  b200_map(const b200_map&) = delete;
  b200_map& operator=(const b200_map&) = delete;

  /// counting_densehash_map::insert(std::vector<Key>&) — distributed_densehash_map.hpp:2714-2899
  size_t insert(std::vector<Key>& keys) {
    check((kc_ ? kmi_dist_insert : kmi_insert)(h_, reinterpret_cast<const uint64_t*>(keys.data()), nullptr, keys.size()));
    return keys.size();
  }
  /// insert(std::vector<std::pair<Key,T>>&): COUNT adds the tuple's count to the key's entry (reduction insert with
  /// std::plus, distributed_densehash_map.hpp:2463-2482; KmerCountTupleParser emits count 1, kmer_parser.hpp:941,966);
  /// POSITION stores every (key, id) — :2329-2384
  size_t insert(std::vector<value_type>& tuples) {
    std::vector<uint64_t> k(tuples.size() * nw()), v(tuples.size());
    for (size_t i = 0; i < tuples.size(); ++i) {
      std::memcpy(&k[i * nw()], tuples[i].first.getData(), sizeof(Key));
      uint64_t id = 0;
      std::memcpy(&id, &tuples[i].second, sizeof(T) < 8 ? sizeof(T) : 8);
      v[i] = id;
    }
    check((kc_ ? kmi_dist_insert : kmi_insert)(h_, k.data(), v.data(), tuples.size()));
    return tuples.size();
  }
  /// find: COUNT -> (key, frequency) for hits; POSITION -> every (key, id) of every hit — :327-448. As in the reference the
  /// query vector is transformed in place (canonical keys, distributed_map_base.hpp:286-289) and results carry those keys.
  std::vector<value_type> find(std::vector<Key>& keys) const {
    std::vector<uint64_t> off, vals;
    lookup(keys, off, vals);
    std::vector<value_type> out;
    out.reserve(vals.size());
    for (size_t i = 0; i < keys.size(); ++i)
      for (uint64_t j = off[i]; j < off[i + 1]; ++j) out.emplace_back(keys[i], from_u64(vals[j]));
    return out;
  }
  /// find(keys, sorted_input, pred) — kmer_index.hpp:159: hits whose (key, value) satisfies the predicate
  template <typename Pred>
  std::vector<value_type> find(std::vector<Key>& keys, bool /*sorted_input*/, const Pred& pred) const {
    std::vector<value_type> out = find(keys);
    out.erase(std::remove_if(out.begin(), out.end(), [&](const value_type& v) { return !pred(v); }), out.end());
    return out;
  }
  /// find(pred) — kmer_index.hpp:171: every local entry that satisfies the predicate
  template <typename Pred>
  std::vector<value_type> find(const Pred& pred) const {
    std::vector<value_type> out = to_vector();
    out.erase(std::remove_if(out.begin(), out.end(), [&](const value_type& v) { return !pred(v); }), out.end());
    return out;
  }
  /// find(pred) with a device-side value predicate: only the matching entries are read back
  std::vector<value_type> find(const value_pred& pred) const {
    if (Mode != KMI_MODE_COUNT) return find<value_pred>(pred);
    uint64_t n = 0;
    check(kmi_count_if(h_, pred.op, pred.x, &n));
    std::vector<uint64_t> k(n * nw() + 1), v(n + 1);
    size_t got = 0;
    check(kmi_dump_if(h_, pred.op, pred.x, k.data(), v.data(), n, &got));
    std::vector<value_type> out;
    out.reserve(got);
    for (size_t i = 0; i < got; ++i) {
      Key key;
      std::memcpy(key.getDataRef(), &k[i * nw()], sizeof(Key));
      out.emplace_back(key, from_u64(v[i]));
    }
    return out;
  }
  /// number of local entries that satisfy a device-side value predicate (no entry leaves the GPU)
  size_t count_if(const value_pred& pred) const {
    uint64_t n = 0;
    check(kmi_count_if(h_, pred.op, pred.x, &n));
    return (size_t)n;
  }
  /// erase(pred) with a device-side value predicate: one sweep of the table
  size_t erase(const value_pred& pred) {
    if (Mode != KMI_MODE_COUNT) return erase<value_pred>(pred);
    uint64_t n = 0;
    check(kmi_erase_if(h_, pred.op, pred.x, &n));
    return (size_t)n;
  }
  /// count: one (key, size_t) per query, zeros included — LocalCount, :234-266
  std::vector<std::pair<Key, size_t>> count(std::vector<Key>& keys) const {
    transform_input(keys);
    std::vector<uint64_t> cnt(keys.size());
    check((kc_ ? kmi_dist_count : kmi_count)(h_, reinterpret_cast<const uint64_t*>(keys.data()), keys.size(), cnt.data()));
    std::vector<std::pair<Key, size_t>> out;
    out.reserve(keys.size());
    for (size_t i = 0; i < keys.size(); ++i) out.emplace_back(keys[i], (size_t)cnt[i]);
    return out;
  }
  /// count(keys, sorted_input, pred) — kmer_index.hpp:177: per query, how many of its entries satisfy the predicate
  template <typename Pred>
  std::vector<std::pair<Key, size_t>> count(std::vector<Key>& keys, bool /*sorted_input*/, const Pred& pred) const {
    std::vector<uint64_t> off, vals;
    lookup(keys, off, vals);
    std::vector<std::pair<Key, size_t>> out;
    out.reserve(keys.size());
    for (size_t i = 0; i < keys.size(); ++i) {
      size_t c = 0;
      for (uint64_t j = off[i]; j < off[i + 1]; ++j) c += pred(value_type(keys[i], from_u64(vals[j]))) ? 1 : 0;
      out.emplace_back(keys[i], c);
    }
    return out;
  }
  /// count(pred) — kmer_index.hpp:182: for every local key, how many of its entries satisfy the predicate
  template <typename Pred>
  std::vector<std::pair<Key, size_t>> count(const Pred& pred) const {
    std::vector<value_type> all = to_vector();
    std::sort(all.begin(), all.end(), [](const value_type& a, const value_type& b) { return a.first < b.first; });
    std::vector<std::pair<Key, size_t>> out;
    for (size_t i = 0; i < all.size(); ++i) {
      if (i == 0 || !(all[i].first == all[i - 1].first)) out.emplace_back(all[i].first, 0);
      out.back().second += pred(all[i]) ? 1 : 0;
    }
    return out;
  }
  /// erase(keys): every entry of the keys — :1602-1663
  size_t erase(std::vector<Key>& keys) {
    transform_input(keys);
    uint64_t before = local_size();
    check((kc_ ? kmi_dist_erase : kmi_erase)(h_, reinterpret_cast<const uint64_t*>(keys.data()), keys.size()));
    return (size_t)(before - local_size());
  }
  /// erase(keys, sorted_input, pred) — kmer_index.hpp:188: the entries of the keys that satisfy the predicate (for the
  /// multimap, single values leave a key: densehash_map.hpp:2267-2328). The key's entries are read, the key is erased,
  /// and the survivors go back in.
  template <typename Pred>
  size_t erase(std::vector<Key>& keys, bool /*sorted_input*/, const Pred& pred) {
    transform_input(keys);
    std::sort(keys.begin(), keys.end());
    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
    std::vector<uint64_t> off, vals;
    lookup(keys, off, vals);
    std::vector<Key> gone;
    std::vector<value_type> back;
    size_t removed = 0;
    for (size_t i = 0; i < keys.size(); ++i) {
      size_t hit = 0;
      for (uint64_t j = off[i]; j < off[i + 1]; ++j) hit += pred(value_type(keys[i], from_u64(vals[j]))) ? 1 : 0;
      if (!hit) continue;
      removed += hit;
      gone.push_back(keys[i]);
      for (uint64_t j = off[i]; j < off[i + 1]; ++j) {
        value_type v(keys[i], from_u64(vals[j]));
        if (!pred(v)) back.push_back(v);
      }
    }
    if (!gone.empty()) check(kmi_erase(h_, reinterpret_cast<const uint64_t*>(gone.data()), gone.size()));
    if (!back.empty()) insert(back);
    return removed;
  }
  /// erase(pred) — kmer_index.hpp:193: every local entry that satisfies the predicate
  template <typename Pred>
  size_t erase(const Pred& pred) {
    std::vector<value_type> all = to_vector();
    std::vector<Key> keys;
    keys.reserve(all.size());
    for (const value_type& v : all) keys.push_back(v.first);
    return erase(keys, false, pred);
  }
  void reserve(size_t n) { check(kmi_reserve(h_, n)); }
  void clear() { check(kmi_clear(h_)); }
  size_t local_size() const { uint64_t n = 0; check(kmi_local_size(h_, &n)); return (size_t)n; }
  /// collective (allreduce of the local sizes, distributed_map_base.hpp:243) when the map spans several ranks
  size_t size() const {
    if (!kc_) return local_size();
    uint64_t n = 0;
    check(kmi_dist_size(h_, &n));
    return (size_t)n;
  }
  bool local_empty() const { return local_size() == 0; }
  float get_multiplicity() const {
    if (Mode == KMI_MODE_COUNT) return 1.0f;
    uint64_t keys = 0;
    check(kmi_num_keys(h_, &keys));
    return keys ? (float)local_size() / (float)keys : 1.0f;
  }
  /// to_vector(): every entry, unordered
  void to_vector(std::vector<value_type>& out) const {
    size_t n = local_size(), got = 0;
    std::vector<uint64_t> k(n * nw() + 1), v(n + 1);
    check(kmi_dump(h_, k.data(), v.data(), n, &got));
    out.clear();
    out.reserve(got);
    for (size_t i = 0; i < got; ++i) {
      Key key;
      std::memcpy(key.getDataRef(), &k[i * nw()], sizeof(Key));
      out.emplace_back(key, from_u64(v[i]));
    }
  }
  std::vector<value_type> to_vector() const { std::vector<value_type> o; to_vector(o); return o; }
  /// cbegin()/cend() (kmer_index.hpp:379-383) over a snapshot
  const_iterator cbegin() const { to_vector(snapshot_); return snapshot_.cbegin(); }
  const_iterator cend() const { return snapshot_.cend(); }

  /// Index::build_posix<FASTQParser|FASTAParser>(filename, comm) — kmer_index.hpp:231-260,332-372. A map that spans several
  /// ranks is built collectively: every rank passes the same file and takes its block of it (partitioned_file::read_file,
  /// file.hpp:1342-1431 -> kmi_dist_build_file); returns the k-mers generated from this rank's share
  uint64_t build_file(const std::string& filename) {
    uint64_t kmers = 0;
    check((kc_ ? kmi_dist_build_file : kmi_build_file)(h_, filename.c_str(), 0, &kmers));   // wrong extension -> std::invalid_argument, as the reference
    return kmers;
  }
  /// this rank's record-aligned block of the reads (kmi::fastq_cuts); sharded maps exchange the k-mers by owner
  uint64_t build_fastq(const uint8_t* bytes, size_t n, uint64_t file_offset = 0) {
    uint64_t kmers = 0;
    check((kc_ ? kmi_dist_build_fastq : kmi_build_fastq)(h_, bytes, n, file_offset, &kmers));
    return kmers;
  }
  uint64_t build_fasta(const uint8_t* bytes, size_t n) {
    uint64_t kmers = 0;
    check(kmi_build_fasta(h_, bytes, n, &kmers));
    return kmers;
  }
  /// this shard's entries to / from a file (kmi_save / kmi_load; load adds to what the map holds)
  void save(const std::string& filename) const { check(kmi_save(h_, filename.c_str())); }
  size_t load(const std::string& filename) { uint64_t n = 0; check(kmi_load(h_, filename.c_str(), &n)); return (size_t)n; }
  kmi_index* handle() const { return h_; }
  const comm& get_comm() const { return comm_; }

 private:
  static constexpr size_t nw() { return sizeof(Key) / 8; }
  /// distributed_map_base::transform_input (distributed_map_base.hpp:286-289): queries become canonical in place
  void transform_input(std::vector<Key>& keys) const {
    if (canonical_) for (Key& k : keys) canonical_words(k.getDataRef(), K);
  }
  void create(const map_options& o) {
    kmi_config cfg;
    cfg.k = (int32_t)K; cfg.mode = Mode; cfg.canonical = o.canonical ? 1 : 0; cfg.hash = o.hash; cfg.nsplit = o.nsplit ? 1 : 0;
    cfg.device = o.device; cfg.rank = comm_.rank(); cfg.nranks = comm_.size(); cfg.reserve = o.reserve;
    check(kmi_create(&cfg, &h_));
    if (o.dist_trans) check(kmi_set_param(h_, KMI_PARAM_DIST_TRANS, o.dist_trans));
    if (o.bulk_load_keys && Mode == KMI_MODE_COUNT) check(kmi_set_param(h_, KMI_PARAM_DEFER_INSERT, (int64_t)o.bulk_load_keys));
  }
  /// per query key i its values vals[off[i] .. off[i+1]): COUNT 0 or 1 value (the frequency), POSITION every id
  void lookup(std::vector<Key>& keys, std::vector<uint64_t>& off, std::vector<uint64_t>& vals) const {
    transform_input(keys);
    const uint64_t* kp = reinterpret_cast<const uint64_t*>(keys.data());
    off.assign(keys.size() + 1, 0);
    if (Mode == KMI_MODE_COUNT) {
      std::vector<uint64_t> v(keys.size());
      std::vector<uint8_t> found(keys.size());
      check((kc_ ? kmi_dist_find : kmi_find)(h_, kp, keys.size(), v.data(), found.data()));
      vals.clear();
      for (size_t i = 0; i < keys.size(); ++i) {
        if (found[i]) vals.push_back(v[i]);
        off[i + 1] = vals.size();
      }
    } else {
      std::vector<uint64_t> cnt(keys.size());
      check((kc_ ? kmi_dist_count : kmi_count)(h_, kp, keys.size(), cnt.data()));
      uint64_t total = 0;
      for (uint64_t c : cnt) total += c;
      vals.assign(total ? total : 1, 0);
      check((kc_ ? kmi_dist_find_positions : kmi_find_positions)(h_, kp, keys.size(), off.data(), vals.data(), total));
      vals.resize(total);
    }
  }
  static T from_u64(uint64_t v) { T t; std::memset(&t, 0, sizeof(T)); std::memcpy(&t, &v, sizeof(T) < 8 ? sizeof(T) : 8); return t; }
  comm comm_;
  kmi_comm* kc_ = nullptr;      // the map spans comm_.size() > 1 ranks: batch operations are the collective kmi_dist_* calls
  bool canonical_ = true;
  kmi_index* h_ = nullptr;
  mutable std::vector<value_type> snapshot_;
};

/// The sort-based MapType: dsc::counting_sorted_map (Mode COUNT) / dsc::sorted_multimap (Mode POSITION) of
/// reference src/containers/distributed_sorted_map.hpp — what BenchmarkKmerIndex.cpp:206-219 picks for pMAP == SORTED — over
/// kmi_sorted_* (device radix sort + run-length reduction, binary-search queries). Same member shapes as b200_map for what
/// Index<> forwards (insert / find / count / size / local_size / to_vector / reserve); ONE rank only, as the C layer.
template <typename Key, typename T, int Mode>
class b200_sorted_map {
  static_assert(sizeof(Key) == 8 || sizeof(Key) == 16, "Key must be 1 or 2 uint64_t words");

 public:
  using key_type = Key;
  using mapped_type = T;
  using value_type = std::pair<Key, T>;
  using const_iterator = typename std::vector<value_type>::const_iterator;
  static constexpr unsigned K = Key::size;

  template <typename Comm>
  explicit b200_sorted_map(const Comm& c, const map_options& o = map_options()) : canonical_(o.canonical) {
    if (c.size() != 1) throw std::invalid_argument("b200_sorted_map spans one rank");
    create(o);
  }
  explicit b200_sorted_map(const map_options& o = map_options()) : canonical_(o.canonical) { create(o); }
  ~b200_sorted_map() { kmi_sorted_destroy(h_); }
  b200_sorted_map(const b200_sorted_map&) = delete;
  b200_sorted_map& operator=(const b200_sorted_map&) = delete;

  /// sorted_map_base::insert (distributed_sorted_map.hpp:1186-1300): appended; sorted and reduced by the next rehash / query
  size_t insert(std::vector<Key>& keys) {
    if (Mode != KMI_MODE_COUNT) throw std::invalid_argument("a sorted multimap inserts (key, id) tuples");
    check(kmi_sorted_insert(h_, reinterpret_cast<const uint64_t*>(keys.data()), nullptr, keys.size()));
    return keys.size();
  }
  size_t insert(std::vector<value_type>& tuples) {
    std::vector<uint64_t> k(tuples.size() * nw()), v(tuples.size());
    for (size_t i = 0; i < tuples.size(); ++i) {
      std::memcpy(&k[i * nw()], tuples[i].first.getData(), sizeof(Key));
      uint64_t id = 0;
      std::memcpy(&id, &tuples[i].second, sizeof(T) < 8 ? sizeof(T) : 8);
      v[i] = id;
    }
    check(kmi_sorted_insert(h_, k.data(), v.data(), tuples.size()));
    return tuples.size();
  }
  size_t build_fastq(const uint8_t* bytes, size_t n, uint64_t file_offset = 0) {
    uint64_t m = 0;
    check(kmi_sorted_build_fastq(h_, bytes, n, file_offset, &m));
    return (size_t)m;
  }
  size_t build_fasta(const uint8_t* bytes, size_t n) {
    uint64_t m = 0;
    check(kmi_sorted_build_fasta(h_, bytes, n, &m));
    return (size_t)m;
  }
  /// local_reduction + sort (distributed_sorted_map.hpp:1521-1540,1601-1700 at one rank)
  void rehash() { check(kmi_sorted_rehash(h_)); }
  void reserve(size_t) {}
  std::vector<value_type> find(std::vector<Key>& keys) const {
    transform_input(keys);
    const uint64_t* kp = reinterpret_cast<const uint64_t*>(keys.data());
    std::vector<value_type> out;
    if (Mode == KMI_MODE_COUNT) {
      std::vector<uint64_t> v(keys.size());
      std::vector<uint8_t> found(keys.size());
      check(kmi_sorted_find(h_, kp, keys.size(), v.data(), found.data()));
      for (size_t i = 0; i < keys.size(); ++i) if (found[i]) out.emplace_back(keys[i], from_u64(v[i]));
    } else {
      std::vector<uint64_t> cnt(keys.size()), off(keys.size() + 1, 0);
      check(kmi_sorted_count(h_, kp, keys.size(), cnt.data()));
      uint64_t total = 0;
      for (uint64_t c : cnt) total += c;
      std::vector<uint64_t> vals(total ? total : 1);
      check(kmi_sorted_find_positions(h_, kp, keys.size(), off.data(), vals.data(), total));
      for (size_t i = 0; i < keys.size(); ++i)
        for (uint64_t j = off[i]; j < off[i + 1]; ++j) out.emplace_back(keys[i], from_u64(vals[j]));
    }
    return out;
  }
  std::vector<std::pair<Key, size_t>> count(std::vector<Key>& keys) const {
    transform_input(keys);
    std::vector<uint64_t> cnt(keys.size());
    check(kmi_sorted_count(h_, reinterpret_cast<const uint64_t*>(keys.data()), keys.size(), cnt.data()));
    std::vector<std::pair<Key, size_t>> out;
    out.reserve(keys.size());
    for (size_t i = 0; i < keys.size(); ++i) out.emplace_back(keys[i], (size_t)cnt[i]);
    return out;
  }
  size_t local_size() const {
    uint64_t n = 0;
    check(kmi_sorted_local_size(h_, &n));
    return (size_t)n;
  }
  size_t size() const { return local_size(); }
  /// the container itself: entries in key order (a sorted_map's vector after rehash)
  std::vector<value_type> to_vector() const {
    const size_t n = local_size();
    std::vector<uint64_t> k(n * nw() + 1), v(n + 1);
    size_t got = 0;
    check(kmi_sorted_dump(h_, k.data(), v.data(), n, &got));
    std::vector<value_type> out(got);
    for (size_t i = 0; i < got; ++i) {
      std::memcpy(out[i].first.getDataRef(), &k[i * nw()], sizeof(Key));
      out[i].second = from_u64(v[i]);
    }
    return out;
  }
  void to_vector(std::vector<value_type>& out) const { out = to_vector(); }
  kmi_sorted* handle() const { return h_; }

 private:
  static constexpr size_t nw() { return sizeof(Key) / 8; }
  void transform_input(std::vector<Key>& keys) const {
    if (canonical_) for (Key& k : keys) canonical_words(k.getDataRef(), K);
  }
  void create(const map_options& o) {
    kmi_config cfg;
    cfg.k = (int32_t)K; cfg.mode = Mode; cfg.canonical = o.canonical ? 1 : 0; cfg.hash = o.hash; cfg.nsplit = o.nsplit ? 1 : 0;
    cfg.device = o.device; cfg.rank = 0; cfg.nranks = 1; cfg.reserve = 0;
    check(kmi_sorted_create(&cfg, &h_));
  }
  static T from_u64(uint64_t v) { T t; std::memset(&t, 0, sizeof(T)); std::memcpy(&t, &v, sizeof(T) < 8 ? sizeof(T) : 8); return t; }
  bool canonical_ = true;
  kmi_sorted* h_ = nullptr;
};

/// PositionQualityIndex's map (kmer_index.hpp:406, BenchmarkKmerIndex.cpp:131-137): k-mer -> (position id,
/// quality score), built from FASTQ bytes by kmi_posqual_* (the scores are KmerPositionQualityTupleParser's,
/// src/io/kmer_parser.hpp:578-760). mapped_type = std::pair<uint64_t, float>, the reference's KmerInfoType with the id as its
/// raw 64-bit word. One rank.
template <typename Key>
class b200_posqual_map {
 public:
  using key_type = Key;
  using mapped_type = std::pair<uint64_t, float>;
  using value_type = std::pair<Key, mapped_type>;
  static constexpr unsigned K = Key::size;

  explicit b200_posqual_map(const map_options& o = map_options()) : canonical_(o.canonical) {
    kmi_config cfg;
    cfg.k = (int32_t)K; cfg.mode = KMI_MODE_POSITION; cfg.canonical = o.canonical ? 1 : 0; cfg.hash = o.hash; cfg.nsplit = o.nsplit ? 1 : 0;
    cfg.device = o.device; cfg.rank = 0; cfg.nranks = 1; cfg.reserve = o.reserve;
    check(kmi_posqual_create(&cfg, &h_));
  }
  ~b200_posqual_map() { kmi_posqual_destroy(h_); }
  b200_posqual_map(const b200_posqual_map&) = delete;
  b200_posqual_map& operator=(const b200_posqual_map&) = delete;

  size_t build_fastq(const uint8_t* bytes, size_t n, uint64_t file_offset = 0) {
    uint64_t m = 0;
    check(kmi_posqual_build_fastq(h_, bytes, n, file_offset, &m));
    return (size_t)m;
  }
  std::vector<value_type> find(std::vector<Key>& keys) const {
    if (canonical_) for (Key& k : keys) canonical_words(k.getDataRef(), K);
    const uint64_t* kp = reinterpret_cast<const uint64_t*>(keys.data());
    std::vector<uint64_t> cnt(keys.size()), off(keys.size() + 1, 0);
    check(kmi_posqual_count(h_, kp, keys.size(), cnt.data()));
    uint64_t total = 0;
    for (uint64_t c : cnt) total += c;
    std::vector<uint64_t> ids(total ? total : 1);
    std::vector<float> q(total ? total : 1);
    check(kmi_posqual_find(h_, kp, keys.size(), off.data(), ids.data(), q.data(), total));
    std::vector<value_type> out;
    out.reserve(total);
    for (size_t i = 0; i < keys.size(); ++i)
      for (uint64_t j = off[i]; j < off[i + 1]; ++j) out.emplace_back(keys[i], mapped_type(ids[j], q[j]));
    return out;
  }
  size_t local_size() const {
    uint64_t n = 0;
    check(kmi_posqual_local_size(h_, &n));
    return (size_t)n;
  }
  size_t size() const { return local_size(); }
  std::vector<value_type> to_vector() const {
    const size_t n = local_size();
    std::vector<uint64_t> k(n * (sizeof(Key) / 8) + 1), ids(n + 1);
    std::vector<float> q(n + 1);
    size_t got = 0;
    check(kmi_posqual_dump(h_, k.data(), ids.data(), q.data(), n, &got));
    std::vector<value_type> out(got);
    for (size_t i = 0; i < got; ++i) {
      std::memcpy(out[i].first.getDataRef(), &k[i * (sizeof(Key) / 8)], sizeof(Key));
      out[i].second = mapped_type(ids[i], q[i]);
    }
    return out;
  }
  kmi_posqual* handle() const { return h_; }

 private:
  bool canonical_ = true;
  kmi_posqual* h_ = nullptr;
};

/// standalone aliases mirroring bliss::index::kmer::CountIndex / PositionIndex (kmer_index.hpp:402-409)
template <unsigned K> using CountIndex = b200_map<Kmer<K>, uint32_t, KMI_MODE_COUNT>;
template <unsigned K> using PositionIndex = b200_map<Kmer<K>, uint64_t, KMI_MODE_POSITION>;

}  // namespace kmi
#endif  // KMI_INDEX_HPP_
