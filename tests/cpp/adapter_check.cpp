// Drives the C++ adapter (include/kmi_index.hpp) the way the reference's container tests drive its maps
// (src/containers/test/test_densehash_map.cpp:100-762: every operation checked against a std:: container), for the member
// set Index<MapType,KmerParser> forwards to (kmer_index.hpp:134-194): insert / find / count / erase, each also with a
// predicate, and to_vector / size. Needs a GPU; exits 0 when every check holds.
#include <cstdio>
#include <map>
#include <random>
#include <string>
#include <vector>

#include "kmi_index.hpp"

// -DKMI_CHECK_BLISS: the key type is the REFERENCE's bliss::common::Kmer<K, DNA, uint64_t> (compiled against
// /root/reference/src where it lies; oracle/Makefile builds that flavour into oracle/_ref/adapter_check_bliss), otherwise the
// self-contained kmi::Kmer<K>. Everything below is written against the members both have.
#ifdef KMI_CHECK_BLISS
#include "common/alphabets.hpp"
#include "common/kmer.hpp"
#include "common/kmer_transform.hpp"
template <unsigned K> using KeyT = bliss::common::Kmer<K, bliss::common::DNA, uint64_t>;
template <typename Key> static Key canon(const Key& k) { return bliss::kmer::transform::lex_less<Key>()(k); }
struct ref_like_comm { int rank() const { return 0; } int size() const { return 1; } };   // stands in for mxx::comm (rank()/size())
#define KEY_FLAVOUR "bliss::common::Kmer"
#else
template <unsigned K> using KeyT = kmi::Kmer<K>;
template <typename Key> static Key canon(const Key& k) { return k.canonical(); }
typedef kmi::comm ref_like_comm;
#define KEY_FLAVOUR "kmi::Kmer"
#endif

#define CHECK(c)                                                     \
  do {                                                               \
    if (!(c)) {                                                      \
      std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #c);       \
      return 1;                                                      \
    }                                                                \
  } while (0)

template <unsigned K>
static KeyT<K> random_kmer(std::mt19937_64& rng) {
  KeyT<K> k;
  for (unsigned i = 0; i < K; ++i) k.nextFromChar((uint8_t)(rng() & 3));
  return k;
}

template <typename Pair>
static void sort_pairs(std::vector<Pair>& v) {
  std::sort(v.begin(), v.end(), [](const Pair& a, const Pair& b) { return a.first < b.first || (a.first == b.first && a.second < b.second); });
}

template <unsigned K>
static int count_checks() {
  using Key = KeyT<K>;
  using Map = kmi::b200_map<Key, uint32_t, KMI_MODE_COUNT>;
  using V = typename Map::value_type;
  std::mt19937_64 rng(K);
  std::vector<Key> pool;
  for (int i = 0; i < 3000; ++i) pool.push_back(random_kmer<K>(rng));
  std::vector<Key> input;
  for (int i = 0; i < 20000; ++i) {
    Key k = pool[(rng() % 3000) * (rng() % 3000) / 3000];   // skewed multiplicities
    input.push_back((rng() & 1) ? k : k.reverse_complement());
  }
  std::map<Key, uint32_t> want;
  for (const Key& k : input) want[canon(k)]++;
  kmi::map_options o;
  o.hash = KMI_HASH_FARM;
  Map idx(ref_like_comm(), o);
  idx.insert(input);
  CHECK(idx.local_size() == want.size() && idx.size() == want.size());
  CHECK(idx.get_multiplicity() == 1.0f);
  std::vector<V> all = idx.to_vector();
  sort_pairs(all);
  CHECK(all.size() == want.size());
  {
    size_t i = 0;
    for (const auto& kv : want) { CHECK(all[i].first == kv.first && all[i].second == kv.second); ++i; }
  }
  // (key, count) tuples add their count (reduction insert), also onto existing entries and for new keys
  {
    std::vector<V> tuples;
    for (int i = 0; i < 1500; ++i) {
      Key k = (i % 3) ? pool[rng() % 3000] : random_kmer<K>(rng);
      uint32_t c = (uint32_t)(rng() % 1000);
      tuples.emplace_back(k.reverse_complement(), c);
      want[canon(k)] += c;
    }
    tuples.emplace_back(pool[0], 0xFFFFFFF0u);       // std::plus<uint32_t> wraps
    want[canon(pool[0])] += 0xFFFFFFF0u;
    idx.insert(tuples);
    CHECK(idx.local_size() == want.size());
    all = idx.to_vector();
    sort_pairs(all);
    size_t i = 0;
    for (const auto& kv : want) { CHECK(all[i].first == kv.first && all[i].second == kv.second); ++i; }
  }
  // queries: every pool key in a random orientation + keys that are (almost surely) absent
  std::vector<Key> q;
  for (const Key& k : pool) q.push_back((rng() & 1) ? k : k.reverse_complement());
  for (int i = 0; i < 500; ++i) q.push_back(random_kmer<K>(rng));
  {
    std::vector<Key> qq = q;
    std::vector<V> got = idx.find(qq);
    size_t hits = 0;
    for (const Key& k : q) hits += want.count(canon(k));
    CHECK(got.size() == hits);
    for (const V& v : got) CHECK(want.count(v.first) && want[v.first] == v.second);   // canonical keys come back
    qq = q;
    auto cnt = idx.count(qq);
    CHECK(cnt.size() == q.size());
    for (size_t i = 0; i < q.size(); ++i) CHECK(cnt[i].first == canon(q[i]) && cnt[i].second == want.count(canon(q[i])));
  }
  auto often = [](const V& v) { return v.second >= 3; };
  {
    std::vector<Key> qq = q;
    std::vector<V> got = idx.find(qq, false, often);
    size_t hits = 0;
    for (const Key& k : q) { auto it = want.find(canon(k)); hits += (it != want.end() && it->second >= 3) ? 1 : 0; }
    CHECK(got.size() == hits);
    for (const V& v : got) CHECK(v.second >= 3 && want[v.first] == v.second);
    std::vector<V> every = idx.find(often);
    size_t n3 = 0;
    for (const auto& kv : want) n3 += kv.second >= 3 ? 1 : 0;
    CHECK(every.size() == n3);
    qq = q;
    auto cnt = idx.count(qq, false, often);
    for (size_t i = 0; i < q.size(); ++i) {
      auto it = want.find(canon(q[i]));
      CHECK(cnt[i].second == ((it != want.end() && it->second >= 3) ? 1u : 0u));
    }
    auto per_key = idx.count(often);
    CHECK(per_key.size() == want.size());
    size_t s = 0;
    for (const auto& kc : per_key) s += kc.second;
    CHECK(s == n3);
  }
  // the same predicates evaluated on the device (kmi::count_at_least & co.): only matching entries are read back
  {
    std::vector<V> dev = idx.find(kmi::count_at_least(3));
    std::vector<V> host = idx.find(often);
    sort_pairs(dev); sort_pairs(host);
    CHECK(dev.size() == host.size());
    for (size_t i = 0; i < dev.size(); ++i) CHECK(dev[i].first == host[i].first && dev[i].second == host[i].second);
    CHECK(idx.count_if(kmi::count_at_least(3)) == dev.size());
    size_t ones = 0, le2 = 0;
    for (const auto& kv : want) { ones += kv.second == 1 ? 1 : 0; le2 += kv.second <= 2 ? 1 : 0; }
    CHECK(idx.count_if(kmi::count_equal(1)) == ones && idx.count_if(kmi::count_at_most(2)) == le2);
  }
  // erase with a predicate on half of the queries, then by predicate over the whole map, then plain
  {
    std::vector<Key> half(q.begin(), q.begin() + q.size() / 2);
    size_t expect = 0;
    for (const Key& k : half) {
      auto it = want.find(canon(k));
      if (it != want.end() && it->second >= 3) { want.erase(it); ++expect; }
    }
    CHECK(idx.erase(half, false, often) == expect);
    CHECK(idx.local_size() == want.size());
    auto single = [](const V& v) { return v.second == 1; };
    expect = 0;
    for (auto it = want.begin(); it != want.end();) {
      if (it->second == 1) { it = want.erase(it); ++expect; } else ++it;
    }
    CHECK(idx.erase(kmi::count_equal(1)) == expect);      // device-side sweep
    CHECK(idx.erase(single) == 0);                        // nothing left for the host-side form
    CHECK(idx.local_size() == want.size());
    std::vector<Key> some(q.begin() + q.size() / 2, q.begin() + q.size() / 2 + 300);
    expect = 0;
    for (const Key& k : some) expect += want.erase(canon(k));
    CHECK(idx.erase(some) == expect);
    all = idx.to_vector();
    sort_pairs(all);
    CHECK(all.size() == want.size());
    size_t i = 0;
    for (const auto& kv : want) { CHECK(all[i].first == kv.first && all[i].second == kv.second); ++i; }
  }
  return 0;
}

template <unsigned K>
static int position_checks() {
  using Key = KeyT<K>;
  using Map = kmi::b200_map<Key, uint64_t, KMI_MODE_POSITION>;
  using V = typename Map::value_type;
  std::mt19937_64 rng(100 + K);
  std::vector<Key> pool;
  for (int i = 0; i < 2000; ++i) pool.push_back(random_kmer<K>(rng));
  std::vector<V> tuples;
  for (int i = 0; i < 12000; ++i) {
    Key k = pool[(rng() % 2000) * (rng() % 2000) / 2000];
    tuples.emplace_back((rng() & 1) ? k : k.reverse_complement(), (uint64_t)(rng() % 64));   // few ids: (key, id) repeats happen
  }
  std::vector<V> want;
  for (const V& t : tuples) want.emplace_back(canon(t.first), t.second);
  sort_pairs(want);
  Map idx;
  idx.insert(tuples);
  CHECK(idx.local_size() == want.size());       // duplicates are all kept (densehash_map.hpp:2190-2249)
  std::vector<V> all = idx.to_vector();
  sort_pairs(all);
  CHECK(all == want);
  std::map<Key, std::vector<uint64_t>> by_key;
  for (const V& t : want) by_key[t.first].push_back(t.second);
  CHECK(idx.get_multiplicity() == (float)want.size() / (float)by_key.size());
  std::vector<Key> q;
  for (int i = 0; i < 500; ++i) q.push_back(pool[rng() % 2000].reverse_complement());
  for (int i = 0; i < 100; ++i) q.push_back(random_kmer<K>(rng));
  {
    std::vector<Key> qq = q;
    std::vector<V> got = idx.find(qq), expect;
    for (const Key& k : q) {
      auto it = by_key.find(canon(k));
      if (it != by_key.end()) for (uint64_t v : it->second) expect.emplace_back(it->first, v);
    }
    sort_pairs(got); sort_pairs(expect);
    CHECK(got == expect);
    qq = q;
    auto cnt = idx.count(qq);
    for (size_t i = 0; i < q.size(); ++i) {
      auto it = by_key.find(canon(q[i]));
      CHECK(cnt[i].second == (it == by_key.end() ? 0 : it->second.size()));
    }
    auto odd = [](const V& v) { return (v.second & 1) != 0; };
    qq = q;
    auto codd = idx.count(qq, false, odd);
    for (size_t i = 0; i < q.size(); ++i) {
      auto it = by_key.find(canon(q[i]));
      size_t c = 0;
      if (it != by_key.end()) for (uint64_t v : it->second) c += v & 1;
      CHECK(codd[i].second == c);
    }
    // erase the odd ids of the queried keys: single values leave their key
    size_t removed = 0;
    std::vector<Key> uq;
    for (const Key& k : q) uq.push_back(canon(k));
    std::sort(uq.begin(), uq.end());
    uq.erase(std::unique(uq.begin(), uq.end()), uq.end());
    for (const Key& k : uq) {
      auto it = by_key.find(k);
      if (it == by_key.end()) continue;
      std::vector<uint64_t> keep;
      for (uint64_t v : it->second) { if (v & 1) ++removed; else keep.push_back(v); }
      if (keep.empty()) by_key.erase(it); else it->second.swap(keep);
    }
    qq = q;
    CHECK(idx.erase(qq, false, odd) == removed);
    want.clear();
    for (const auto& kv : by_key) for (uint64_t v : kv.second) want.emplace_back(kv.first, v);
    sort_pairs(want);
    all = idx.to_vector();
    sort_pairs(all);
    CHECK(all == want);
    // then every id below 8 anywhere in the map
    auto low = [](const V& v) { return v.second < 8; };
    removed = 0;
    for (const V& t : want) removed += t.second < 8 ? 1 : 0;
    CHECK(idx.erase(low) == removed);
    want.erase(std::remove_if(want.begin(), want.end(), low), want.end());
    all = idx.to_vector();
    sort_pairs(all);
    CHECK(all == want);
    CHECK(idx.find(low).empty());
  }
  return 0;
}

// the sort-based MapType (dsc::counting_sorted_map / sorted_multimap's member shapes) against std::map, and the
// PositionQualityIndex map: ids equal the position map's, scores in (0, 1] for 'I' qualities
template <unsigned K>
static int sorted_checks() {
  using Key = KeyT<K>;
  using Map = kmi::b200_sorted_map<Key, uint32_t, KMI_MODE_COUNT>;
  using V = typename Map::value_type;
  std::mt19937_64 rng(100 + K);
  std::vector<Key> pool;
  for (int i = 0; i < 2000; ++i) pool.push_back(random_kmer<K>(rng));
  std::vector<Key> input;
  std::map<Key, uint32_t> want;
  for (int i = 0; i < 15000; ++i) {
    Key k = pool[(rng() % 2000) * (rng() % 2000) / 2000];
    if (rng() & 1) k = k.reverse_complement();
    input.push_back(k);
    ++want[canon(k)];
  }
  Map m{ref_like_comm()};
  std::vector<Key> in2 = input;
  CHECK(m.insert(in2) == input.size());
  CHECK(m.size() == want.size());
  std::vector<V> all = m.to_vector();                      // in key order, as the reference's sorted vector
  CHECK(all.size() == want.size());
  size_t i = 0;
  for (const auto& kv : want) { CHECK(all[i].first == kv.first && all[i].second == kv.second); ++i; }
  std::vector<Key> q(pool.begin(), pool.begin() + 300);
  for (int j = 0; j < 50; ++j) q.push_back(random_kmer<K>(rng));
  std::vector<Key> qq = q;
  auto cnt = m.count(qq);
  qq = q;
  auto found = m.find(qq);
  size_t hits = 0;
  for (size_t j = 0; j < q.size(); ++j) {
    auto it = want.find(canon(q[j]));
    CHECK(cnt[j].second == (it != want.end() ? 1u : 0u));
    hits += it != want.end();
  }
  CHECK(found.size() == hits);
  for (const V& f : found) CHECK(want.at(f.first) == f.second);
  std::vector<V> more;
  more.emplace_back(all[0].first, 7u);
  m.insert(more);
  all = m.to_vector();
  CHECK(all[0].second == want.begin()->second + 7u);
  return 0;
}

static int posqual_checks() {
  using Key = KeyT<21>;
  std::mt19937_64 rng(5);
  std::string fq;
  const char* acgt = "ACGT";
  for (int r = 0; r < 200; ++r) {
    std::string seq;
    for (int b = 0; b < 100; ++b) seq += acgt[rng() & 3];
    fq += "@r" + std::to_string(r) + "\n" + seq + "\n+\n" + std::string(100, 'I') + "\n";
  }
  kmi::b200_posqual_map<Key> pq;
  kmi::b200_map<Key, uint64_t, KMI_MODE_POSITION> pos{ref_like_comm()};
  const size_t n = pq.build_fastq(reinterpret_cast<const uint8_t*>(fq.data()), fq.size());
  CHECK(n == 200u * 80u && pq.size() == n);
  CHECK(pos.build_fastq(reinterpret_cast<const uint8_t*>(fq.data()), fq.size()) == n);
  auto a = pq.to_vector();
  auto b = pos.to_vector();
  std::vector<std::pair<Key, uint64_t>> ai;
  for (const auto& t : a) {
    ai.emplace_back(t.first, t.second.first);
    CHECK(t.second.second > 0.99f && t.second.second <= 1.0f);   // 21 bases of Phred 40: (1 - 1e-4)^21
  }
  sort_pairs(ai);
  sort_pairs(b);
  CHECK(ai == b);
  std::vector<Key> q;
  for (size_t j = 0; j < a.size(); j += 97) q.push_back(a[j].first.reverse_complement());
  auto f = pq.find(q);
  CHECK(f.size() >= q.size());
  return 0;
}

int main() {
  if (sorted_checks<31>()) return 1;
  if (sorted_checks<47>()) return 1;
  if (posqual_checks()) return 1;
  if (count_checks<21>()) return 1;
  if (count_checks<31>()) return 1;
  if (count_checks<63>()) return 1;
  if (position_checks<21>()) return 1;
  if (position_checks<35>()) return 1;
  std::printf("adapter ok (%s)\n", KEY_FLAVOUR);
  return 0;
}
