// TEST INFRASTRUCTURE — oracle tier T1: the REFERENCE's real file reader, record parsers, k-mer parsers and
// distributed maps, compiled from where they lie under /root/reference against the single-rank mpi.h of oracle/mpishim,
// run at p = 1. Writes binary dumps that tests/golden/make_golden_t1.py turns into committed golden vectors:
//   ref_t1 fastq_pos  K file  ->  (k-mer words, ShortSequenceKmerId) of KmerPositionTupleParser over FASTQParser, file order
//   ref_t1 fasta_pos  K file  ->  (k-mer words, LongSequenceKmerId) over FASTAParser
//   ref_t1 count      K file  ->  sorted (canonical k-mer, count) of CountIndex<counting_densehash_map, lex_less, farm> built by
//                                  Index::build_posix / insert — the reference's whole path, end to end
//   ref_t1 first_record 0 file ->  FASTQParser::find_first_record over a grid of search ranges + the p-rank block boundaries
//   ref_t1 adapter    K file  ->  the reference's Index<> template instantiated with kmi::b200_map as MapType (compile-time
//                                  proof of the drop-in; needs a GPU to run)
// Nothing here is copied: it only includes and drives reference headers.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>

#include "bliss-config.hpp"
#include <mxx/env.hpp>
#include <mxx/comm.hpp>
#include "common/alphabets.hpp"
#include "common/kmer.hpp"
#include "common/kmer_transform.hpp"
#include "index/kmer_index.hpp"
#include "index/kmer_hash.hpp"
#include "io/kmer_file_helper.hpp"
#include "io/kmer_parser.hpp"
#include "io/fastq_loader.hpp"
#include "io/fasta_loader.hpp"
#include "io/sequence_iterator.hpp"
#include "containers/distributed_densehash_map.hpp"
#ifdef KMI_T1_ADAPTER
#include "kmi_index.hpp"
#endif

using bliss::common::DNA;
template <unsigned K> using KM = bliss::common::Kmer<K, DNA, uint64_t>;

static void put(FILE* f, const void* p, size_t n) { if (fwrite(p, 1, n, f) != n) { perror("fwrite"); exit(2); } }

template <unsigned K, typename IdType, template <typename> class SeqParser>
static int dump_positions(const std::string& file, const char* out, const mxx::comm& comm) {
  using Kmer = KM<K>;
  using Tuple = std::pair<Kmer, IdType>;
  std::vector<Tuple> tuples;
  ::bliss::io::KmerFileHelper::read_file_posix<::bliss::index::kmer::KmerPositionTupleParser<Tuple>, SeqParser, bliss::io::SequencesIterator>(file, tuples, comm);
  FILE* f = fopen(out, "wb");
  if (!f) return 2;
  const uint64_t n = tuples.size(), nw = Kmer::nWords;
  put(f, &n, 8); put(f, &nw, 8);
  for (const Tuple& t : tuples) { put(f, t.first.getData(), 8 * nw); static_assert(sizeof(IdType) == 8, "id is one word"); uint64_t id; std::memcpy(&id, &t.second, 8); put(f, &id, 8); /* the raw id word, as it travels */ }
  fclose(f);
  std::printf("%s: %llu tuples\n", out, (unsigned long long)n);
  return 0;
}

// the reference's CountIndex exactly as test/benchmark/BenchmarkKmerIndex.cpp:245-306 instantiates it for
// -DpKmerStore=CANONICAL -DpMAP=DENSEHASH -DpINDEX=COUNT -DpDistHash=FARM -DpStoreHash=FARM
template <unsigned K> struct RefCount {
  using Kmer = KM<K>;
  template <typename Key> using DistHash = ::bliss::index::kmer::DistHashFarm<Key>;
  template <typename Key> using StoreHash = ::bliss::index::kmer::StoreHashFarm<Key>;
  template <typename Key> using MapParams = ::bliss::index::kmer::CanonicalHashMapParams<Key, DistHash, StoreHash>;
  using SpecialKeys = ::bliss::kmer::hash::sparsehash::special_keys<Kmer, true>;
  using MapType = ::dsc::counting_densehash_map<Kmer, uint32_t, MapParams, SpecialKeys>;
  using IndexType = ::bliss::index::kmer::CountIndex<MapType>;
};

template <unsigned K, typename IndexType>
static int run_count(const std::string& file, const char* out, const mxx::comm& comm) {
  using Kmer = KM<K>;
  IndexType idx(comm);
  // Index::build_posix (kmer_index.hpp:231-260,332-372): read_file_posix + insert
  idx.template build_posix<bliss::io::FASTQParser, bliss::io::SequencesIterator>(file, comm);
  std::vector<std::pair<Kmer, uint32_t>> all;
  idx.get_map().to_vector(all);
  std::sort(all.begin(), all.end(), [](const std::pair<Kmer, uint32_t>& a, const std::pair<Kmer, uint32_t>& b) { return a.first < b.first; });
  FILE* f = fopen(out, "wb");
  if (!f) return 2;
  const uint64_t n = all.size(), nw = Kmer::nWords;
  put(f, &n, 8); put(f, &nw, 8);
  for (const auto& t : all) { put(f, t.first.getData(), 8 * nw); const uint64_t c = t.second; put(f, &c, 8); }
  fclose(f);
  // the batch queries of the same facade, on the index's own keys in the other orientation
  std::vector<Kmer> q;
  for (size_t i = 0; i < all.size() && i < 500; ++i) q.push_back(all[i].first.reverse_complement());
  auto found = idx.find(q);
  auto counted = idx.count(q);
  size_t hits = 0;
  for (const auto& c : counted) hits += c.second;
  std::printf("%s: %llu distinct k-mers, size() %llu, find %zu / count %zu of %zu queries\n", out, (unsigned long long)n,
              (unsigned long long)idx.size(), found.size(), hits, q.size());
  return (found.size() == q.size() && hits == q.size() && idx.size() == n) ? 0 : 3;
}

#ifdef KMI_T1_ADAPTER
// the same Index<> template of the reference, the map type swapped for the B200 one: this is the drop-in
template <unsigned K> struct B200Count {
  using MapType = kmi::b200_map<KM<K>, uint32_t, KMI_MODE_COUNT>;
  using IndexType = ::bliss::index::kmer::CountIndex<MapType>;
};
#endif

// the reference's FASTQParser::find_first_record (fastq_loader.hpp:269-357) on its own, over the byte ranges a partitioned read
// hands it: for every cut c of a grid over the file, the search range [c, c + window) clipped to the file, the file being the
// parent and in memory whole. One row per call: c, end of the search range, result (or ~0 when the reference throws: no FASTQ
// markers in a range that does not reach the parent's end). Also the block boundaries of p ranks as partitioned_file<..,
// FASTQParser>::read_file computes them (file.hpp:1342-1431): find_first_record over [size * r / p, size).
static int dump_first_record(const std::string& file, const char* out) {
  std::vector<unsigned char> data;
  {
    FILE* f = fopen(file.c_str(), "rb");
    if (!f) return 2;
    unsigned char buf[65536];
    size_t got;
    while ((got = fread(buf, 1, sizeof(buf), f)) > 0) data.insert(data.end(), buf, buf + got);
    fclose(f);
  }
  typedef bliss::partition::range<size_t> Range;
  const size_t n = data.size();
  bliss::io::FASTQParser<unsigned char*> parser;
  std::vector<uint64_t> rows;
  auto call = [&](size_t b, size_t e) {
    uint64_t r;
    try { r = parser.find_first_record(data.data(), Range(0, n), Range(0, n), Range(b, e)); }
    catch (const std::exception&) { r = ~0ULL; }
    rows.push_back(b); rows.push_back(e); rows.push_back(r);
  };
  const size_t step = std::max<size_t>(1, n / 400);
  const size_t windows[] = {n, 64, 300, 1000, 5000};
  for (size_t w : windows)
    for (size_t c = 0; c <= n; c += step) call(c, std::min(n, c + w));
  const int parts[] = {2, 3, 8, 64};
  for (int p : parts)
    for (int r = 0; r <= p; ++r) call((size_t)((unsigned __int128)n * r / p), n);
  FILE* f = fopen(out, "wb");
  if (!f) return 2;
  const uint64_t nr = rows.size() / 3, nw = 2;   // same container as the other dumps: nr rows of (nw + 1) words
  put(f, &nr, 8); put(f, &nw, 8);
  put(f, rows.data(), rows.size() * 8);
  fclose(f);
  std::printf("%s: %llu calls of find_first_record\n", out, (unsigned long long)nr);
  return 0;
}

int main(int argc, char** argv) {
  mxx::env e(argc, argv);
  mxx::comm comm;
  if (argc < 5) { std::fprintf(stderr, "usage: ref_t1 fastq_pos|fasta_pos|count K file out\n"); return 1; }
  const std::string what = argv[1], file = argv[3];
  const int K = std::atoi(argv[2]);
  if (what == "first_record") return dump_first_record(file, argv[4]);
  if (what == "fastq_pos") {
    using Id = bliss::common::ShortSequenceKmerId;
    if (K == 21) return dump_positions<21, Id, bliss::io::FASTQParser>(file, argv[4], comm);
    if (K == 31) return dump_positions<31, Id, bliss::io::FASTQParser>(file, argv[4], comm);
    if (K == 35) return dump_positions<35, Id, bliss::io::FASTQParser>(file, argv[4], comm);
  }
  if (what == "fasta_pos") {
    using Id = bliss::common::LongSequenceKmerId;
    if (K == 21) return dump_positions<21, Id, bliss::io::FASTAParser>(file, argv[4], comm);
    if (K == 35) return dump_positions<35, Id, bliss::io::FASTAParser>(file, argv[4], comm);
  }
  if (what == "count") {
    if (K == 21) return run_count<21, RefCount<21>::IndexType>(file, argv[4], comm);
    if (K == 31) return run_count<31, RefCount<31>::IndexType>(file, argv[4], comm);
    if (K == 35) return run_count<35, RefCount<35>::IndexType>(file, argv[4], comm);
  }
#ifdef KMI_T1_ADAPTER
  if (what == "adapter") {
    if (K == 31) return run_count<31, B200Count<31>::IndexType>(file, argv[4], comm);
    if (K == 35) return run_count<35, B200Count<35>::IndexType>(file, argv[4], comm);
  }
#endif
  std::fprintf(stderr, "unsupported %s K=%d\n", what.c_str(), K);
  return 1;
}
