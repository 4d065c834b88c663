// TEST INFRASTRUCTURE — oracle tier T1: the REFERENCE's real file reader, record parsers, k-mer parsers and
// distributed maps, compiled from where they lie under /root/reference against the single-rank mpi.h of oracle/mpishim,
// run at p = 1. Writes binary dumps that tests/golden/make_golden_t1.py turns into committed golden vectors:
//   ref_t1 fastq_pos  K file  ->  (k-mer words, ShortSequenceKmerId) of KmerPositionTupleParser over FASTQParser, file order
//   ref_t1 fasta_pos  K file  ->  (k-mer words, LongSequenceKmerId) over FASTAParser
//   ref_t1 count      K file  ->  sorted (canonical k-mer, count) of CountIndex<counting_densehash_map, lex_less, farm> built by
//                                  Index::build_posix / insert — the reference's whole path, end to end
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

int main(int argc, char** argv) {
  mxx::env e(argc, argv);
  mxx::comm comm;
  if (argc < 5) { std::fprintf(stderr, "usage: ref_t1 fastq_pos|fasta_pos|count K file out\n"); return 1; }
  const std::string what = argv[1], file = argv[3];
  const int K = std::atoi(argv[2]);
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
