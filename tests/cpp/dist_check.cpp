// dist_check — the sharded index driven from C++ with no Python in the loop: one process per GPU (forked here, before any
// CUDA call), NCCL id handed from rank 0 to the others through shared memory, then the collective API of include/kmi.h:
//   kmi_comm_create / kmi_dist_attach
//   kmi_dist_build_fastq   rank r builds from ITS record-aligned block of one synthetic FASTQ file (kmi::fastq_cuts)
//   kmi_dist_size, kmi_local_size, kmi_dump                 each shard == the oracle's k-mers owned by that rank, counts equal
//   kmi_dist_find / kmi_dist_count / kmi_dist_erase         answers in input order, also for keys no rank holds
//   kmi_dist_build_fasta + kmi_dist_find_positions          PositionIndex, rank r = block r of one FASTA file cut at any byte
// The expected values come from the CPU oracle (oracle/libkmi_oracle.so, TEST INFRASTRUCTURE), as the reference's own
// MPI tests compare against a sequential recomputation (src/io/test/mpi_test_distribute.cpp:431-559).
// usage: dist_check [nranks = visible GPUs] [n_reads = 60000] [K = 31 | 63] ; exits 0 when every rank passes.
#include <sys/mman.h>
#include <signal.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>

#include "kmi_index.hpp"

extern "C" {
size_t kmio_synth_record_bytes(uint32_t L);
void kmio_synth_fastq(uint8_t* out, uint64_t first_read, uint64_t n_reads, uint32_t L, uint64_t G, uint64_t seed, uint32_t err_ppm, uint32_t n_ppm);
size_t kmio_synth_fasta(uint8_t* out, uint32_t n_seq, uint64_t seq_len, uint32_t width, uint64_t seed);
long kmio_fastq_kmers(const uint8_t* buf, size_t n, uint64_t file_offset, int K, int canonical, int nsplit, uint64_t* keys, uint64_t* ids, size_t cap, uint64_t* n_reads);
long kmio_fasta_kmers(const uint8_t* buf, size_t n, int K, int canonical, int nsplit, uint64_t* keys, uint64_t* ids, size_t cap, uint64_t* n_seqs);
void kmio_hash(int K, int which, const uint64_t* in, uint64_t* out, size_t n);
}

struct Shared {
  std::atomic<int> id_ready;
  uint8_t id[KMI_COMM_ID_BYTES];
  std::atomic<int> box_ready;            // second hand-over: the adapter's own communicator (kmi::comm_bootstrap)
  uint8_t box[KMI_COMM_ID_BYTES];
};
struct BoxArg { Shared* sh; int rank; };
static void box_bcast(void* buf, int n, void* arg) {
  BoxArg* a = static_cast<BoxArg*>(arg);
  if (a->rank == 0) { std::memcpy(a->sh->box, buf, (size_t)n); a->sh->box_ready.store(1); }
  else { while (!a->sh->box_ready.load()) usleep(1000); std::memcpy(buf, a->sh->box, (size_t)n); }
}

#define CHECK(c)                                                                         \
  do {                                                                                   \
    if (!(c)) { std::printf("rank %d FAIL %s:%d: %s\n", rank, __FILE__, __LINE__, #c); return 1; } \
  } while (0)
#define KOK(call)                                                                                      \
  do {                                                                                                 \
    int r_ = (call);                                                                                   \
    if (r_ != KMI_OK) { std::printf("rank %d FAIL %s:%d: %s -> %d %s\n", rank, __FILE__, __LINE__, #call, r_, kmi_last_error()); return 1; } \
  } while (0)

typedef std::vector<uint64_t> Key;   // nw words, most significant LAST (word 0 = least significant, as in the ABI)
struct KeyLess {
  bool operator()(const Key& a, const Key& b) const {
    for (int i = (int)a.size() - 1; i >= 0; --i) if (a[i] != b[i]) return a[i] < b[i];
    return false;
  }
};

static int run_rank(int rank, int p, size_t n_reads, int K, Shared* sh) {
  const int nw = K <= 32 ? 1 : 2;
  // ---- the whole synthetic file, identical on every rank (the oracle needs all of it; a rank only BUILDS from its block)
  const size_t rec = kmio_synth_record_bytes(150);
  std::vector<uint8_t> file(n_reads * rec);
  kmio_synth_fastq(file.data(), 0, n_reads, 150, n_reads * 5, 21, 5000, 300);
  const std::vector<size_t> cuts = kmi::fastq_cuts(file.data(), file.size(), p);
  // ---- communicator
  if (rank == 0) {
    KOK(kmi_comm_unique_id(sh->id));
    sh->id_ready.store(1);
  } else {
    while (!sh->id_ready.load()) usleep(1000);
  }
  int ndev = 0;
  KOK(kmi_device_count(&ndev));
  const int device = rank % (ndev > 0 ? ndev : 1);
  kmi_comm* comm = nullptr;
  KOK(kmi_comm_create(sh->id, p, rank, device, &comm));
  CHECK(kmi_comm_rank(comm) == rank && kmi_comm_size(comm) == p);
  // ---- expected contents: every canonical k-mer of the file, by owner = farm dist hash % p
  std::vector<uint64_t> keys(file.size() / 2 * nw + 2);
  uint64_t nr = 0;
  const long total = kmio_fastq_kmers(file.data(), file.size(), 0, K, 1, 0, keys.data(), nullptr, file.size() / 2, &nr);
  CHECK(total > 0 && nr == n_reads);
  std::vector<uint64_t> dh((size_t)total);
  kmio_hash(K, 3, keys.data(), dh.data(), (size_t)total);
  std::map<Key, uint32_t, KeyLess> mine;
  for (long i = 0; i < total; ++i) {
    if ((int)(dh[i] % (uint64_t)p) != rank) continue;
    mine[Key(keys.begin() + i * nw, keys.begin() + (i + 1) * nw)]++;
  }
  // ---- sharded CountIndex from this rank's block (host bytes: H2D pipelined inside the library)
  kmi_config cfg;
  cfg.k = K; cfg.mode = KMI_MODE_COUNT; cfg.canonical = 1; cfg.hash = KMI_HASH_FARM; cfg.nsplit = 0; cfg.device = device;
  cfg.rank = rank; cfg.nranks = p; cfg.reserve = 0;
  kmi_index* idx = nullptr;
  KOK(kmi_create(&cfg, &idx));
  KOK(kmi_dist_attach(idx, comm));
  for (int round = 0; round < 2; ++round) {                  // twice: the receive buffers are reused
    KOK(kmi_clear(idx));
    uint64_t produced = 0;
    KOK(kmi_dist_build_fastq(idx, file.data() + cuts[rank], cuts[rank + 1] - cuts[rank], cuts[rank], &produced));
    uint64_t sum[2] = {produced, 0};
    KOK(kmi_comm_allreduce_u64(comm, sum, 1, 0));
    CHECK(sum[0] == (uint64_t)total);
    uint64_t local = 0, global = 0;
    KOK(kmi_local_size(idx, &local));
    KOK(kmi_dist_size(idx, &global));
    CHECK(local == mine.size());
    uint64_t s2[1] = {(uint64_t)mine.size()};
    KOK(kmi_comm_allreduce_u64(comm, s2, 1, 0));
    CHECK(global == s2[0]);
    std::vector<uint64_t> dk(local * nw + 1), dv(local + 1);
    size_t got = 0;
    KOK(kmi_dump(idx, dk.data(), dv.data(), local, &got));
    CHECK(got == local);
    for (size_t i = 0; i < got; ++i) {
      auto it = mine.find(Key(dk.begin() + i * nw, dk.begin() + (i + 1) * nw));
      CHECK(it != mine.end() && it->second == dv[i]);
    }
  }
  int32_t path = 0;
  float ms[4];
  uint64_t n_recv = 0;
  KOK(kmi_dist_last_stats(idx, ms, &n_recv, &path));
  if (rank == 0) std::printf("count build: path %d (1 = NVLink push, 2 = NCCL all-to-all), %.2f ms, rank 0 received %llu keys\n", path, ms[0], (unsigned long long)n_recv);
  // ---- queries from every rank: its own block's first k-mers in either strand + keys nobody holds
  {
    std::vector<uint64_t> q;
    const size_t nq = std::min<size_t>(5000, (size_t)total);
    const size_t start = (size_t)total / p * rank;
    for (size_t i = 0; i < nq; ++i) for (int w = 0; w < nw; ++w) q.push_back(keys[((start + i * 7) % (size_t)total) * nw + w]);
    for (size_t i = 0; i < 1000; ++i) { q.push_back((0x1234567ULL * (i + 1) * 2654435761ULL) & ((1ULL << 40) - 1)); if (nw == 2) q.push_back(i & 0xFFFF); }
    const size_t n = q.size() / nw;
    std::vector<uint64_t> vals(n), cnt(n);
    std::vector<uint8_t> found(n);
    KOK(kmi_dist_find(idx, q.data(), n, vals.data(), found.data()));
    KOK(kmi_dist_count(idx, q.data(), n, cnt.data()));
    // expected from the whole file: count of the canonical key anywhere
    std::map<Key, uint32_t, KeyLess> all;
    for (long i = 0; i < total; ++i) all[Key(keys.begin() + i * nw, keys.begin() + (i + 1) * nw)]++;
    for (size_t i = 0; i < n; ++i) {
      Key k(q.begin() + i * nw, q.begin() + (i + 1) * nw);
      kmi::canonical_words(k.data(), (unsigned)K);
      auto it = all.find(k);
      const uint64_t want = it == all.end() ? 0 : it->second;
      CHECK(found[i] == (want ? 1 : 0) && vals[i] == want && cnt[i] == (want ? 1u : 0u));
    }
    // erase what rank 0 asked for (collective: the others pass nothing), then nobody finds those keys
    std::vector<uint64_t> eq;
    for (size_t i = 0; i < 1000; ++i) for (int w = 0; w < nw; ++w) eq.push_back(keys[((i * 7) % (size_t)total) * nw + w]);   // rank 0's first queries
    uint64_t before = 0, after = 0;
    KOK(kmi_dist_size(idx, &before));
    KOK(kmi_dist_erase(idx, rank == 0 ? eq.data() : nullptr, rank == 0 ? 1000 : 0));
    KOK(kmi_dist_size(idx, &after));
    uint64_t tag = 1000 + (uint64_t)rank;
    std::vector<uint64_t> ag(p);
    KOK(kmi_comm_allgather_u64(comm, &tag, 1, ag.data()));
    for (int r = 0; r < p; ++r) CHECK(ag[r] == 1000 + (uint64_t)r);
    // every rank probes the keys rank 0 erased (its query list is rebuilt deterministically here)
    std::vector<uint64_t> q0;
    for (size_t i = 0; i < 1000; ++i) for (int w = 0; w < nw; ++w) q0.push_back(keys[((i * 7) % (size_t)total) * nw + w]);
    std::vector<uint64_t> c0(1000);
    KOK(kmi_dist_count(idx, q0.data(), 1000, c0.data()));
    for (size_t i = 0; i < 1000; ++i) CHECK(c0[i] == 0);
    std::map<Key, int, KeyLess> erased;
    for (size_t i = 0; i < 1000; ++i) { Key k(q0.begin() + i * nw, q0.begin() + (i + 1) * nw); kmi::canonical_words(k.data(), (unsigned)K); erased[k] = 1; }
    CHECK(before - after == erased.size());
  }
  kmi_destroy(idx);
  // ---- PositionIndex: block r of one FASTA file, cut at arbitrary bytes; ids are global
  {
    const int KP = 21;
    const size_t fn = kmio_synth_fasta(nullptr, 5, 40000, 70, 9);
    std::vector<uint8_t> fa(fn);
    kmio_synth_fasta(fa.data(), 5, 40000, 70, 9);
    std::vector<uint64_t> fk(fn + 2), fi(fn + 2);
    uint64_t ns = 0;
    const long ft = kmio_fasta_kmers(fa.data(), fn, KP, 1, 0, fk.data(), fi.data(), fn, &ns);
    CHECK(ft > 0 && ns == 5);
    std::vector<uint64_t> fdh((size_t)ft);
    kmio_hash(KP, 1, fk.data(), fdh.data(), (size_t)ft);     // murmur dist hash
    std::multimap<uint64_t, uint64_t> want;
    for (long i = 0; i < ft; ++i) if ((int)(fdh[i] % (uint64_t)p) == rank) want.insert(std::make_pair(fk[i], fi[i]));
    kmi_config pc = cfg;
    pc.k = KP; pc.mode = KMI_MODE_POSITION; pc.hash = KMI_HASH_MURMUR;
    kmi_index* pidx = nullptr;
    KOK(kmi_create(&pc, &pidx));
    KOK(kmi_dist_attach(pidx, comm));
    const uint64_t b0 = fn * (uint64_t)rank / p + (rank ? 13 : 0), b1 = rank + 1 == p ? fn : fn * (uint64_t)(rank + 1) / p + 13;   // mid-line cuts
    uint64_t halo = 0;
    KOK(kmi_fasta_halo_end(fa.data(), 0, fn, b1, KP, &halo));
    kmi_fasta_part part;
    part.file_offset = b0; part.n_valid = b1 - b0; part.bol = (b0 == 0 || fa[b0 - 1] == '\n') ? 1 : 0; part.reserved = 0;
    uint64_t produced = 0;
    KOK(kmi_dist_build_fasta(pidx, fa.data() + b0, halo - b0, &part, &produced));
    uint64_t tot[1] = {produced};
    KOK(kmi_comm_allreduce_u64(comm, tot, 1, 0));
    CHECK(tot[0] == (uint64_t)ft);
    uint64_t local = 0;
    KOK(kmi_local_size(pidx, &local));
    CHECK(local == want.size());
    std::vector<uint64_t> dk(local + 1), dv(local + 1);
    size_t got = 0;
    KOK(kmi_dump(pidx, dk.data(), dv.data(), local, &got));
    std::vector<std::pair<uint64_t, uint64_t>> a, b(want.begin(), want.end());
    for (size_t i = 0; i < got; ++i) a.push_back(std::make_pair(dk[i], dv[i]));
    std::sort(a.begin(), a.end()); std::sort(b.begin(), b.end());
    CHECK(a == b);
    // find_positions from every rank on k-mers of the whole file + absent keys
    std::vector<uint64_t> q;
    for (size_t i = 0; i < 3000; ++i) q.push_back(fk[(i * 31 + rank * 17) % (size_t)ft]);
    for (size_t i = 0; i < 500; ++i) q.push_back((i * 0x9E3779B97F4A7C15ULL) & ((1ULL << 42) - 1));
    std::vector<uint64_t> cnt(q.size()), off(q.size() + 1);
    KOK(kmi_dist_count(pidx, q.data(), q.size(), cnt.data()));
    uint64_t cap = 0;
    for (uint64_t c : cnt) cap += c;
    std::vector<uint64_t> ids(cap + 1);
    KOK(kmi_dist_find_positions(pidx, q.data(), q.size(), off.data(), ids.data(), cap));
    std::multimap<uint64_t, uint64_t> allpos;
    for (long i = 0; i < ft; ++i) allpos.insert(std::make_pair(fk[i], fi[i]));
    for (size_t i = 0; i < q.size(); ++i) {
      uint64_t k = q[i];
      kmi::canonical_words(&k, KP);
      auto range = allpos.equal_range(k);
      std::vector<uint64_t> e, g(ids.begin() + off[i], ids.begin() + off[i + 1]);
      for (auto it = range.first; it != range.second; ++it) e.push_back(it->second);
      std::sort(e.begin(), e.end()); std::sort(g.begin(), g.end());
      CHECK(e == g && cnt[i] == e.size());
    }
    kmi_destroy(pidx);
  }
  // ---- the C++ adapter over all ranks: kmi::b200_map constructed from a p-rank communicator routes to the collective calls
  if (K <= 31) {
    BoxArg arg = {sh, rank};
    kmi::map_options o;
    o.hash = KMI_HASH_FARM; o.device = device;
    try {
      kmi::b200_map<kmi::Kmer<31>, uint32_t, KMI_MODE_COUNT> map(kmi::comm(rank, p, box_bcast, &arg), o);
      if (K == 31) {
        const uint64_t produced = map.build_fastq(file.data() + cuts[rank], cuts[rank + 1] - cuts[rank], cuts[rank]);
        uint64_t s1[1] = {produced};
        KOK(kmi_comm_allreduce_u64(comm, s1, 1, 0));
        CHECK(s1[0] == (uint64_t)total);
        CHECK(map.local_size() == mine.size());
        uint64_t s2[1] = {(uint64_t)mine.size()};
        KOK(kmi_comm_allreduce_u64(comm, s2, 1, 0));
        CHECK(map.size() == s2[0]);
        std::vector<kmi::Kmer<31>> q(200);
        for (size_t i = 0; i < q.size(); ++i) q[i].getDataRef()[0] = keys[(i * 11 + (size_t)rank) % (size_t)total];
        auto hits = map.find(q);                           // every query is a k-mer of the file: all found, with its global count
        CHECK(hits.size() == q.size());
        std::map<uint64_t, uint32_t> all1;
        for (long i = 0; i < total; ++i) all1[keys[i]]++;
        for (const auto& h : hits) CHECK(all1[h.first.getData()[0]] == h.second);
      }
    } catch (const std::exception& ex) {
      std::printf("rank %d FAIL adapter: %s\n", rank, ex.what());
      return 1;
    }
  }
  KOK(kmi_comm_barrier(comm));
  kmi_comm_destroy(comm);
  if (rank == 0) std::printf("dist ok: %d ranks, K=%d, %zu reads\n", p, K, n_reads);
  return 0;
}

int main(int argc, char** argv) {
  setvbuf(stdout, nullptr, _IONBF, 0);                      // the ranks leave with _exit
  int p = argc > 1 ? std::atoi(argv[1]) : 0;
  const size_t n_reads = argc > 2 ? (size_t)std::atoll(argv[2]) : 60000;
  const int K = argc > 3 ? std::atoi(argv[3]) : 31;
  if (p <= 0) {
    // count the devices in a child so that the parent stays free of CUDA state before it forks the ranks
    int fd[2];
    if (pipe(fd) != 0) return 2;
    pid_t c = fork();
    if (c == 0) { int n = 0; kmi_device_count(&n); if (write(fd[1], &n, sizeof(n)) != sizeof(n)) _exit(1); _exit(0); }
    int n = 0;
    if (read(fd[0], &n, sizeof(n)) != sizeof(n)) n = 0;
    waitpid(c, nullptr, 0);
    p = n > 0 ? n : 1;
  }
  Shared* sh = static_cast<Shared*>(mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0));
  if (sh == MAP_FAILED) return 2;
  new (sh) Shared();
  sh->id_ready.store(0);
  sh->box_ready.store(0);
  std::vector<pid_t> kids;
  for (int r = 0; r < p; ++r) {
    pid_t c = fork();
    if (c == 0) _exit(run_rank(r, p, n_reads, K, sh));
    kids.push_back(c);
  }
  // a rank that fails leaves its peers inside a collective: stop them (exact pids) instead of waiting forever
  int bad = 0;
  for (size_t left = kids.size(); left > 0; --left) {
    int st = 0;
    const pid_t c = waitpid(-1, &st, 0);
    if (c < 0) break;
    if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) {
      bad = 1;
      for (pid_t k : kids) if (k != c) kill(k, SIGKILL);
    }
  }
  return bad;
}
