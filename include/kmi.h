/* kmi.h — C ABI of the B200-native k-mer index construction path (libkmi.so).
 *
 * Drop-in boundary for ParBLiSS/kmerind's `bliss::index::kmer::Index<MapType, KmerParser>`
 * (reference src/index/kmer_index.hpp:98-394) on the build / count / find / erase path. The
 * reference has no FFI: its boundary is a C++ template, so the replacement has two faces — this
 * C ABI, and the header-only C++ adapter include/kmi_index.hpp that gives it the reference's
 * member set (CountIndex / PositionIndex). INTEGRATION.md shows the reference-side binding.
 *
 * Conventions
 *  - A k-mer is `nw = (K <= 32 ? 1 : 2)` consecutive uint64_t, word 0 least significant, first
 *    base in the most significant used bits: exactly `bliss::common::Kmer<K, DNA, uint64_t>::data`
 *    (reference src/common/kmer.hpp:76-100,177), so `std::vector<Kmer>` can be passed by
 *    reinterpret_cast. K must leave >= 2 pad bits (K <= 31 or 33 <= K <= 63), as the reference's
 *    unsplit dense-hash maps require for their sentinels (src/index/kmer_hash.hpp:396-414).
 *  - Every function returns KMI_OK (0) or a negative kmi_status; nothing throws across the ABI.
 *    kmi_last_error() returns a thread-local message for the last failure.
 *  - `*_dev` entry points take DEVICE pointers (inputs already resident in HBM); the others take
 *    HOST pointers and do their own host<->device copies. Work is issued on the index's stream
 *    (default: the legacy default stream 0); host-pointer calls are blocking, `*_dev` calls are
 *    stream-ordered unless they return a count to the host.
 *  - There is no CPU fallback: without a CUDA device every compute call fails with KMI_ERR_CUDA.
 */
#ifndef KMI_H_
#define KMI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  KMI_OK = 0,
  KMI_ERR_ARG = -1,          /* std::invalid_argument in the reference */
  KMI_ERR_CUDA = -2,         /* device/runtime failure, or no device */
  KMI_ERR_FORMAT = -3,       /* malformed FASTQ: std::logic_error (fastq_loader.hpp:393,419,435) */
  KMI_ERR_CAPACITY = -4,     /* caller-provided output buffer too small */
  KMI_ERR_UNSUPPORTED = -5,  /* K / mode combination not instantiated */
  KMI_ERR_NOMEM = -6
} kmi_status;

typedef enum { KMI_MODE_COUNT = 0, KMI_MODE_POSITION = 1 } kmi_mode;

/* Hash families of reference src/index/kmer_hash.hpp:243-311. The store hash (Prefix=false:
 * murmur h[0] / farm seed 42) places a key in the local table, the dist hash (Prefix=true:
 * murmur h[1] / farm seed 83) picks the owning rank: owner = dist % nranks
 * (src/containers/distributed_densehash_map.hpp:148-170). FARM_DEBUG is farmhash as the
 * reference's default RelWithDebInfo build computes it (DebugTweak, no NDEBUG).
 * IDENTITY and STD are the remaining two flavours of the reference's benchmark matrix (kmer_hash.hpp:157-232,
 * BenchmarkKmerIndex.cpp:164-193): identity = the key's low 64 bits (store) / its top ceilLog2(nranks) bits (dist,
 * Kmer::getPrefix); cpp_std = XOR of (std::hash<size_t>(word) << 1) over the words, shifted down to ceilLog2(nranks) + ...
 * bits for the dist flavour exactly as cpp_std<KMER, true> does. Both make the owner a PREFIX of the key, so shards are as
 * skewed as the keys are (the reference's too); their sharded builds always take the partition + all-to-all path. */
typedef enum { KMI_HASH_MURMUR = 0, KMI_HASH_FARM = 1, KMI_HASH_FARM_DEBUG = 2, KMI_HASH_IDENTITY = 3, KMI_HASH_STD = 4 } kmi_hash_kind;

typedef struct {
  int32_t k;            /* k-mer length */
  int32_t mode;         /* kmi_mode */
  int32_t canonical;    /* 1: lex_less input transform (CanonicalHashMapParams), 0: identity */
  int32_t hash;         /* kmi_hash_kind, used for both dist and store hash */
  int32_t nsplit;       /* 0: N -> A (default parser), 1: NSplitSequencesIterator semantics */
  int32_t device;       /* CUDA device ordinal */
  int32_t rank;         /* this index' shard (GPU ordinal within the job) */
  int32_t nranks;       /* number of shards; owner(key) = dist_hash(key) % nranks */
  uint64_t reserve;     /* expected distinct keys (0 = grow on demand, like the reference) */
} kmi_config;

typedef struct kmi_index kmi_index;

/* ---- lifetime ------------------------------------------------------------------------------ */
/* Index::Index(comm)  — kmer_index.hpp:115 */
int kmi_create(const kmi_config* cfg, kmi_index** out);
void kmi_destroy(kmi_index* idx);
const char* kmi_last_error(void);
/* CUDA devices visible to this process (0 and KMI_ERR_CUDA without a driver) */
int kmi_device_count(int* n);
/* run subsequent work of this index on `cuda_stream` (a cudaStream_t); NULL = default stream */
int kmi_set_stream(kmi_index* idx, void* cuda_stream);
/* Tuning knobs without a reference counterpart. KMI_PARAM_BUILD_PATH: how k-mers reach the table —
 * 0 auto (by input size), 1 direct atomics from the parser, 2 through the 256 hash-prefix buckets
 * (table segments stay L2-resident). KMI_PARAM_SEGMENT_INSERT (bucketed path): 1 (default) =
 * second-level scatter + table segments updated in shared memory, 0 = L2-resident global atomics.
 * Results are identical on every path.
 * KMI_PARAM_SATURATE (COUNT index): 1 = counts stick at 2^32 - 1 instead of wrapping — saturating_counting_densehash_map,
 * distributed_densehash_map.hpp:2904-3075 (default 0: std::plus<uint32_t>, :2475); such an index inserts on path 1.
 * KMI_PARAM_DIST_TRANS (non-canonical index): the DistTrans of SingleStrandHashMapParams<Key, DistHash, StoreHash, DistTrans>
 * (BenchmarkKmerIndex.cpp:152-161,246-249) — owner = dist_hash(T(key)) % nranks with T = 0 identity (default), 1 lex_less,
 * 2 xor_rev_comp (kmer_transform.hpp:91-107), so both strands of a k-mer meet on one rank while the stored key stays as
 * given. Enters kmi_partition_dev and every sharded operation built on it.
 * KMI_PARAM_DEFER_INSERT (COUNT index; bulk load): value = staging budget in keys, 0 = off (default). The reference inserts
 * a whole file's k-mers in ONE Index::insert (kmer_index.hpp:201-225, 332-372); a GPU build far larger than one bucket round
 * (2^31 k-mers) otherwise runs the shared-memory segment insert — a read + write of the WHOLE table — once per round or per
 * call. In bulk-load mode the bucketed builds (kmi_build_fastq_dev of >= 8 MiB, kmi_dist_build_fastq*) stop after the fine
 * scatter, which appends to a staging area laid out for `value` keys (clamped to 60 % of the free device memory); the single
 * segment insert runs when the area is full, when the table must grow, when anything reads or changes the table (find, count,
 * erase, dump, size, checksum, save ...) or on kmi_flush. Results are identical; only the number of table sweeps changes.
 * KMI_PARAM_MAX_LOAD: percentage (10..70, default 70 = dense_hash_map's max_load_factor, densehash_map.hpp:969) of the
 * power-of-two table that keys may fill before it grows. Linear probing at 0.7 walks 2.2 slots per hit and 6 per miss; an
 * index built to be QUERIED (PositionIndex + batched find / count) answers ~2x faster at 50, for twice the table memory in the
 * unlucky case. Set it before kmi_reserve / the first build. Not observable in results. */
typedef enum { KMI_PARAM_BUILD_PATH = 1, KMI_PARAM_SEGMENT_INSERT = 2, KMI_PARAM_SATURATE = 3, KMI_PARAM_DIST_TRANS = 4,
               KMI_PARAM_DEFER_INSERT = 5, KMI_PARAM_MAX_LOAD = 6 } kmi_param;
int kmi_set_param(kmi_index* idx, int param, int64_t value);
/* insert whatever bulk-load mode has staged and wait for it; reports a sticky table error like the size readers do */
int kmi_flush(kmi_index* idx);
/* bulk-load accounting: keys staged right now, deferred segment inserts (table sweeps) run since creation */
int kmi_bulk_stats(kmi_index* idx, uint64_t* staged_keys, uint64_t* flushes);
/* map_base::reserve — src/containers/distributed_map_base.hpp:149-302 */
int kmi_reserve(kmi_index* idx, uint64_t n_distinct);
/* remove every entry, keep the allocation */
int kmi_clear(kmi_index* idx);

/* ---- build from file bytes (Index::build_posix<FASTQParser|FASTAParser>, kmer_index.hpp:332-372;
 *      KmerFileHelper::read_file_posix + Index::insert, kmer_file_helper.hpp:622-633) ------------
 * `bytes` must start at a record boundary (FASTQ: an '@' header line; FASTA: anywhere in line
 * start position) and hold whole records; `file_offset` is the offset of bytes[0] in the file
 * (only used for position ids: a POSITION-mode index built from FASTQ stores ShortSequenceKmerId values, from FASTA
 * LongSequenceKmerId values). Every k-mer of the block is inserted into THIS index regardless
 * of owner (nranks == 1 semantics); the sharded build is kmi_dist_build_*. n_kmers (optional) receives the number of k-mer occurrences generated. */
int kmi_build_fastq(kmi_index* idx, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_build_fasta(kmi_index* idx, const uint8_t* bytes, size_t n, uint64_t* n_kmers);
int kmi_build_fastq_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_build_fasta_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t* n_kmers);

/* ---- a6: one block of a partitioned FASTA file (KmerParser::get_valid_iterator_range, kmer_parser.hpp:114-157; the
 * FASTA tail rule of KmerFileHelper::read_block, kmer_file_helper.hpp:153-173; the cross-rank part of
 * FASTAParser::init_parser, fasta_loader.hpp:202-483). A block may begin and end at ANY byte. The `n` bytes passed are
 * the block's own bytes [0, n_valid) followed by halo bytes: a k-mer is emitted iff its first base lies in the block,
 * and it may run K-1 bases into the halo, so the halo must hold K-1 non-EOL bytes or reach the end of the file
 * (kmi_fasta_halo_end). What precedes the block enters as three facts: how many header runs began before it
 * (sequence id = that count, minus one when the file opens with a header line), whether the last line start before it
 * began a header line, and whether its first byte starts a line. kmi_fasta_block_summary_dev computes a block's own
 * contribution so that ranks can combine them with one exclusive scan (kmi::fasta_chain in kmi_index.hpp):
 *   out[0] the block holds a line start, out[1] its first line start begins a header, out[2] its last one does,
 *   out[3] header runs starting inside it, counted as if the line before the block were no header (subtract one when
 *   out[1] is set and the previous block hands over a header line).
 * block == NULL means the whole file (kmi_build_fasta / kmi_parse_fasta_dev are exactly that). */
typedef struct {
  uint64_t file_offset;         /* file byte of bytes[0] (position ids carry file offsets) */
  uint64_t n_valid;             /* bytes owned by the block; the rest of the buffer is halo */
  uint32_t header_runs_before;  /* header runs that start before the block */
  uint32_t flags;               /* KMI_FASTA_* */
} kmi_fasta_block;
#define KMI_FASTA_BOL 1u                       /* bytes[0] starts a line (file start, or the previous byte is '\n') */
#define KMI_FASTA_PREV_HEADER 2u               /* the last line start before the block begins a header line */
#define KMI_FASTA_FILE_STARTS_WITH_HEADER 4u   /* byte 0 of the FILE is '>' or ';' */
int kmi_fasta_block_summary_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n_valid, int bol, uint32_t* out /* 4 */);
int kmi_build_fasta_block(kmi_index* idx, const uint8_t* bytes, size_t n, const kmi_fasta_block* block, uint64_t* n_kmers);
int kmi_build_fasta_block_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, const kmi_fasta_block* block, uint64_t* n_kmers);
/* host: end of the halo of a block ending at file byte valid_end: the first offset by which K-1 non-EOL bytes follow
 * valid_end, or n_file. `bytes` holds the file bytes [mem_offset, mem_offset + n). */
int kmi_fasta_halo_end(const uint8_t* bytes, uint64_t mem_offset, size_t n, uint64_t valid_end, int k, uint64_t* halo_end);

/* ---- f1: build from a file (Index::build_posix<FASTQParser|FASTAParser>, kmer_index.hpp:231-260,332-372, with the
 * posix_file reader of src/io/file.hpp:552-900). format 0 = by extension (.fastq/.fq, .fasta/.fa; anything else is
 * KMI_ERR_ARG "input filename extension is not supported." as kmer_index.hpp:245-254 throws), 1 = FASTQ, 2 = FASTA. The
 * file is read with pread into two pinned buffers by a reader thread while the GPU builds from the other buffer: FASTQ
 * chunks end at a record boundary (kmi_fastq_find_first_record), FASTA chunks are blocks with a K-1 halo (kmi_fasta_block);
 * position ids carry file offsets. Chunk size 1 GiB (environment KMI_FILE_CHUNK_BYTES overrides it). */
int kmi_build_file(kmi_index* idx, const char* path, int format, uint64_t* n_kmers);
/* a1: rank `rank`'s block of ONE file shared by `nranks` ranks — partitioned_file<posix_file, Parser>::read_file with the
 * BlockPartitioner (src/io/file.hpp:1239-1251,1342-1431): the equal byte ranges [size * r / p, size * (r + 1) / p); for FASTQ
 * every boundary is moved forward to the first whole record at or after it (FASTQParser::find_first_record on a window of
 * the file, fastq_loader.hpp:269-357,696-873), so blocks hold whole records and tile the file; a FASTA block is the byte
 * range as it is (kmi_fasta_block / kmi_fasta_part take any cut). Host only: reads a few KiB around the two boundaries.
 * format as in kmi_build_file. file_size is optional. */
int kmi_file_block(const char* path, int format, int rank, int nranks, uint64_t* begin, uint64_t* end, uint64_t* file_size);

/* ---- f2: persistence. The reference keeps no index on disk; this is the wire format of one shard: a 64-byte header
 * ("KMIDX001", k, mode, words per key, canonical, entry count) followed by the entries of to_vector()
 * (distributed_densehash_map.hpp:1103-1113) sorted by (key, value): all keys (nw little-endian uint64_t each), then all
 * values as uint64_t (COUNT: frequency; POSITION: one entry per (key, id)). kmi_load ADDS a file's entries to the index
 * (counts are summed, ids appended), so shards saved by several ranks can be merged into one index; K and mode must
 * match (KMI_ERR_ARG). */
int kmi_save(kmi_index* idx, const char* path);
int kmi_load(kmi_index* idx, const char* path, uint64_t* n_entries);

/* ---- a2: record-aligned blocks (FASTQParser::find_first_record, fastq_loader.hpp:269-357,696-873) ---------------
 * Host side, as in the reference (it looks at four lines). `bytes` holds file bytes [mem_offset, mem_offset + n); the
 * search starts at search_begin (dropping a partial first line unless search_begin == parent_begin) and matches the first
 * characters of four consecutive non-empty lines against @?+? / ?@?+ / +?@? / ?+?@. *offset = file offset of the '@' of
 * the first whole record, or the end of the search range when it holds no complete record. KMI_ERR_FORMAT when no marker
 * pattern is found in a range that is not the parent's tail (the reference throws IOException). Use it to cut a file
 * into the blocks kmi_build_fastq / kmi_parse_fastq_dev take: block r = [first_record(cut r), first_record(cut r+1)). */
int kmi_fastq_find_first_record(const uint8_t* bytes, uint64_t mem_offset, size_t n, uint64_t search_begin,
                                uint64_t search_end, uint64_t parent_begin, uint64_t parent_end, uint64_t* offset);

/* ---- stages, device resident ------------------------------------- */
/* a2-a10: record boundaries + 2-bit encode + sliding window (+ lex_less if cfg.canonical).
 * Writes up to `cap` keys (nw words each) and, if d_ids != NULL, their position ids
 * (FASTA: LongSequenceKmerId = seq_id << 40 | file byte of the first base, sequence.hpp:180-236;
 *  FASTQ: ShortSequenceKmerId = (file offset of the record's '@' & (2^40 - 1)) << 16, plus the offset of the first base
 *  inside the record, sequence.hpp:103-160, kmer_parser.hpp:378-399, with file offset = file_offset + offset in d_bytes).
 * Order of the output is unspecified. */
int kmi_parse_fastq_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset,
                        uint64_t* d_keys, uint64_t* d_ids, size_t cap, uint64_t* n_kmers);
int kmi_parse_fasta_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n,
                        uint64_t* d_keys, uint64_t* d_ids, size_t cap, uint64_t* n_kmers);
int kmi_parse_fasta_block_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, const kmi_fasta_block* block,
                              uint64_t* d_keys, uint64_t* d_ids, size_t cap, uint64_t* n_kmers);
/* a10: out = canonical(in) (lex_less, kmer_transform.hpp:109-126); in == out allowed */
int kmi_canonicalize_dev(kmi_index* idx, const uint64_t* d_keys, uint64_t* d_out, size_t n);
/* a11/a12: which = 0 store hash, 1 dist hash, of the keys AS GIVEN (no transform) */
int kmi_hash_dev(kmi_index* idx, int which, const uint64_t* d_keys, uint64_t* d_hash, size_t n);
/* a13/a14: imxx::distribute's local half (incremental_mxx.hpp:273-364,596-640): stable
 * counting-sort of n keys (+ optional 8-byte values) by owner = dist_hash % nranks.
 * h_counts[nranks] receives the per-owner counts (the send counts of the all-to-all). */
int kmi_partition_dev(kmi_index* idx, const uint64_t* d_keys, const uint64_t* d_vals, size_t n,
                      uint64_t* d_keys_out, uint64_t* d_vals_out, uint64_t* h_counts);

/* kmi_partition_dev for callers that cannot allocate. NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). Same stable
 * counting-sort by owner, same outputs, but the scratch (per-chunk owner histogram + scan sums) is one file-scope
 * __device__ array of KMI_PARTITION_NOALLOC_MAX_HIST + ~1100 32-bit words (4.2 MB) that every GPU loading the library holds
 * until exit, and the per-owner counts are written to DEVICE memory (d_counts[nranks]) with no stream synchronisation.
 * The call performs no cudaMalloc / cudaFree and grows no buffer of the index; work is queued on the index's stream.
 * Limit: ceil(n / 2048) * nranks <= KMI_PARTITION_NOALLOC_MAX_HIST, i.e. n <= kmi_partition_noalloc_max_keys(nranks)
 * (268 M keys at 8 ranks); beyond it the call returns KMI_ERR_ARG and touches nothing. All calls on one GPU share the
 * scratch: calls that overlap on the device (different streams / indexes / threads) race — serialise them. The first call per
 * device looks the array's address up (cudaGetSymbolAddress); make that call once before starting a graph capture. */
#define KMI_PARTITION_NOALLOC_MAX_HIST (1u << 20)
size_t kmi_partition_noalloc_max_keys(int nranks);
int kmi_partition_noalloc_dev(kmi_index* idx, const uint64_t* d_keys, const uint64_t* d_vals, size_t n,
                              uint64_t* d_keys_out, uint64_t* d_vals_out, uint64_t* d_counts);

/* ---- insert (Index::insert, kmer_index.hpp:201-225) ---------------------------------------- */
/* COUNT mode: vals == NULL: each key counts 1 (counting_densehash_map::insert(vector<Key>&),
 * distributed_densehash_map.hpp:2714-2899); vals != NULL: (key, count) tuples, vals[i] is ADDED to the entry of keys[i]
 * (reduction_densehash_map::insert with std::plus<uint32_t>, :2463-2482,2565-2628: the sum wraps at 2^32). POSITION mode: vals[i] is the id of keys[i]
 * (densehash_multimap::insert, :2329-2384; duplicates kept). Keys are canonicalised first when
 * cfg.canonical (transform_input, distributed_map_base.hpp:286-301); the caller's arrays are not
 * modified. */
int kmi_insert(kmi_index* idx, const uint64_t* keys, const uint64_t* vals, size_t n);
int kmi_insert_dev(kmi_index* idx, const uint64_t* d_keys, const uint64_t* d_vals, size_t n);

/* ---- batched query (kmer_index.hpp:132-194; distributed_densehash_map.hpp:1122-1216,327-448) -
 * Results are in INPUT order (the reference returns them grouped by answering rank; the adapter
 * documents the difference). Queries are canonicalised first when cfg.canonical.
 * count: counts[i] = 0/1 presence for a COUNT index (dense_hash_map::count), multiplicity for a
 *        POSITION index (LocalCount, :234-266).
 * find (COUNT): vals[i] = stored frequency, found[i] = 0/1 (LocalFind, :1769-1798).
 * find (POSITION): CSR answer — offsets[n+1] into vals; cap = capacity of vals (LocalFind :2094-2133). */
int kmi_count(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* counts);
int kmi_find(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* vals, uint8_t* found);
int kmi_find_positions(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* offsets,
                       uint64_t* vals, size_t cap);
int kmi_erase(kmi_index* idx, const uint64_t* keys, size_t n);
int kmi_count_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_counts);
int kmi_find_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_vals, uint8_t* d_found);

/* ---- accessors ----------------------------------------------------------------------------- */
/* local_size(): distinct keys (COUNT) / stored tuples (POSITION) — kmer_index.hpp:386-391 */
int kmi_local_size(kmi_index* idx, uint64_t* n);
/* distinct keys in either mode */
int kmi_num_keys(kmi_index* idx, uint64_t* n);
/* to_vector(): every (key, value) entry, unordered. COUNT: value = frequency. POSITION: one
 * entry per stored (key, id) tuple. n_out receives the number written (<= cap). */
int kmi_dump(kmi_index* idx, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out);
/* the same entries ordered by (key, value) — Kmer::operator< on the key, most significant word first — sorted on the
 * device (the radix sort of the sorted back end below). This is the order kmi_save writes. */
int kmi_dump_sorted(kmi_index* idx, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out);
/* ---- f2: find(pred) / count(pred) / erase(pred) over the local map (kmer_index.hpp:156-194;
 * distributed_densehash_map.hpp:1666-1686,1835-1903) for predicates on the stored VALUE, evaluated on the device: one
 * compare fused into the table sweep instead of a dump of the whole table to the host. COUNT indices (value = frequency):
 * "k-mers seen at least c times", "singletons" (the error k-mers), ... kmi_count_if sizes the buffers of kmi_dump_if. */
typedef enum { KMI_PRED_GE = 0, KMI_PRED_LE = 1, KMI_PRED_EQ = 2 } kmi_pred_op;
int kmi_count_if(kmi_index* idx, int op, uint64_t value, uint64_t* n);
int kmi_dump_if(kmi_index* idx, int op, uint64_t value, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out);
int kmi_erase_if(kmi_index* idx, int op, uint64_t value, uint64_t* n_removed);
/* order-independent 64-bit checksums of the contents: sum over entries of mix(key)*value and of
 * value (mod 2^64) — used to compare builds at sizes where a dump diff is impractical */
int kmi_checksum(kmi_index* idx, uint64_t* key_value_sum, uint64_t* value_sum);
/* table geometry for reporting: slots, bytes per slot, device bytes held */
int kmi_table_info(kmi_index* idx, uint64_t* slots, uint64_t* slot_bytes, uint64_t* device_bytes);
/* number of kernels this index has launched since creation (bench.py's gpu_launches) */
int kmi_launch_count(kmi_index* idx, uint64_t* n);
/* device time (ms) of the stages of the last build_* call, in the reference's BL_BENCH stage
 * names order: [0] scan (record boundaries), [1] parse+insert (fused) or parse, [2] insert */
int kmi_last_stage_ms(kmi_index* idx, float* ms3);
/* device time (ms, CUDA events on the index's stream) of the kernels of the last build_fastq* call:
 * [0] whole build, [1] parse (+ bucket scatter), [2] fine scatter, [3] table insert (shared-memory
 * segments or L2 atomics), [4] overflow / spill insert. [2]-[4] are 0 on the direct path. */
int kmi_last_kernel_ms(kmi_index* idx, float* ms5);

/* ---- the sharded index: one process per GPU, owner(key) = dist_hash(key) % nranks -------------------------------------
 * Replaces the collective half of the reference's distributed maps (every rank calls every operation, also with empty
 * input: distributed_densehash_map.hpp:2567-2573): imxx::distribute (src/io/incremental_mxx.hpp:1039-1109: bucket by
 * KeyToRank, mxx::all2all of the counts, mxx::all2allv of the elements) + the local operation + the way back for queries.
 *
 * kmi_comm wraps one NCCL communicator (libnccl.so.2 is loaded with dlopen on first use). Bootstrap as NCCL wants it: rank 0
 * calls kmi_comm_unique_id and hands the KMI_COMM_ID_BYTES bytes to every rank by any means (MPI_Bcast in the reference's
 * world, a pipe, torch.distributed), then every rank calls kmi_comm_create. kmi_dist_attach binds an index created with the
 * same rank / nranks / device to the communicator (collective; it also maps every rank's receive buffer with CUDA IPC when
 * the ranks share an NVLink domain). After that:
 *   kmi_dist_build_fastq[_dev]  rank r passes ITS record-aligned block of the reads (kmi_fastq_find_first_record cuts them).
 *       COUNT index, <= 8 ranks, peer mapping available: the parser's 256 hash-prefix buckets are pushed by k_owner_push
 *       straight into the owners' HBM over NVLink while the next round is parsed, and every owner starts its table build at
 *       the fine scatter (no pass over received keys, no NCCL for the keys). Otherwise, and for POSITION indices: parse ->
 *       stable partition (kmi_partition_dev) -> NCCL all-to-all -> insert. Skewed keys that overflow a receive region make
 *       every rank redo the build on the second path (decided by one all-reduce: no rank is left alone in a collective).
 *   kmi_dist_build_fasta[_dev]  rank r passes block r of ONE file (kmi_fasta_part; blocks in rank order, cut at any byte,
 *       followed by the K-1 halo of kmi_fasta_halo_end) or, with part == NULL, a whole file of its own. The cross-rank part
 *       of FASTAParser::init_parser (fasta_loader.hpp:202-483) is one all-gather of the block summaries.
 *   kmi_dist_insert / count / find / find_positions / erase   the batch operations of kmer_index.hpp:132-225 on keys held
 *       by any rank; answers arrive in the caller's input order. kmi_dist_size = map_base::size (distributed_map_base.hpp:243).
 * kmi_local_size / kmi_dump / kmi_checksum / kmi_save / ... keep working on this rank's shard. */
typedef struct kmi_comm kmi_comm;
#define KMI_COMM_ID_BYTES 128
int kmi_comm_unique_id(uint8_t* id_out /* KMI_COMM_ID_BYTES */);
int kmi_comm_create(const uint8_t* id, int nranks, int rank, int device, kmi_comm** out);
void kmi_comm_destroy(kmi_comm* comm);
int kmi_comm_rank(const kmi_comm* comm);
int kmi_comm_size(const kmi_comm* comm);
int kmi_comm_barrier(kmi_comm* comm);
/* small host-value collectives for callers that have no other transport: op 0 sum, 1 max, 2 min */
int kmi_comm_allreduce_u64(kmi_comm* comm, uint64_t* vals, int n, int op);
int kmi_comm_allgather_u64(kmi_comm* comm, const uint64_t* mine, int n, uint64_t* all /* nranks * n */);

typedef struct {
  uint64_t file_offset;   /* file byte of the block's first byte */
  uint64_t n_valid;       /* bytes the rank owns; the rest of the buffer is halo */
  int32_t bol;            /* the block's first byte starts a line */
  int32_t reserved;
} kmi_fasta_part;
int kmi_dist_attach(kmi_index* idx, kmi_comm* comm);
int kmi_dist_build_fastq(kmi_index* idx, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_dist_build_fastq_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_dist_build_fasta(kmi_index* idx, const uint8_t* bytes, size_t n, const kmi_fasta_part* part, uint64_t* n_kmers);
int kmi_dist_build_fasta_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, const kmi_fasta_part* part, uint64_t* n_kmers);
/* Index::build_posix<FASTQParser|FASTAParser>(filename, comm) at p ranks (kmer_index.hpp:231-260,332-372 over
 * partitioned_file::read_file, file.hpp:1342-1431): collective; every rank passes the SAME path, takes its block of the
 * file (kmi_file_block with the communicator's rank / size), reads it — FASTQ: the pinned slice ring's threads pread it
 * straight into pinned slices while the sharded build runs; FASTA: block + K-1 halo, then kmi_dist_build_fasta with the block
 * as kmi_fasta_part — and the k-mers go to their owners. n_kmers: k-mers generated from THIS rank's block. A rank that cannot
 * read the file makes every rank return the error (one all-reduce before the build). */
int kmi_dist_build_file(kmi_index* idx, const char* path, int format, uint64_t* n_kmers);
int kmi_dist_insert(kmi_index* idx, const uint64_t* keys, const uint64_t* vals, size_t n);
int kmi_dist_insert_dev(kmi_index* idx, const uint64_t* d_keys, const uint64_t* d_vals, size_t n);
int kmi_dist_count(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* counts);
int kmi_dist_find(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* vals, uint8_t* found);
int kmi_dist_find_positions(kmi_index* idx, const uint64_t* keys, size_t n, uint64_t* offsets, uint64_t* vals, size_t cap);
int kmi_dist_erase(kmi_index* idx, const uint64_t* keys, size_t n);
/* the same batch queries with the keys already in HBM and the answers left there (input order) */
int kmi_dist_count_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_counts);
int kmi_dist_find_dev(kmi_index* idx, const uint64_t* d_keys, size_t n, uint64_t* d_vals, uint8_t* d_found);
int kmi_dist_size(kmi_index* idx, uint64_t* n);
/* last sharded build: device ms [0] whole, [1] parse + push / exchange, [2] insert of the received keys, [3] unused; keys this
 * rank received; path 1 = NVLink push, 2 = NCCL all-to-all */
int kmi_dist_last_stats(kmi_index* idx, float* ms4, uint64_t* n_received, int32_t* path);

/* ---- f4: the sort-based back end — dsc::counting_sorted_map (COUNT) / dsc::sorted_multimap (POSITION) of reference
 * src/containers/distributed_sorted_map.hpp, the alternative MapType of the same Index<> facade (BenchmarkKmerIndex.cpp:
 * 206-219, pMAP == SORTED). The reference keeps a std::vector per rank: insert() appends, rehash() sorts and reduces
 * (local_reduction, :1521-1540,2560-2600), count()/find() are binary searches (:228-480). Here the entries live in HBM:
 * build / insert append canonical keys (+ ids), kmi_sorted_rehash is one device LSD radix sort (8-bit digits, stable) over
 * old + new entries followed by a run-length reduction (COUNT: counts summed as uint32_t, wrapping like the mapped type),
 * queries are one binary search per key and call rehash first when entries are pending. kmi_sorted_dump returns the
 * entries IN KEY ORDER (POSITION: ordered by (key, id)). ONE SHARD ONLY: the reference's cross-rank half — sample sort
 * with splitters, :1544-1700 — is not built, so cfg.nranks must be 1 (KMI_ERR_UNSUPPORTED otherwise).
 * Same conventions as the hash index: keys are canonicalised first when cfg.canonical, host pointers unless `_dev`. */
typedef struct kmi_sorted kmi_sorted;
int kmi_sorted_create(const kmi_config* cfg, kmi_sorted** out);
void kmi_sorted_destroy(kmi_sorted* s);
int kmi_sorted_build_fastq(kmi_sorted* s, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_sorted_build_fasta(kmi_sorted* s, const uint8_t* bytes, size_t n, uint64_t* n_kmers);
int kmi_sorted_build_fastq_dev(kmi_sorted* s, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_sorted_build_fasta_dev(kmi_sorted* s, const uint8_t* d_bytes, size_t n, uint64_t* n_kmers);
/* COUNT: vals == NULL counts 1 per key, else vals[i] is added; POSITION: vals[i] is the id of keys[i] */
int kmi_sorted_insert(kmi_sorted* s, const uint64_t* keys, const uint64_t* vals, size_t n);
int kmi_sorted_insert_dev(kmi_sorted* s, const uint64_t* d_keys, const uint64_t* d_vals, size_t n);
int kmi_sorted_rehash(kmi_sorted* s);
int kmi_sorted_local_size(kmi_sorted* s, uint64_t* n);
/* counts[i]: 0/1 (COUNT) or the multiplicity (POSITION) */
int kmi_sorted_count(kmi_sorted* s, const uint64_t* keys, size_t n, uint64_t* counts);
int kmi_sorted_find(kmi_sorted* s, const uint64_t* keys, size_t n, uint64_t* vals, uint8_t* found);
int kmi_sorted_find_positions(kmi_sorted* s, const uint64_t* keys, size_t n, uint64_t* offsets, uint64_t* vals, size_t cap);
int kmi_sorted_dump(kmi_sorted* s, uint64_t* keys, uint64_t* vals, size_t cap, size_t* n_out);
/* device ms of the last build's parse [0] and of the last rehash: sort [1], reduce [2]; kernels launched so far */
int kmi_sorted_stats(kmi_sorted* s, float* ms3, uint64_t* launches);
/* the radix sort by itself: n entries of idx's key width in d_keys (+ optional 8-byte values in d_vals) are sorted by
 * (key, low val_bits bits of the value); the *_tmp arrays are scratch of the same sizes and *sorted_in_tmp tells which pair
 * holds the result (the passes ping-pong). Stream-ordered on the index's stream. */
int kmi_sort_pairs_dev(kmi_index* idx, uint64_t* d_keys, uint64_t* d_vals, uint64_t* d_keys_tmp, uint64_t* d_vals_tmp, size_t n,
                       int val_bits, int* sorted_in_tmp);

/* ---- f4: PositionQualityIndex — k-mer -> (position id, quality score): Index<MapType, KmerPositionQualityTupleParser<...>>
 * (kmer_index.hpp:406; src/io/kmer_parser.hpp:578-760) with the score of QualityScoreSlidingWindow over
 * Illumina18QualityScoreCodec<float> (src/index/quality_score_iterator.hpp:60-140, quality_scores.hpp:60-170,529): Phred+33,
 * per base log2(1 - 10^(-q/10)), per k-mer exp2 of the window's running float sum, 0 when a base of the window has a score
 * outside 1..93. THE ONLY FLOATING-POINT PATH: the running sum is replayed in the reference's operation order (bit-identical
 * float adds / subtracts), the final exp2f is the device's — tolerance 1e-6 relative (tests/test_quality.py).
 * kmi_fastq_quality_dev is the stage: d_ids are the ShortSequenceKmerId values kmi_parse_fastq_dev produced over the same
 * d_bytes / file_offset (any order); d_quals[i] receives the score of the k-mer d_ids[i] names. With cfg.nsplit the window
 * restarts after every N, as the sub-sequences of NSplitSequencesIterator do. kmi_quality_lut returns the 96-entry table.
 * kmi_posqual is the index (one shard): a POSITION-mode index over the same parser that also keeps every tuple's score. */
int kmi_quality_lut(float* out96);
int kmi_fastq_quality_dev(kmi_index* idx, const uint8_t* d_bytes, size_t n, uint64_t file_offset, const uint64_t* d_ids, size_t n_ids,
                          float* d_quals);
typedef struct kmi_posqual kmi_posqual;
int kmi_posqual_create(const kmi_config* cfg, kmi_posqual** out);
void kmi_posqual_destroy(kmi_posqual* pq);
int kmi_posqual_build_fastq(kmi_posqual* pq, const uint8_t* bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_posqual_build_fastq_dev(kmi_posqual* pq, const uint8_t* d_bytes, size_t n, uint64_t file_offset, uint64_t* n_kmers);
int kmi_posqual_local_size(kmi_posqual* pq, uint64_t* n);
int kmi_posqual_count(kmi_posqual* pq, const uint64_t* keys, size_t n, uint64_t* counts);
/* CSR answer like kmi_find_positions, with the score of every (k-mer, id) hit in quals */
int kmi_posqual_find(kmi_posqual* pq, const uint64_t* keys, size_t n, uint64_t* offsets, uint64_t* ids, float* quals, size_t cap);
int kmi_posqual_erase(kmi_posqual* pq, const uint64_t* keys, size_t n);
/* to_vector(): one entry per stored tuple, unordered */
int kmi_posqual_dump(kmi_posqual* pq, uint64_t* keys, uint64_t* ids, float* quals, size_t cap, size_t* n_out);
int kmi_posqual_launch_count(kmi_posqual* pq, uint64_t* n);

/* ---- synthetic workload (same bytes as oracle/kmi_oracle.c kmio_synth_fastq) - */
size_t kmi_synth_record_bytes(uint32_t read_len);
int kmi_synth_fastq_dev(int device, uint8_t* d_out, uint64_t first_read, uint64_t n_reads,
                        uint32_t read_len, uint64_t genome_len, uint64_t seed, uint32_t err_ppm,
                        uint32_t n_ppm);

/* config 5's inputs: bytes [file_offset, file_offset + n_bytes) of a synthetic FASTA genome (n_seq sequences of seq_len bases,
 * `width` bases per line, headers ">chrNN"; same bytes as oracle kmio_synth_fasta), and query k-mers: even ones are K bases
 * of that genome at a random place and strand, odd ones uniform random */
size_t kmi_synth_fasta_bytes(uint32_t n_seq, uint64_t seq_len, uint32_t width);
int kmi_synth_fasta_dev(int device, uint8_t* d_out, uint64_t file_offset, uint64_t n_bytes, uint32_t n_seq, uint64_t seq_len,
                        uint32_t width, uint64_t seed);
int kmi_synth_queries_dev(int device, uint64_t* d_keys, uint64_t first, uint64_t n, int k, uint32_t n_seq, uint64_t seq_len,
                          uint64_t genome_seed, uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif /* KMI_H_ */
