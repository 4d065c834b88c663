// kmi_ops.cuh — host-side launchers, one table of function pointers per (K, hash).
// kmi_ops_inst.cu is compiled once per supported K (-DKMI_K=...), so that the API translation
// unit carries no templates and the per-K objects build in parallel.
#pragma once
#include <algorithm>
#include <cstdlib>
#include "kmi_parse.cuh"
#include "kmi_kernels.cuh"
#include "kmi_fasta.cuh"

namespace kmi {

struct Ops {
  int K, NW, hash;
  // parse (FASTQ). sink_kind 0: key stream, 1: straight into the table, 2: hash-prefix buckets
  void (*fastq_kmers)(cudaStream_t, int sink_kind, const uint8_t* bytes, u64 n, u64 n_tiles, const u32* tile_lines,
                      const u32* block_sums, KeySink sink, TableView table, BucketSink bsink, u64* n_kmers, u32* perr,
                      bool canon, bool nsplit, u64 tile_begin, u32 nranks);
  void (*scatter_keys)(cudaStream_t, const u64* keys, u64 n, BucketSink bsink, u32* perr, bool canon);
  void (*hll_buckets)(cudaStream_t, BucketSink bsink);   // HLL registers from the sampled buckets (table sizing)
  void (*insert_buckets)(cudaStream_t, TableView, BucketSink bsink, u64 n_hint);
  // second level (one-word keys only; null for K > 32): fine scatter, shared-memory segment insert
  void (*scatter_fine)(cudaStream_t, BucketSink bsink, FineSink fsink, u32* perr, u64 n_hint);
  int (*insert_segments)(cudaStream_t, TableView, FineSink fsink, BucketSink bsink, u32 log_seg, u32 n_segs, u32* perr, bool fresh);
  void (*insert_overflow)(cudaStream_t, TableView, BucketSink bsink);
  // sharded build: coarse buckets -> owners' receive buffers (peer memory), owner = dist hash % p
  void (*owner_push)(cudaStream_t, BucketSink bsink, PushSink psink, u64 n_hint);
  // parse (FASTA) to a key(+id) stream; wide_halo selects the 1 KiB halo variant
  void (*fasta_kmers)(cudaStream_t, bool wide_halo, const uint8_t* bytes, u64 n, u64 n_tiles, const u32* tile_enc_scanned,
                      const u32* enc_block_max, const u32* tile_runs, const u32* runs_block_sums, KeySink sink, u64* n_kmers,
                      u32* perr, bool canon, bool nsplit, u64 tile_begin, FaBlock blk);
  void (*table_clear)(cudaStream_t, void* slots, u64 n);
  void (*insert_keys)(cudaStream_t, TableView, const u64* keys, u64 n, bool canon, const u64* vals);
  void (*rehash)(cudaStream_t, const void* old_slots, u64 n_old, TableView);
  void (*lookup)(cudaStream_t, TableView, const u64* keys, u64 n, u64* vals, uint8_t* found, u64* slot_idx, bool canon);
  void (*erase)(cudaStream_t, TableView, const u64* keys, u64 n, u64* vals_removed, bool canon);
  void (*dump)(cudaStream_t, const void* slots, u64 n_slots, u64* keys, u64* vals, u64 cap, u64* cursor, int op, u64 x);
  void (*erase_if)(cudaStream_t, void* slots, u64 n_slots, u64* size, u64* removed, int op, u64 x);
  void (*checksum)(cudaStream_t, const void* slots, u64 n_slots, u64* out2);
  void (*canonicalize)(cudaStream_t, const u64* in, u64* out, u64 n);
  void (*hash_keys)(cudaStream_t, int which, const u64* in, u64* out, u64 n, u32 nranks);
  void (*part_hist)(cudaStream_t, const u64* keys, u64 n, u32 p, u32* hist, u64 n_chunks, u32 dist_trans);
  void (*part_scatter)(cudaStream_t, const u64* keys, const u64* vals, u64 n, u32 p, const u32* hist_scanned,
                       const u32* hist_block_sums, u64 n_chunks, u64* keys_out, u64* vals_out, u32 dist_trans);
  void (*csr_scan_l1)(cudaStream_t, const void* slots, u64 n_slots, u64* offsets, u64* block_sums);
  void (*csr_fill)(cudaStream_t, TableView, const u64* keys, const u64* ids, u64 n, const u64* offsets, u64* values);
  void (*log_filter)(cudaStream_t, TableView, const u64* keys, const u64* ids, u64 n, u64* keys_out, u64* ids_out,
                     u64* cursor);
  void (*zero_aux)(cudaStream_t, void* slots, u64 n);
};

static inline unsigned grid_for(u64 n, unsigned block, unsigned max_blocks = 148u * 16u) {
  u64 g = (n + block - 1) / block;
  if (g < 1) g = 1;
  if (g > max_blocks) g = max_blocks;
  return (unsigned)g;
}

template <int K, int HASH>
struct OpsImpl {
  static constexpr int NW = (K <= 32 ? 1 : 2);
  typedef Slot<NW> S;
  static void fastq_kmers(cudaStream_t st, int sink_kind, const uint8_t* bytes, u64 n, u64 n_tiles, const u32* tile_lines,
                          const u32* block_sums, KeySink sink, TableView table, BucketSink bsink, u64* n_kmers, u32* perr,
                          bool canon, bool nsplit, u64 tile_begin, u32 nranks) {
    if (n_tiles == 0) return;
    if (sink_kind == 0)
      k_fastq_kmers<K, HASH, 0><<<(unsigned)n_tiles, TILE_THREADS, 0, st>>>(bytes, n, tile_lines, block_sums, sink, table, bsink,
                                                                          n_kmers, perr, canon, nsplit, tile_begin, nranks);
    else if (sink_kind == 1)
      k_fastq_kmers<K, HASH, 1><<<(unsigned)n_tiles, TILE_THREADS, 0, st>>>(bytes, n, tile_lines, block_sums, sink, table, bsink,
                                                                          n_kmers, perr, canon, nsplit, tile_begin, nranks);
    else
      k_fastq_kmers<K, HASH, 2><<<(unsigned)n_tiles, TILE_THREADS, 0, st>>>(bytes, n, tile_lines, block_sums, sink, table, bsink,
                                                                          n_kmers, perr, canon, nsplit, tile_begin, nranks);
  }
  static void scatter_keys(cudaStream_t st, const u64* keys, u64 n, BucketSink bsink, u32* perr, bool canon) {
    if (n) k_scatter_keys<K, HASH><<<grid_for(n, TILE_THREADS * BktRound<NW>::R, 148u * 8u * 16u), TILE_THREADS, 0, st>>>(keys, n, bsink, perr, canon);
  }
  static void hll_buckets(cudaStream_t st, BucketSink bsink) {
    k_hll_buckets<K><<<dim3(148u * 2u, HLL_SAMPLE_BUCKETS * (N_BUCKETS >> bsink.log_pre) * bsink.sub + 1), 256, 0, st>>>(bsink);
  }
  static void insert_buckets(cudaStream_t st, TableView t, BucketSink bsink, u64 n_hint) {
    k_insert_buckets<K, HASH><<<grid_for(n_hint ? n_hint : 1, 256, 148u * 8u), 256, 0, st>>>(t, bsink);
  }
  static void fasta_kmers(cudaStream_t st, bool wide_halo, const uint8_t* bytes, u64 n, u64 n_tiles, const u32* tile_enc_scanned,
                          const u32* enc_block_max, const u32* tile_runs, const u32* runs_block_sums, KeySink sink,
                          u64* n_kmers, u32* perr, bool canon, bool nsplit, u64 tile_begin, FaBlock blk) {
    if (n_tiles == 0) return;
    if (!wide_halo)
      k_fasta_kmers<K, HASH, 4><<<(unsigned)n_tiles, TILE_THREADS, 0, st>>>(bytes, n, tile_enc_scanned, enc_block_max, tile_runs,
                                                                          runs_block_sums, sink, n_kmers, perr, canon, nsplit,
                                                                          tile_begin, blk);
    else
      k_fasta_kmers<K, HASH, 64><<<(unsigned)n_tiles, TILE_THREADS, 0, st>>>(bytes, n, tile_enc_scanned, enc_block_max, tile_runs,
                                                                           runs_block_sums, sink, n_kmers, perr, canon, nsplit,
                                                                           tile_begin, blk);
  }
  static void table_clear(cudaStream_t st, void* slots, u64 n) {
    k_table_clear<NW><<<grid_for(n, 256), 256, 0, st>>>(reinterpret_cast<S*>(slots), n);
  }
  static void insert_keys(cudaStream_t st, TableView t, const u64* keys, u64 n, bool canon, const u64* vals) {
    if (n) k_insert_keys<K, HASH><<<grid_for(n, 256, 148u * 32u), 256, 0, st>>>(t, keys, n, canon, vals);
  }
  static void rehash(cudaStream_t st, const void* old_slots, u64 n_old, TableView t) {
    k_rehash<K, HASH><<<grid_for(n_old, 256, 148u * 32u), 256, 0, st>>>(reinterpret_cast<const S*>(old_slots), n_old, t);
  }
  static void scatter_fine(cudaStream_t st, BucketSink bsink, FineSink fsink, u32* perr, u64 n_hint) {
    if constexpr (NW == 1) {
      const size_t smem = scatter_fine_smem(fsink.log_f);
      // (the kernel also holds ~9 KB of static shared memory: opt in as soon as the sum passes the 48 KB default)
      if (smem > 36 * 1024) cudaFuncSetAttribute(k_scatter_fine<K, HASH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      // ~20 chunks per CTA: enough for the register prefetch to pay, short enough for the block scheduler to even out
      // the SMs (measured: 4.13 ms with one resident wave of CTAs, 3.50 with 32 waves, 4.09 with one chunk per CTA)
      k_scatter_fine<K, HASH><<<grid_for(n_hint ? n_hint : 1, FINE_ROUND, 148u * KMI_FINE_MINB * 32u), TILE_THREADS, smem, st>>>(bsink, fsink, perr);
    } else {
      const size_t smem = scatter_fine2_smem(fsink.log_f);
      if (smem > 36 * 1024) cudaFuncSetAttribute(k_scatter_fine2<K, HASH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      k_scatter_fine2<K, HASH><<<grid_for(n_hint ? n_hint : 1, FINE_ROUND2, 148u * 4u * 32u), TILE_THREADS, smem, st>>>(bsink, fsink, perr);
    }
  }
  static int insert_segments(cudaStream_t st, TableView t, FineSink fsink, BucketSink bsink, u32 log_seg, u32 n_segs, u32* perr, bool fresh) {
    if constexpr (NW == 1) {
      const size_t smem = segment_smem(log_seg);
      if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_insert_segments<K, HASH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
      }
      const unsigned grid = n_segs;   // one CTA per segment: the block scheduler balances the SMs (4.60 ms vs 4.85 at 64 CTAs per SM)
      k_insert_segments<K, HASH><<<grid, SEG_THREADS, smem, st>>>(t, fsink, bsink, log_seg, n_segs, perr, fresh ? 1u : 0u);
    } else {
      const size_t smem = segment2_smem(log_seg);
      if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_insert_segments2<K, HASH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
      }
      k_insert_segments2<K, HASH><<<n_segs, 384, smem, st>>>(t, fsink, bsink, log_seg, n_segs, perr, fresh ? 1u : 0u);
    }
    return 0;
  }
  static void owner_push(cudaStream_t st, BucketSink bsink, PushSink psink, u64 n_hint) {
    // ~16 chunks per CTA: enough for the key prefetch to pay, short enough for the block scheduler to even out the SMs
    k_owner_push<K, HASH><<<grid_for(n_hint ? n_hint / 16 + 1 : 1, PUSH_ROUND, 148u * 3u * 8u), TILE_THREADS, 0, st>>>(bsink, psink);
  }
  static void insert_overflow(cudaStream_t st, TableView t, BucketSink bsink) {
    k_insert_overflow<K, HASH><<<148u * 4u, 256, 0, st>>>(t, bsink);
  }
  static void lookup(cudaStream_t st, TableView t, const u64* keys, u64 n, u64* vals, uint8_t* found, u64* slot_idx,
                     bool canon) {
    if (n) k_lookup<K, HASH><<<grid_for(n, 256, 148u * 32u), 256, 0, st>>>(t, keys, n, vals, found, slot_idx, canon);
  }
  static void erase(cudaStream_t st, TableView t, const u64* keys, u64 n, u64* vals_removed, bool canon) {
    if (n) k_erase<K, HASH><<<grid_for(n, 256), 256, 0, st>>>(t, keys, n, vals_removed, canon);
  }
  static void dump(cudaStream_t st, const void* slots, u64 n_slots, u64* keys, u64* vals, u64 cap, u64* cursor, int op, u64 x) {
    k_dump<NW><<<grid_for(n_slots, 256), 256, 0, st>>>(reinterpret_cast<const S*>(slots), n_slots, keys, vals, cap, cursor, op, x);
  }
  static void erase_if(cudaStream_t st, void* slots, u64 n_slots, u64* size, u64* removed, int op, u64 x) {
    k_erase_if<NW><<<grid_for(n_slots, 256), 256, 0, st>>>(reinterpret_cast<S*>(slots), n_slots, size, removed, op, x);
  }
  static void checksum(cudaStream_t st, const void* slots, u64 n_slots, u64* out2) {
    k_checksum<NW><<<grid_for(n_slots, 256), 256, 0, st>>>(reinterpret_cast<const S*>(slots), n_slots, out2);
  }
  static void canonicalize(cudaStream_t st, const u64* in, u64* out, u64 n) {
    if (n) k_canonicalize<K><<<grid_for(n, 256), 256, 0, st>>>(in, out, n);
  }
  static void hash_keys(cudaStream_t st, int which, const u64* in, u64* out, u64 n, u32 nranks) {
    if (n) k_hash<K, HASH><<<grid_for(n, 256), 256, 0, st>>>(which, in, out, n, ceil_log2_u32(nranks ? nranks : 1u));
  }
  static void part_hist(cudaStream_t st, const u64* keys, u64 n, u32 p, u32* hist, u64 n_chunks, u32 dist_trans) {
    if (n_chunks) k_part_hist<K, HASH><<<(unsigned)n_chunks, PART_THREADS, 0, st>>>(keys, n, p, hist, n_chunks, dist_trans);
  }
  static void part_scatter(cudaStream_t st, const u64* keys, const u64* vals, u64 n, u32 p, const u32* hist_scanned,
                           const u32* hist_block_sums, u64 n_chunks, u64* keys_out, u64* vals_out, u32 dist_trans) {
    if (n_chunks)
      k_part_scatter<K, HASH><<<(unsigned)n_chunks, PART_THREADS, 0, st>>>(keys, vals, n, p, hist_scanned, hist_block_sums,
                                                                          n_chunks, keys_out, vals_out, dist_trans);
  }
  static void csr_scan_l1(cudaStream_t st, const void* slots, u64 n_slots, u64* offsets, u64* block_sums) {
    k_csr_scan_l1<NW><<<(unsigned)((n_slots + 1023) / 1024), 1024, 0, st>>>(reinterpret_cast<const S*>(slots), n_slots,
                                                                           offsets, block_sums);
  }
  static void csr_fill(cudaStream_t st, TableView t, const u64* keys, const u64* ids, u64 n, const u64* offsets, u64* values) {
    if (n) k_csr_fill<K, HASH><<<grid_for(n, 256, 148u * 32u), 256, 0, st>>>(t, keys, ids, n, offsets, values);
  }
  static void log_filter(cudaStream_t st, TableView t, const u64* keys, const u64* ids, u64 n, u64* keys_out, u64* ids_out,
                         u64* cursor) {
    if (n) k_log_filter<K, HASH><<<grid_for(n, 256), 256, 0, st>>>(t, keys, ids, n, keys_out, ids_out, cursor);
  }
  static void zero_aux(cudaStream_t st, void* slots, u64 n) {
    k_zero_aux<NW><<<grid_for(n, 256), 256, 0, st>>>(reinterpret_cast<S*>(slots), n);
  }
  static Ops make() {
    Ops o;
    o.K = K; o.NW = NW; o.hash = HASH;
    o.fastq_kmers = fastq_kmers; o.scatter_keys = scatter_keys; o.insert_buckets = insert_buckets; o.hll_buckets = hll_buckets;
    o.scatter_fine = scatter_fine; o.insert_segments = insert_segments; o.insert_overflow = insert_overflow; o.owner_push = owner_push; o.fasta_kmers = fasta_kmers; o.table_clear = table_clear; o.insert_keys = insert_keys; o.rehash = rehash;
    o.lookup = lookup; o.erase = erase; o.dump = dump; o.erase_if = erase_if; o.checksum = checksum; o.canonicalize = canonicalize;
    o.hash_keys = hash_keys; o.part_hist = part_hist; o.part_scatter = part_scatter; o.csr_scan_l1 = csr_scan_l1;
    o.csr_fill = csr_fill; o.log_filter = log_filter; o.zero_aux = zero_aux;
    return o;
  }
  // The hash kind only enters the kernels that evaluate the DIST hash (kmi_hash_dev, the partition by owner, the NVLink
  // push): the identity / cpp_std flavours reuse every other launcher of this instantiation and swap those in. They get no
  // owner_push — a prefix of the key as the owner is far too skewed for fixed-capacity receive regions, so their sharded
  // builds take the partition + all-to-all path.
  template <int H2> static Ops make_with_dist_hash() {
    Ops o = make();
    o.hash = H2;
    o.hash_keys = OpsImpl<K, H2>::hash_keys; o.part_hist = OpsImpl<K, H2>::part_hist; o.part_scatter = OpsImpl<K, H2>::part_scatter;
    o.owner_push = nullptr;
    return o;
  }
};

}  // namespace kmi
