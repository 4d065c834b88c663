/* TEST INFRASTRUCTURE — not product code. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this.
 *
 * Plain-C CPU restatement of kmerind's k-mer index construction hot path, one
 * function per reference row, each citing the reference file:line it follows (paths relative to
 * the ParBLiSS/kmerind checkout).
 *
 * Parity pinning: tests/test_oracle_vs_ref.py checks every arithmetic function here against the
 * reference's own compiled headers (oracle/_ref, built by oracle/Makefile from /root/reference),
 * against the reference's golden windows (src/common/test/test_kmer.cpp:393-458) and against the
 * known-answer hash values recorded in SURVEY.md §8c. The record walkers (FASTQ/FASTA) cannot be
 * compiled from the reference without mpi.h; they are pinned by the reference's golden
 * (reads, k-mers, file size) tables of mpi_test_fastq_seq_parse.cpp:446-459 and
 * mpi_test_fasta_seq_parse.cpp:398-403 via tests/golden/.
 *
 * Key layout everywhere: nw = (K<=32 ? 1 : 2) consecutive uint64_t per k-mer, word 0 least
 * significant, first base in the most significant used bits (src/common/kmer.hpp:76-100,177).
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#define KMIO_OK 0
#define KMIO_ERR_NO_AT (-1)      /* fastq_loader.hpp:393,419 "missing @ on first line" */
#define KMIO_ERR_NO_PLUS (-2)    /* fastq_loader.hpp:435 "missing + on third line" */
#define KMIO_ERR_LEN (-3)        /* fastq_loader.hpp:460 "seq and qual differ in length" */
#define KMIO_ERR_CAP (-4)

/* ---- a7: DNA::FROM_ASCII, src/common/alphabets.hpp:139-157 — A/a 0, C/c 1, G/g 2, T/t 3, else 0 */
uint8_t kmio_from_ascii(uint8_t c) {
  switch (c) {
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 0;
  }
}
/* NotEOL, src/utils/file_utils.hpp:43-54 */
static inline int is_eol(uint8_t c) { return c == '\n' || c == '\r'; }
/* NCharFilter, src/io/filtered_sequence_iterator.hpp:429-436 */
static inline int is_n(uint8_t c) { return c == 'N' || c == 'n'; }

int kmio_nwords(int K) { return K <= 32 ? 1 : 2; }

/* ---- a8: Kmer::nextFromChar, src/common/kmer.hpp:731-741,1419-1430 + do_sanitize :1454-1460 */
static inline void roll(uint64_t* w, int nw, int K, uint8_t code) {
  if (nw == 1) {
    uint64_t mask = (K == 32) ? ~0ULL : ((1ULL << (2 * K)) - 1);
    w[0] = ((w[0] << 2) | code) & mask;
  } else {
    uint64_t himask = (K == 64) ? ~0ULL : ((1ULL << (2 * (K - 32))) - 1);
    w[1] = ((w[1] << 2) | (w[0] >> 62)) & himask;
    w[0] = (w[0] << 2) | code;
  }
}

/* ---- a10: Kmer::reverse_complement, src/common/kmer.hpp:1118-1123,1726-1742 — stated per
 * character as the reference's own test helper does (kmer_reverse_helper.hpp). */
void kmio_revcomp(int K, const uint64_t* in, uint64_t* out, size_t n) {
  int nw = kmio_nwords(K);
  for (size_t i = 0; i < n; ++i) {
    uint64_t r[2] = {0, 0};
    const uint64_t* x = in + i * nw;
    for (int j = 0; j < K; ++j) {
      uint64_t c = (x[(2 * j) >> 6] >> ((2 * j) & 63)) & 3;
      int d = 2 * (K - 1 - j);
      r[d >> 6] |= (3 - c) << (d & 63);
    }
    for (int j = 0; j < nw; ++j) out[i * nw + j] = r[j];
  }
}
/* operator<, src/common/kmer.hpp:820-823 → bit_less, most significant word first */
static inline int key_less(const uint64_t* a, const uint64_t* b, int nw) {
  for (int j = nw - 1; j >= 0; --j) {
    if (a[j] != b[j]) return a[j] < b[j];
  }
  return 0;
}
/* lex_less, src/common/kmer_transform.hpp:109-126: (x < rc) ? x : rc */
void kmio_canonical(int K, const uint64_t* in, uint64_t* out, size_t n) {
  int nw = kmio_nwords(K);
  uint64_t rc[2];
  for (size_t i = 0; i < n; ++i) {
    kmio_revcomp(K, in + i * nw, rc, 1);
    const uint64_t* pick = key_less(in + i * nw, rc, nw) ? in + i * nw : rc;
    uint64_t t0 = pick[0], t1 = nw > 1 ? pick[1] : 0;
    out[i * nw] = t0;
    if (nw > 1) out[i * nw + 1] = t1;
  }
}

/* ---- a11: MurmurHash3_x64_128, ext/smhasher/MurmurHash3.cpp:255-332 (fmix64 :81-90), for the
 * key lengths the path uses (len = (2K+7)/8 <= 16), seed 42 (src/index/kmer_hash.hpp:243-276). */
static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
static inline uint64_t fmix64(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33;
  return k;
}
static void murmur3_x64_128(const uint8_t* data, int len, uint32_t seed, uint64_t out[2]) {
  const uint64_t c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
  uint64_t h1 = seed, h2 = seed;
  int nblocks = len / 16;
  for (int i = 0; i < nblocks; ++i) {
    uint64_t k1, k2;
    memcpy(&k1, data + 16 * i, 8); memcpy(&k2, data + 16 * i + 8, 8);
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
    h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
    k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
    h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
  }
  const uint8_t* tail = data + nblocks * 16;
  int t = len & 15;
  uint64_t k1 = 0, k2 = 0;
  for (int j = t - 1; j >= 8; --j) k2 ^= (uint64_t)tail[j] << (8 * (j - 8));
  if (t > 8) { k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2; }
  for (int j = (t > 8 ? 8 : t) - 1; j >= 0; --j) k1 ^= (uint64_t)tail[j] << (8 * j);
  if (t > 0) { k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1; }
  h1 ^= (uint64_t)len; h2 ^= (uint64_t)len;
  h1 += h2; h2 += h1;
  h1 = fmix64(h1); h2 = fmix64(h2);
  h1 += h2; h2 += h1;
  out[0] = h1; out[1] = h2;
}

/* ---- a12: farmhashna::Hash64WithSeed for len <= 16, ext/farmhash/src/farmhash.cc:523-529,
 * HashLen0to16 :388-410, HashLen16 :378-386, Hash128to64 farmhash.h:129-138, DebugTweak :334-343 */
static const uint64_t FK0 = 0xc3a5c85c97cb3127ULL, FK1 = 0xb492b66fbe98f273ULL, FK2 = 0x9ae16a3b2f90404fULL;
static inline uint64_t rotr64(uint64_t x, int r) { return r == 0 ? x : (x >> r) | (x << (64 - r)); }
static inline uint64_t hashlen16_mul(uint64_t u, uint64_t v, uint64_t mul) {
  uint64_t a = (u ^ v) * mul; a ^= a >> 47;
  uint64_t b = (v ^ a) * mul; b ^= b >> 47; b *= mul;
  return b;
}
static uint64_t farm_len0to16(const uint8_t* s, size_t len) {
  if (len >= 8) {
    uint64_t mul = FK2 + len * 2, a, b;
    memcpy(&a, s, 8); memcpy(&b, s + len - 8, 8);
    a += FK2;
    uint64_t c = rotr64(b, 37) * mul + a;
    uint64_t d = (rotr64(a, 25) + b) * mul;
    return hashlen16_mul(c, d, mul);
  }
  if (len >= 4) {
    uint64_t mul = FK2 + len * 2;
    uint32_t a32, b32;
    memcpy(&a32, s, 4); memcpy(&b32, s + len - 4, 4);
    uint64_t a = a32;
    return hashlen16_mul(len + (a << 3), b32, mul);
  }
  if (len > 0) {
    uint8_t a = s[0], b = s[len >> 1], c = s[len - 1];
    uint32_t y = (uint32_t)a + ((uint32_t)b << 8);
    uint32_t z = (uint32_t)len + ((uint32_t)c << 2);
    uint64_t v = (y * FK2) ^ (z * FK0);
    v ^= v >> 47;
    return v * FK2;
  }
  return FK2;
}
static uint64_t farm64_seed(const uint8_t* s, size_t len, uint64_t seed, int debug_tweak) {
  /* Hash64WithSeeds(s,len,k2,seed) = HashLen16(Hash64(s,len) - k2, seed) = Hash128to64(lo,hi) */
  uint64_t h = hashlen16_mul(farm_len0to16(s, len) - FK2, seed, 0x9ddfea08eb382d69ULL);
  if (debug_tweak) h = ~__builtin_bswap64(h * FK1);
  return h;
}

/* which: 0 murmur store h[0], 1 murmur dist h[1] (kmer_hash.hpp:257-274);
 *        2 farm store seed 42, 3 farm dist seed (42<<1)-1 = 83 (kmer_hash.hpp:304-309);
 *        4/5 = 2/3 as built WITHOUT NDEBUG (DebugTweak). */
void kmio_hash(int K, int which, const uint64_t* in, uint64_t* out, size_t n) {
  int nw = kmio_nwords(K);
  int nbytes = (2 * K + 7) / 8;
  for (size_t i = 0; i < n; ++i) {
    const uint8_t* p = (const uint8_t*)(in + i * nw);
    if (which <= 1) {
      uint64_t h[2];
      murmur3_x64_128(p, nbytes, 42, h);
      out[i] = h[which];
    } else {
      int dist = (which & 1);
      out[i] = farm64_seed(p, (size_t)nbytes, dist ? 83 : 42, which >= 4);
    }
  }
}
/* f3: the identity and cpp_std flavours (src/index/kmer_hash.hpp:157-232), built the way KeyToRank builds its DistHash:
 * prefix bits = ceilLog2(comm_size) (distributed_densehash_map.hpp:153-156).
 * kind 0 identity store = Kmer::getSuffix(min(nBits,64)) (kmer.hpp:1245-1249: data[0] masked to the low bits);
 *      1 identity dist  = Kmer::getPrefix(min(nBits,pbits)) (kmer.hpp:1203-1221: top word >> (64 - bits - padBits), or,
 *        when the prefix is longer than the top word's used bits, the whole k-mer >> (nBits - bits), low `bits` bits);
 *      2 cpp_std store  = XOR over the 64-bit words of (std::hash<size_t>(w) << 1), std::hash<size_t> = identity (libstdc++);
 *      3 cpp_std dist   = that >> (min(nBits,64) - min(pbits,nBits)) (kmer_hash.hpp:170,196-199). */
void kmio_hash_prefix(int K, int kind, unsigned pbits, const uint64_t* in, uint64_t* out, size_t n) {
  int nw = kmio_nwords(K);
  unsigned nbits = 2u * (unsigned)K, pad = 64u * (unsigned)nw - nbits;
  unsigned bits = pbits < nbits ? pbits : nbits;
  for (size_t i = 0; i < n; ++i) {
    const uint64_t* w = in + i * nw;
    if (kind == 0) {
      out[i] = (nw == 1 && nbits < 64) ? (w[0] & ((1ULL << nbits) - 1ULL)) : w[0];
    } else if (kind == 1) {
      if (!bits) out[i] = 0ULL;
      else if (pad + bits <= 64u) out[i] = w[nw - 1] >> (64u - bits - pad);
      else {                                      /* the prefix runs into the low word: shift right by nBits - bits, keep `bits` bits */
        unsigned sh = nbits - bits;               /* 0 < sh < 64 */
        out[i] = ((w[0] >> sh) | (w[1] << (64u - sh))) & (bits >= 64u ? ~0ULL : ((1ULL << bits) - 1ULL));
      }
    } else {
      uint64_t h = 0;
      for (int j = 0; j < nw; ++j) h ^= (w[j] << 1);
      if (kind == 3) {
        unsigned sh = (nbits < 64u ? nbits : 64u) - bits;
        h = sh >= 64u ? 0ULL : (h >> sh);   /* the reference's shift by 64 is undefined; only reachable at p = 1, owner 0 */
      }
      out[i] = h;
    }
  }
}
/* a13: KeyToRank, src/containers/distributed_densehash_map.hpp:148-170: dist_hash % p */
void kmio_owner(const uint64_t* dist_hash, size_t n, uint32_t p, uint32_t* out) {
  for (size_t i = 0; i < n; ++i) out[i] = (uint32_t)(dist_hash[i] % p);
}

/* ---- a8/a9/a5/a6: k-mers of one sequence [b,e) of the file buffer.
 * KmerSlidingWindow over NotEOL-filtered, ASCII2<DNA>-mapped chars
 * (src/common/kmer_iterators.hpp:82-100; src/io/kmer_parser.hpp:963-966,1080). EOL bytes inside
 * [b,e) are skipped (FASTA multi-line). nsplit: NSplitSequencesIterator semantics
 * (src/io/filtered_sequence_iterator.hpp:225-236,429-440) == no window may contain N/n.
 * id of a k-mer = id_base + raw byte offset of its first base (kmer_parser.hpp:378-399).
 * Returns the number emitted, or KMIO_ERR_CAP. */
static long emit_seq(const uint8_t* buf, size_t b, size_t e, int K, int nw, int nsplit,
                     uint64_t id_base, uint64_t id_origin, uint64_t* keys, uint64_t* ids,
                     size_t have, size_t cap) {
  uint64_t w[2] = {0, 0};
  size_t* first = (size_t*)malloc(sizeof(size_t) * (size_t)K); /* ring of raw offsets */
  int run = 0;
  long m = 0;
  size_t nb = 0;
  for (size_t i = b; i < e; ++i) {
    uint8_t c = buf[i];
    if (is_eol(c)) continue;
    if (nsplit && is_n(c)) { run = 0; continue; }
    roll(w, nw, K, kmio_from_ascii(c));
    first[nb % (size_t)K] = i; ++nb;
    if (++run >= K) {
      if (have + (size_t)m >= cap) { free(first); return KMIO_ERR_CAP; }
      size_t at = have + (size_t)m;
      if (keys) { keys[at * nw] = w[0]; if (nw > 1) keys[at * nw + 1] = w[1]; }
      if (ids) ids[at] = id_base + (first[(nb - (size_t)K) % (size_t)K] - id_origin);
      ++m;
    }
  }
  free(first);
  return m;
}

/* ---- a3: SequentialFASTQParser::get_next_record, src/io/fastq_loader.hpp:389-470 with
 * findNonEOL/findEOL of src/io/file_loader.hpp:146-172, driven over the whole buffer as
 * SequencesIterator does (p = 1: valid range == file, kmer_parser.hpp:114-157 clips nothing).
 * file_offset = offset of buf[0] in the file (for ids). Position id = ShortSequenceKmerId:
 * (record_start << 16) | offset_in_record (src/common/sequence.hpp:103-160).
 * seq_off (optional, cap_seq+... ) receives [begin,end) of every sequence line. */
long kmio_fastq_kmers(const uint8_t* buf, size_t n, uint64_t file_offset, int K, int canonical,
                      int nsplit, uint64_t* keys, uint64_t* ids, size_t cap, uint64_t* n_reads) {
  int nw = kmio_nwords(K);
  size_t i = 0, total = 0;
  uint64_t reads = 0;
  while (i < n && is_eol(buf[i])) ++i;
  while (i < n) {
    if (buf[i] != '@') return KMIO_ERR_NO_AT;
    size_t rec = i;
    while (i < n && !is_eol(buf[i])) ++i;           /* line 1 */
    while (i < n && is_eol(buf[i])) ++i;
    size_t sb = i;
    while (i < n && !is_eol(buf[i])) ++i;           /* line 2 */
    size_t se = i;
    while (i < n && is_eol(buf[i])) ++i;
    if (i < n && buf[i] != '+') return KMIO_ERR_NO_PLUS;
    while (i < n && !is_eol(buf[i])) ++i;           /* line 3 */
    while (i < n && is_eol(buf[i])) ++i;
    size_t qb = i;
    while (i < n && !is_eol(buf[i])) ++i;           /* line 4 */
    size_t qe = i;
    while (i < n && is_eol(buf[i])) ++i;
    if (sb != se && qb != qe && (se - sb) != (qe - qb)) return KMIO_ERR_LEN;
    /* truncated record (sb==se or qb==qe): warned and returned as is (:453-459) */
    ++reads;
    uint64_t idb = ((file_offset + rec) & 0xFFFFFFFFFFULL) << 16;
    long m = emit_seq(buf, sb, se, K, nw, nsplit, idb, rec, keys, ids, total, cap);
    if (m < 0) return m;
    total += (size_t)m;
  }
  if (canonical && keys) kmio_canonical(K, keys, keys, total);
  if (n_reads) *n_reads = reads;
  return (long)total;
}

/* ---- f4: k-mer quality scores of the PositionQualityIndex (KmerPositionQualityTupleParser, src/io/kmer_parser.hpp:578-760;
 * QualityScoreSlidingWindow, src/index/quality_score_iterator.hpp:60-140; Illumina18QualityScoreCodec<float> =
 * QualityScoreCodec<float, 33, 126, 0>, src/index/quality_scores.hpp:60-170,529). One float per k-mer, in the order
 * kmio_fastq_kmers emits its tuples: DecodeLUT[c - 33] = log2(1 - 10^(-q/10)) (the table's literals, reproduced with log1p and
 * rounded to float: checked against the reference's own table in tests/test_oracle.py); a base is "correct" when its value
 * lies strictly between LUT[0] (lowest) and LUT[95] (0.0), i.e. Phred 1..93; the window keeps a running float sum, adding
 * the entering and subtracting the leaving value IN THAT ORDER OF OPERATIONS (init sums the first K left to right; next()
 * subtracts the old value, then adds the new one), and a count of incorrect bases; value = 0 if any base of the window is
 * incorrect, else exp2f(sum). With nsplit every N-free run restarts the window (the sub-sequence's qual_begin is advanced
 * by the same distance as its seq_begin, kmer_parser.hpp:650-654). A record without a quality line yields 0. */
#include <math.h>
static float kmio_qlut[96];
static int kmio_qlut_ready = 0;
void kmio_quality_lut(float* out96) {
  if (!kmio_qlut_ready) {
    kmio_qlut[0] = -3.402823466e+38F;                                   /* std::numeric_limits<float>::lowest() */
    for (int q = 1; q <= 93; ++q) kmio_qlut[q] = (float)(log1p(-pow(10.0, -(double)q / 10.0)) / log(2.0));
    kmio_qlut[94] = kmio_qlut[95] = 0.0f;
    kmio_qlut_ready = 1;
  }
  if (out96) memcpy(out96, kmio_qlut, sizeof(kmio_qlut));
}
static float kmio_qdecode(uint8_t c, int* ok) {
  if (c < 33 || c > 128) { *ok = 0; return 0.0f; }                     /* outside the table: the reference reads out of bounds */
  float v = kmio_qlut[c - 33];
  *ok = (v > kmio_qlut[0]) && (v < kmio_qlut[95]);
  return v;
}
/* windows of one run: bases [b, e) of the sequence line, their quality bytes at q[0 .. e-b) */
static long kmio_quality_run(const uint8_t* q, size_t len, size_t q_avail, int K, float* out, size_t have, size_t cap) {
  if (len < (size_t)K) return 0;
  long m = 0;
  float sum = 0.0f;
  int bad = 0;
  for (size_t t = 0; t + (size_t)K <= len; ++t) {
    if (t == 0) {
      for (int j = 0; j < K; ++j) { int ok; float v = ((size_t)j < q_avail) ? kmio_qdecode(q[j], &ok) : (ok = 0, 0.0f); if (ok) sum += v; else ++bad; }
    } else {
      int ok_o, ok_n;
      float o = (t - 1 < q_avail) ? kmio_qdecode(q[t - 1], &ok_o) : (ok_o = 0, 0.0f);
      float v = (t - 1 + (size_t)K < q_avail) ? kmio_qdecode(q[t - 1 + (size_t)K], &ok_n) : (ok_n = 0, 0.0f);
      if (ok_o) sum -= o; else --bad;
      if (ok_n) sum += v; else ++bad;
    }
    if (have + (size_t)m >= cap) return KMIO_ERR_CAP;
    out[have + (size_t)m] = bad > 0 ? 0.0f : exp2f(sum);
    ++m;
  }
  return m;
}
long kmio_fastq_quality(const uint8_t* buf, size_t n, int K, int nsplit, float* quals, size_t cap) {
  kmio_quality_lut(NULL);
  size_t i = 0, total = 0;
  while (i < n && is_eol(buf[i])) ++i;
  while (i < n) {
    if (buf[i] != '@') return KMIO_ERR_NO_AT;
    while (i < n && !is_eol(buf[i])) ++i;
    while (i < n && is_eol(buf[i])) ++i;
    size_t sb = i;
    while (i < n && !is_eol(buf[i])) ++i;
    size_t se = i;
    while (i < n && is_eol(buf[i])) ++i;
    if (i < n && buf[i] != '+') return KMIO_ERR_NO_PLUS;
    while (i < n && !is_eol(buf[i])) ++i;
    while (i < n && is_eol(buf[i])) ++i;
    size_t qb = i;
    while (i < n && !is_eol(buf[i])) ++i;
    size_t qe = i;
    while (i < n && is_eol(buf[i])) ++i;
    if (sb != se && qb != qe && (se - sb) != (qe - qb)) return KMIO_ERR_LEN;
    size_t b = sb;
    while (b < se) {
      size_t e = se;
      if (nsplit) {
        while (b < se && is_n(buf[b])) ++b;
        e = b;
        while (e < se && !is_n(buf[e])) ++e;
      }
      size_t qoff = b - sb, qav = (qe - qb) > qoff ? (qe - qb) - qoff : 0;
      long m = kmio_quality_run(buf + qb + qoff, e - b, qav, K, quals, total, cap);
      if (m < 0) return m;
      total += (size_t)m;
      b = e;
    }
  }
  return (long)total;
}

/* ---- a2: FASTQParser::find_first_record (serial), src/io/fastq_loader.hpp:269-357: first record start inside the
 * search range [s, e) of a file whose bytes [mem, mem + n) are in buf; parent = [ps, pe) is the enclosing
 * record-aligned range. Walks to the first line start, takes the first characters of four consecutive non-empty
 * lines and votes @?+? / ?@?+ / +?@? / ?+?@ (:337-340). Returns 0 and the offset (e when the range holds no complete
 * record and ends with the parent), -1 when no marker pattern is found inside a range that is not the parent's tail
 * (the reference throws IOException there, :342-357). */
int kmio_fastq_find_first_record(const uint8_t* buf, uint64_t mem, size_t n, uint64_t s, uint64_t e, uint64_t ps,
                                 uint64_t pe, uint64_t* out) {
  uint64_t ts = s > mem ? s : mem, te = e < mem + n ? e : mem + n;
  uint64_t i = ts;
  uint8_t first[4] = {0, 0, 0, 0};
  uint64_t off[4];
  if (te < ts) te = ts;
  off[0] = off[1] = off[2] = off[3] = te;
  if (i == te) { *out = te; return 0; }
  if (ts > ps && !is_eol(buf[i - mem])) { while (i < te && !is_eol(buf[i - mem])) ++i; }   /* findEOL */
  if (i == te) { *out = te; return 0; }
  while (i < te && is_eol(buf[i - mem])) ++i;                                              /* findNonEOL */
  if (i == te) { *out = te; return 0; }
  first[0] = buf[i - mem]; off[0] = i;
  for (int j = 1; j < 4; ++j) {
    while (i < te && !is_eol(buf[i - mem])) ++i;
    while (i < te && is_eol(buf[i - mem])) ++i;
    off[j] = i;
    if (i != te) first[j] = buf[i - mem];
  }
  if (i == te) { *out = te; return 0; }
  if (first[0] == '@' && first[2] == '+') { *out = off[0]; return 0; }
  if (first[1] == '@' && first[3] == '+') { *out = off[1]; return 0; }
  if (first[0] == '+' && first[2] == '@') { *out = off[2]; return 0; }
  if (first[1] == '+' && first[3] == '@') { *out = off[3]; return 0; }
  if (te == pe) { *out = te; return 0; }
  return -1;
}

/* ---- a4: FASTAParser::init_parser (serial), src/io/fasta_loader.hpp:485-616, and
 * get_next_record :618-723 at p = 1. Line starts follow '\n' only; a line is a header iff it
 * starts with '>' or ';'; runs of same-kind lines are merged; a sequence is everything from the
 * first non-header line start to the next header line start (or file end). seq ids count from 0
 * when the file starts with a header, from 1 otherwise (:571-590, the k/2 arithmetic).
 * Position id = LongSequenceKmerId: (seq_id << 40) | file byte of first base
 * (src/common/sequence.hpp:180-236). */
long kmio_fasta_kmers(const uint8_t* buf, size_t n, int K, int canonical, int nsplit,
                      uint64_t* keys, uint64_t* ids, size_t cap, uint64_t* n_seqs) {
  int nw = kmio_nwords(K);
  size_t total = 0;
  uint64_t seqs = 0;
  if (n == 0) { if (n_seqs) *n_seqs = 0; return 0; }
  /* walk the merged (line-start, kind) list without materialising it */
  size_t k = 0;           /* index in the unique'd list */
  int prev_kind = -1;
  size_t pos = 0, pos2 = 0;
  int have_pos = 0, have_pos2 = 0;
  int first_is_header = (buf[0] == '>' || buf[0] == ';');
  for (size_t i = 0; i <= n; ++i) {
    int kind;
    if (i == n) kind = 1;                                   /* sentinel (parentRange.end, 1) */
    else if (i == 0 || buf[i - 1] == '\n') kind = (buf[i] == '>' || buf[i] == ';') ? 1 : 0;
    else continue;
    if (kind == prev_kind) continue;                        /* std::unique on .second */
    prev_kind = kind;
    size_t idx = k++;
    if (idx == 0 && !first_is_header) continue;             /* skip a leading non-header run */
    if (kind == 1) {
      if (have_pos && have_pos2) {
        /* sequence (pos, pos2, i, id) ; id = (list index of pos2)/2 */
        size_t list_idx_pos2 = idx - 1;
        uint64_t sid = list_idx_pos2 / 2;
        ++seqs;
        uint64_t idb = ((sid & 0xFFFF) << 40);
        long m = emit_seq(buf, pos2, i, K, nw, nsplit, idb, 0, keys, ids, total, cap);
        if (m < 0) return m;
        total += (size_t)m;
      }
      pos = i; have_pos = 1; have_pos2 = 0;
    } else {
      pos2 = i; have_pos2 = have_pos;
    }
  }
  (void)pos;
  if (canonical && keys) kmio_canonical(K, keys, keys, total);
  if (n_seqs) *n_seqs = seqs;
  return (long)total;
}

/* ---- a15: reduction/counting map contents, distributed_densehash_map.hpp:2463-2482: one entry
 * per distinct key, value = number of occurrences as uint32_t (std::plus, wraps at 2^32).
 * Produces the sorted dump directly: keys ascending as multi-word unsigned integers. In place. */
static int g_nw = 1;
static int cmp_keys(const void* a, const void* b) {
  const uint64_t* x = (const uint64_t*)a; const uint64_t* y = (const uint64_t*)b;
  for (int j = g_nw - 1; j >= 0; --j) if (x[j] != y[j]) return x[j] < y[j] ? -1 : 1;
  return 0;
}
size_t kmio_count_sorted(int K, uint64_t* keys, size_t n, uint32_t* counts) {
  int nw = kmio_nwords(K);
  g_nw = nw;
  qsort(keys, n, sizeof(uint64_t) * (size_t)nw, cmp_keys);
  size_t m = 0;
  for (size_t i = 0; i < n;) {
    size_t j = i + 1;
    while (j < n && cmp_keys(keys + i * nw, keys + j * nw) == 0) ++j;
    for (int w = 0; w < nw; ++w) keys[m * nw + w] = keys[i * nw + w];
    counts[m] = (uint32_t)(j - i);
    ++m; i = j;
  }
  return m;
}
/* a17: multimap contents = every (key,id) tuple kept; sorted dump by (key, id). In place. */
typedef struct { uint64_t k[2]; uint64_t id; } kmio_tuple;
static int cmp_tuple(const void* a, const void* b) {
  const kmio_tuple* x = (const kmio_tuple*)a; const kmio_tuple* y = (const kmio_tuple*)b;
  if (x->k[1] != y->k[1]) return x->k[1] < y->k[1] ? -1 : 1;
  if (x->k[0] != y->k[0]) return x->k[0] < y->k[0] ? -1 : 1;
  if (x->id != y->id) return x->id < y->id ? -1 : 1;
  return 0;
}
void kmio_sort_tuples(int K, uint64_t* keys, uint64_t* ids, size_t n) {
  int nw = kmio_nwords(K);
  kmio_tuple* t = (kmio_tuple*)malloc(sizeof(kmio_tuple) * (n ? n : 1));
  for (size_t i = 0; i < n; ++i) {
    t[i].k[0] = keys[i * nw]; t[i].k[1] = nw > 1 ? keys[i * nw + 1] : 0; t[i].id = ids[i];
  }
  qsort(t, n, sizeof(kmio_tuple), cmp_tuple);
  for (size_t i = 0; i < n; ++i) {
    keys[i * nw] = t[i].k[0]; if (nw > 1) keys[i * nw + 1] = t[i].k[1]; ids[i] = t[i].id;
  }
  free(t);
}

/* ---- Synthetic reads (this repo's own generator, identical on host and device:
 * kmerind_b200/csrc/kmi_synth.cuh). Counter-based splitmix64; nothing is materialised but the text. */
static inline uint64_t sm64(uint64_t x) {
  uint64_t z = x + 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
#define SYN_HDR 13
static inline uint8_t syn_genome_base(uint64_t seed, uint64_t i) {
  return (uint8_t)((sm64(seed * 0x100000001b3ULL + (i >> 5)) >> (2 * (i & 31))) & 3);
}
/* byte `o` (0..rec-1) of record `r`; L = read length, G = genome length, err_ppm / n_ppm per base */
static inline uint8_t syn_byte(uint64_t r, uint32_t o, uint32_t L, uint64_t G, uint64_t seed,
                               uint32_t err_ppm, uint32_t n_ppm) {
  if (o < SYN_HDR) {
    if (o == 0) return '@';
    if (o == 1) return 'S';
    if (o == SYN_HDR - 1) return '\n';
    uint64_t d = r;
    for (uint32_t k = SYN_HDR - 2; k > o; --k) d /= 10;
    return (uint8_t)('0' + d % 10);
  }
  o -= SYN_HDR;
  if (o < L) {
    uint64_t h = sm64((seed + 1) * 0x9e3779b97f4a7c15ULL + r);
    uint64_t start = (h >> 1) % (G - L + 1);
    int rev = (int)(h & 1);
    uint8_t b = rev ? (uint8_t)(3 - syn_genome_base(seed, start + (L - 1 - o)))
                    : syn_genome_base(seed, start + o);
    uint64_t e = sm64((seed + 2) * 0xd1b54a32d192ed03ULL + r * L + o);
    if ((uint32_t)(e % 1000000u) < err_ppm) b = (uint8_t)((b + 1 + ((e >> 40) % 3)) & 3);
    if ((uint32_t)((e >> 20) % 1000000u) < n_ppm) return 'N';
    return (uint8_t)"ACGT"[b];
  }
  o -= L;
  if (o == 0) return '\n';
  if (o == 1) return '+';
  if (o == 2) return '\n';
  o -= 3;
  if (o < L) return 'I';
  return '\n';
}
size_t kmio_synth_record_bytes(uint32_t L) { return SYN_HDR + 2 * (size_t)L + 4; }
void kmio_synth_fastq(uint8_t* out, uint64_t first_read, uint64_t n_reads, uint32_t L, uint64_t G,
                      uint64_t seed, uint32_t err_ppm, uint32_t n_ppm) {
  size_t rec = kmio_synth_record_bytes(L);
  for (uint64_t r = 0; r < n_reads; ++r)
    for (uint32_t o = 0; o < rec; ++o)
      out[r * rec + o] = syn_byte(first_read + r, o, L, G, seed, err_ppm, n_ppm);
}
/* synthetic FASTA genome (config 5): n_seq sequences of seq_len bases, `width` bases per line,
 * header ">chrNN\n" (7 bytes). Returns bytes written (call with out==NULL to size). */
size_t kmio_synth_fasta(uint8_t* out, uint32_t n_seq, uint64_t seq_len, uint32_t width, uint64_t seed) {
  size_t w = 0;
  for (uint32_t s = 0; s < n_seq; ++s) {
    const char hdr[7] = {'>', 'c', 'h', 'r', (char)('0' + (s / 10) % 10), (char)('0' + s % 10), '\n'};
    for (int j = 0; j < 7; ++j) { if (out) out[w] = (uint8_t)hdr[j]; ++w; }
    for (uint64_t i = 0; i < seq_len; ++i) {
      if (out) out[w] = (uint8_t)"ACGT"[syn_genome_base(seed + 977 * (s + 1), i)];
      ++w;
      if ((i + 1) % width == 0 || i + 1 == seq_len) { if (out) out[w] = '\n'; ++w; }
    }
  }
  return w;
}

/* ---- CPU baseline ("port"): the reference's single-rank CountIndex build restated end to end —
 * walk records (a3), roll the window (a8), lex_less (a10), store hash (a11/a12), insert into an
 * open-addressing table with google::dense_hash_map's parameters: power-of-two buckets starting
 * at 128 (densehash_map.hpp:960), triangular probing (sparsehash internal/densehashtable.h:110),
 * grow x2 at load 0.7 (densehash_map.hpp:969-970), count += 1 on a hit
 * (distributed_densehash_map.hpp:2463-2482). K <= 32 only. Returns distinct keys. */
typedef struct { uint64_t key; uint32_t cnt; uint32_t pad; } kmio_slot;
static void tbl_insert(kmio_slot* t, size_t cap, uint64_t key, uint32_t add, int hk, size_t* sz) {
  uint64_t h; kmio_hash(31, hk, &key, &h, 1);
  size_t b = (size_t)h & (cap - 1), probes = 0;
  for (;;) {
    if (t[b].key == 0x8000000000000000ULL) { t[b].key = key; t[b].cnt = add; ++*sz; return; }
    if (t[b].key == key) { t[b].cnt += add; return; }
    ++probes; b = (b + probes) & (cap - 1);
  }
}
size_t kmio_count_build_port(const uint8_t* buf, size_t n, int K, int hash_kind /*0 murmur,2 farm*/,
                             uint64_t* out_keys, uint32_t* out_counts, size_t out_cap, long* n_kmers) {
  if (K > 32) return 0;
  size_t cap = 128, sz = 0;
  kmio_slot* t = (kmio_slot*)malloc(cap * sizeof(kmio_slot));
  for (size_t i = 0; i < cap; ++i) { t[i].key = 0x8000000000000000ULL; t[i].cnt = 0; }
  size_t i = 0; long total = 0;
  uint64_t mask = (K == 32) ? ~0ULL : ((1ULL << (2 * K)) - 1);
  while (i < n && is_eol(buf[i])) ++i;
  while (i < n) {
    if (buf[i] != '@') break;
    while (i < n && !is_eol(buf[i])) ++i;
    while (i < n && is_eol(buf[i])) ++i;
    size_t sb = i; while (i < n && !is_eol(buf[i])) ++i; size_t se = i;
    while (i < n && is_eol(buf[i])) ++i;
    while (i < n && !is_eol(buf[i])) ++i;
    while (i < n && is_eol(buf[i])) ++i;
    while (i < n && !is_eol(buf[i])) ++i;
    while (i < n && is_eol(buf[i])) ++i;
    uint64_t w = 0;
    for (size_t j = sb; j < se; ++j) {
      w = ((w << 2) | kmio_from_ascii(buf[j])) & mask;
      if (j + 1 - sb >= (size_t)K) {
        uint64_t rc; kmio_revcomp(K, &w, &rc, 1);
        uint64_t c = w < rc ? w : rc;
        if ((double)(sz + 1) > 0.7 * (double)cap) {
          size_t ncap = cap * 2;
          kmio_slot* nt = (kmio_slot*)malloc(ncap * sizeof(kmio_slot));
          for (size_t q = 0; q < ncap; ++q) { nt[q].key = 0x8000000000000000ULL; nt[q].cnt = 0; }
          size_t nsz = 0;
          for (size_t q = 0; q < cap; ++q)
            if (t[q].key != 0x8000000000000000ULL) tbl_insert(nt, ncap, t[q].key, t[q].cnt, hash_kind, &nsz);
          free(t); t = nt; cap = ncap;
        }
        tbl_insert(t, cap, c, 1, hash_kind, &sz);
        ++total;
      }
    }
  }
  if (out_keys) {
    size_t wr = 0;
    for (size_t q = 0; q < cap && wr < out_cap; ++q)
      if (t[q].key != 0x8000000000000000ULL) { out_keys[wr] = t[q].key; out_counts[wr] = t[q].cnt; ++wr; }
  }
  free(t);
  if (n_kmers) *n_kmers = total;
  return sz;
}
