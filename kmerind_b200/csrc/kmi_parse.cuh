// kmi_parse.cuh — record-boundary finding, 2-bit encoding and k-mer generation (sm_100a).
//
// Replaces, for whole record blocks resident in HBM, the reference's per-byte iterator stack:
//   SequentialFASTQParser::get_next_record      src/io/fastq_loader.hpp:389-470       (a3)
//   NotEOL filter + ASCII2<DNA>                 src/utils/file_utils.hpp:43-54,
//                                               src/common/alphabets.hpp:139-157      (a7)
//   KmerSlidingWindow / Kmer::nextFromChar      src/common/kmer_iterators.hpp:82-100,
//                                               src/common/kmer.hpp:731-741           (a8)
//   NSplitSequencesIterator                     src/io/filtered_sequence_iterator.hpp:225-236 (a5)
//   lex_less                                    src/common/kmer_transform.hpp:109-126 (a10)
//
// Layout: the text is cut into 4 KiB tiles of 256 16-byte groups; one thread owns one group and
// loads it with a single coalesced 128-bit load. FASTQ structure is recovered exactly as the
// reference walks it — a record is four consecutive NON-EMPTY lines (findNonEOL skips every
// '\n'/'\r', so blank lines vanish) — by numbering line starts globally: pass A counts line
// starts per tile, a scan turns that into each tile's first line number, pass B classifies
// every byte (line number mod 4 == 1 -> sequence line). Per tile the 2-bit codes (32 bits per
// group, first base most significant) and a one-bit-per-byte "is a base of a sequence line"
// flag live in shared memory; a k-mer starts at byte p iff flags p..p+K-1 are all set, and its
// value is a funnel-shift of the packed codes. Valid start positions are compacted into a
// shared work list so that the expensive part (reverse complement, min, hash, sink) runs with
// all lanes busy and the key stream is written coalesced.
#pragma once
#include "kmi_bucket.cuh"

namespace kmi {

// ---- byte-parallel classification of one 16-byte group ------------------------------------
// zb: 0x80 in every byte of x that is zero (exact, no carries between bytes); nib80 gathers those four
// flag bits into a nibble, byte 0 -> bit 0.
__device__ __forceinline__ u32 zb(u32 x) { return ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu); }
__device__ __forceinline__ u32 nib80(u32 z) { return ((z >> 7) * 0x01020408u) >> 24; }
__device__ __forceinline__ u32 mask16(u32 a, u32 b, u32 c, u32 d) {
  return nib80(a) | (nib80(b) << 4) | (nib80(c) << 8) | (nib80(d) << 12);
}
__device__ __forceinline__ u32 eq4(u32 w, u32 c4) { return zb(w ^ c4); }
// 2-bit codes of 4 bytes packed into 8 bits, byte 0 in the two most significant bits.
// DNA::FROM_ASCII: A/a 0, C/c 1, G/g 2, T/t 3, everything else 0.
__device__ __forceinline__ u32 codes4(u32 w) {
  const u32 f = w & 0xDFDFDFDFu;  // fold case: only 'c','g','t' fold onto 'C','G','T'
  const u32 zc = eq4(f, 0x43434343u), zg = eq4(f, 0x47474747u), zt = eq4(f, 0x54545454u);
  const u32 c = ((zc | zt) >> 7) | ((zg | zt) >> 6);
  return (c * 0x40100401u) >> 24;
}
struct Group {
  u32 codes;  // 16 bases, first base in bits 31..30
  u32 eol;    // bit b: byte b is '\n' or '\r'
  u32 nl;     // bit b: byte b is '\n'
  u32 nmask;  // bit b: byte b is 'N' or 'n'
};
__device__ __forceinline__ u32 eol16(const uint4& v) {
  return mask16(eq4(v.x, 0x0A0A0A0Au) | eq4(v.x, 0x0D0D0D0Du), eq4(v.y, 0x0A0A0A0Au) | eq4(v.y, 0x0D0D0D0Du),
                eq4(v.z, 0x0A0A0A0Au) | eq4(v.z, 0x0D0D0D0Du), eq4(v.w, 0x0A0A0A0Au) | eq4(v.w, 0x0D0D0D0Du));
}
__device__ __forceinline__ u32 codes16(const uint4& v) {
  return (codes4(v.x) << 24) | (codes4(v.y) << 16) | (codes4(v.z) << 8) | codes4(v.w);
}
__device__ __forceinline__ u32 nmask16(const uint4& v) {
  return mask16(eq4(v.x & 0xDFDFDFDFu, 0x4E4E4E4Eu), eq4(v.y & 0xDFDFDFDFu, 0x4E4E4E4Eu),
                eq4(v.z & 0xDFDFDFDFu, 0x4E4E4E4Eu), eq4(v.w & 0xDFDFDFDFu, 0x4E4E4E4Eu));
}
__device__ __forceinline__ Group classify(const uint4& v) {
  Group g;
  g.codes = codes16(v);
  g.nl = mask16(eq4(v.x, 0x0A0A0A0Au), eq4(v.y, 0x0A0A0A0Au), eq4(v.z, 0x0A0A0A0Au), eq4(v.w, 0x0A0A0A0Au));
  g.eol = g.nl | mask16(eq4(v.x, 0x0D0D0D0Du), eq4(v.y, 0x0D0D0D0Du), eq4(v.z, 0x0D0D0D0Du), eq4(v.w, 0x0D0D0D0Du));
  g.nmask = nmask16(v);
  return g;
}
// 16 bytes at offset `off` of a 16-byte aligned buffer of n bytes; bytes past the end read as '\n'
__device__ __forceinline__ uint4 load_group(const uint8_t* __restrict__ bytes, u64 off, u64 n) {
  if (off + GROUP_BYTES <= n) {
    return __ldg(reinterpret_cast<const uint4*>(bytes + off));
  }
  u32 w[4] = {0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au, 0x0A0A0A0Au};
  for (int b = 0; b < GROUP_BYTES; ++b)
    if (off + b < n) w[b >> 2] = (w[b >> 2] & ~(0xFFu << (8 * (b & 3)))) | ((u32)bytes[off + b] << (8 * (b & 3)));
  return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ u32 byte_of(const uint4& v, int b) {
  u32 w = (b < 4) ? v.x : (b < 8) ? v.y : (b < 12) ? v.z : v.w;
  return (w >> (8 * (b & 3))) & 0xFFu;
}
// is the byte before offset `off` an EOL (file start counts as one)
__device__ __forceinline__ bool prev_is_eol(const uint8_t* __restrict__ bytes, u64 off, u64 n) {
  if (off == 0 || off > n) return true;
  uint8_t c = __ldg(bytes + off - 1);
  return c == '\n' || c == '\r';
}
// FASTQ line starts: a non-EOL byte preceded by an EOL byte (blank lines vanish, as findNonEOL
// skips every EOL char: reference src/io/file_loader.hpp:146-152)
__device__ __forceinline__ u32 line_starts(u32 eol, bool prev_eol) {
  return (~eol) & ((eol << 1) | (prev_eol ? 1u : 0u)) & 0xFFFFu;
}

#ifdef KMI_SHARED_KERNELS  // non-template kernels: compiled once, in the API translation unit
// ---- pass A: line starts per 512-byte CHUNK (= one warp's 32 groups in the parser) -----------------------------------
// One WARP per 4 KiB tile, no block barrier: the eight 512-byte rows (= chunks) of the tile are requested together
// (eight 128-bit loads in flight per lane), the "previous byte is an EOL" bit of a group comes from its left
// neighbour's mask by shuffle, and the eight chunk totals go out as two 16-byte stores. Warps stride over tiles.
// Per-chunk (not per-tile) counts let the parser number its lines with a warp scan instead of a block scan.
static constexpr int COUNT_THREADS = 256;
__global__ void __launch_bounds__(COUNT_THREADS) k_fastq_count_lines(const uint8_t* __restrict__ bytes, u64 n,
                                                                      u32* __restrict__ chunk_lines, const u32 tile_begin,
                                                                      const u32 n_tiles) {
  constexpr int ROWS = TILE_CHUNKS;   // 8
  const int lane = threadIdx.x & 31;
  const u32 warps = gridDim.x * (COUNT_THREADS / 32);
  for (u32 w = blockIdx.x * (COUNT_THREADS / 32) + (threadIdx.x >> 5); w < n_tiles; w += warps) {
    const u64 tile = (u64)tile_begin + w;
    const u64 base = tile * TILE_BYTES + (u64)lane * GROUP_BYTES;
    u32 e[ROWS];
#pragma unroll
    for (int j = 0; j < ROWS; ++j) e[j] = eol16(load_group(bytes, base + (u64)j * CHUNK_BYTES, n));   // past the end: all EOL
    u32 prev_last = prev_is_eol(bytes, tile * TILE_BYTES, n) ? 0x8000u : 0u;   // mask whose bit 15 = byte before the row
    u32 pk[ROWS / 2];                                      // two rows per word: a row holds <= 512 line starts
#pragma unroll
    for (int j = 0; j < ROWS; ++j) {
      u32 left = __shfl_up_sync(0xffffffffu, e[j], 1);
      if (lane == 0) left = prev_last;
      const u32 c = __popc(line_starts(e[j], (left >> 15) & 1u));
      if (j & 1) pk[j >> 1] |= c << 16; else pk[j >> 1] = c;
      prev_last = __shfl_sync(0xffffffffu, e[j], 31);
    }
#pragma unroll
    for (int q = 0; q < ROWS / 2; ++q) {
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) pk[q] += __shfl_xor_sync(0xffffffffu, pk[q], d);
    }
    if (lane < 2) {
      const u32 a = lane ? pk[2] : pk[0], b = lane ? pk[3] : pk[1];
      reinterpret_cast<uint4*>(chunk_lines + tile * TILE_CHUNKS)[lane] = make_uint4(a & 0xFFFFu, a >> 16, b & 0xFFFFu, b >> 16);
    }
  }
}

// ---- record starts for FASTQ position ids (ShortSequenceKmerId, src/common/sequence.hpp:103-160) ----------
// tile_rec[tile] = 1 + offset of the '@' line (line index 0 mod 4) that is open at the tile's first byte, i.e. the
// most recent header-line start strictly before the tile; 0 when there is none. With F line starts before the tile,
// that is the j-th line start walking backwards, j = F mod 4 (4 when F mod 4 == 0). One warp per tile walks back in
// 512-byte windows (a record header is normally found in the first one; long reads take more).
__global__ void __launch_bounds__(COUNT_THREADS) k_fastq_rec_carry(const uint8_t* __restrict__ bytes, u64 n,
                                                                    const u32* __restrict__ tile_lines,
                                                                    const u32* __restrict__ block_sums,
                                                                    u64* __restrict__ tile_rec, const u32 n_tiles) {
  const int lane = threadIdx.x & 31;
  const u32 warps = gridDim.x * (COUNT_THREADS / 32);
  for (u32 tile = blockIdx.x * (COUNT_THREADS / 32) + (threadIdx.x >> 5); tile < n_tiles; tile += warps) {
    const u32 first = tile_lines[(u64)tile * TILE_CHUNKS] + block_sums[((u64)tile * TILE_CHUNKS) >> 10];   // per-chunk prefixes: the tile's first chunk
    u32 j = first & 3u;
    if (j == 0) j = 4;
    u64 end = (u64)tile * TILE_BYTES, found = 0;
    while (end > 0) {                                       // warp-uniform
      const u64 start = end >= 512 ? end - 512 : 0;         // 16-byte aligned: end is
      const u64 off = start + (u64)lane * GROUP_BYTES;
      u32 e = 0xFFFFu;
      if (off < end) e = eol16(load_group(bytes, off, n));
      u32 left = __shfl_up_sync(0xffffffffu, e, 1);
      bool prev = ((left >> 15) & 1u) != 0;
      if (lane == 0) prev = prev_is_eol(bytes, off, n);
      const u32 ls = (off < end) ? line_starts(e, prev) : 0u;
      u32 suf = __popc(ls);                                 // line starts in lanes >= this one (closest to the tile last)
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const u32 o = __shfl_down_sync(0xffffffffu, suf, d);
        if (lane + d < 32) suf += o;
      }
      const u32 total = __shfl_sync(0xffffffffu, suf, 0);
      if (total >= j) {
        const u32 have = __ballot_sync(0xffffffffu, suf >= j);   // lanes 0..L
        const int L = 31 - __clz((int)have);
        const u32 suf_next = __shfl_sync(0xffffffffu, suf, min(L + 1, 31));
        if (lane == L) {
          u32 k = j - ((L == 31) ? 0u : suf_next);          // the k-th last set bit of this lane's mask
          u32 m = ls;
          while (k > 1) { m &= ~(1u << (31 - __clz((int)m))); --k; }
          found = off + (u64)(31 - __clz((int)m)) + 1;
        }
        found = __shfl_sync(0xffffffffu, found, L);
        break;
      }
      j -= total;
      end = start;
    }
    if (lane == 0) tile_rec[tile] = found;
  }
}

// exclusive scan of the per-tile counts (modulo 2^32: only the value mod 4 matters downstream)
__global__ void __launch_bounds__(1024) k_scan_tiles(u32* __restrict__ v, u64 n, u32* __restrict__ carry_io = nullptr) {
  __shared__ u32 tmp[33];
  __shared__ u32 carry_s;
  if (threadIdx.x == 0) carry_s = carry_io ? *carry_io : 0;  // running total of earlier rounds (pipelined builds)
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (u64 base = 0; base < n; base += 1024) {
    u64 i = base + threadIdx.x;
    u32 x = (i < n) ? v[i] : 0;
    u32 inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u32 o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) tmp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      u32 w = tmp[lane], winc = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        u32 o = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += o;
      }
      tmp[lane] = winc - w;
      if (lane == 31) tmp[32] = winc;
    }
    __syncthreads();
    u32 carry = carry_s;
    if (i < n) v[i] = carry + tmp[warp] + inc - x;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + tmp[32];
    __syncthreads();
  }
  if (threadIdx.x == 0 && carry_io) *carry_io = carry_s;
}

// level-1 of the tile scan: each block scans 1024 tile counts in place (exclusive) and writes
// its total; k_scan_tiles then scans the block totals; consumers add the two.
__global__ void __launch_bounds__(1024) k_scan_l1(u32* __restrict__ v, u64 n, u32* __restrict__ block_sums) {
  __shared__ u32 tmp[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u64 i = (u64)blockIdx.x * 1024 + threadIdx.x;
  u32 x = (i < n) ? v[i] : 0;
  u32 inc = x;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    u32 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    u32 w = tmp[lane], winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u32 o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += o;
    }
    tmp[lane] = winc - w;
    if (lane == 31) tmp[32] = winc;
  }
  __syncthreads();
  if (i < n) v[i] = tmp[warp] + inc - x;
  if (threadIdx.x == 0) block_sums[blockIdx.x] = tmp[32];
}

#endif  // KMI_SHARED_KERNELS

// ---- k-mer extraction from the packed code stream --------------------------------------------
// c[0..4]: code words of groups g..g+4 (first base most significant); j: start base in group g.
template <int K> __device__ __forceinline__ Key<(K <= 32 ? 1 : 2)> extract_kmer(const u32* c, int j) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  const int s = 2 * j;  // 0..30
  Key<NW> k;
  if (NW == 1) {
    u32 w0 = __funnelshift_l(c[1], c[0], s), w1 = __funnelshift_l(c[2], c[1], s);
    u64 top = ((u64)w0 << 32) | w1;
    k.w[0] = top >> (64 - 2 * K);
  } else {
    u32 w0 = __funnelshift_l(c[1], c[0], s), w1 = __funnelshift_l(c[2], c[1], s);
    u32 w2 = __funnelshift_l(c[3], c[2], s), w3 = __funnelshift_l(c[4], c[3], s);
    u64 hi = ((u64)w0 << 32) | w1, lo = ((u64)w2 << 32) | w3;
    constexpr int S = 128 - 2 * K;  // 2..62
    k.w[0] = (lo >> S) | (hi << (64 - S));
    k.w[NW - 1] = hi >> S;
  }
  return k;
}

// bit j of the result: flag bits j..j+K-1 are all set. f[0..4]: 16-bit flag words of groups g..g+4.
template <int K> __device__ __forceinline__ u32 window_all_set(const u32* f) {
  if (K <= 33) {  // needs bits 0..K+14 <= 47
    u64 F = (u64)f[0] | ((u64)f[1] << 16) | ((u64)f[2] << 32) | ((u64)f[3] << 48);
    u64 A = F, R = ~0ULL;
    int r = 0;
#pragma unroll
    for (int len = 1; len <= K; len <<= 1) {
      if (K & len) { R &= (A >> r); r += len; }
      A &= (A >> len);
    }
    return (u32)(R & 0xFFFFu);
  } else {
    unsigned __int128 F = (unsigned __int128)((u64)f[0] | ((u64)f[1] << 16) | ((u64)f[2] << 32) | ((u64)f[3] << 48)) |
                          ((unsigned __int128)f[4] << 64);
    unsigned __int128 A = F, R = ~(unsigned __int128)0;
    int r = 0;
#pragma unroll
    for (int len = 1; len <= K; len <<= 1) {
      if (K & len) { R &= (A >> r); r += len; }
      A &= (A >> len);
    }
    return (u32)((u64)R & 0xFFFFu);
  }
}

// ---- rolling extraction: R consecutive start positions from one aligned block of code words -------------
// W[0..N): code words, first base in the top bits of W[0]. window_at returns the K bases that start at bit
// offset `bitoff` (a compile-time constant after unrolling: the word picks and shifts fold to immediates).
template <int N> __device__ __forceinline__ u32 word_at(const u32 (&W)[N], int i) { return (i >= 0 && i < N) ? W[i] : 0u; }
template <int K, int N> __device__ __forceinline__ Key<(K <= 32 ? 1 : 2)> window_at(const u32 (&W)[N], const int bitoff) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  const int ws = bitoff >> 5, sh = bitoff & 31;
  Key<NW> k;
  const u32 w0 = __funnelshift_l(word_at<N>(W, ws + 1), word_at<N>(W, ws), sh);
  const u32 w1 = __funnelshift_l(word_at<N>(W, ws + 2), word_at<N>(W, ws + 1), sh);
  if (NW == 1) {
    k.w[0] = (((u64)w0 << 32) | w1) >> (64 - 2 * K);
  } else {
    const u32 w2 = __funnelshift_l(word_at<N>(W, ws + 3), word_at<N>(W, ws + 2), sh);
    const u32 w3 = __funnelshift_l(word_at<N>(W, ws + 4), word_at<N>(W, ws + 3), sh);
    const u64 hi = ((u64)w0 << 32) | w1, lo = ((u64)w2 << 32) | w3;
    constexpr int S = 128 - 2 * K;  // 2..62
    k.w[0] = (lo >> S) | (hi << (64 - S));
    k.w[NW - 1] = hi >> S;
  }
  return k;
}
// reverse complement of 16 bases held in one code word (first base in the top bits)
__device__ __forceinline__ u32 revcomp16(u32 x) {
  const u32 r = __brev(x);
  return ~(((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1));
}
// Entry = R consecutive candidate start positions (R = keys per thread per bucket round): the thread that owns
// an entry loads the code words once and rolls the window over them, forward and reverse-complement strand
// (the reverse complement of the whole block is built once, then each k-mer's is a window of it: lex_less,
// kmer_transform.hpp:109-126, without a per-k-mer bit reversal).
template <int K> struct Roll {
  static constexpr int NW = (K <= 32 ? 1 : 2);
  static constexpr int R = BktRound<NW>::R;            // 8 (one word) / 4 (two words)
  static constexpr int EPG = GROUP_BYTES / R;           // entries per 16-byte group
  static constexpr int NA = (R - 1 + K + 15) / 16;      // aligned code words an entry needs
  static constexpr int LOG_EPG = (EPG == 2 ? 1 : 2);
  u32 A[NA], RB[NA];
  // s_codes: the tile's code words; entry id = group * EPG + sub
  __device__ __forceinline__ void load(const u32* s_codes, u32 id, bool canon) {
    const u32 g = id >> LOG_EPG, sh = (id & (EPG - 1)) * (2 * R);
    u32 c[NA + 1];
#pragma unroll
    for (int w = 0; w <= NA; ++w) c[w] = s_codes[min(g + w, (u32)TILE_GROUPS)];
#pragma unroll
    for (int w = 0; w < NA; ++w) A[w] = __funnelshift_l(c[w + 1], c[w], sh);
    if (canon) {
#pragma unroll
      for (int w = 0; w < NA; ++w) RB[w] = revcomp16(A[NA - 1 - w]);
    }
  }
  // k-mer that starts at position i (0 <= i < R, compile-time) of the entry
  __device__ __forceinline__ Key<NW> get(const int i, bool canon) const {
    Key<NW> f = window_at<K, NA>(A, 2 * i);
    if (canon) {
      const Key<NW> rc = window_at<K, NA>(RB, 2 * (16 * NA - K - i));
      if (!key_less<NW>(f, rc)) f = rc;
    }
    return f;
  }
};

// ---- sinks: what happens to a generated k-mer ---------------------------------------------------
struct KeySink {     // dense key (+id) stream
  u64* keys; u64* ids; u64 cap; u64* cursor;
  u64 file_offset = 0;            // offset of bytes[0] in the file (FASTQ position ids)
  const u64* tile_rec = nullptr;  // k_fastq_rec_carry output, needed when ids != nullptr
};
// per-group record starts inside the tile (position-id variant of the FASTQ parser only)
template <bool ON> struct RecSmem { u32 unused; };
template <> struct RecSmem<true> { u32 hdr[TILE_THREADS]; u32 before[TILE_THREADS]; u32 tmp[TILE_THREADS / 32]; };


// ---- pass B: FASTQ tile -> k-mers ------------------------------------------------------------------
// SINK 0: write canonical (or forward) keys to sink.keys (block-contiguous, order of blocks free)
// SINK 1: insert straight into the count table (val += 1), no key stream at all
// SINK 2: scatter into the 256 hash-prefix buckets of kmi_bucket.cuh (the default build path)
#ifndef KMI_PARSE_MINB
#define KMI_PARSE_MINB 5      // CTAs per SM the parser is compiled for (48 registers; measured 5.59 ms at 5, 5.61 at 6, 5.71 at 7 and 8)
#endif
template <int K, int HASH, int SINK>
__global__ void __launch_bounds__(TILE_THREADS, KMI_PARSE_MINB)
k_fastq_kmers(const uint8_t* __restrict__ bytes, u64 n, const u32* __restrict__ tile_lines,
              const u32* __restrict__ block_sums, KeySink sink, TableView table, BucketSink bsink,
              u64* __restrict__ n_kmers, u32* __restrict__ perr, const bool CANON, const bool NSPLIT,
              const u64 tile_begin, const u32 nranks) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  __shared__ BucketSmem<(SINK >= 2 ? NW : 0)> s_bkt;
  __shared__ u32 s_codes[TILE_GROUPS + 1];
  __shared__ u32 s_flags[TILE_GROUPS + 1];
  __shared__ u32 s_aux[HALO_GROUPS * 2];       // halo: line-start masks / popcounts
  __shared__ u32 s_ent[TILE_THREADS * (GROUP_BYTES / BktRound<(K <= 32 ? 1 : 2)>::R)];   // entry list (P4)
  __shared__ u64 s_base;
  __shared__ u32 s_cnt;                           // entries << 16 | k-mers of the tile (P4)
  __shared__ RecSmem<(SINK == 0)> s_rec;

  const int t = threadIdx.x;
  const u64 tile = tile_begin + blockIdx.x;
  const u64 tile_off = tile * TILE_BYTES;
  const u64 off = tile_off + (u64)t * GROUP_BYTES;

  // -- P1: load + classify own group (and, for the first HALO_GROUPS threads, a halo group)
  // line starts before this warp's chunk (and, for the halo lanes, before the next tile); needed in P3: in flight with the text
  const u64 chunk = tile * TILE_CHUNKS + (u64)(t >> 5);
  const u32 chunk_first_a = __ldg(tile_lines + chunk), chunk_first_b = __ldg(block_sums + (chunk >> 10));
  u32 halo_first_a = 0, halo_first_b = 0;
  if (t < HALO_GROUPS && tile_off + TILE_BYTES < n) {
    halo_first_a = __ldg(tile_lines + (tile + 1) * TILE_CHUNKS);
    halo_first_b = __ldg(block_sums + (((tile + 1) * TILE_CHUNKS) >> 10));
  }
  uint4 v = load_group(bytes, off, n);
  Group g;                                            // FASTQ needs no separate '\n' mask, and the N mask only when splitting
  g.codes = codes16(v); g.eol = eol16(v); g.nl = 0; g.nmask = NSPLIT ? nmask16(v) : 0u;
  bool prev_eol;                                      // byte before the group: the left neighbour's last byte
  {
    const u32 left = __shfl_up_sync(0xffffffffu, v.w >> 24, 1);
    prev_eol = (left == '\n' || left == '\r');
    if ((t & 31) == 0) prev_eol = prev_is_eol(bytes, off, n);
    if (off > n) prev_eol = true;
  }
  u32 ls = (off < n) ? line_starts(g.eol, prev_eol) : 0u;
  s_codes[t] = g.codes;
  u32 hls = 0, heol = 0, hn = 0;
  if (t < HALO_GROUPS) {
    u64 hoff = tile_off + TILE_BYTES + (u64)t * GROUP_BYTES;
    uint4 hv = load_group(bytes, hoff, n);
    Group hg;
    hg.codes = codes16(hv); hg.eol = eol16(hv); hg.nl = 0; hg.nmask = NSPLIT ? nmask16(hv) : 0u;
    hls = (hoff < n) ? line_starts(hg.eol, prev_is_eol(bytes, hoff, n)) : 0u;
    heol = hg.eol; hn = hg.nmask;
    s_codes[TILE_THREADS + t] = hg.codes;
    s_aux[t] = __popc(hls);
  }
  if (t == 0) { s_codes[TILE_GROUPS] = 0; s_cnt = 0; }
  if (SINK >= 2) reinterpret_cast<BucketSmem<NW>&>(s_bkt).hist[t] = 0;   // bucket_init, under the barriers of P2-P4

  // -- P2: line number at the start of every group. A warp is one 512-byte chunk and pass A counted the line starts before
  // every chunk, so this is a warp scan on top of the chunk's prefix; the halo continues at the next tile's first chunk
  // (a tile whose halo lies beyond the text needs no line numbers there: those bytes read as EOLs).
  u32 before;
  {
    const u32 own = __popc(ls);
    u32 inc = own;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const u32 o = __shfl_up_sync(0xffffffffu, inc, d);
      if ((t & 31) >= d) inc += o;
    }
    before = inc - own;
  }
  const u32 chunk_first = chunk_first_a + chunk_first_b;   // line starts before this warp's chunk
  __syncwarp();                                           // s_aux of the halo lanes (all in warp 0) is visible to them

  // -- P3: per-byte "base of a sequence line" flags; header checks at line starts
  {
    u32 idx = chunk_first + before;  // number of line starts before this group
    u32 flags;
    u32 hdr = 0;                    // 1 + byte of the last header-line ('@', line 0 mod 4) start in this group
    if (ls == 0) {
      // every byte of the group is on line (idx-1); sequence lines are 1 mod 4
      flags = (idx != 0 && ((idx - 1) & 3u) == 1u) ? (~g.eol & 0xFFFFu) : 0u;
    } else {
      // walk the line starts of the group: bytes [lo, hi) lie on line (idx-1); a start at hi opens line idx
      flags = 0;
      u32 err = 0, m = ls, lo = 0;
      for (;;) {
        const u32 hi = m ? (u32)__ffs(m) - 1u : (u32)GROUP_BYTES;
        if ((idx & 3u) == 2u) flags |= (1u << hi) - (1u << lo);
        if (!m) break;
        const u32 c = byte_of(v, (int)hi);
        if ((idx & 3u) == 0u && c != '@') err |= PERR_NO_AT;
        if ((idx & 3u) == 2u && c != '+') err |= PERR_NO_PLUS;
        if ((idx & 3u) == 0u) hdr = hi + 1u;
        ++idx; lo = hi; m &= m - 1;
      }
      flags &= ~g.eol;
      if (err) atomicOr(perr, err);
    }
    if (NSPLIT) flags &= ~g.nmask;
    s_flags[t] = flags & 0xFFFFu;
    if (t < HALO_GROUPS) {
      u32 hidx = halo_first_a + halo_first_b;               // line starts before the next tile
      for (int q = 0; q < t; ++q) hidx += s_aux[q];
      u32 hflags = 0;
      {
        u32 m = hls, lo = 0;
        for (;;) {
          const u32 hi = m ? (u32)__ffs(m) - 1u : (u32)GROUP_BYTES;
          if ((hidx & 3u) == 2u) hflags |= (1u << hi) - (1u << lo);
          if (!m) break;
          ++hidx; lo = hi; m &= m - 1;
        }
      }
      hflags &= ~heol;
      if (NSPLIT) hflags &= ~hn;
      s_flags[TILE_THREADS + t] = hflags & 0xFFFFu;
    }
    if (t == 0) s_flags[TILE_GROUPS] = 0;
    if (SINK == 0) {
      if (sink.ids) {               // position ids: latest header start before every group (block max-scan, 1-based)
        RecSmem<true>& rs = reinterpret_cast<RecSmem<true>&>(s_rec);
        const int lane = t & 31, warp = t >> 5;
        const u32 x = hdr ? (u32)t * GROUP_BYTES + hdr : 0u;
        u32 inc = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const u32 o = __shfl_up_sync(0xffffffffu, inc, d);
          if (lane >= d) inc = max(inc, o);
        }
        if (lane == 31) rs.tmp[warp] = inc;
        rs.hdr[t] = x;
        __syncthreads();
        u32 ex = __shfl_up_sync(0xffffffffu, inc, 1);
        if (lane == 0) ex = 0;
        for (int w = 0; w < warp; ++w) ex = max(ex, rs.tmp[w]);
        rs.before[t] = ex;
      }
    }
  }
  __syncthreads();

  // -- P4: valid k-mer starts -> dense list of ENTRIES (R consecutive start positions with at least one valid)
  using RollT = Roll<K>;
  constexpr int R = RollT::R, EPG = RollT::EPG;
  u32 f5[5];
#pragma unroll
  for (int q = 0; q < 5; ++q) f5[q] = s_flags[min(t + q, TILE_GROUPS)];
  const u32 valid = window_all_set<K>(f5);
  u32 ne = 0;
#pragma unroll
  for (int e = 0; e < EPG; ++e) ne += ((valid >> (e * R)) & ((1u << R) - 1u)) ? 1u : 0u;
  // dense positions: warp scan of (entries : k-mers), one shared-memory atomic per warp for the warp's base
  u32 pre;
  {
    const u32 own = (ne << 16) | (u32)__popc(valid);
    u32 inc = own;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const u32 o = __shfl_up_sync(0xffffffffu, inc, d);
      if ((t & 31) >= d) inc += o;
    }
    u32 wbase = 0;
    if ((t & 31) == 31 && inc) wbase = atomicAdd(&s_cnt, inc);
    pre = __shfl_sync(0xffffffffu, wbase, 31) + inc - own;
  }
  {
    u32 er = pre >> 16, kp = pre & 0xFFFFu;
#pragma unroll
    for (int e = 0; e < EPG; ++e) {
      const u32 v = (valid >> (e * R)) & ((1u << R) - 1u);
      if (v) { s_ent[er++] = (u32)(t * EPG + e) | (v << 10) | (kp << 18); kp += __popc(v); }   // id : mask : k-mers before it
    }
  }
  __syncthreads();
  const u32 total_e = s_cnt >> 16, total = s_cnt & 0xFFFFu;   // entries : k-mers of the tile (<= 4096)
  if (t == 0 && total) {
    if (SINK == 0) s_base = atomicAdd(sink.cursor, (u64)total);
    atomicAdd(n_kmers, (u64)total);
  }
  if (SINK == 0) __syncthreads();                         // s_base

  // -- P5: one entry per thread: load its code words once, roll the window over the R positions
  if (SINK >= 2) {
    BucketSmem<NW>& sm = reinterpret_cast<BucketSmem<NW>&>(s_bkt);     // histogram zeroed in P1
    for (u32 e0 = 0; e0 < total_e; e0 += TILE_THREADS) {     // block-uniform trip count
      Key<NW> key[R];
      u32 bkt[R];
#pragma unroll
      for (int q = 0; q < R; ++q) bkt[q] = BKT_NONE;
      if (e0 + t < total_e) {
        const u32 ent = s_ent[e0 + t];
        const u32 v = (ent >> 10) & 0xFFu;
        RollT roll;
        roll.load(s_codes, ent & 1023u, CANON);
#pragma unroll
        for (int q = 0; q < R; ++q) {
          key[q] = roll.get(q, CANON);
          bkt[q] = ((v >> q) & 1u) ? bucket_of(slot_hash<K>(key[q])) : BKT_NONE;
        }
      }
      bucket_round<NW>(bsink, sm, key, bkt, perr);
    }
    return;
  }
  u32 n_new = 0;
  for (u32 e = t; e < total_e; e += TILE_THREADS) {
    const u32 ent = s_ent[e];
    const u32 v = (ent >> 10) & 0xFFu;
    u64 at = s_base + (ent >> 18);
    RollT roll;
    roll.load(s_codes, ent & 1023u, CANON);
    u64 id0 = 0;                    // ShortSequenceKmerId of start position 0 of the entry
    if (SINK == 0) {
      if (sink.ids) {
        const RecSmem<true>& rs = reinterpret_cast<const RecSmem<true>&>(s_rec);
        const u32 g = (ent & 1023u) >> RollT::LOG_EPG, pos0 = g * GROUP_BYTES + (ent & (EPG - 1)) * R;
        const u32 first = pos0 + (u32)__ffs((int)v) - 1u, own = rs.hdr[g];
        u64 rec;                    // file-buffer offset of the record's '@'
        if (own && own - 1u <= first) rec = tile_off + own - 1u;
        else if (rs.before[g]) rec = tile_off + rs.before[g] - 1u;
        else rec = sink.tile_rec[tile] - 1u;                 // record opened in an earlier tile (0 - 1 if none: malformed, flagged by PERR_NO_AT)
        // offset of the k-mer inside its record must fit 16 bits (ShortSequenceKmerId::operator+= throws otherwise,
        // src/common/sequence.hpp:177-183)
        if (tile_off + pos0 + (u32)(31 - __clz((int)v)) - rec > 0xFFFFull) atomicOr(perr, PERR_ID_RANGE);
        id0 = (((sink.file_offset + rec) & 0xFFFFFFFFFFULL) << 16) + (tile_off + pos0 - rec);
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (!((v >> q) & 1u)) continue;
      const Key<NW> key = roll.get(q, CANON);
      if (SINK == 0) {
        if (at < sink.cap) {
#pragma unroll
          for (int w = 0; w < NW; ++w) sink.keys[at * NW + w] = key.w[w];
          if (sink.ids) sink.ids[at] = id0 + (u64)q;
        } else {
          atomicOr(perr, PERR_OVERFLOW);
        }
        ++at;
      } else {
        bool was_new;
        u64 slot = table_upsert<NW>(table, key, slot_hash<K>(key), was_new);
        if (slot != ~0ULL) {
          table_add<NW>(table, slot, 1u);
          n_new += was_new ? 1u : 0u;
        }
      }
    }
  }
  if (SINK == 1) {
    // one size update per warp
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) n_new += __shfl_xor_sync(0xffffffffu, n_new, d);
    if ((t & 31) == 0 && n_new) atomicAdd(table.size, (u64)n_new);
  }
}

}  // namespace kmi
