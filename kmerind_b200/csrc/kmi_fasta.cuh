// kmi_fasta.cuh — FASTA record structure and k-mer/position generation (sm_100a).
//
// Replaces FASTAParser::init_parser / get_next_record (reference src/io/fasta_loader.hpp:485-616,
// 618-723) and KmerPositionTupleParser (src/io/kmer_parser.hpp:360-412,465-567) for a whole file
// resident in HBM (p = 1 semantics: every sequence is complete).
//
// Reference semantics restated: a line starts at byte 0 and after every '\n' (only '\n'); a line
// is a header iff it starts with '>' or ';'; consecutive lines of the same kind merge into runs; a
// leading non-header run is ignored; every non-header run that follows a header run is one
// sequence, whose k-mers slide over its non-EOL ('\n','\r') bytes ACROSS line breaks; sequence id =
// (index of the run in the merged list) / 2, i.e. header runs seen so far, minus one when the file
// starts with a header. Position id (LongSequenceKmerId, src/common/sequence.hpp:180-236) =
// (seq_id & 0xFFFF) << 40 | file byte of the k-mer's first base.
//
// GPU formulation: per 16-byte group, a mask of line starts and of header line starts. "Kind of
// the line in progress" is a last-writer value -> exclusive MAX-scan of (position<<1 | kind);
// "header runs so far" is a SUM-scan of header-run starts (header line start whose previous line
// is not a header). Both scans run over tiles (global, 2 levels) and again over groups inside a
// tile. Bases are then compacted inside the tile (EOLs and header bytes dropped) into a packed
// 2-bit stream in shared memory with their raw offsets, so a k-mer is again a funnel shift; a
// window is valid iff its K bases are "good" (N-split mode) and its first and last base carry the
// same header-run count (no header in between).
#pragma once
#include "kmi_parse.cuh"

namespace kmi {

__device__ __forceinline__ u32 block_excl_scan_max(u32 v, u32* tmp, u32& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u32 inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    u32 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc = max(inc, o);
  }
  u32 excl = __shfl_up_sync(0xffffffffu, inc, 1);
  if (lane == 0) excl = 0;
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    u32 w = (lane < TILE_THREADS / 32) ? tmp[lane] : 0;
    u32 winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u32 o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc = max(winc, o);
    }
    u32 wex = __shfl_up_sync(0xffffffffu, winc, 1);
    if (lane == 0) wex = 0;
    if (lane < TILE_THREADS / 32) tmp[lane] = wex;
    if (lane == TILE_THREADS / 32 - 1) tmp[TILE_THREADS / 32] = winc;
  }
  __syncthreads();
  total = tmp[TILE_THREADS / 32];
  u32 r = max(tmp[warp], excl);
  __syncthreads();
  return r;
}

// line structure of one group. prev_nl: the byte before the group is '\n' (or the group starts the file)
struct FaGroup {
  u32 ls;    // line starts
  u32 hdr;   // line starts whose first byte is '>' or ';'
};
__device__ __forceinline__ FaGroup fasta_lines(const uint4& v, u32 nl, bool prev_nl) {
  FaGroup f;
  f.ls = ((nl << 1) | (prev_nl ? 1u : 0u)) & 0xFFFFu;
  u32 h = mask16(eq4(v.x, 0x3E3E3E3Eu) | eq4(v.x, 0x3B3B3B3Bu), eq4(v.y, 0x3E3E3E3Eu) | eq4(v.y, 0x3B3B3B3Bu),
                 eq4(v.z, 0x3E3E3E3Eu) | eq4(v.z, 0x3B3B3B3Bu), eq4(v.w, 0x3E3E3E3Eu) | eq4(v.w, 0x3B3B3B3Bu));
  f.hdr = f.ls & h;
  return f;
}
__device__ __forceinline__ bool prev_is_nl(const uint8_t* __restrict__ bytes, u64 off, bool bol) {
  return off == 0 ? bol : __ldg(bytes + off - 1) == '\n';
}
// One block of a FASTA file (rank r's byte range of a partitioned file, or one chunk of a streamed one). The text
// passed holds the block's own bytes [0, n_valid) followed by halo bytes up to n: k-mers are emitted for first bases
// inside [0, n_valid) and may run K-1 bases into the halo (KmerParser::get_valid_iterator_range, kmer_parser.hpp:114-157;
// FASTA tail rule kmer_file_helper.hpp:153-173). What the bytes before the block contribute is three numbers.
struct FaBlock {
  u64 file_offset;      // file byte of bytes[0]
  u64 n_valid;          // bytes owned by the block
  u32 runs0;            // header runs that start before the block
  u32 kind0;            // 1: the last line start before the block begins a header line
  u32 bol;              // 1: bytes[0] starts a line (file start or previous byte '\n')
  u32 first_is_header;  // 1: the FILE starts with a header line (sequence id = header runs so far - this)
};
// header-run starts inside a group: header line starts whose previous line start (inside the
// group, else `prev_kind`) is a non-header
__device__ __forceinline__ u32 header_run_starts(u32 ls, u32 hdr, u32 prev_kind) {
  u32 out = 0, kind = prev_kind, m = ls;
  while (m) {
    int b = __ffs(m) - 1;
    m &= m - 1;
    u32 k = (hdr >> b) & 1u;
    if (k && !kind) out |= 1u << b;
    kind = k;
  }
  return out;
}
// per-group / per-tile summary for the max-scan: 0 = no line start, else ((pos+1) << 1) | kind of the last one
__device__ __forceinline__ u32 enc_last(u32 ls, u32 hdr, u32 pos) {
  if (!ls) return 0;
  int b = 31 - __clz(ls);
  return ((pos + 1) << 1) | ((hdr >> b) & 1u);
}

#ifdef KMI_SHARED_KERNELS
// pass A1: per tile: enc (for the max scan); (internal header-run starts << 2) | has << 1 | kind of first line start
__global__ void __launch_bounds__(TILE_THREADS)
k_fasta_tiles(const uint8_t* __restrict__ bytes, u64 n, const u64 n_valid, const u32 bol, u32* __restrict__ tile_enc,
              u32* __restrict__ tile_first_t, u64* __restrict__ first_ls) {
  __shared__ u32 tmp[TILE_THREADS / 32 + 1];
  __shared__ u32 s_first;
  const int t = threadIdx.x;
  if (t == 0) s_first = 0xFFFFFFFFu;
  u64 off = (u64)blockIdx.x * TILE_BYTES + (u64)t * GROUP_BYTES;
  FaGroup f; f.ls = 0; f.hdr = 0;
  if (off < n_valid) {
    uint4 v = load_group(bytes, off, n);
    Group g = classify(v);
    f = fasta_lines(v, g.nl, prev_is_nl(bytes, off, bol != 0));
    // line starts at or past the end of the block are not this block's
    if (off + GROUP_BYTES > n_valid) { u32 keep = (1u << (n_valid - off)) - 1; f.ls &= keep; f.hdr &= keep; }
  }
  u32 total_enc;
  u32 prev = block_excl_scan_max(enc_last(f.ls, f.hdr, (u32)t), tmp, total_enc);
  // inside the tile the kind before the first line start is unknown: count it as "non-header"
  // only when a previous line start exists in the tile; the tile-boundary case is added later
  u32 hrs = header_run_starts(f.ls, f.hdr, prev ? (prev & 1u) : 1u);
  if (f.ls) atomicMin(&s_first, ((u32)t << 1) | ((f.hdr >> (__ffs(f.ls) - 1)) & 1u));
  u32 total_t;
  block_excl_scan(__popc(hrs), tmp, total_t);
  if (t == 0) {
    tile_enc[blockIdx.x] = total_enc ? ((((u32)blockIdx.x + 1) << 1) | (total_enc & 1u)) : 0u;
    u32 first_kind = (s_first == 0xFFFFFFFFu) ? 0u : (s_first & 1u);
    tile_first_t[blockIdx.x] = (total_t << 2) | (total_enc ? 2u : 0u) | first_kind;
    if (blockIdx.x == 0) tile_enc[gridDim.x] = 0;   // sentinel: its exclusive scan value is the whole block's
    // first line start of the block (tile << 1 | kind), for the block summary
    const u64 mine = ((u64)blockIdx.x << 1) | first_kind;
    if (total_enc && mine < *first_ls) atomicMin(reinterpret_cast<unsigned long long*>(first_ls), (unsigned long long)mine);
  }
}
// two-level exclusive max scan (same shape as k_scan_l1 / k_scan_tiles)
__global__ void __launch_bounds__(1024) k_maxscan_l1(u32* __restrict__ v, u64 n, u32* __restrict__ block_max) {
  __shared__ u32 tmp[33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u64 i = (u64)blockIdx.x * 1024 + threadIdx.x;
  u32 x = (i < n) ? v[i] : 0, inc = x;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    u32 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc = max(inc, o);
  }
  u32 excl = __shfl_up_sync(0xffffffffu, inc, 1);
  if (lane == 0) excl = 0;
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    u32 w = tmp[lane], winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u32 o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc = max(winc, o);
    }
    u32 wex = __shfl_up_sync(0xffffffffu, winc, 1);
    if (lane == 0) wex = 0;
    tmp[lane] = wex;
    if (lane == 31) tmp[32] = winc;
  }
  __syncthreads();
  if (i < n) v[i] = max(tmp[warp], excl);
  if (threadIdx.x == 0) block_max[blockIdx.x] = tmp[32];
}
__global__ void __launch_bounds__(32) k_maxscan_l2(u32* __restrict__ v, u64 n) {
  // block maxima are few (n_tiles / 1024): one warp walks them
  u32 carry = 0;
  for (u64 base = 0; base < n; base += 32) {
    u64 i = base + threadIdx.x;
    u32 x = (i < n) ? v[i] : 0, inc = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u32 o = __shfl_up_sync(0xffffffffu, inc, d);
      if ((int)threadIdx.x >= d) inc = max(inc, o);
    }
    u32 excl = __shfl_up_sync(0xffffffffu, inc, 1);
    if (threadIdx.x == 0) excl = 0;
    if (i < n) v[i] = max(carry, excl);
    carry = max(carry, __shfl_sync(0xffffffffu, inc, 31));
  }
}
// header-run starts per tile, now that the kind of the line entering each tile is known
__global__ void __launch_bounds__(256)
k_fasta_tile_starts(const u32* __restrict__ tile_enc_scanned, const u32* __restrict__ block_max, const u32* __restrict__ tile_first_t,
                    u64 n_tiles, const u32 kind0, u32* __restrict__ starts) {
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i <= n_tiles; i += (u64)gridDim.x * blockDim.x) {
    if (i == n_tiles) { starts[i] = 0; continue; }   // sentinel
    u32 prev = max(tile_enc_scanned[i], block_max[i >> 10]);
    u32 pk = prev ? (prev & 1u) : kind0;
    u32 ft = tile_first_t[i];
    u32 add = ((ft & 2u) && (ft & 1u) && !pk) ? 1u : 0u;  // has a line start, first is a header, entering line is not
    starts[i] = (ft >> 2) + add;
  }
}
#endif  // KMI_SHARED_KERNELS

// ---- pass B: FASTA tile -> (k-mer, position id) ---------------------------------------------------------
// FH: halo groups after the tile (FH*16 raw bytes must hold K-1 bases of a continuing sequence)
template <int K, int HASH, int FH>
__global__ void __launch_bounds__(TILE_THREADS)
k_fasta_kmers(const uint8_t* __restrict__ bytes, u64 n, const u32* __restrict__ tile_enc_scanned, const u32* __restrict__ enc_block_max,
              const u32* __restrict__ tile_runs, const u32* __restrict__ runs_block_sums, KeySink sink, u64* __restrict__ n_kmers,
              u32* __restrict__ perr, const bool CANON, const bool NSPLIT, const u64 tile_begin, const FaBlock blk) {
  constexpr int NW = (K <= 32 ? 1 : 2);
  constexpr int TG = TILE_THREADS + FH;          // raw groups incl. halo
  constexpr int MAXC = TG * GROUP_BYTES;         // compacted bases upper bound
  constexpr int CW = MAXC / 16 + 6;              // packed words (+ zero tail)
  __shared__ u32 s_codes[TG];                    // raw 2-bit codes per group
  __shared__ unsigned short s_ls[TG], s_hdr[TG], s_eol[TG], s_nm[TG], s_hrs[TG], s_flags[TG];
  __shared__ u32 s_runs[TG];
  __shared__ unsigned char s_kind[TG];
  __shared__ u32 s_cbase[TG];                    // compacted index of the group's first base
  __shared__ u32 s_cc[CW], s_good[CW];
  __shared__ unsigned short s_raw[MAXC];
  __shared__ unsigned short s_list[TILE_BYTES];
  __shared__ u32 s_tmp[TILE_THREADS / 32 + 1];
  __shared__ u64 s_base;

  const int t = threadIdx.x;
  const u64 tile = tile_begin + blockIdx.x;
  const u64 tile_off = tile * TILE_BYTES;
  const u32 first_is_header = blk.first_is_header;

  // -- P1: classify all raw groups
  for (int g = t; g < TG; g += TILE_THREADS) {
    u64 off = tile_off + (u64)g * GROUP_BYTES;
    uint4 v = load_group(bytes, min(off, n), n);
    Group gr = classify(v);
    FaGroup f; f.ls = 0; f.hdr = 0;
    if (off < n) {
      f = fasta_lines(v, gr.nl, prev_is_nl(bytes, off, blk.bol != 0));
      if (off + GROUP_BYTES > n) { u32 keep = (1u << (n - off)) - 1; f.ls &= keep; f.hdr &= keep; }
    }
    s_codes[g] = gr.codes; s_ls[g] = f.ls; s_hdr[g] = f.hdr; s_eol[g] = gr.eol; s_nm[g] = gr.nmask;
  }
  for (int w = t; w < CW; w += TILE_THREADS) { s_cc[w] = 0; s_good[w] = 0; }
  __syncthreads();

  // -- P2: kind of the line in progress and header-run count at the start of every group
  const u32 tprev = max(tile_enc_scanned[tile], enc_block_max[tile >> 10]);
  const u32 tile_kind = tprev ? (tprev & 1u) : blk.kind0;   // no line start before the tile: what the block inherits
  const u32 tile_run0 = blk.runs0 + tile_runs[tile] + runs_block_sums[tile >> 10];
  u32 main_enc_total, main_runs_total;
  {
    u32 ls = s_ls[t], hdr = s_hdr[t];
    u32 prev = block_excl_scan_max(enc_last(ls, hdr, (u32)t), s_tmp, main_enc_total);
    u32 kind_in = prev ? (prev & 1u) : tile_kind;
    u32 hrs = header_run_starts(ls, hdr, kind_in);
    u32 before = block_excl_scan(__popc(hrs), s_tmp, main_runs_total);
    s_kind[t] = (unsigned char)kind_in; s_hrs[t] = (unsigned short)hrs; s_runs[t] = tile_run0 + before;
  }
  {
    const bool h = t < FH;
    u32 ls = h ? s_ls[TILE_THREADS + t] : 0u, hdr = h ? s_hdr[TILE_THREADS + t] : 0u;
    u32 dummy;
    u32 prev = block_excl_scan_max(h ? enc_last(ls, hdr, (u32)t) : 0u, s_tmp, dummy);
    u32 kind_in = prev ? (prev & 1u) : (main_enc_total ? (main_enc_total & 1u) : tile_kind);
    u32 hrs = header_run_starts(ls, hdr, kind_in);
    u32 before = block_excl_scan(h ? __popc(hrs) : 0u, s_tmp, dummy);
    if (h) {
      s_kind[TILE_THREADS + t] = (unsigned char)kind_in; s_hrs[TILE_THREADS + t] = (unsigned short)hrs;
      s_runs[TILE_THREADS + t] = tile_run0 + main_runs_total + before;
    }
  }
  __syncthreads();

  // -- P3: base flags, compaction
  auto group_flags = [&](int g) -> u32 {
    u32 ls = s_ls[g], hdr = s_hdr[g], hrs = s_hrs[g], kind = s_kind[g], runs = s_runs[g], flags = 0;
    if (ls == 0) return (kind == 0 && runs >= 1) ? (~(u32)s_eol[g] & 0xFFFFu) : 0u;
#pragma unroll
    for (int b = 0; b < GROUP_BYTES; ++b) {
      if ((ls >> b) & 1u) { kind = (hdr >> b) & 1u; runs += (hrs >> b) & 1u; }
      if (kind == 0 && runs >= 1) flags |= 1u << b;
    }
    return flags & ~(u32)s_eol[g] & 0xFFFFu;
  };
  u32 fl_main = group_flags(t), fl_halo = (t < FH) ? group_flags(TILE_THREADS + t) : 0u;
  u32 m_owned, m_halo;
  u32 cb_main = block_excl_scan(__popc(fl_main), s_tmp, m_owned);
  u32 cb_halo = block_excl_scan(__popc(fl_halo), s_tmp, m_halo);
  const u32 M = m_owned + m_halo;
  auto compact = [&](int g, u32 flags, u32 c) {
    s_cbase[g] = c; s_flags[g] = (unsigned short)flags;
    u32 codes = s_codes[g], nm = s_nm[g];
    while (flags) {
      int b = __ffs(flags) - 1;
      flags &= flags - 1;
      u32 code = (codes >> (30 - 2 * b)) & 3u;
      if (code) atomicOr(&s_cc[c >> 4], code << (30 - 2 * (c & 15)));
      if (!(NSPLIT && ((nm >> b) & 1u))) atomicOr(&s_good[c >> 4], 1u << (c & 15));
      s_raw[c] = (unsigned short)(g * GROUP_BYTES + b);
      ++c;
    }
  };
  compact(t, fl_main, cb_main);
  if (t < FH) compact(TILE_THREADS + t, fl_halo, m_owned + cb_halo);
  __syncthreads();

  // header-run count of the base at raw tile offset r
  auto runs_at = [&](u32 r) -> u32 { return s_runs[r >> 4] + __popc((u32)s_hrs[r >> 4] & ((2u << (r & 15)) - 1u)); };

  // halo sufficiency: the last owned base lacks K-1 followers although the text and its sequence go on
  if (t == 0 && m_owned > 0 && tile_off + (u64)TG * GROUP_BYTES < n && m_halo < (u32)(K - 1)) {
    u32 runs_end = s_runs[TG - 1] + __popc((u32)s_hrs[TG - 1]);
    u32 kind_end = s_ls[TG - 1] ? ((s_hdr[TG - 1] >> (31 - __clz((u32)s_ls[TG - 1]))) & 1u) : s_kind[TG - 1];
    if (runs_end == runs_at(s_raw[m_owned - 1]) && kind_end == 0) atomicOr(perr, PERR_HALO);
  }

  // -- P4: valid starts among this thread's 16 compacted positions
  u32 valid = 0;
  {
    // good bits are stored 16 per word (bit c&15 of word c>>4), like the raw flag words
    u32 f5[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) f5[q] = s_good[t + q] & 0xFFFFu;
    u32 cand = window_all_set<K>(f5);
    const u32 c0 = (u32)t * 16u;
    while (cand) {
      int j = __ffs(cand) - 1;
      cand &= cand - 1;
      u32 c = c0 + j;
      if (c >= m_owned || c + K > M) continue;
      if (tile_off + s_raw[c] >= blk.n_valid) continue;   // first base in the halo: the next block's k-mer
      if (runs_at(s_raw[c]) == runs_at(s_raw[c + K - 1])) valid |= 1u << j;
    }
  }
  u32 total;
  u32 rank = block_excl_scan(__popc(valid), s_tmp, total);
  {
    u32 m = valid;
    while (m) {
      int j = __ffs(m) - 1;
      m &= m - 1;
      s_list[rank++] = (unsigned short)(t * 16 + j);
    }
  }
  if (t == 0) {
    u64 base = 0;
    if (total) { base = atomicAdd(sink.cursor, (u64)total); atomicAdd(n_kmers, (u64)total); }
    s_base = base;
  }
  __syncthreads();

  // -- P5: dense loop
  for (u32 r = t; r < total; r += TILE_THREADS) {
    const u32 c = s_list[r];
    const int gi = c >> 4, j = c & 15;
    u32 c5[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) c5[q] = s_cc[gi + q];
    Key<NW> key = extract_kmer<K>(c5, j);
    if (CANON) key = canonical<K>(key);
    const u64 at = s_base + r;
    if (at < sink.cap) {
#pragma unroll
      for (int w = 0; w < NW; ++w) sink.keys[at * NW + w] = key.w[w];
      if (sink.ids) {
        const u32 raw = s_raw[c];
        const u64 sid = (u64)((runs_at(raw) - first_is_header) & 0xFFFFu);
        sink.ids[at] = (sid << 40) | (blk.file_offset + tile_off + raw);
      }
    } else {
      atomicOr(perr, PERR_OVERFLOW);
    }
  }
}

}  // namespace kmi
