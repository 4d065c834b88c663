// kmi_block.cuh — tile geometry, error bits and block/warp-level helpers shared by every kernel file.
#pragma once
#include "kmi_table.cuh"

namespace kmi {

static constexpr int TILE_THREADS = 256;
static constexpr int GROUP_BYTES = 16;
static constexpr int TILE_BYTES = TILE_THREADS * GROUP_BYTES;  // 4096
static constexpr int HALO_GROUPS = 4;                          // 64 bytes >= K-1 for K <= 63
static constexpr int TILE_GROUPS = TILE_THREADS + HALO_GROUPS;
static constexpr int CHUNK_BYTES = 32 * GROUP_BYTES;            // 512: what one warp of the parser owns; line starts are counted per chunk
static constexpr int TILE_CHUNKS = TILE_BYTES / CHUNK_BYTES;    // 8

// error bits reported by the parse kernels
static constexpr u32 PERR_NO_AT = 1u;    // fastq_loader.hpp:393,419
static constexpr u32 PERR_NO_PLUS = 2u;  // fastq_loader.hpp:435
static constexpr u32 PERR_OVERFLOW = 4u; // output capacity exceeded
static constexpr u32 PERR_HALO = 8u;     // FASTA: K-1 bases do not fit the halo (retry wider)
static constexpr u32 PERR_ID_RANGE = 32u; // FASTQ position id: offset inside the record >= 2^16 (sequence.hpp:177-183)

template <int NW> __device__ __forceinline__ Key<NW> load_key(const u64* __restrict__ keys, u64 i) {
  Key<NW> k;
  if (NW == 1) {
    k.w[0] = __ldg(keys + i);
  } else {
    const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2*>(keys) + i);
    k.w[0] = v.x; k.w[NW - 1] = v.y;
  }
  return k;
}
template <int NW> __device__ __forceinline__ void store_key(u64* __restrict__ keys, u64 i, const Key<NW>& k) {
  if (NW == 1) keys[i] = k.w[0];
  else reinterpret_cast<ulonglong2*>(keys)[i] = make_ulonglong2(k.w[0], k.w[NW - 1]);
}
__device__ __forceinline__ u32 warp_sum(u32 v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

// block-wide exclusive scan of one u32 per thread (TILE_THREADS threads); returns the exclusive
// prefix, total in `total`. `tmp` is TILE_THREADS/32 + 1 words of shared memory.
__device__ __forceinline__ u32 block_excl_scan(u32 v, u32* tmp, u32& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  u32 inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    u32 o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) tmp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    u32 w = (lane < TILE_THREADS / 32) ? tmp[lane] : 0;
    u32 winc = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      u32 o = __shfl_up_sync(0xffffffffu, winc, d);
      if (lane >= d) winc += o;
    }
    if (lane < TILE_THREADS / 32) tmp[lane] = winc - w;
    if (lane == TILE_THREADS / 32 - 1) tmp[TILE_THREADS / 32] = winc;
  }
  __syncthreads();
  total = tmp[TILE_THREADS / 32];
  u32 r = tmp[warp] + inc - v;
  __syncthreads();
  return r;
}

}  // namespace kmi
