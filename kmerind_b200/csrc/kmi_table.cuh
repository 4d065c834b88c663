// kmi_table.cuh — open-addressing k-mer table in HBM and its kernels (sm_100a).
// Replaces the local half of dsc::counting_densehash_map / densehash_multimap
// (reference src/containers/densehash_map.hpp:928-1229,2015-2400 over google::dense_hash_map):
// power-of-two slots, the reference's store hash picks the home slot (its top log2(slots) bits), linear probing inside HBM
// sectors (one 16-byte slot per K<=31 key, one 32-byte slot per K<=63 key, so a probe touches a
// single 32-byte sector), claim by 64/128-bit CAS on the key, count by a fire-and-forget RED.
#pragma once
#include "kmi_device.cuh"

namespace kmi {

template <int NW> struct Slot;
template <> struct alignas(16) Slot<1> { u64 key; u32 val; u32 aux; };
template <> struct alignas(32) Slot<2> { u64 lo; u64 hi; u32 val; u32 aux; u64 pad; };

static constexpr u32 TERR_FULL = 1u, TERR_SPILL = 2u;

struct TableView {
  void* slots;
  u64 mask;        // slots - 1
  u32 shift;       // 64 - log2(slots): home slot = store_hash >> shift (TOP bits, so that a hash
                   // prefix names a contiguous table segment for any capacity: kmi_bucket.cuh)
  u64* size;       // device counter: occupied (live) keys
  u32 saturate;    // 1: counts stick at 2^32 - 1 instead of wrapping (saturating_counting_densehash_map,
                   // distributed_densehash_map.hpp:2904-3075); 0: std::plus<uint32_t> wraps (:2475)
  u32* error;      // device flags, sticky until kmi_clear: TERR_FULL probe limit hit (table full), TERR_SPILL a segment insert
                   // dropped a key because the spill list was exhausted (kmi_bucket.cuh)
};

// ---- raw slot access -----------------------------------------------------------------------
__device__ __forceinline__ u64 ld_cg_u64(const u64* p) {
  u64 v; asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p)); return v;
}
__device__ __forceinline__ void ld_cg_u64x2(const void* p, u64& a, u64& b) {
  asm volatile("ld.global.cg.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p));
}
__device__ __forceinline__ void cas128(void* addr, u64 clo, u64 chi, u64 vlo, u64 vhi, u64& olo, u64& ohi) {
  asm volatile(
      "{\n .reg .b128 c, v, o;\n mov.b128 c, {%2, %3};\n mov.b128 v, {%4, %5};\n"
      " atom.global.relaxed.gpu.cas.b128 o, [%6], c, v;\n mov.b128 {%0, %1}, o;\n}\n"
      : "=l"(olo), "=l"(ohi) : "l"(clo), "l"(chi), "l"(vlo), "l"(vhi), "l"(addr) : "memory");
}

template <int NW> __device__ __forceinline__ Key<NW> slot_key(const Slot<NW>* s);
template <> __device__ __forceinline__ Key<1> slot_key<1>(const Slot<1>* s) {
  Key<1> k; k.w[0] = ld_cg_u64(&s->key); return k;
}
template <> __device__ __forceinline__ Key<2> slot_key<2>(const Slot<2>* s) {
  Key<2> k; ld_cg_u64x2(s, k.w[0], k.w[1]); return k;
}
template <int NW> __device__ __forceinline__ bool is_empty(const Key<NW>& k) { return k.w[NW - 1] == KEY_EMPTY_TOP; }
template <int NW> __device__ __forceinline__ bool is_deleted(const Key<NW>& k) { return k.w[NW - 1] == KEY_DELETED_TOP; }
template <int NW> __device__ __forceinline__ bool is_live(const Key<NW>& k) { return (k.w[NW - 1] >> 63) == 0; }

// try to claim an EMPTY slot for `key`: true when this thread's CAS installed it; `seen` is what
// the slot held before the CAS
template <int NW> __device__ __forceinline__ bool slot_claim(Slot<NW>* s, const Key<NW>& key, Key<NW>& seen);
template <> __device__ __forceinline__ bool slot_claim<1>(Slot<1>* s, const Key<1>& key, Key<1>& seen) {
  seen.w[0] = atomicCAS(&s->key, KEY_EMPTY_TOP, key.w[0]);
  return seen.w[0] == KEY_EMPTY_TOP;
}
template <> __device__ __forceinline__ bool slot_claim<2>(Slot<2>* s, const Key<2>& key, Key<2>& seen) {
  cas128(s, 0ULL, KEY_EMPTY_TOP, key.w[0], key.w[1], seen.w[0], seen.w[1]);
  return seen.w[1] == KEY_EMPTY_TOP;
}

// ---- find-or-insert: returns the slot index of `key`, claiming a slot when absent.
// was_new tells whether this call created the entry. Returns ~0 when the table is full.
template <int NW>
__device__ __forceinline__ u64 table_upsert(const TableView& t, const Key<NW>& key, u64 hash, bool& was_new) {
  Slot<NW>* slots = reinterpret_cast<Slot<NW>*>(t.slots);
  u64 i = hash >> t.shift;
  was_new = false;
  for (u64 probes = 0; probes <= t.mask; ++probes) {
    Key<NW> cur = slot_key<NW>(slots + i);
    if (key_eq<NW>(cur, key)) return i;
    if (is_empty<NW>(cur)) {
      Key<NW> seen;
      if (slot_claim<NW>(slots + i, key, seen)) { was_new = true; return i; }
      if (key_eq<NW>(seen, key)) return i;  // a racing twin installed the same key first
    }
    i = (i + 1) & t.mask;
  }
  atomicOr(t.error, TERR_FULL);
  return ~0ULL;
}

// val += add. Saturating mode: an add that wraps is followed by a store of the maximum; any later add either lands before
// that store (and is overwritten by it) or after it (and wraps itself, storing the maximum again), so the final value is
// the maximum whatever the interleaving.
template <int NW> __device__ __forceinline__ void table_add(const TableView& t, u64 slot, u32 add) {
  u32* const v = &reinterpret_cast<Slot<NW>*>(t.slots)[slot].val;
  const u32 old = atomicAdd(v, add);
  if (t.saturate && old + add < old) atomicExch(v, 0xFFFFFFFFu);
}

// read-only probe: slot index or ~0
template <int NW>
__device__ __forceinline__ u64 table_find(const TableView& t, const Key<NW>& key, u64 hash) {
  const Slot<NW>* slots = reinterpret_cast<const Slot<NW>*>(t.slots);
  u64 i = hash >> t.shift;
  for (u64 probes = 0; probes <= t.mask; ++probes) {
    Key<NW> cur = slot_key<NW>(slots + i);
    if (key_eq<NW>(cur, key)) return i;
    if (is_empty<NW>(cur)) return ~0ULL;
    i = (i + 1) & t.mask;
  }
  return ~0ULL;
}

// read-only probe that also returns the value, for the batched queries. One-word keys: a slot is 16 bytes, so an aligned
// PAIR of slots is one 32-byte sector — both are requested together (two 128-bit loads in flight, one sector, one latency)
// and tested in probe order, which halves the dependent round trips of a long probe sequence (tables sit at load 0.35-0.7;
// at 0.7 a miss walks 6 slots on average). Two-word keys: a slot is its own sector; key and value come with one 32-byte read.
template <int NW>
__device__ __forceinline__ u64 table_find_val(const TableView& t, const Key<NW>& key, u64 hash, u32& val) {
  val = 0;
  u64 i = hash >> t.shift;
  if (NW == 1) {
    const char* base = reinterpret_cast<const char*>(t.slots);
    u64 k0, v0, k1, v1;
    u64 probes = 0;
    if (i & 1) {                                          // odd home slot: alone in its sector's second half
      ld_cg_u64x2(base + i * 16, k0, v0);
      if (k0 == key.w[0]) { val = (u32)v0; return i; }
      if (k0 == KEY_EMPTY_TOP) return ~0ULL;
      i = (i + 1) & t.mask; probes = 1;
    }
    for (; probes <= t.mask; probes += 2) {               // i is even: slots i, i + 1 share a sector (mask is odd)
      ld_cg_u64x2(base + i * 16, k0, v0);
      ld_cg_u64x2(base + i * 16 + 16, k1, v1);
      if (k0 == key.w[0]) { val = (u32)v0; return i; }
      if (k0 == KEY_EMPTY_TOP) return ~0ULL;
      if (k1 == key.w[0]) { val = (u32)v1; return i + 1; }
      if (k1 == KEY_EMPTY_TOP) return ~0ULL;
      i = (i + 2) & t.mask;
    }
    return ~0ULL;
  } else {
    const Slot<NW>* slots = reinterpret_cast<const Slot<NW>*>(t.slots);
    for (u64 probes = 0; probes <= t.mask; ++probes) {
      u64 lo, hi, vw, pad;
      asm volatile("ld.global.cg.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(lo), "=l"(hi), "=l"(vw), "=l"(pad) : "l"(slots + i));
      if (lo == key.w[0] && hi == key.w[NW - 1]) { val = (u32)vw; return i; }
      if (hi == KEY_EMPTY_TOP) return ~0ULL;
      i = (i + 1) & t.mask;
    }
    return ~0ULL;
  }
}

}  // namespace kmi
