"""GPU parity of the sort-based back end (kmi_sorted_*): dsc::counting_sorted_map / dsc::sorted_multimap of
reference src/containers/distributed_sorted_map.hpp, one shard. A sorted map's contents are the same set of (key, value)
entries the dense-hash map holds (the reference's own tests compare both against std::map-like containers,
src/containers/test/test_sorted_map.cpp), so the checker is the oracle's sorted count dump / sorted tuple list, which
tests/test_oracle.py pins to reference-executed vectors. All calls go through the C ABI."""
import numpy as np
import pytest

import kmerind_b200 as kb
import oracle_util as ou

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64).reshape(-1)).cuda()


def host_u64(t):
    return t.cpu().numpy().view(np.uint64)


def np_order(keys, vals, nw):
    if nw == 2:
        return np.lexsort((vals, keys[:, 0], keys[:, 1]))
    return np.lexsort((vals, keys))


@pytest.mark.parametrize("k", (15, 31, 47, 63))
def test_radix_sort_pairs_sorted_and_stable(k):
    """kmi_sort_pairs_dev: keys in Kmer::operator< order (most significant word first); with val_bits = 0 equal keys keep
    their input order (LSD passes are stable); with val_bits > 0 ties are ordered by the value."""
    nw = ou.nwords(k)
    rng = np.random.default_rng(k)
    base = ou.shape_keys(np.ascontiguousarray(ou.random_keys(k, 40000, seed=k)).reshape(-1), k)
    keys = base[rng.integers(0, len(base), 150001)]            # plenty of duplicates
    n = len(keys)
    with kb.Index(k=k) as idx:
        for val_bits, vals in ((0, np.arange(n, dtype=np.uint64)), (40, rng.integers(0, 1 << 40, n, dtype=np.uint64)), (None, None)):
            dk, dkt = dev(keys), torch.zeros(n * nw, dtype=torch.int64, device="cuda")
            if vals is None:
                in_tmp = idx.sort_pairs_dev(dk.data_ptr(), None, dkt.data_ptr(), None, n, 0)
                got_k = ou.shape_keys(host_u64(dkt if in_tmp else dk).copy(), k)
                order = np_order(keys, np.zeros(n, dtype=np.uint64), nw)
                assert np.array_equal(got_k, keys[order])
                continue
            dv, dvt = dev(vals), torch.zeros(n, dtype=torch.int64, device="cuda")
            in_tmp = idx.sort_pairs_dev(dk.data_ptr(), dv.data_ptr(), dkt.data_ptr(), dvt.data_ptr(), n, val_bits)
            torch.cuda.synchronize()
            got_k = ou.shape_keys(host_u64(dkt if in_tmp else dk).copy(), k)
            got_v = host_u64(dvt if in_tmp else dv)
            order = np_order(keys, vals, nw)                   # for val_bits = 0 the values are the input positions: stability
            assert np.array_equal(got_k, keys[order]), (k, val_bits)
            assert np.array_equal(got_v, vals[order]), (k, val_bits)


@pytest.mark.parametrize("k", (21, 31, 63))
@pytest.mark.parametrize("name", ("natural.fastq", "natural.withN.fastq", "test.medium.fastq"))
def test_counting_sorted_map_equals_oracle(name, k):
    buf = ou.read_fixture(name)
    for nsplit in (False, True):
        keys, _, _ = ou.fastq_kmers(buf, k, canonical=True, nsplit=nsplit)
        wk, wc = ou.count_sorted(keys, k)
        with kb.SortedMap(k, nsplit=nsplit) as sm:
            assert sm.build_fastq(buf) == len(keys)
            assert sm.local_size() == len(wk)
            gk, gc = sm.dump()                                  # already in key order
            assert np.array_equal(gk, wk) and np.array_equal(gc, wc), (name, k, nsplit)
            # count / find: present keys by either strand, and absent ones
            q = np.concatenate([np.ascontiguousarray(ou.revcomp(wk[::3], k)).reshape(-1), np.ascontiguousarray(ou.random_keys(k, 500, seed=7)).reshape(-1)])
            cq = ou.shape_keys(np.ascontiguousarray(ou.canonical(q, k)).reshape(-1), k)
            lut = {tuple(np.atleast_1d(a)): int(c) for a, c in zip(wk, wc)}
            want = np.array([lut.get(tuple(np.atleast_1d(a)), 0) for a in cq], dtype=np.uint64)
            assert np.array_equal(sm.count(q), (want > 0).astype(np.uint64))
            vals, found = sm.find(q)
            assert np.array_equal(found, want > 0) and np.array_equal(vals, want)
            # a second build of the same reads doubles every count; (key, count) tuples add
            sm.build_fastq(buf)
            gk, gc = sm.dump()
            assert np.array_equal(gk, wk) and np.array_equal(gc, 2 * wc)
            sm.insert(wk[:100], np.full(min(100, len(wk)), 5, dtype=np.uint64))
            gk, gc = sm.dump()
            want2 = 2 * wc
            want2[:100] += 5
            assert np.array_equal(gk, wk) and np.array_equal(gc, want2)


def test_counting_sorted_map_wraps_like_uint32():
    k = 31
    key = np.ascontiguousarray(ou.canonical(ou.random_keys(k, 1, seed=1), k)).reshape(-1)
    with kb.SortedMap(k) as sm:
        sm.insert(key, np.array([0xFFFFFFFF], dtype=np.uint64))
        sm.insert(key, np.array([3], dtype=np.uint64))
        _, c = sm.dump()
        assert c.tolist() == [2]            # std::plus<uint32_t>: distributed_sorted_map.hpp:2560-2600 reduces in the mapped type


@pytest.mark.parametrize("k,fasta,name", ((21, True, "test2.fasta"), (21, True, "test.medium.fasta"), (35, True, "test.fasta"),
                                          (31, False, "natural.fastq"), (63, False, "test.unitiq1.fastq")))
def test_sorted_multimap_equals_oracle(k, fasta, name):
    buf = ou.read_fixture(name)
    if fasta:
        keys, ids, _ = ou.fasta_kmers(buf, k)
    else:
        keys, ids, _ = ou.fastq_kmers(buf, k, want_ids=True)
    sk, si = ou.sort_tuples(keys, ids, k)
    with kb.SortedMap(k, mode=kb.MODE_POSITION) as sm:
        n = sm.build_fasta(buf) if fasta else sm.build_fastq(buf)
        assert n == len(keys) and sm.local_size() == len(keys)       # duplicates kept
        gk, gi = sm.dump()
        assert np.array_equal(gk, sk) and np.array_equal(gi, si)
        nw = ou.nwords(k)
        assert len(sk) > 0
        rows = sk.reshape(len(sk), -1)
        uk, first, cnt = np.unique(rows, axis=0, return_index=True, return_counts=True)
        if nw == 2:                                                  # np.unique sorts by column 0 (low word) first: re-derive from sk order
            change = np.ones(len(rows), dtype=bool)
            change[1:] = np.any(rows[1:] != rows[:-1], axis=1)
            first = np.flatnonzero(change)
            cnt = np.diff(np.append(first, len(rows)))
            uk = rows[first]
        pick = np.arange(0, len(uk), max(1, len(uk) // 400))
        q = np.concatenate([np.ascontiguousarray(ou.revcomp(np.ascontiguousarray(uk[pick]).reshape(-1), k)).reshape(-1),
                            np.ascontiguousarray(ou.random_keys(k, 100, seed=3)).reshape(-1)])
        want_cnt = np.concatenate([cnt[pick], np.zeros(100, dtype=np.int64)]).astype(np.uint64)
        # (a random key is absent with overwhelming probability for K >= 21)
        assert np.array_equal(sm.count(q), want_cnt)
        off, vals = sm.find_positions(q)
        assert np.array_equal(np.diff(off), want_cnt)
        for j, u in enumerate(pick):
            assert np.array_equal(vals[int(off[j]):int(off[j + 1])], si[first[u]:first[u] + cnt[u]])   # ids ascending per key


def test_sorted_map_agrees_with_hash_index_on_synthetic_reads():
    """The bench workload's shape (30x synthetic reads): the sort-based map, the dense-hash index and the oracle hold the
    same entries; kmi_dump_sorted (device sort of the hash table's entries) equals the host-sorted dump."""
    n_reads, G = 80000, 400000
    buf = ou.synth_fastq(0, n_reads, 150, G, 1, 5000, 0)
    keys, _, _ = ou.fastq_kmers(buf, 31)
    wk, wc = ou.count_sorted(keys, 31)
    t = torch.from_numpy(buf).cuda()
    with kb.SortedMap(31) as sm:
        assert sm.build_fastq_dev(t.data_ptr(), buf.size) == len(keys)
        gk, gc = sm.dump()
        ms, launches = sm.stats()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
        assert launches > 0 and ms[1] > 0
    with kb.CountIndex(31) as idx:
        idx.build_fastq(buf)
        hk, hc = idx.dump_sorted()
        assert np.array_equal(hk, wk) and np.array_equal(hc, wc)
    with kb.CountIndex(63) as idx:
        idx.build_fastq(buf[:20000 * 317])
        a, b = idx.dump_sorted()
        c, d = idx.sorted_dump()
        assert np.array_equal(a, c) and np.array_equal(b, d)
    fa = ou.synth_fasta(3, 100000, 80, 5)
    with kb.PositionIndex(21) as idx:
        idx.build_fasta(fa)
        a, b = idx.dump_sorted()
        c, d = idx.sorted_dump()
        assert np.array_equal(a, c) and np.array_equal(b, d)


def test_sorted_map_refuses_shards_and_empty_is_fine():
    lib = kb.load_library()
    import ctypes
    cfg = kb.KmiConfig(31, 0, 1, 0, 0, 0, 0, 2, 0)
    h = ctypes.c_void_p()
    assert lib.kmi_sorted_create(ctypes.byref(cfg), ctypes.byref(h)) == -5      # KMI_ERR_UNSUPPORTED: one shard only
    with kb.SortedMap(31) as sm:
        assert sm.local_size() == 0
        assert sm.count(np.array([5], dtype=np.uint64)).tolist() == [0]
        k, v = sm.dump()
        assert len(k) == 0 and len(v) == 0
