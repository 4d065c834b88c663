"""Bulk-load mode (KMI_PARAM_DEFER_INSERT): the fine buckets of several rounds / calls are staged and inserted with one
table sweep. Observable results must be exactly those of the immediate insert — the reference's Index::insert of a whole
file's k-mers (kmer_index.hpp:201-225) — whatever triggers the flush: a reader, a full staging area, a growing table,
kmi_flush or kmi_clear. Checked against the oracle's sorted count dump, through the C ABI."""
import numpy as np
import pytest

import kmerind_b200 as kb
import oracle_util as ou

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

READS = 30000          # 9.5 MB of FASTQ per call: above the 8 MiB threshold of the bucketed build
REC = 317


def pieces(n_calls, genome, seed=3):
    return [ou.synth_fastq(i * READS, READS, 150, genome, seed, 5000, 0) for i in range(n_calls)]


def oracle_dump(bufs, k):
    keys = np.concatenate([np.ascontiguousarray(ou.fastq_kmers(b, k)[0]).reshape(-1) for b in bufs])
    return ou.count_sorted(keys, k)


@pytest.mark.parametrize("k", (31, 63))
@pytest.mark.parametrize("budget", (1 << 30, 9_000_000, 3_000_000))
def test_staged_builds_equal_oracle(k, budget):
    """budget 2^30: everything is staged until the dump; 9e6: the staging area fills after two calls (3.6e6 k-mers each);
    3e6: smaller than one call twice over -> nothing is staged, every call commits at once. Same contents every time."""
    bufs = pieces(4, 2_000_000)
    wk, wc = oracle_dump(bufs, k)
    with kb.CountIndex(k, hash=kb.HASH_FARM, reserve=8_000_000) as idx:    # (reserved: a growing table would flush first)
        idx.set_defer_insert(budget)
        total = 0
        for b in bufs:
            t = torch.from_numpy(b).cuda()
            total += idx.build_fastq_dev(t.data_ptr(), b.size)
        assert total == 4 * READS * (150 - k + 1)
        staged, flushes = idx.bulk_stats()
        per_call = READS * (150 - k + 1)
        if budget == 1 << 30:
            assert (staged, flushes) == (total, 0)       # nothing has touched the table yet
        elif budget == 9_000_000:
            assert flushes >= 1 and 0 < staged <= 2 * per_call
        else:
            assert (staged, flushes) == (0, 0)           # every call committed at once
        gk, gc = idx.sorted_dump()                       # the dump sees every staged key
        assert idx.bulk_stats()[0] == 0
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
        assert idx.checksum()[1] == total


def test_readers_between_calls_and_growth():
    """no reserve and a large genome: the table grows while keys are staged (flush, then rehash); queries and size readers
    in between see everything built so far; host-pointer builds and key inserts mix in"""
    k = 31
    bufs = pieces(5, 40_000_000, seed=9)                # mostly distinct k-mers: the table has to grow call after call
    with kb.CountIndex(k, hash=kb.HASH_FARM) as idx:
        idx.set_defer_insert(1 << 30)
        seen = []
        for i, b in enumerate(bufs):
            t = torch.from_numpy(b).cuda()
            idx.build_fastq_dev(t.data_ptr(), b.size)
            seen.append(b)
            if i in (1, 3):
                wk, wc = oracle_dump(seen, k)
                assert idx.local_size() == len(wk)
                q = wk[:: max(1, len(wk) // 5000)]
                vals, found = idx.find(q)
                assert found.all() and np.array_equal(vals, wc[:: max(1, len(wk) // 5000)])
        extra = ou.random_keys(k, 1000, seed=4)
        idx.insert(extra)                                # direct insert on top of staged keys
        idx.build_fastq(bufs[0])                         # host-pointer build (pipelined rounds: commits at once)
        keys = np.concatenate([np.ascontiguousarray(ou.fastq_kmers(b, k)[0]).reshape(-1) for b in bufs + [bufs[0]]] +
                              [np.ascontiguousarray(ou.canonical(extra, k)).reshape(-1)])
        wk, wc = ou.count_sorted(keys, k)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)


def test_flush_clear_and_switching_off():
    k = 31
    bufs = pieces(2, 1_000_000, seed=5)
    wk, wc = oracle_dump(bufs[:1], k)
    with kb.CountIndex(k, hash=kb.HASH_FARM, reserve=4_000_000) as idx:
        idx.set_defer_insert(1 << 28)
        t0 = torch.from_numpy(bufs[0]).cuda()
        idx.build_fastq_dev(t0.data_ptr(), bufs[0].size)
        idx.flush()
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
        idx.build_fastq_dev(t0.data_ptr(), bufs[0].size)          # staged ...
        idx.clear()                                                # ... and dropped with everything else
        assert idx.local_size() == 0
        idx.build_fastq_dev(t0.data_ptr(), bufs[0].size)
        idx.set_defer_insert(0)                                    # switching off settles what is staged
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
        t1 = torch.from_numpy(bufs[1]).cuda()
        idx.build_fastq_dev(t1.data_ptr(), bufs[1].size)
        wk2, wc2 = oracle_dump(bufs, k)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk2) and np.array_equal(gc, wc2)


def test_skewed_keys_fall_back_without_losing_staged_keys():
    """one k-mer far beyond the overflow region's capacity (poly-A reads) on top of staged keys: the round is rolled back
    and committed the regular way, after what was staged before it"""
    k = 31
    normal = pieces(1, 1_000_000, seed=6)[0]
    rec = b"@S0000000000\n" + b"A" * 150 + b"\n+\n" + b"I" * 150 + b"\n"
    poly = np.frombuffer(rec * READS, dtype=np.uint8).copy()
    wk, wc = oracle_dump([normal, poly], k)
    with kb.CountIndex(k, hash=kb.HASH_FARM) as idx:
        idx.set_defer_insert(1 << 28)
        for b in (normal, poly):
            t = torch.from_numpy(b).cuda()
            idx.build_fastq_dev(t.data_ptr(), b.size)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
