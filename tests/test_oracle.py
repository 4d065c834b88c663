"""CPU tests: the oracle (oracle/kmi_oracle.c) against the reference's golden vectors.

Pins, in order: (1) the reference's own golden window table (src/common/test/test_kmer.cpp:416-458),
(2) the reference's golden (records, k-mers, bytes) tables for every fixture
(src/io/test/mpi_test_fastq_seq_parse.cpp:446-459,834-845,1420-1432 and
mpi_test_fasta_seq_parse.cpp:398-403,764-769,1255-1260; K = 35), (3) values produced by executing
the reference's headers (tests/golden/ref_kat.json, made by tests/golden/make_golden.py), (4) the
live reference shim oracle/_ref when it is present.
"""
import json
import os

import numpy as np
import pytest

import oracle_util as ou

KAT = json.load(open(os.path.join(ou.ROOT, "tests", "golden", "ref_kat.json")))
KS = (15, 21, 31, 35, 47, 63)
kb_hash_identity, kb_hash_std = 3, 4   # kmi_hash_kind (include/kmi.h)


def unhex(lst, k):
    return ou.shape_keys(np.array([int(x, 16) for x in lst], dtype=np.uint64), k)


def test_golden_windows_test_kmer_cpp():
    # test_kmer.cpp:416-458: 64 2-bit chars of 0xabbacafebabe1234deadbeef01c0ffee; window i of 32
    # chars == kmer_ex[i]; a K-mer starting at i holds the top 2K bits of that window.
    val = 0xabbacafebabe1234deadbeef01c0ffee
    kmer_ex = [0xabbacafebabe1234, 0xaeeb2bfaeaf848d3, 0xbbacafebabe1234d, 0xeeb2bfaeaf848d37, 0xbacafebabe1234de,
               0xeb2bfaeaf848d37a, 0xacafebabe1234dea, 0xb2bfaeaf848d37ab, 0xcafebabe1234dead, 0x2bfaeaf848d37ab6,
               0xafebabe1234deadb, 0xbfaeaf848d37ab6f, 0xfebabe1234deadbe, 0xfaeaf848d37ab6fb, 0xebabe1234deadbee,
               0xaeaf848d37ab6fbb, 0xbabe1234deadbeef, 0xeaf848d37ab6fbbc, 0xabe1234deadbeef0, 0xaf848d37ab6fbbc0,
               0xbe1234deadbeef01, 0xf848d37ab6fbbc07, 0xe1234deadbeef01c, 0x848d37ab6fbbc070, 0x1234deadbeef01c0,
               0x48d37ab6fbbc0703, 0x234deadbeef01c0f, 0x8d37ab6fbbc0703f, 0x34deadbeef01c0ff, 0xd37ab6fbbc0703ff,
               0x4deadbeef01c0ffe, 0x37ab6fbbc0703ffb, 0xdeadbeef01c0ffee]
    codes = [(val >> (126 - 2 * i)) & 3 for i in range(64)]
    assert codes[:8] == [2, 2, 2, 3, 2, 3, 2, 2]  # kmer_data of the reference test
    text = bytes(b"ACGT"[c] for c in codes)
    fq = np.frombuffer(b"@r\n" + text + b"\n+\n" + b"I" * 64 + b"\n", dtype=np.uint8)
    for k in (31, 28, 13, 4, 1):
        keys, _, reads = ou.fastq_kmers(fq, k, canonical=False)
        assert reads == 1 and len(keys) == 64 - k + 1
        for i in range(33):
            assert int(keys[i]) == kmer_ex[i] >> (64 - 2 * k), (k, i)
    keys, _, _ = ou.fastq_kmers(fq, 63, canonical=False)
    for i in range(2):
        want = (val >> (128 - 2 * i - 126)) & ((1 << 126) - 1)
        assert int(keys[i, 0]) == want & (2**64 - 1) and int(keys[i, 1]) == want >> 64


FASTQ_GOLDEN = [  # (records, 35-mers, 35-mers under N-split, bytes, file)
    (243, 434, 434, 27580, "natural.fastq"), (250, 500, 490, 29250, "natural.withN.fastq"),
    (7, 182, 182, 939, "test.debruijn.small.fastq"), (1, 26, 26, 134, "test.debruijn.tiny.fastq"),
    (140, 3640, 3640, 18761, "test.medium.fastq"), (7, 182, 182, 939, "test.small.fastq"),
    (1, 16552, 16552, 33194, "test.unitiq1.fastq"), (1, 3527, 3527, 7144, "test.unitiq1.short2.fastq"),
    (1, 7367, 7367, 14824, "test.unitiq1.short.fastq"), (1, 14603, 14603, 29296, "test.unitiq2.fastq"),
    (2, 31155, 31155, 62490, "test.unitiqs.fastq")]
FASTA_GOLDEN = [
    (243, 434, 434, 13790, "natural.fasta"), (250, 500, 490, 14625, "natural.withN.fasta"), (6, 246, 246, 940, "test2.fasta"),
    (4, 64, 64, 512, "test.fasta"), (5000, 335000, 334994, 1092580, "test.medium.fasta"), (2, 31155, 31155, 31244, "test.unitiqs.fasta")]


@pytest.mark.parametrize("recs,kmers,kmers_split,size,name", FASTQ_GOLDEN)
def test_fastq_golden_table(recs, kmers, kmers_split, size, name):
    buf = ou.read_fixture(name)
    assert buf.size == size
    keys, _, reads = ou.fastq_kmers(buf, 35, canonical=False)
    assert (reads, len(keys)) == (recs, kmers)
    keys, _, _ = ou.fastq_kmers(buf, 35, canonical=False, nsplit=True)
    assert len(keys) == kmers_split


@pytest.mark.parametrize("recs,kmers,kmers_split,size,name", FASTA_GOLDEN)
def test_fasta_golden_table(recs, kmers, kmers_split, size, name):
    buf = ou.read_fixture(name)
    assert buf.size == size
    keys, ids, seqs = ou.fasta_kmers(buf, 35, canonical=False)
    assert (seqs, len(keys)) == (recs, kmers)
    keys2, ids2, _ = ou.fasta_kmers(buf, 35, canonical=False, nsplit=True)
    assert len(keys2) == kmers_split
    # mpi_test_fasta_seq_parse.cpp:130-150: every k-mer equals the file bytes at its position
    text = bytes(buf)
    for j in (0, len(keys) // 2, len(keys) - 1):
        pos = int(ids[j]) & ((1 << 40) - 1)
        bases = bytes(c for c in text[pos:pos + 200] if c not in b"\r\n")[:35]
        want = 0
        for c in bases:
            want = (want << 2) | {65: 0, 67: 1, 71: 2, 84: 3, 97: 0, 99: 1, 103: 2, 116: 3}.get(c, 0)
        assert int(keys[j, 0]) | (int(keys[j, 1]) << 64) == want


@pytest.mark.parametrize("k", KS)
def test_oracle_matches_reference_vectors(k):
    e = KAT["k"][str(k)]
    seq = np.frombuffer(KAT["seq"].encode(), dtype=np.uint8)
    fq = np.concatenate([np.frombuffer(b"@x\n", dtype=np.uint8), seq, np.frombuffer(b"\n+\n", dtype=np.uint8),
                         np.full(seq.size, ord("I"), dtype=np.uint8), np.frombuffer(b"\n", dtype=np.uint8)])
    win, _, _ = ou.fastq_kmers(fq, k, canonical=False)
    keys = unhex(e["keys"], k)
    assert len(win) == e["n_windows"]
    assert np.array_equal(win, keys[:len(win)])
    assert np.array_equal(ou.revcomp(keys, k), unhex(e["revcomp"], k))
    can = ou.canonical(keys, k)
    assert np.array_equal(can, unhex(e["canonical"], k))
    for name, which in (("murmur_store", 0), ("murmur_dist", 1), ("farm_store", 2), ("farm_dist", 3),
                        ("farm_debug_store", 4), ("farm_debug_dist", 5)):
        want = np.array([int(x, 16) for x in e["hash_of_canonical"][name]], dtype=np.uint64)
        assert np.array_equal(ou.hashes(can, k, which), want), name
    # sentinels are the top word 0x8000.. / 0xC000.., others 0 (kmer_hash.hpp:408-414)
    s = unhex(e["sentinels"], k).reshape(2, -1)
    assert int(s[0, -1]) == 0x8000000000000000 and int(s[1, -1]) == 0xC000000000000000 and not s[:, :-1].any()


def test_survey_known_answers():
    # SURVEY.md §8c probe values (reference headers, g++ -O3)
    s = b"ACGTACGTTTGACCAGTACGATCGATCGGGATCCAAATTTGGGCCCAAACGTACGTAGCTAGCTAGCATCG"
    fq = np.frombuffer(b"@x\n" + s + b"\n+\n" + b"I" * len(s) + b"\n", dtype=np.uint8)
    k31, _, _ = ou.fastq_kmers(fq, 31, canonical=False)
    assert int(k31[0]) == 0x06c6fe14b18d8da8
    assert int(ou.revcomp(k31[:1], 31)[0]) == 0x3563636c7ad01b1b
    c = ou.canonical(k31[:1], 31)
    assert int(c[0]) == 0x06c6fe14b18d8da8
    want = {0: 0x02c378a1e090fadc, 1: 0xa9172f0e7a1e50ea, 2: 0xb4fa159c1a1b9a24, 3: 0x333e3d3aa9d08394,
            4: 0xd3b983f0cb4f1571, 5: 0x83fc12a37c2fef46}
    for which, v in want.items():
        assert int(ou.hashes(c, 31, which)[0]) == v
    k63, _, _ = ou.fastq_kmers(fq, 63, canonical=True)
    assert (int(k63[0, 1]), int(k63[0, 0])) == (0x06c6fe14b18d8da8, 0xd40fea5406c6c9c9)
    want = {0: 0x6ee88ef30aaaa1af, 1: 0xe0a334424a045bb4, 2: 0xdb93cecf1ff1144a, 3: 0x51f3bd39c78535c4,
            4: 0xc1ee95c979004246, 5: 0xf390f50d74637ad9}
    for which, v in want.items():
        assert int(ou.hashes(k63[:1], 63, which)[0]) == v


@pytest.mark.parametrize("k", (15, 21, 31, 35))
@pytest.mark.parametrize("fx", ("test.small.fastq", "natural.fastq", "natural.withN.fastq"))
def test_count_dump_matches_reference_dense_hash_map(k, fx):
    e = KAT["k"][str(k)]["count"][fx]
    keys, _, _ = ou.fastq_kmers(ou.read_fixture(fx), k, canonical=True)
    dk, dc = ou.count_sorted(keys, k)
    assert len(dk) == e["distinct"]
    assert np.array_equal(dk, unhex(e["keys"], k))
    assert dc.tolist() == e["counts"]
    # per-rank sizes at p ranks: KeyToRank = dist_hash % p (distributed_densehash_map.hpp:148-170)
    for hname, hk in (("murmur", 0), ("farm", 1)):
        dist = ou.hashes(dk, k, ou.hash_which(hk, True))
        for p in (1, 2, 3):
            sizes = np.bincount((dist % np.uint64(p)).astype(np.int64), minlength=p).tolist()
            assert sizes == e["rank_sizes"]["%s_p%d" % (hname, p)]


def test_port_baseline_agrees_with_sorted_count():
    buf = ou.read_fixture("test.medium.fastq")
    keys, _, _ = ou.fastq_kmers(buf, 31, canonical=True)
    dk, dc = ou.count_sorted(keys, 31)
    ok = np.zeros(len(keys), dtype=np.uint64); oc = np.zeros(len(keys), dtype=np.uint32)
    import ctypes
    nk = ctypes.c_long(0)
    d = ou.oracle().kmio_count_build_port(ou._ptr(buf), buf.size, 31, 2, ou._ptr(ok), ou._ptr(oc), len(keys), ctypes.byref(nk))
    assert d == len(dk) and nk.value == len(keys)
    order = np.argsort(ok[:d])
    assert np.array_equal(ok[:d][order], dk) and np.array_equal(oc[:d][order].astype(np.uint64), dc)


def test_fastq_malformed_raises():
    with pytest.raises(ValueError):
        ou.fastq_kmers(np.frombuffer(b"ACGT\nACGT\n+\nIIII\n", dtype=np.uint8), 3)
    with pytest.raises(ValueError):
        ou.fastq_kmers(np.frombuffer(b"@a\nACGT\n-\nIIII\n", dtype=np.uint8), 3)


def test_edge_cases():
    assert len(ou.fastq_kmers(np.zeros(0, dtype=np.uint8), 31)[0]) == 0
    # read shorter than K, CRLF line ends, blank lines, no trailing newline
    fq = np.frombuffer(b"\n\n@a\r\nACGTAC\r\n+\r\nIIIIII\r\n\r\n@b\nACG\n+\nIII", dtype=np.uint8)
    keys, ids, reads = ou.fastq_kmers(fq, 4, canonical=False, want_ids=True)
    assert reads == 2 and len(keys) == 3
    assert [int(x) for x in keys] == [0b00011011, 0b01101100, 0b10110001]
    # ShortSequenceKmerId: record offset << 16 | offset in record (sequence.hpp:103-160)
    assert [int(x) for x in ids] == [(2 << 16) | 4, (2 << 16) | 5, (2 << 16) | 6]
    fa = np.frombuffer(b"ACGT\n>s1\nAC\nGT\r\nA\n;c\n>s2\nTTTT", dtype=np.uint8)
    keys, ids, seqs = ou.fasta_kmers(fa, 4, canonical=False)
    assert seqs == 2 and [int(x) for x in keys] == [0b00011011, 0b01101100, 0xFF]
    # leading non-header run is skipped and shifts the ids by one (fasta_loader.hpp:571-590)
    assert [int(x) >> 40 for x in ids] == [1, 1, 2]
    assert [int(x) & 0xFFFFFFFFFF for x in ids] == [9, 10, 12 + 0, ][:2] + [int(ids[2]) & 0xFFFFFFFFFF]


@pytest.mark.skipif(ou.ref_lib() is None, reason="oracle/_ref not built (no /root/reference on this box)")
@pytest.mark.parametrize("k", KS)
def test_oracle_vs_live_reference(k):
    ref, dbg = ou.ref_lib("ndebug"), ou.ref_lib("debug")
    keys = np.ascontiguousarray(ou.random_keys(k, 20000, seed=100 + k))
    n = keys.size // ou.nwords(k)
    a = np.zeros_like(keys)
    ref.kmiref_revcomp(k, ou._ptr(keys), ou._ptr(a), n)
    assert np.array_equal(a, ou.revcomp(keys, k))
    ref.kmiref_canonical(k, ou._ptr(keys), ou._ptr(a), n)
    assert np.array_equal(a, ou.canonical(keys, k))
    h = np.zeros(n, dtype=np.uint64)
    for which in range(4):
        ref.kmiref_hash(k, which, ou._ptr(keys), ou._ptr(h), n)
        assert np.array_equal(h, ou.hashes(keys, k, which))
    for which in (2, 3):
        dbg.kmiref_hash(k, which, ou._ptr(keys), ou._ptr(h), n)
        assert np.array_equal(h, ou.hashes(keys, k, which + 2))
    for c in range(256):
        assert ref.kmiref_from_ascii(c) == ou.oracle().kmio_from_ascii(c)


# ---- oracle tier T1: the reference's REAL parsers and index, executed at p = 1 (tests/golden/make_golden_t1.py) -------------
import hashlib

T1 = json.load(open(os.path.join(ou.ROOT, "tests", "golden", "ref_t1.json")))["entries"] if os.path.exists(
    os.path.join(ou.ROOT, "tests", "golden", "ref_t1.json")) else []


def _rows(keys, last, k):
    keys = np.ascontiguousarray(keys, dtype=np.uint64).reshape(len(last), -1)
    return np.concatenate([keys, np.asarray(last, dtype=np.uint64).reshape(-1, 1)], axis=1)


@pytest.mark.parametrize("e", T1, ids=["%s-k%d-%s" % (e["mode"], e["k"], e["fixture"]) for e in T1])
def test_oracle_equals_reference_executed_vectors(e):
    """Every (k-mer, position id) tuple the oracle's walkers emit — in file order — and every (canonical k-mer, count) entry of
    its count index equal what the reference's own FASTQParser / FASTAParser / KmerPositionTupleParser / CountIndex code
    produced when it was run over the same fixture (SHA-256 of the row dump)."""
    assert "error" not in e, e
    buf = ou.read_fixture(e["fixture"])
    k = e["k"]
    if e["mode"] == "fastq_pos":
        keys, ids, _ = ou.fastq_kmers(buf, k, canonical=False, want_ids=True)
        rows = _rows(keys, ids, k)
    elif e["mode"] == "fasta_pos":
        keys, ids, _ = ou.fasta_kmers(buf, k, canonical=False)
        rows = _rows(keys, ids, k)
    else:
        keys, _, _ = ou.fastq_kmers(buf, k, canonical=True)
        dk, dc = ou.count_sorted(keys, k)
        rows = _rows(dk, dc, k)
    assert rows.shape[0] == e["rows"]
    assert hashlib.sha256(np.ascontiguousarray(rows).tobytes()).hexdigest() == e["sha256"]


# ---- f3: identity / cpp_std hash flavours (reference src/index/kmer_hash.hpp:157-232) -----------------------------------------
KAT2 = json.load(open(os.path.join(ou.ROOT, "tests", "golden", "ref_kat_hash2.json")))


@pytest.mark.parametrize("k", KS)
def test_identity_and_std_hash_match_reference_executed_vectors(k):
    """The C oracle's restatement against values computed by the reference's own identity<> / cpp_std<> classes, built
    with ceilLog2(p) prefix bits as KeyToRank builds them (tests/golden/make_golden_hash2.py)."""
    can = np.array([int(x, 16) for x in KAT["k"][str(k)]["canonical"]], dtype=np.uint64)
    e = KAT2["k"][str(k)]
    for kind, name in ((kb_hash_identity, "identity"), (kb_hash_std, "std")):
        want = np.array([int(x, 16) for x in e[name + "_store"]], dtype=np.uint64)
        assert np.array_equal(ou.hashes_prefix(can, k, kind, False), want), (k, name)
        for p in KAT2["ranks"]:
            want = np.array([int(x, 16) for x in e["%s_dist_p%d" % (name, p)]], dtype=np.uint64)
            assert np.array_equal(ou.hashes_prefix(can, k, kind, True, p), want), (k, name, p)
            if name == "identity":
                assert int(want.max()) < (1 << ou.ceil_log2(p))   # a prefix of ceilLog2(p) bits


def test_identity_std_vs_live_reference():
    ref = ou.ref_lib("ndebug")
    if ref is None or not hasattr(ref, "kmiref_hash_prefix"):
        pytest.skip("oracle/_ref not built")
    for k in KS:
        keys = np.ascontiguousarray(ou.canonical(ou.random_keys(k, 20000, seed=100 + k), k)).reshape(-1)
        n = keys.size // ou.nwords(k)
        for kind in (3, 4):
            for dist, pbits_p in ((False, 1), (True, 2), (True, 5), (True, 8), (True, 256)):
                got = ou.hashes_prefix(keys, k, kind, dist, pbits_p)
                want = np.zeros(n, dtype=np.uint64)
                ref.kmiref_hash_prefix(k, (kind - 3) * 2 + (1 if dist else 0), ou.ceil_log2(pbits_p), ou._ptr(keys), ou._ptr(want), n)
                assert np.array_equal(got, want), (k, kind, dist, pbits_p)


# ---- f4: k-mer quality scores (reference src/index/quality_score_iterator.hpp, quality_scores.hpp) ---------------------------
QKAT = json.load(open(os.path.join(ou.ROOT, "tests", "golden", "ref_quality.json")))


def test_quality_lut_matches_reference_table():
    want = np.array(QKAT["lut_bits"], dtype=np.uint32).view(np.float32)
    assert np.array_equal(ou.quality_lut().view(np.uint32), want.view(np.uint32))     # bit for bit


@pytest.mark.parametrize("k", KS)
def test_quality_scores_match_reference_executed_vectors(k):
    """The C oracle's rolling window against the reference's own QualityScoreGenerationIterator<.., Illumina18<float>>,
    float bit patterns (same glibc exp2f on both sides, so even the last step is exact here)."""
    for name, lines in QKAT["lines"].items():
        for ln, want_bits in zip(lines, QKAT["scores"][str(k)][name]):
            q = ln.encode("latin1")
            rec = np.frombuffer(b"@r\n" + b"A" * len(q) + b"\n+\n" + q + b"\n", dtype=np.uint8).copy()
            got = ou.fastq_quality(rec, k)
            assert np.array_equal(got.view(np.uint32), np.array(want_bits, dtype=np.uint32)), (k, name)


def test_quality_order_and_nsplit_follow_the_tuple_stream():
    """scores come in the order fastq_kmers emits tuples; with nsplit every N-free run restarts the window"""
    buf = ou.read_fixture("natural.withN.fastq")
    for nsplit in (False, True):
        keys, ids, _ = ou.fastq_kmers(buf, 21, nsplit=nsplit, want_ids=True)
        q = ou.fastq_quality(buf, 21, nsplit=nsplit)
        assert len(q) == len(keys)
        assert np.all((q >= 0) & (q <= 1))
