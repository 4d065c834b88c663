"""GPU parity tests: every CUDA stage, through the C ABI, against the CPU oracle and the golden
vectors produced by the reference's own code. Bit-exact (integer work): no tolerances anywhere."""
import json
import os

import numpy as np
import pytest

import kmerind_b200 as kb
import oracle_util as ou

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

KAT = json.load(open(os.path.join(ou.ROOT, "tests", "golden", "ref_kat.json")))
KS = (15, 21, 31, 35, 47, 63)
FASTQ = ["natural.fastq", "natural.withN.fastq", "test.debruijn.small.fastq", "test.debruijn.tiny.fastq",
         "test.medium.fastq", "test.small.fastq", "test.unitiq1.fastq", "test.unitiq1.short2.fastq",
         "test.unitiq1.short.fastq", "test.unitiq2.fastq", "test.unitiqs.fastq"]
FASTA = ["natural.fasta", "natural.withN.fasta", "test2.fasta", "test.fasta", "test.medium.fasta", "test.unitiqs.fasta"]


def dev(a):
    """numpy -> CUDA tensor (torch is only the device-memory plumbing here)."""
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint64:
        return torch.from_numpy(a.view(np.int64)).cuda()
    return torch.from_numpy(a).cuda()


def host_u64(t):
    return t.cpu().numpy().view(np.uint64)


def dev_text(buf):
    t = torch.empty(buf.size + 64, dtype=torch.uint8, device="cuda")
    t[:buf.size] = torch.from_numpy(np.array(buf, dtype=np.uint8, copy=True)).cuda()
    return t


def sort_keys(keys, k):
    keys = ou.shape_keys(np.ascontiguousarray(keys).reshape(-1), k)
    if ou.nwords(k) == 2:
        return keys[np.lexsort((keys[:, 0], keys[:, 1]))]
    return np.sort(keys)


def gpu_parse(buf, k, fasta=False, canonical=True, nsplit=False, want_ids=False, hash=kb.HASH_MURMUR, file_offset=0):
    nw = ou.nwords(k)
    with kb.Index(k=k, canonical=canonical, nsplit=nsplit, hash=hash) as idx:
        text = dev_text(buf)
        cap = max(int(buf.size), 1)
        keys = torch.zeros(cap * nw, dtype=torch.int64, device="cuda")
        ids = torch.zeros(cap, dtype=torch.int64, device="cuda") if want_ids else None
        if fasta:
            n = idx.parse_fasta_dev(text.data_ptr(), buf.size, keys.data_ptr(), cap, ids.data_ptr() if want_ids else None)
        else:
            n = idx.parse_fastq_dev(text.data_ptr(), buf.size, keys.data_ptr(), cap, ids.data_ptr() if want_ids else None,
                                    file_offset)
        torch.cuda.synchronize()
        k_host = ou.shape_keys(host_u64(keys)[:n * nw].copy(), k)
        return k_host, (host_u64(ids)[:n].copy() if want_ids else None)


@pytest.mark.parametrize("k", KS)
@pytest.mark.parametrize("hash_kind", (kb.HASH_MURMUR, kb.HASH_FARM, kb.HASH_FARM_DEBUG))
def test_hash_and_canonical_bit_exact(k, hash_kind):
    nw = ou.nwords(k)
    e = KAT["k"][str(k)]
    golden = ou.shape_keys(np.array([int(x, 16) for x in e["keys"]], dtype=np.uint64), k)
    keys = np.concatenate([golden.reshape(-1), np.ascontiguousarray(ou.random_keys(k, 200000, seed=k)).reshape(-1)])
    n = keys.size // nw
    with kb.Index(k=k, hash=hash_kind) as idx:
        d_in = dev(keys)
        d_can = torch.zeros_like(d_in)
        idx.canonicalize_dev(d_in.data_ptr(), d_can.data_ptr(), n)
        can = host_u64(d_can)
        assert np.array_equal(can, ou.canonical(keys, k).reshape(-1))
        ng = golden.size // nw
        assert np.array_equal(can[:golden.size], np.array([int(x, 16) for x in e["canonical"]], dtype=np.uint64))
        names = {kb.HASH_MURMUR: "murmur", kb.HASH_FARM: "farm", kb.HASH_FARM_DEBUG: "farm_debug"}[hash_kind]
        for which, suffix in ((0, "store"), (1, "dist")):
            d_h = torch.zeros(n, dtype=torch.int64, device="cuda")
            idx.hash_dev(which, d_can.data_ptr(), d_h.data_ptr(), n)
            h = host_u64(d_h)
            assert np.array_equal(h, ou.hashes(can, k, ou.hash_which(hash_kind, which == 1)))
            want = np.array([int(x, 16) for x in e["hash_of_canonical"]["%s_%s" % (names, suffix)]], dtype=np.uint64)
            assert np.array_equal(h[:ng], want)  # values computed by the reference's own code


@pytest.mark.parametrize("k", KS)
@pytest.mark.parametrize("hash_kind", (kb.HASH_IDENTITY, kb.HASH_STD))
def test_identity_and_std_hash_bit_exact(k, hash_kind):
    """f3: the identity / cpp_std flavours (kmer_hash.hpp:157-232). The dist hash carries ceilLog2(nranks) prefix bits, as
    KeyToRank builds it; values of the reference's own classes from tests/golden/ref_kat_hash2.json."""
    kat2 = json.load(open(os.path.join(ou.ROOT, "tests", "golden", "ref_kat_hash2.json")))
    nw = ou.nwords(k)
    can_g = np.array([int(x, 16) for x in KAT["k"][str(k)]["canonical"]], dtype=np.uint64)
    ng = can_g.size // nw
    keys = np.concatenate([can_g, np.ascontiguousarray(ou.canonical(ou.random_keys(k, 100000, seed=3 * k), k)).reshape(-1)])
    n = keys.size // nw
    name = {kb.HASH_IDENTITY: "identity", kb.HASH_STD: "std"}[hash_kind]
    for p in (1, 2, 3, 8, 64, 200):
        with kb.Index(k=k, hash=hash_kind, nranks=p) as idx:
            d_in = dev(keys)
            for which in (0, 1):
                if which == 1 and p == 1 and hash_kind == kb.HASH_STD and k > 32:
                    continue   # the reference shifts a 64-bit word by 64 there (undefined); every key is rank 0's anyway
                d_h = torch.zeros(n, dtype=torch.int64, device="cuda")
                idx.hash_dev(which, d_in.data_ptr(), d_h.data_ptr(), n)
                h = host_u64(d_h)
                assert np.array_equal(h, ou.hashes_prefix(keys, k, hash_kind, which == 1, p)), (k, name, p, which)
                if which == 0 or p in kat2["ranks"]:
                    want = kat2["k"][str(k)][name + ("_store" if which == 0 else "_dist_p%d" % p)]
                    assert np.array_equal(h[:ng], np.array([int(x, 16) for x in want], dtype=np.uint64))
            if p > 1:   # KeyToRank with these flavours: stable partition by (prefix hash % p)
                vals = np.arange(n, dtype=np.uint64)
                dk, dv = dev(keys), dev(vals)
                ok, ov = torch.zeros_like(dk), torch.zeros_like(dv)
                counts = idx.partition_dev(dk.data_ptr(), dv.data_ptr(), n, ok.data_ptr(), ov.data_ptr(), p)
                torch.cuda.synchronize()
                owner = (ou.hashes_prefix(keys, k, hash_kind, True, p) % np.uint64(p)).astype(np.int64)
                assert counts.tolist() == np.bincount(owner, minlength=p).tolist()
                order = np.argsort(owner, kind="stable")
                assert np.array_equal(host_u64(ov), vals[order])
                assert np.array_equal(host_u64(ok).reshape(n, nw), keys.reshape(n, nw)[order])


@pytest.mark.parametrize("name", FASTQ)
@pytest.mark.parametrize("k", (15, 31, 35, 63))
def test_fastq_parse_matches_oracle(name, k):
    buf = ou.read_fixture(name)
    for nsplit in (False, True):
        want, _, _ = ou.fastq_kmers(buf, k, canonical=True, nsplit=nsplit)
        got, _ = gpu_parse(buf, k, nsplit=nsplit)
        assert len(got) == len(want), (name, k, nsplit)
        assert np.array_equal(sort_keys(got, k), sort_keys(want, k))
    want, _, _ = ou.fastq_kmers(buf, k, canonical=False)
    got, _ = gpu_parse(buf, k, canonical=False)
    assert np.array_equal(sort_keys(got, k), sort_keys(want, k))


@pytest.mark.parametrize("name", FASTA)
@pytest.mark.parametrize("k", (21, 35, 63))
def test_fasta_parse_matches_oracle(name, k):
    buf = ou.read_fixture(name)
    for nsplit in (False, True):
        wk, wi, _ = ou.fasta_kmers(buf, k, canonical=True, nsplit=nsplit)
        gk, gi = gpu_parse(buf, k, fasta=True, nsplit=nsplit, want_ids=True)
        assert len(gk) == len(wk), (name, k, nsplit)
        a, b = ou.sort_tuples(gk, gi, k)
        c, d = ou.sort_tuples(wk, wi, k)
        assert np.array_equal(a, c) and np.array_equal(b, d)


EDGE_FASTQ = [
    b"", b"\n\n", b"@a\nACGT\n+\nIIII", b"@a\nACGT\n+\nIIII\n", b"\r\n@a\r\nACGTACGTAC\r\n+\r\nIIIIIIIIII\r\n\r\n\r\n@b\nAC\n+\nII\n",
    b"@a\nacgtnNACGTacgtRYKM\n+a\n@@@@@@@@@@@@@@@@@@\n@b\nTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTT\n+\n++++++++++++++++++++++++++++++++++++++++\n",
    b"@q\n" + b"ACGT" * 5000 + b"\n+\n" + b"@" * 20000 + b"\n",
]


@pytest.mark.parametrize("i", range(len(EDGE_FASTQ)))
def test_fastq_edge_cases(i):
    buf = np.frombuffer(EDGE_FASTQ[i], dtype=np.uint8)
    for k in (15, 31, 63):
        for nsplit in (False, True):
            want, _, _ = ou.fastq_kmers(buf, k, nsplit=nsplit)
            got, _ = gpu_parse(buf, k, nsplit=nsplit)
            assert np.array_equal(sort_keys(got, k), sort_keys(want, k)), (i, k, nsplit)


def test_fastq_ragged_random_records():
    # ragged read lengths straddling tile (4096 B) and group (16 B) boundaries, blank lines, CRLF
    rng = np.random.default_rng(7)
    parts = []
    for r in range(3000):
        L = int(rng.integers(1, 400))
        seq = bytes(rng.choice(np.frombuffer(b"ACGTNacgtn", dtype=np.uint8), size=L, p=[.24, .24, .24, .24, .01, .005, .005, .005, .005, .01]))
        eol = b"\r\n" if r % 7 == 0 else b"\n"
        parts.append(b"@r%d" % r + eol + seq + eol + b"+" + eol + b"@" * L + eol + (b"\n" if r % 11 == 0 else b""))
    buf = np.frombuffer(b"".join(parts), dtype=np.uint8)
    for k in (21, 31, 47, 63):
        for nsplit in (False, True):
            want, _, _ = ou.fastq_kmers(buf, k, nsplit=nsplit)
            got, _ = gpu_parse(buf, k, nsplit=nsplit)
            assert np.array_equal(sort_keys(got, k), sort_keys(want, k)), (k, nsplit)


def _ragged_fastq(seed, n_rec, max_len):
    rng = np.random.default_rng(seed)
    parts = []
    for r in range(n_rec):
        L = int(rng.integers(1, max_len))
        seq = bytes(rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), size=L, p=[.245, .245, .245, .245, .02]))
        eol = b"\r\n" if r % 7 == 0 else b"\n"
        name = b"@" if r % 13 == 0 else b"@read%d some description" % r      # 1-byte headers: '@' and bases in one 16-byte group
        parts.append(name + eol + seq + eol + b"+" + eol + b"@" * L + eol + (b"\n" if r % 11 == 0 else b""))
    return np.frombuffer(b"".join(parts), dtype=np.uint8)


@pytest.mark.parametrize("case", ("fixtures", "edge", "ragged", "long_reads"))
def test_fastq_position_ids_match_oracle(case):
    """ShortSequenceKmerId = (file offset of the record's '@' << 16) + offset in the record (sequence.hpp:103-160,
    kmer_parser.hpp:378-399): the (k-mer, id) multiset equals the oracle's, also when the '@' lies in an earlier tile."""
    if case == "fixtures":
        bufs = [ou.read_fixture(n) for n in FASTQ]
    elif case == "edge":
        bufs = [np.frombuffer(b, dtype=np.uint8) for b in EDGE_FASTQ]
    elif case == "ragged":
        bufs = [_ragged_fastq(5, 3000, 400)]
    else:
        bufs = [_ragged_fastq(6, 40, 30000)]      # reads of up to 30 kbp: records span several 4 KiB tiles
    for buf in bufs:
        for k, nsplit, off in ((15, False, 0), (31, True, 12345678901), (63, False, 77)):
            wk, wi, _ = ou.fastq_kmers(buf, k, nsplit=nsplit, want_ids=True, file_offset=off)
            gk, gi = gpu_parse(buf, k, nsplit=nsplit, want_ids=True, file_offset=off)
            assert len(gk) == len(wk)
            a, b = ou.sort_tuples(wk, wi, k), ou.sort_tuples(gk, gi, k)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (case, k, nsplit)


def test_fastq_position_index_equals_oracle():
    buf = _ragged_fastq(8, 2000, 300)
    k = 21
    wk, wi, _ = ou.fastq_kmers(buf, k, want_ids=True, file_offset=4096)
    swk, swi = ou.sort_tuples(wk, wi, k)
    for host in (True, False):
        with kb.PositionIndex(k) as idx:
            if host:
                n = idx.build_fastq(buf, file_offset=4096)
            else:
                text = dev_text(buf)
                n = idx.build_fastq_dev(text.data_ptr(), buf.size, file_offset=4096)
            assert n == len(wk) and idx.local_size() == len(wk)
            gk, gi = idx.sorted_dump()
            assert np.array_equal(gk, swk) and np.array_equal(gi, swi)
            dk, dc = ou.count_sorted(wk, k)
            sel = dk[:: max(1, len(dk) // 200)]
            off, vals = idx.find_positions(sel)
            assert np.diff(off.astype(np.int64)).tolist() == dc[:: max(1, len(dk) // 200)].tolist()


def test_fastq_malformed_is_reported():
    for bad in (b"ACGT\nACGT\n+\nIIII\n", b"@a\nACGT\n-\nIIII\n"):
        buf = np.frombuffer(bad, dtype=np.uint8)
        with pytest.raises(ValueError):
            ou.fastq_kmers(buf, 3)
        with pytest.raises(kb.KmiFormatError):
            gpu_parse(buf, 15)


def _fasta_blocks_gpu(buf, k, cuts, nsplit=False):
    """Parse a FASTA file as len(cuts)-1 blocks cut at arbitrary bytes, the way ranks of a partitioned file do: block
    summaries on the GPU, chained on the host, each block parsed from its own bytes + halo."""
    nw = ou.nwords(k)
    parts = len(cuts) - 1
    out = []
    with kb.Index(k=k, nsplit=nsplit) as idx:
        texts, summ, bol = [], [], []
        for r in range(parts):
            b, e = cuts[r], cuts[r + 1]
            he = kb.fasta_halo_end(buf, e, k)
            texts.append((dev_text(buf[b:he]), he - b))
            bol.append(b == 0 or buf[b - 1] == 0x0A)
            summ.append(idx.fasta_block_summary_dev(texts[r][0].data_ptr(), e - b, bol[r]))
        fsh = buf.size > 0 and buf[0] in (0x3E, 0x3B)
        chain = kb.fasta_chain(summ, bol, fsh)
        for r in range(parts):
            b, e = cuts[r], cuts[r + 1]
            assert (chain[r][0], (chain[r][1] >> 1) & 1, bool(chain[r][1] & 1)) == ou.fasta_prefix_state(buf, b), (r, cuts)
            cap = max(e - b, 1)
            keys = torch.zeros(cap * nw, dtype=torch.int64, device="cuda")
            ids = torch.zeros(cap, dtype=torch.int64, device="cuda")
            n = idx.parse_fasta_dev(texts[r][0].data_ptr(), texts[r][1], keys.data_ptr(), cap, ids.data_ptr(),
                                    block=(b, e - b, chain[r][0], chain[r][1]))
            torch.cuda.synchronize()
            out.append((ou.shape_keys(host_u64(keys)[:n * nw].copy(), k), host_u64(ids)[:n].copy()))
    return out


@pytest.mark.parametrize("name,k", [("test2.fasta", 21), ("test.fasta", 35), ("natural.withN.fasta", 21), ("test.medium.fasta", 21),
                                    ("test.unitiqs.fasta", 63)])
def test_fasta_blocks_match_oracle(name, k):
    """Row a6: every block of a file cut at arbitrary bytes (inside headers, inside lines, empty blocks) emits exactly the
    reference's tuples for that valid range — global sequence ids, file-offset positions, windows running K-1 bases on."""
    buf = ou.read_fixture(name)
    rng = np.random.default_rng(len(name) + k)
    for parts in (2, 3, 7):
        cuts = [0] + sorted(int(x) for x in rng.integers(0, buf.size + 1, size=parts - 1)) + [int(buf.size)]
        if parts == 3:
            cuts[2] = cuts[1]   # an empty block in the middle
        for nsplit in ((False, True) if parts == 3 else (False,)):
            got = _fasta_blocks_gpu(buf, k, cuts, nsplit)
            total = 0
            for r in range(parts):
                wk, wi = ou.fasta_block_kmers(buf, k, cuts[r], cuts[r + 1], nsplit=nsplit)
                a, b = ou.sort_tuples(got[r][0], got[r][1], k)
                c, d = ou.sort_tuples(wk, wi, k)
                assert len(got[r][0]) == len(wk), (name, parts, r, cuts)
                assert np.array_equal(a, c) and np.array_equal(b, d), (name, parts, r, cuts)
                total += len(wk)
            assert total == len(ou.fasta_kmers(buf, k, nsplit=nsplit)[0])


def test_fasta_blocks_edge_cuts():
    """cuts right after '>' / inside a header run / before an EOL, on a file with comment lines, CRLF and blank lines"""
    c = (b"ACGT\n>s1 desc\n;more\nACGTACGTTG\nGTACGTAAGT\r\nACGTACGTACGTAAC\n\n;c\n>s2\nTTTTTGTTTTTTTCTTTTTTTTATTTTT\n>s3\n"
         + b"ACGTTGCA" * 1200 + b"\n")
    buf = np.frombuffer(c, dtype=np.uint8)
    k = 15
    whole = len(ou.fasta_kmers(buf, k)[0])
    for cuts in ([0, 6, 7, 16, 22, int(buf.size)], [0, 5, 5, 5, int(buf.size)], [0, 23, 34, 35, 46, 64, 70, 75, 5000, 5001, int(buf.size)]):
        got = _fasta_blocks_gpu(buf, k, cuts)
        n = 0
        for r in range(len(cuts) - 1):
            wk, wi = ou.fasta_block_kmers(buf, k, cuts[r], cuts[r + 1])
            a, b = ou.sort_tuples(got[r][0], got[r][1], k)
            c2, d = ou.sort_tuples(wk, wi, k)
            assert np.array_equal(a, c2) and np.array_equal(b, d), (cuts, r)
            n += len(wk)
        assert n == whole


@pytest.mark.parametrize("name,k,chunk", [("test.medium.fasta", 21, 4096), ("test.medium.fasta", 21, 40000), ("test.unitiqs.fasta", 63, 4096),
                                          ("natural.withN.fasta", 21, 4096)])
def test_fasta_host_build_in_chunks(name, k, chunk, monkeypatch):
    """kmi_build_fasta from a host pointer streams the text in chunks and builds each chunk as a block of the file (halo = the
    next chunk, inherited header runs / line kind chained): the position index equals the oracle's whole-file tuples."""
    buf = ou.read_fixture(name)
    wk, wi, _ = ou.fasta_kmers(buf, k)
    swk, swi = ou.sort_tuples(wk, wi, k)
    monkeypatch.setenv("KMI_FASTA_CHUNK_BYTES", str(chunk))
    assert buf.size > 2 * chunk
    with kb.PositionIndex(k) as idx:
        assert idx.build_fasta(buf) == len(wk)
        gk, gi = idx.sorted_dump()
        assert np.array_equal(gk, swk) and np.array_equal(gi, swi)
    # the same file as two host blocks (a rank's share each), every one streamed in chunks
    cut = int(buf.size) // 3 + 5
    he = kb.fasta_halo_end(buf, cut, k)
    pre = ou.fasta_prefix_state(buf, cut)
    fsh = kb.FASTA_FILE_STARTS_WITH_HEADER if buf[0] in (0x3E, 0x3B) else 0
    with kb.PositionIndex(k) as idx:
        n0 = idx.build_fasta(buf[:he], block=(0, cut, 0, kb.FASTA_BOL | fsh))
        n1 = idx.build_fasta(buf[cut:], block=(cut, buf.size - cut, pre[0], (kb.FASTA_BOL if pre[2] else 0) | (kb.FASTA_PREV_HEADER if pre[1] else 0) | fsh))
        assert n0 + n1 == len(wk)
        gk, gi = idx.sorted_dump()
        assert np.array_equal(gk, swk) and np.array_equal(gi, swi)


def test_fasta_edge_cases():
    cases = [b"", b"ACGT\n", b">h\n", b">h", b"ACGT\n>s1\nACGTAC\nGTACGT\r\nACGTACGTACGTAAC\n;c\n>s2\nTTTTTTTTTTTTTTTTTTTTTTTTT",
             b">a\n" + b"ACGTTGCA" * 700 + b"\n>b\n\n\n" + b"GATTACA" * 3 + b"\n" + b"C" * 5000,
             b">x\n" + b"\n".join(b"ACGTA"[:1 + (i % 5)] for i in range(4000)) + b"\n"]
    for i, c in enumerate(cases):
        buf = np.frombuffer(c, dtype=np.uint8)
        for k in (15, 21, 63):
            wk, wi, _ = ou.fasta_kmers(buf, k)
            gk, gi = gpu_parse(buf, k, fasta=True, want_ids=True)
            a, b = ou.sort_tuples(gk, gi, k)
            c2, d = ou.sort_tuples(wk, wi, k)
            assert np.array_equal(a, c2) and np.array_equal(b, d), (i, k)


@pytest.mark.parametrize("k", (15, 21, 31, 35))
@pytest.mark.parametrize("fx", ("test.small.fastq", "natural.fastq", "natural.withN.fastq"))
def test_count_index_equals_reference_dump(k, fx):
    """Sorted (canonical k-mer, count) dump == what the reference's Kmer + lex_less + dense_hash_map produce."""
    e = KAT["k"][str(k)]["count"][fx]
    buf = ou.read_fixture(fx)
    for hash_kind in (kb.HASH_MURMUR, kb.HASH_FARM, kb.HASH_FARM_DEBUG):
        with kb.CountIndex(k, hash=hash_kind) as idx:
            n = idx.build_fastq(buf)
            keys, counts = idx.sorted_dump()
            assert idx.size() == e["distinct"]
            assert np.array_equal(keys, ou.shape_keys(np.array([int(x, 16) for x in e["keys"]], dtype=np.uint64), k))
            assert counts.tolist() == e["counts"]
            assert n == sum(e["counts"])


@pytest.mark.parametrize("name", ["test.medium.fastq", "test.unitiqs.fastq", "natural.withN.fastq"])
@pytest.mark.parametrize("k", (31, 63))
@pytest.mark.parametrize("path", (1, 2))
def test_count_index_fused_build_equals_oracle(name, k, path):
    """path 1: direct atomics from the parser; path 2: through the hash-prefix buckets."""
    buf = ou.read_fixture(name)
    for nsplit in (False, True):
        keys, _, _ = ou.fastq_kmers(buf, k, nsplit=nsplit)
        wk, wc = ou.count_sorted(keys, k)
        with kb.CountIndex(k, nsplit=nsplit) as idx:  # no reserve: exercises growth + rehash
            idx.set_build_path(path)
            n = idx.build_fastq(buf)
            gk, gc = idx.sorted_dump()
            assert n == len(keys)
            assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
            # insert(vector<Kmer>) path gives the same contents, and doubles every count when repeated
            with kb.CountIndex(k, nsplit=nsplit) as idx2:
                idx2.set_build_path(path)
                fwd, _, _ = ou.fastq_kmers(buf, k, canonical=False, nsplit=nsplit)
                idx2.insert(fwd)
                assert idx2.checksum() == idx.checksum()
                idx2.insert(fwd)
                k2, c2 = idx2.sorted_dump()
                assert np.array_equal(k2, wk) and np.array_equal(c2, 2 * wc)


def test_count_find_erase_semantics():
    k = 31
    buf = ou.read_fixture("test.medium.fastq")
    keys, _, _ = ou.fastq_kmers(buf, k)
    wk, wc = ou.count_sorted(keys, k)
    absent = ou.canonical(ou.random_keys(k, 1000, seed=5), k)
    absent = absent[~np.isin(absent, wk)]
    with kb.CountIndex(k) as idx:
        idx.build_fastq(buf)
        q = np.concatenate([wk, absent, ou.revcomp(wk, k)])  # queries are canonicalised like inserts
        cnt = idx.count(q)
        assert cnt[:len(wk)].tolist() == [1] * len(wk)          # dense_hash_map::count is 0/1
        assert not cnt[len(wk):len(wk) + len(absent)].any()
        assert cnt[len(wk) + len(absent):].tolist() == [1] * len(wk)
        vals, found = idx.find(q)
        assert np.array_equal(vals[:len(wk)], wc) and found[:len(wk)].all()
        assert not found[len(wk):len(wk) + len(absent)].any()
        assert np.array_equal(vals[len(wk) + len(absent):], wc)
        # erase half, duplicates in the erase batch are harmless
        gone = wk[::2]
        idx.erase(np.concatenate([gone, gone, absent]))
        assert idx.size() == len(wk) - len(gone)
        vals, found = idx.find(wk)
        assert not found[::2].any() and found[1::2].all()
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk[1::2]) and np.array_equal(gc, wc[1::2])
        idx.insert(gone[:5])
        assert idx.size() == len(wk) - len(gone) + 5
        assert idx.find(gone[:5])[0].tolist() == [1] * 5


@pytest.mark.parametrize("name,k", [("test2.fasta", 21), ("test.medium.fasta", 21), ("test.unitiqs.fasta", 63), ("natural.withN.fasta", 15)])
def test_position_index_equals_oracle(name, k):
    buf = ou.read_fixture(name)
    wk, wi, _ = ou.fasta_kmers(buf, k)
    swk, swi = ou.sort_tuples(wk, wi, k)
    with kb.PositionIndex(k) as idx:
        n = idx.build_fasta(buf)
        assert n == len(wk) and idx.local_size() == len(wk)
        gk, gi = idx.sorted_dump()
        assert np.array_equal(gk, swk) and np.array_equal(gi, swi)
        dk, dc = ou.count_sorted(wk, k)
        assert idx.num_keys() == len(dk)
        # count = multiplicity; find = every position of the key
        sel = dk[:: max(1, len(dk) // 500)]
        absent = ou.canonical(ou.random_keys(k, 50, seed=3), k)
        q = np.concatenate([sel.reshape(-1), np.ascontiguousarray(absent).reshape(-1)])
        q = ou.shape_keys(q, k)
        cnt = idx.count(q)
        mult = dict()
        for kk, c in zip(map(lambda r: tuple(np.atleast_1d(r).tolist()), dk), dc):
            mult[kk] = int(c)
        want_cnt = [mult.get(tuple(np.atleast_1d(r).tolist()), 0) for r in q]
        assert cnt.tolist() == want_cnt
        off, vals = idx.find_positions(q)
        assert np.diff(off.astype(np.int64)).tolist() == want_cnt
        pos_of = {}
        for kk, i in zip(map(lambda r: tuple(np.atleast_1d(r).tolist()), swk), swi):
            pos_of.setdefault(kk, []).append(int(i))
        for j, r in enumerate(q):
            got = sorted(int(x) for x in vals[int(off[j]):int(off[j + 1])])
            assert got == pos_of.get(tuple(np.atleast_1d(r).tolist()), [])
        # erase drops every tuple of the key
        idx.erase(sel[:3])
        assert idx.local_size() == len(wk) - sum(want_cnt[:3])
        assert idx.count(sel[:3]).tolist() == [0, 0, 0]


@pytest.mark.parametrize("k", (21, 31, 63))
@pytest.mark.parametrize("p", (1, 2, 3, 8))
def test_partition_is_stable_and_by_owner(k, p):
    nw = ou.nwords(k)
    keys = np.ascontiguousarray(ou.canonical(ou.random_keys(k, 100003, seed=11 + k), k)).reshape(-1)
    n = keys.size // nw
    vals = np.arange(n, dtype=np.uint64)
    for hash_kind in (kb.HASH_MURMUR, kb.HASH_FARM):
        with kb.Index(k=k, hash=hash_kind, nranks=p) as idx:
            dk, dv = dev(keys), dev(vals)
            ok, ov = torch.zeros_like(dk), torch.zeros_like(dv)
            counts = idx.partition_dev(dk.data_ptr(), dv.data_ptr(), n, ok.data_ptr(), ov.data_ptr(), p)
            torch.cuda.synchronize()
            owner = (ou.hashes(keys, k, ou.hash_which(hash_kind, True)) % np.uint64(p)).astype(np.int64)
            assert counts.tolist() == np.bincount(owner, minlength=p).tolist()
            order = np.argsort(owner, kind="stable")  # imxx::distribute keeps the sender's order per bucket
            assert np.array_equal(host_u64(ov), vals[order])
            assert np.array_equal(host_u64(ok).reshape(n, nw), keys.reshape(n, nw)[order])


@pytest.mark.parametrize("k", (21, 31, 63))
def test_partition_dist_transform(k):
    """f3: DistTrans of SingleStrandHashMapParams (BenchmarkKmerIndex.cpp:152-161): owner = dist_hash(T(key)) % p with
    T = lex_less / xor_rev_comp (kmer_transform.hpp:91-126); the keys themselves travel as given, both strands to one rank."""
    nw = ou.nwords(k)
    p = 5
    fw = np.ascontiguousarray(ou.random_keys(k, 50001, seed=77 + k)).reshape(-1)
    keys = np.concatenate([fw, np.ascontiguousarray(ou.revcomp(fw, k)).reshape(-1)])   # every k-mer with its other strand
    n = keys.size // nw
    vals = np.arange(n, dtype=np.uint64)
    rc = np.ascontiguousarray(ou.revcomp(keys, k)).reshape(-1)
    hashed = {0: keys, 1: np.ascontiguousarray(ou.canonical(keys, k)).reshape(-1), 2: keys ^ rc}
    with kb.Index(k=k, hash=kb.HASH_FARM, nranks=p, canonical=False) as idx:
        for trans in (0, 1, 2, 0):
            idx.set_dist_trans(trans)
            dk, dv = dev(keys), dev(vals)
            ok, ov = torch.zeros_like(dk), torch.zeros_like(dv)
            counts = idx.partition_dev(dk.data_ptr(), dv.data_ptr(), n, ok.data_ptr(), ov.data_ptr(), p)
            torch.cuda.synchronize()
            owner = (ou.hashes(hashed[trans], k, ou.hash_which(kb.HASH_FARM, True)) % np.uint64(p)).astype(np.int64)
            if trans:
                assert np.array_equal(owner[:n // 2], owner[n // 2:])   # the two strands meet
            assert counts.tolist() == np.bincount(owner, minlength=p).tolist()
            order = np.argsort(owner, kind="stable")
            assert np.array_equal(host_u64(ov), vals[order])
            assert np.array_equal(host_u64(ok).reshape(n, nw), keys.reshape(n, nw)[order])
    with kb.Index(k=k, hash=kb.HASH_FARM, nranks=p) as idx:   # canonical index: xor is refused, lex_less is a no-op
        with pytest.raises(kb.KmiError):
            idx.set_dist_trans(2)
        idx.set_dist_trans(1)


@pytest.mark.parametrize("k,mode", ((31, kb.MODE_COUNT), (63, kb.MODE_COUNT), (21, kb.MODE_POSITION)))
def test_large_query_batches_equal_small_batches_and_oracle(k, mode):
    """kmi_find_dev / kmi_count_dev on a batch of 2^21 queries against a table beyond the L2 (the pair-probing lookup: two
    16-byte slots = one sector per step, value in the same load): same answers, in input order, as many small batches — and
    as the oracle's counts. The index is also rebuilt at a 50 % fill limit (KMI_PARAM_MAX_LOAD): answers do not change."""
    nw = ou.nwords(k)
    buf = ou.synth_fastq(0, 60000, 150, 3_000_000, 21, 5000, 0)
    keys, ids, _ = ou.fastq_kmers(buf, k, want_ids=(mode == kb.MODE_POSITION))
    wk, wc = ou.count_sorted(keys, k)
    rng = np.random.default_rng(k)
    nq = (1 << 21) + 12345
    present = wk[rng.integers(0, len(wk), nq // 2)]
    present = ou.shape_keys(np.ascontiguousarray(ou.revcomp(np.ascontiguousarray(present).reshape(-1), k)).reshape(-1), k)   # the other strand
    absent = ou.shape_keys(np.ascontiguousarray(ou.random_keys(k, nq - nq // 2, seed=99)).reshape(-1), k)
    q = np.concatenate([present, absent])
    q = q[rng.permutation(nq)]
    with kb.Index(k=k, mode=mode, hash=kb.HASH_FARM, reserve=9_000_000) as idx:     # 2^24 slots: 268 / 537 MB, beyond the L2
        if mode == kb.MODE_COUNT:
            idx.build_fastq(buf)
        else:
            idx.insert(keys, ids)
        assert idx.table_info()["slots"] * idx.table_info()["slot_bytes"] > (96 << 20)
        dq = dev(q)
        dv = torch.zeros(nq, dtype=torch.int64, device="cuda")
        df = torch.zeros(nq, dtype=torch.uint8, device="cuda")
        dc = torch.zeros(nq, dtype=torch.int64, device="cuda")
        idx.find_dev(dq.data_ptr(), nq, dv.data_ptr(), df.data_ptr())            # one large batch
        idx.count_dev(dq.data_ptr(), nq, dc.data_ptr())
        torch.cuda.synchronize()
        big_v, big_f, big_c = host_u64(dv).copy(), df.cpu().numpy().copy(), host_u64(dc).copy()
        dv.zero_(); df.zero_(); dc.zero_()
        step = 1 << 19                                                             # small batches
        for a in range(0, nq, step):
            m = min(step, nq - a)
            idx.find_dev(dq.data_ptr() + a * nw * 8, m, dv.data_ptr() + a * 8, df.data_ptr() + a)
            idx.count_dev(dq.data_ptr() + a * nw * 8, m, dc.data_ptr() + a * 8)
        torch.cuda.synchronize()
        assert np.array_equal(big_v, host_u64(dv)) and np.array_equal(big_f, df.cpu().numpy()) and np.array_equal(big_c, host_u64(dc))
        # against the oracle: frequency (COUNT) / multiplicity (POSITION) of the canonical query
        cq = ou.shape_keys(np.ascontiguousarray(ou.canonical(np.ascontiguousarray(q).reshape(-1), k)).reshape(-1), k)
        if nw == 1:
            pos = np.searchsorted(wk, cq)
            pos[pos >= len(wk)] = 0
            hit = wk[pos] == cq
            want = np.where(hit, wc[pos], 0).astype(np.uint64)
            assert np.array_equal(big_f.astype(bool), hit)
            assert np.array_equal(big_v, want)
            assert np.array_equal(big_c, want if mode == kb.MODE_POSITION else hit.astype(np.uint64))
        else:
            assert int(big_f.sum()) == nq // 2
    with kb.Index(k=k, mode=mode, hash=kb.HASH_FARM) as idx:                       # the same index at a 50 % fill limit
        idx.set_max_load(50)
        if mode == kb.MODE_COUNT:
            idx.build_fastq(buf)
        else:
            idx.insert(keys, ids)
        info = idx.table_info()
        assert len(wk) <= 0.5 * info["slots"]
        dq = dev(q)
        dv = torch.zeros(nq, dtype=torch.int64, device="cuda")
        df = torch.zeros(nq, dtype=torch.uint8, device="cuda")
        idx.find_dev(dq.data_ptr(), nq, dv.data_ptr(), df.data_ptr())
        torch.cuda.synchronize()
        assert np.array_equal(host_u64(dv), big_v) and np.array_equal(df.cpu().numpy(), big_f)


def test_synthetic_generator_matches_host_twin():
    rec = kb.synth_record_bytes(150)
    assert rec == 317
    n_reads = 5000
    for (first, seed, n_ppm) in ((0, 1, 0), (123456789, 9, 1000)):
        t = torch.zeros(n_reads * rec, dtype=torch.uint8, device="cuda")
        kb.synth_fastq_dev(0, t.data_ptr(), first, n_reads, 150, 250000, seed, 5000, n_ppm)
        torch.cuda.synchronize()
        want = ou.synth_fastq(first, n_reads, 150, 250000, seed, 5000, n_ppm)
        assert np.array_equal(t.cpu().numpy(), want)


@pytest.mark.parametrize("path", (0, 1, 2, 3))
def test_synthetic_build_parity_and_properties(path):
    """A 30x synthetic slice (the bench workload's shape, smaller): dump equality with the oracle,
    then size-independent properties: sum of counts == k-mers, chunked build == one-shot build."""
    n_reads, G = 60000, 300000
    buf = ou.synth_fastq(0, n_reads, 150, G, 1, 5000, 0)
    keys, _, reads = ou.fastq_kmers(buf, 31)
    assert reads == n_reads and len(keys) == n_reads * 120
    wk, wc = ou.count_sorted(keys, 31)
    seg = path != 3   # path 3: bucketed with L2 atomics instead of shared-memory segments
    path = min(path, 2)
    with kb.CountIndex(31, hash=kb.HASH_FARM, reserve=len(wk)) as idx:
        idx.set_build_path(path)
        idx.set_segment_insert(seg)
        assert idx.build_fastq(buf) == len(keys)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
        ck = idx.checksum()
        assert ck[1] == len(keys)
        assert idx.table_info()["slots"] * 0.7 >= len(wk)
    with kb.CountIndex(31, hash=kb.HASH_FARM) as idx:  # two record-aligned halves, growth on demand
        idx.set_build_path(path)
        idx.set_segment_insert(seg)
        half = (n_reads // 2) * 317
        idx.build_fastq(buf[:half])
        idx.build_fastq(buf[half:])
        assert idx.checksum() == ck


@pytest.mark.parametrize("path", (1, 2))
def test_k63_count_index_on_synthetic(path):
    buf = ou.synth_fastq(1000, 20000, 150, 100000, 3, 5000, 500)
    for nsplit in (False, True):
        keys, _, _ = ou.fastq_kmers(buf, 63, nsplit=nsplit)
        wk, wc = ou.count_sorted(keys, 63)
        with kb.CountIndex(63, nsplit=nsplit) as idx:
            idx.set_build_path(path)
            assert idx.build_fastq(buf) == len(keys)
            gk, gc = idx.sorted_dump()
            assert np.array_equal(gk, wk) and np.array_equal(gc, wc)


def test_lazy_clear_is_invisible():
    """kmi_clear does not write the table (the next bucketed build initialises it in shared memory); every other access
    must still see an empty table."""
    buf = ou.synth_fastq(0, 20000, 150, 100000, 9, 5000, 0)
    keys, _, _ = ou.fastq_kmers(buf, 31)
    wk, wc = ou.count_sorted(keys, 31)
    q = wk[:2000]
    with kb.CountIndex(31) as idx:
        idx.set_build_path(2)
        assert idx.build_fastq(buf) == len(keys)
        idx.clear()
        assert idx.local_size() == 0
        vals, found = idx.find(q)                       # lookup on the lazily cleared table
        assert not found.any() and not vals.any()
        assert idx.count(q).tolist() == [0] * len(q)
        k0, c0 = idx.sorted_dump()
        assert len(k0) == 0 and idx.checksum() == (0, 0)
        idx.insert(q[:100])                             # direct insert after the reads above
        assert idx.local_size() == 100 and idx.find(q[:100])[1].all()
        idx.clear()
        idx.erase(q[:10])                               # erase on an empty (virgin) table
        idx.insert(q[:50])                              # small insert straight after a clear
        assert idx.local_size() == 50
        idx.clear()
        idx.reserve(4 * len(wk))                        # growth of a virgin table: nothing to rehash
        assert idx.build_fastq(buf) == len(keys)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)


def test_two_fine_passes_on_a_very_large_table():
    """A table of 2^30 slots has 2^18 shared-memory segments = 1024 fine buckets per coarse bucket: the fine scatter
    then runs as two passes (4 mid buckets, then 256 fine buckets per group of 256 mid regions). Same contents."""
    free, _ = torch.cuda.mem_get_info()
    if free < 24 << 30:
        pytest.skip("needs 24 GB of free device memory")
    buf = ou.synth_fastq(0, 30000, 150, 200000, 5, 5000, 0)
    keys, _, _ = ou.fastq_kmers(buf, 31)
    wk, wc = ou.count_sorted(keys, 31)
    with kb.CountIndex(31, hash=kb.HASH_FARM, reserve=700_000_000) as idx:
        idx.set_build_path(2)
        assert idx.table_info()["slots"] == 1 << 30
        assert idx.build_fastq(buf) == len(keys)
        assert idx.build_fastq(buf) == len(keys)           # second build: segments are loaded, not initialised
        assert idx.local_size() == len(wk)
        assert idx.checksum()[1] == 2 * len(keys)
        q = wk[:: max(1, len(wk) // 5000)]
        vals, found = idx.find(q)
        assert found.all() and np.array_equal(vals, 2 * wc[:: max(1, len(wk) // 5000)].astype(np.uint64))


@pytest.mark.parametrize("frac", (0.1, 0.97))
def test_bucketed_build_survives_key_skew(frac):
    """A heavy-hitter key (poly-A reads) overflows its hash-prefix bucket: at 10 % the spill goes to the
    overflow region, at 97 % that is exhausted too and the round is redone with direct atomics. Either
    way the index equals the oracle's."""
    n_reads = 40000
    buf = ou.synth_fastq(0, n_reads, 150, 200000, 7, 5000, 0).copy()
    recs = buf.reshape(n_reads, 317)
    rng = np.random.default_rng(5)
    hot = rng.random(n_reads) < frac
    recs[hot, 13:163] = ord("A")
    keys, _, _ = ou.fastq_kmers(buf, 31)
    wk, wc = ou.count_sorted(keys, 31)
    with kb.CountIndex(31, hash=kb.HASH_FARM) as idx:
        idx.set_build_path(2)
        assert idx.build_fastq(buf) == len(keys)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
    with kb.CountIndex(31, hash=kb.HASH_FARM) as idx:  # the same through insert(vector<Kmer>)
        idx.set_build_path(2)
        idx.insert(keys)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)


def test_cpp_adapter_against_std_containers(tmp_path):
    """The header-only MapType (include/kmi_index.hpp) — insert / find / count / erase, each with and without a predicate,
    to_vector, size — against std::map brute force, in C++, the way the reference tests its maps
    (src/containers/test/test_densehash_map.cpp)."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "adapter_check")
    subprocess.check_call(["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(root, "include"),
                           os.path.join(root, "tests", "cpp", "adapter_check.cpp"), kb.LIB_PATH, "-o", exe,
                           "-Wl,-rpath," + os.path.dirname(kb.LIB_PATH)])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "adapter ok" in r.stdout, r.stdout + r.stderr


def test_cpp_adapter_with_the_reference_key_type():
    """The same checks with Key = bliss::common::Kmer<K, DNA, uint64_t>: the binary was compiled against the reference's
    headers where they lie (oracle/Makefile -> oracle/_ref/adapter_check_bliss; the headers do not travel to the GPU box)."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "oracle", "_ref", "adapter_check_bliss")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/adapter_check_bliss not built (needs /root/reference at build time)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "adapter ok (bliss::common::Kmer)" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("stage_threads", ("3", "0"))
@pytest.mark.parametrize("name,k,chunk", [("natural.fastq", 21, 3000), ("test.medium.fastq", 31, 5000), ("natural.withN.fastq", 15, 1 << 28)])
def test_build_file_fastq_streams_record_aligned_chunks(name, k, chunk, stage_threads, tmp_path, monkeypatch):
    """kmi_build_file: the file is cut into chunks that end at record boundaries — read by the slice ring's threads straight
    into pinned slices (stage_threads > 0: boundaries found in a chunk's tail window), or chunk by chunk into two pinned
    buffers with the rest carried over (stage_threads 0); count index == oracle; position index ids are file offsets of the
    whole file."""
    buf = ou.read_fixture(name)
    path = tmp_path / name
    path.write_bytes(buf.tobytes())
    monkeypatch.setenv("KMI_FILE_CHUNK_BYTES", str(chunk))
    monkeypatch.setenv("KMI_STAGE_THREADS", stage_threads)
    monkeypatch.setenv("KMI_STAGE_SLICE_BYTES", "1024")
    wk, wi, _ = ou.fastq_kmers(buf, k, want_ids=True)
    dk, dc = ou.count_sorted(wk, k)
    with kb.CountIndex(k) as idx:
        assert idx.build_file(str(path)) == len(wk)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, dk) and np.array_equal(gc, dc)
    with kb.PositionIndex(k) as idx:
        assert idx.build_file(str(path), format=1) == len(wk)
        gk, gi = idx.sorted_dump()
        swk, swi = ou.sort_tuples(wk, wi, k)
        assert np.array_equal(gk, swk) and np.array_equal(gi, swi)


@pytest.mark.parametrize("name,k,chunk", [("test.medium.fasta", 21, 50000), ("natural.withN.fasta", 21, 2000), ("test.unitiqs.fasta", 63, 1500),
                                          ("test2.fasta", 21, 1 << 28)])
def test_build_file_fasta_streams_blocks(name, k, chunk, tmp_path, monkeypatch):
    buf = ou.read_fixture(name)
    path = tmp_path / name
    path.write_bytes(buf.tobytes())
    monkeypatch.setenv("KMI_FILE_CHUNK_BYTES", str(chunk))
    wk, wi, _ = ou.fasta_kmers(buf, k)
    swk, swi = ou.sort_tuples(wk, wi, k)
    with kb.PositionIndex(k) as idx:
        assert idx.build_file(str(path)) == len(wk)
        gk, gi = idx.sorted_dump()
        assert np.array_equal(gk, swk) and np.array_equal(gi, swi)


def test_build_file_errors(tmp_path):
    p = tmp_path / "reads.txt"
    p.write_bytes(b"@a\nACGTACGTTGCATGCA\n+\nIIIIIIIIIIIIIIII\n")
    with kb.CountIndex(15) as idx:
        with pytest.raises(kb.KmiError, match="extension is not supported"):
            idx.build_file(str(p))
        with pytest.raises(kb.KmiError, match="cannot open"):
            idx.build_file(str(tmp_path / "missing.fastq"))
        assert idx.build_file(str(p), format=1) == 2
        e = tmp_path / "empty.fastq"
        e.write_bytes(b"")
        assert idx.build_file(str(e)) == 0


@pytest.mark.parametrize("k", (21, 63))
def test_count_tuples_are_added(k):
    """kmi_insert with values in COUNT mode is the reduction insert: the tuple's count is added to the key's entry."""
    keys = ou.random_keys(k, 5000, seed=3)
    keys = np.concatenate([keys, keys[:2000]])
    cnt = np.arange(1, len(keys) + 1, dtype=np.uint64) % np.uint64(977)
    can = ou.shape_keys(np.ascontiguousarray(ou.canonical(keys, k)).reshape(-1), k)
    with kb.CountIndex(k) as idx:
        idx.insert(keys)                       # +1 each
        idx.insert(keys, cnt)                  # + count
        got = idx.count(keys)                  # map semantics: 0/1
        assert got.min() == 1
        vals, found = idx.find(keys)
        assert found.all()
        want = {}
        rows = [tuple(np.atleast_1d(r).tolist()) for r in can]
        for r, c in zip(rows, cnt.tolist()):
            want[r] = want.get(r, 0) + 1 + c
        assert [want[r] & 0xFFFFFFFF for r in rows] == vals.tolist()
        assert idx.local_size() == len(want)


def test_save_and_load_round_trip(tmp_path):
    """kmi_save writes to_vector() sorted by (key, value); kmi_load adds a file's entries: a reloaded index equals the
    original, loading twice doubles the counts, shards merge, a K mismatch is refused."""
    buf = ou.read_fixture("natural.fastq")
    for k in (21, 63):
        p = str(tmp_path / ("count%d.kmi" % k))
        with kb.CountIndex(k) as idx:
            idx.build_fastq(buf)
            dk, dc = idx.sorted_dump()
            idx.save(p)
        raw = np.fromfile(p, dtype=np.uint64)
        nw = ou.nwords(k)
        assert raw[:1].tobytes() == b"KMIDX001" and int(raw[3]) == len(dk)      # header: magic, (k, mode), (nw, canonical), n
        assert np.array_equal(raw[8:8 + len(dk) * nw], np.ascontiguousarray(dk).reshape(-1))   # keys sorted, then values
        assert np.array_equal(raw[8 + len(dk) * nw:], dc.astype(np.uint64))
        with kb.CountIndex(k) as idx:
            assert idx.load(p) == len(dk)
            gk, gc = idx.sorted_dump()
            assert np.array_equal(gk, dk) and np.array_equal(gc, dc)
            idx.load(p)
            gk, gc = idx.sorted_dump()
            assert np.array_equal(gk, dk) and np.array_equal(gc, 2 * dc)
        with kb.CountIndex(31) as idx:
            with pytest.raises(kb.KmiError, match="holds a K="):
                idx.load(p)
    fa = ou.read_fixture("test2.fasta")
    p = str(tmp_path / "pos.kmi")
    with kb.PositionIndex(21) as idx:
        idx.build_fasta(fa)
        dk, di = idx.sorted_dump()
        idx.save(p)
    with kb.PositionIndex(21) as idx:
        assert idx.load(p) == len(dk)
        gk, gi = idx.sorted_dump()
        assert np.array_equal(gk, dk) and np.array_equal(gi, di)
        sel = dk[::7]
        off, vals = idx.find_positions(sel)
        assert int(off[-1]) == sum(int((dk == s).sum()) for s in sel)
    with kb.CountIndex(21) as idx:
        with pytest.raises(kb.KmiError):
            idx.load(p)      # a position file into a count index
        with pytest.raises(kb.KmiError):
            idx.load(str(tmp_path / "missing.kmi"))


def _slot_hash_u64(keys):
    """kmi_device.cuh slot_hash for one-word keys, in numpy (mod 2^64)."""
    with np.errstate(over="ignore"):
        h = keys.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)
        h ^= h >> np.uint64(32)
        h *= np.uint64(0xD6E8FEB86659FD93)
    return h


@pytest.mark.parametrize("n_keys,expect_error", [(3000, False), (40000, True)])
def test_segment_spills_are_exact_or_reported(n_keys, expect_error):
    """Keys whose home slot is one of the last slots of their shared-memory table segment run off the segment end and
    'spill' to the overflow list, which is inserted with global atomics afterwards. A few thousand of them must come out
    exact; so many that the list (n/8 + 4096 entries) is exhausted must be REPORTED (KMI_ERR_CAPACITY from the insert and
    from every later size query: the error is sticky), never silently dropped (round-1 advisor finding)."""
    k = 31
    rng = np.random.default_rng(11)
    cand = rng.integers(0, 1 << 62, size=12_000_000, dtype=np.uint64)
    cand = np.unique(ou.canonical(cand, k))
    h = _slot_hash_u64(cand)
    # home slot of a 2^17-slot table is h >> 47: the last 4 slots of every 1024-slot block are the last slots of every
    # 256-, 512- and 1024-slot segment (and the last 2 of them for a 2^16-slot table)
    sel = cand[((h >> np.uint64(47)) & np.uint64(1023)) >= np.uint64(1020)][:n_keys]
    assert len(sel) == n_keys
    keys = np.repeat(sel, 3)                      # every key three times
    rng.shuffle(keys)
    with kb.CountIndex(k, hash=kb.HASH_FARM) as idx:
        idx.set_build_path(2)
        if expect_error:
            with pytest.raises(kb.KmiError) as e:
                idx.insert(keys)
                idx.local_size()
            assert e.value.code == kb.ERR_CAPACITY
            with pytest.raises(kb.KmiError):
                idx.local_size()                  # sticky until clear()
            idx.clear()
            idx.insert(sel[:100])
            assert idx.local_size() == 100
        else:
            idx.insert(keys)
            gk, gc = idx.sorted_dump()
            assert np.array_equal(gk, np.sort(sel)) and np.all(gc == 3)


def test_erase_insert_churn_reclaims_tombstones():
    """Erased slots stay DELETED until a rehash; the table's load test must count them (dense_hash_map does), otherwise
    erase/insert churn fills the table with tombstones while the host sees a low load (round-1 advisor finding)."""
    k = 31
    with kb.CountIndex(k, hash=kb.HASH_MURMUR) as idx:
        slots0 = None
        for r in range(12):
            keys = ou.canonical(ou.random_keys(k, 60000, seed=100 + r), k)
            idx.insert(keys)
            if slots0 is None:
                slots0 = idx.table_info()["slots"]
            assert idx.local_size() == len(np.unique(keys))
            c = idx.count(keys)
            assert np.all(c == 1)
            idx.erase(keys)
            assert idx.local_size() == 0
        # 12 x 60000 tombstones would not fit the first table (2^17 slots): it was purged, not grown without bound
        assert idx.table_info()["slots"] <= 2 * slots0
        keys = ou.canonical(ou.random_keys(k, 1000, seed=999), k)
        assert np.all(idx.count(keys) == 0)


def test_fastq_position_ids_reject_records_over_64k():
    """ShortSequenceKmerId keeps the k-mer's offset inside its record in 16 bits; the reference throws std::invalid_argument
    beyond that (src/common/sequence.hpp:177-183). A 70 000-base read must be rejected for a position index (KMI_ERR_ARG)
    and still build a count index."""
    rng = np.random.default_rng(3)
    seq = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, 70000)]
    rec = np.concatenate([np.frombuffer(b"@long\n", np.uint8), seq, np.frombuffer(b"\n+\n", np.uint8), np.full(70000, ord("I"), np.uint8),
                          np.frombuffer(b"\n", np.uint8)])
    wk, _, _ = ou.fastq_kmers(rec, 21)
    with kb.CountIndex(21) as idx:
        assert idx.build_fastq(rec) == len(wk)
    with kb.PositionIndex(21) as idx:
        with pytest.raises(kb.KmiError) as e:
            idx.build_fastq(rec)
        assert e.value.code == kb.ERR_ARG


@pytest.mark.parametrize("n_seq,seq_len,width", [(3, 1000, 80), (2, 777, 70), (5, 64, 64)])
def test_synthetic_fasta_and_queries_match_host_twin(n_seq, seq_len, width):
    """kmi_synth_fasta_dev writes the bytes of oracle kmio_synth_fasta for any byte range; the even query k-mers of
    kmi_synth_queries_dev occur in that genome (either strand)."""
    want = ou.synth_fasta(n_seq, seq_len, width, seed=3)
    assert kb.synth_fasta_bytes(n_seq, seq_len, width) == want.size
    for lo, hi in ((0, want.size), (13, want.size - 5), (want.size // 2, want.size // 2 + 100)):
        t = torch.zeros(hi - lo, dtype=torch.uint8, device="cuda")
        kb.synth_fasta_dev(0, t.data_ptr(), lo, hi - lo, n_seq, seq_len, width, seed=3)
        assert np.array_equal(t.cpu().numpy(), want[lo:hi])
    k = 21
    if seq_len >= k:
        keys, _, _ = ou.fasta_kmers(want, k)
        have = set(keys.tolist())
        q = torch.zeros(400, dtype=torch.int64, device="cuda")
        kb.synth_queries_dev(0, q.data_ptr(), 1000, 400, k, n_seq, seq_len, genome_seed=3, seed=4)
        qc = ou.canonical(host_u64(q), k)
        assert all(int(x) in have for x in qc[0::2])
        assert (host_u64(q) < np.uint64(1 << 42)).all()


def test_k63_build_into_a_large_reserved_table_in_batches():
    """Two-word keys, a table of 2^27 slots (256 fine buckets per coarse bucket: the fine scatter's shared memory passes
    the 48 KB default with its static part), two batches into one index: contents == oracle, second batch adds."""
    k, n_reads = 63, 120000
    buf = ou.synth_fastq(0, n_reads, 150, 400000, 5, 5000, 0)
    keys, _, _ = ou.fastq_kmers(buf, k)
    wk, wc = ou.count_sorted(keys, k)
    half = (n_reads // 2) * 317
    with kb.CountIndex(k, hash=kb.HASH_FARM, reserve=90_000_000) as idx:
        assert idx.table_info()["slots"] == 1 << 27
        idx.set_build_path(2)
        n = idx.build_fastq(buf[:half]) + idx.build_fastq(buf[half:])
        assert n == len(keys)
        assert idx.checksum()[1] == len(keys)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)


def test_value_predicates_on_the_device():
    """find(pred) / count(pred) / erase(pred) for frequency thresholds (kmer_index.hpp:156-194) evaluated by the table sweep:
    equal to filtering the oracle's dump; erased entries are gone for queries and size, and the table keeps working."""
    k = 31
    buf = ou.read_fixture("test.medium.fastq")
    keys, _, _ = ou.fastq_kmers(buf, k)
    extra = ou.canonical(ou.random_keys(k, 5000, seed=2), k)
    allk = np.concatenate([keys, extra, extra[:1000]])
    wk, wc = ou.count_sorted(allk, k)
    with kb.CountIndex(k) as idx:
        idx.insert(allk)
        for op, x, sel in (("ge", 3, wc >= 3), ("le", 1, wc <= 1), ("eq", 2, wc == 2), ("ge", 10**9, wc >= 10**9)):
            assert idx.count_if(op, x) == int(sel.sum())
            gk, gv = idx.find_if(op, x)
            o = np.argsort(gk)
            assert np.array_equal(gk[o], wk[sel]) and np.array_equal(gv[o], wc[sel])
        assert idx.erase_if("eq", 1) == int((wc == 1).sum())
        assert idx.local_size() == int((wc != 1).sum())
        assert np.array_equal(idx.count(wk), (wc != 1).astype(np.uint64))
        idx.insert(wk[wc == 1][:10])
        assert idx.local_size() == int((wc != 1).sum()) + min(10, int((wc == 1).sum()))
    with kb.PositionIndex(21) as pidx:
        with pytest.raises(kb.KmiError):
            pidx.count_if("ge", 1)


def test_saturating_counts():
    """std::plus<uint32_t> wraps at 2^32 (distributed_densehash_map.hpp:2475); the saturating map sticks at 2^32 - 1 (:2904-2947)."""
    k = 31
    keys = ou.canonical(ou.random_keys(k, 3, seed=8), k)
    big = np.array([0xFFFFFFF0, 5, 0xFFFFFFFF], dtype=np.uint64)
    more = np.array([0x20, 7, 1], dtype=np.uint64)
    for sat, want in ((False, [0x10, 12, 0]), (True, [0xFFFFFFFF, 12, 0xFFFFFFFF])):
        with kb.CountIndex(k) as idx:
            idx.set_saturate(sat)
            idx.insert(keys, big)
            idx.insert(keys, more)
            vals, found = idx.find(keys)
            assert found.all() and vals.tolist() == want
            if sat:   # plain key inserts on top of a saturated entry stay saturated, also when many threads add at once
                idx.insert(np.repeat(keys[:1], 5000))
                assert idx.find(keys[:1])[0][0] == 0xFFFFFFFF


@pytest.mark.parametrize("k", (17, 27, 33, 55))
def test_kmer_lengths_from_modules(k):
    """bliss::common::Kmer is a template over any K (kmer.hpp:128-139): lengths outside libkmi.so's built-in set come from
    kmods/libkmi_k<K>.so, loaded by kmi_create. Hashes, canonical form, FASTQ parse and a count index against the C oracle."""
    if not os.path.exists(os.path.join(os.path.dirname(kb.LIB_PATH), "kmods", "libkmi_k%d.so" % k)):
        pytest.skip("module for k=%d not built (python -m kmerind_b200.build --k %d)" % (k, k))
    keys = ou.random_keys(k, 4000, seed=k)
    buf = ou.read_fixture("natural.fastq")
    wk, _, _ = ou.fastq_kmers(buf, k)
    dk, dc = ou.count_sorted(wk, k)
    with kb.CountIndex(k, hash=kb.HASH_FARM) as idx:
        d = dev(keys)
        out = torch.empty_like(d)
        idx.canonicalize_dev(d.data_ptr(), out.data_ptr(), 4000)
        assert np.array_equal(host_u64(out).reshape(keys.shape), ou.canonical(keys, k))
        h = torch.empty(4000, dtype=torch.int64, device="cuda")
        for which in (0, 1):
            idx.hash_dev(which, d.data_ptr(), h.data_ptr(), 4000)
            assert np.array_equal(host_u64(h), ou.hashes(keys, k, ou.hash_which(kb.HASH_FARM, which == 1)))
        assert idx.build_fastq(buf) == len(wk)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, dk) and np.array_equal(gc, dc)
    syn = ou.synth_fastq(0, 40000, 150, 200000, 3, 5000, 200)
    sk, _, _ = ou.fastq_kmers(syn, k)
    wk2, wc2 = ou.count_sorted(sk, k)
    with kb.CountIndex(k, hash=kb.HASH_MURMUR) as idx:    # 12.7 MB: the bucketed build
        assert idx.build_fastq(syn) == len(sk)
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, wk2) and np.array_equal(gc, wc2)


@pytest.mark.parametrize("k,fx", [(31, "natural.withN.fastq"), (31, "test.medium.fastq"), (35, "natural.fastq")])
def test_reference_index_template_over_b200_map(k, fx, tmp_path):
    """The drop-in itself: the REFERENCE's bliss::index::kmer::CountIndex<MapType> template (Index::build_posix with its own file
    reader and FASTQ / k-mer parsers, insert, find, count, size: kmer_index.hpp:98-394), compiled from /root/reference with
    MapType = kmi::b200_map<bliss::common::Kmer<K, DNA, uint64_t>, uint32_t, COUNT> over a single-rank mpi.h
    (oracle/Makefile -> oracle/_ref/ref_index_b200), runs on the GPU and dumps the same sorted (k-mer, count) rows as the
    reference's own counting_densehash_map did (tests/golden/ref_t1.json)."""
    import hashlib
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "oracle", "_ref", "ref_index_b200")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/ref_index_b200 not built (needs /root/reference at build time)")
    out = str(tmp_path / "dump.bin")
    r = subprocess.run([exe, "adapter", str(k), os.path.join(ou.DATA, fx), out], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    a = np.fromfile(out, dtype=np.uint64)
    golden = json.load(open(os.path.join(root, "tests", "golden", "ref_t1.json")))["entries"]
    want = [e for e in golden if e["mode"] == "count" and e["k"] == k and e["fixture"] == fx][0]
    assert int(a[0]) == want["rows"]
    assert hashlib.sha256(a[2:].tobytes()).hexdigest() == want["sha256"]
