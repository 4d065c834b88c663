"""Record-aligned cutting of a FASTQ file: kmi_fastq_find_first_record is host code, like the
reference's FASTQParser::find_first_record (fastq_loader.hpp:269-357), so these tests need no GPU. The C ABI function is
compared with the oracle's restatement at every cut point, and both with the true record starts of the fixtures (what the
reference's own partition tests assert: every block starts at an '@' header line, mpi_test_fastq_seq_parse.cpp)."""
import numpy as np
import pytest

import kmerind_b200 as kb
import oracle_util as ou

FASTQ = ["test.small.fastq", "test.medium.fastq", "natural.fastq", "natural.withN.fastq"]


def record_starts(buf):
    """offsets of line 0 (mod 4) starts, counting non-empty lines as the reference's record walk does"""
    b = buf.tobytes()
    starts, i, line, n = [], 0, 0, len(b)
    while i < n:
        while i < n and b[i] in (10, 13):
            i += 1
        if i >= n:
            break
        if line % 4 == 0:
            starts.append(i)
        while i < n and b[i] not in (10, 13):
            i += 1
        line += 1
    return np.array(starts, dtype=np.int64)


@pytest.mark.parametrize("name", FASTQ)
def test_find_first_record_matches_oracle_and_true_starts(name):
    buf = ou.read_fixture(name)
    n = buf.size
    starts = record_starts(buf)
    step = max(1, n // 400)
    for c in list(range(0, n, step)) + [n - 1, n]:
        want = ou.fastq_find_first_record(buf, c)
        got = kb.fastq_find_first_record(buf, c)
        assert got == want, (name, c)
        if got < n:
            assert got in starts and got >= c or c == 0, (name, c, got)
            # it is the FIRST record start whose header line begins at or after the first line start >= c
            nxt = starts[starts >= c]
            assert len(nxt) and got <= nxt[min(1, len(nxt) - 1)], (name, c, got)


def test_find_first_record_windows_and_errors():
    buf = ou.read_fixture("test.small.fastq")
    n = buf.size
    starts = record_starts(buf)
    # bytes in memory are a window of the file: offsets stay file offsets
    lo, hi = int(starts[2]) - 5, int(starts[5]) + 40
    win = buf[lo:hi]
    assert kb.fastq_find_first_record(win, lo, hi, mem_offset=lo, parent=(0, n)) == ou.fastq_find_first_record(win, lo, hi, lo, (0, n)) == int(starts[2])
    # a range too short to hold four lines and not at the parent's end -> end of the range; garbage -> format error
    assert kb.fastq_find_first_record(buf, 3, 20, parent=(0, n)) == 20
    junk = np.frombuffer(b"xxxx\nyyyy\nzzzz\nwwww\nvvvv\nuuuu\n", dtype=np.uint8)
    with pytest.raises(ValueError):
        ou.fastq_find_first_record(junk, 2, 25, 0, (0, 1000))
    with pytest.raises(kb.KmiError):
        kb.fastq_find_first_record(junk, 2, 25, parent=(0, 1000))


@pytest.mark.parametrize("parts", (2, 3, 8))
def test_cuts_partition_the_kmers(parts):
    """blocks cut at kmi_fastq_find_first_record boundaries are whole records: their k-mer multisets add up"""
    buf = ou.read_fixture("natural.fastq")
    cuts = kb.fastq_cuts(buf, parts)
    assert cuts[0] == 0 and cuts[-1] == buf.size and all(a <= b for a, b in zip(cuts, cuts[1:]))
    whole, _, reads = ou.fastq_kmers(buf, 21)
    got, n_reads = [], 0
    for a, b in zip(cuts, cuts[1:]):
        k, _, r = ou.fastq_kmers(buf[a:b], 21)
        got.append(k)
        n_reads += r
    assert n_reads == reads
    assert np.array_equal(np.sort(np.concatenate(got)), np.sort(whole))


# ---- FASTA blocks (row a6): host side of the cross-block scan and of the K-1 tail ------------------------------------
def block_summary(block, bol):
    """numpy restatement of what kmi_fasta_block_summary_dev reports for a block's own bytes"""
    b = block.tobytes()
    starts = ([0] if bol and len(b) else []) + [i + 1 for i in range(len(b) - 1) if b[i] == 0x0A]
    kinds = [1 if b[s] in (0x3E, 0x3B) else 0 for s in starts]
    runs, prev = 0, 0
    for kd in kinds:
        if kd and not prev:
            runs += 1
        prev = kd
    return (1 if starts else 0, kinds[0] if starts else 0, kinds[-1] if starts else 0, runs)


@pytest.mark.parametrize("name", ("test2.fasta", "test.fasta", "natural.withN.fasta"))
def test_fasta_chain_reproduces_the_prefix_state(name):
    buf = ou.read_fixture(name)
    rng = np.random.default_rng(5)
    fsh = buf[0] in (0x3E, 0x3B)
    for parts in (1, 2, 5, 16):
        cuts = [0] + sorted(int(x) for x in rng.integers(0, buf.size + 1, size=parts - 1)) + [int(buf.size)]
        bol = [c == 0 or buf[c - 1] == 0x0A for c in cuts[:-1]]
        summ = [block_summary(buf[a:b], bl) for a, b, bl in zip(cuts, cuts[1:], bol)]
        chain = kb.fasta_chain(summ, bol, fsh)
        for r, c in enumerate(cuts[:-1]):
            runs0, flags = chain[r]
            assert (runs0, (flags >> 1) & 1, bool(flags & 1)) == ou.fasta_prefix_state(buf, c), (name, cuts, r)
            assert bool(flags & kb.FASTA_FILE_STARTS_WITH_HEADER) == bool(fsh)


def test_fasta_halo_end_counts_non_eol_bytes():
    buf = np.frombuffer(b">h\nACGT\r\nAC\n\nGTTT\n>g\nAAAA", dtype=np.uint8)
    n = buf.size
    for k in (1, 2, 5, 8, 40):
        for e in range(n + 1):
            p, need = e, k - 1
            while p < n and need > 0:
                if buf[p] not in (0x0A, 0x0D):
                    need -= 1
                p += 1
            assert kb.fasta_halo_end(buf, e, k) == p, (k, e)
    # a window of the file in memory: offsets stay file offsets
    assert kb.fasta_halo_end(buf[3:], 100 + 4, 3, mem_offset=100) == 100 + 8   # \\r \\n skipped, then A C
    with pytest.raises(kb.KmiError):
        kb.fasta_halo_end(buf, n + 1, 3)


def test_fasta_blocks_oracle_partition():
    """the oracle's per-block tuples (first base in the valid range) partition the whole file's"""
    buf = ou.read_fixture("test.fasta")
    wk, wi, _ = ou.fasta_kmers(buf, 21)
    cuts = [0, 100, 101, 5000, int(buf.size)]
    parts = [ou.fasta_block_kmers(buf, 21, a, b) for a, b in zip(cuts, cuts[1:])]
    assert sum(len(p[0]) for p in parts) == len(wk)
    assert np.array_equal(np.sort(np.concatenate([p[1] for p in parts])), np.sort(wi))


# ---- row a1: one file, p ranks (partitioned_file::read_file + BlockPartitioner, file.hpp:1239-1251,1342-1431) ----------
@pytest.mark.parametrize("name", FASTQ)
@pytest.mark.parametrize("parts", (1, 2, 3, 8, 64))
def test_file_block_fastq_equals_in_memory_cuts(name, parts, tmp_path):
    """kmi_file_block reads only windows of the file around the boundaries; its blocks must be the ones kmi::fastq_cuts /
    kb.fastq_cuts finds with the whole file in memory: whole records, tiling the file, in rank order"""
    buf = ou.read_fixture(name)
    path = tmp_path / name
    path.write_bytes(buf.tobytes())
    cuts = kb.fastq_cuts(buf, parts)
    starts = set(record_starts(buf).tolist()) | {buf.size}
    prev_end = 0
    for r in range(parts):
        b, e, size = kb.file_block(str(path), r, parts)
        assert size == buf.size and (b, e) == (cuts[r], cuts[r + 1]), (name, parts, r)
        assert b == prev_end and b in starts and e in starts
        prev_end = e
    assert prev_end == buf.size


def test_file_block_large_windows_fasta_and_errors(tmp_path):
    # records longer than the first 64 KiB window: the window grows until it holds four lines
    seq = b"ACGT" * 50000
    rec = b"@r\n" + seq + b"\n+\n" + b"I" * len(seq) + b"\n"
    path = tmp_path / "long.fq"
    path.write_bytes(rec * 3)
    assert kb.file_block(str(path), 0, 2)[:2] == (0, 2 * len(rec))
    assert kb.file_block(str(path), 1, 2)[:2] == (2 * len(rec), 3 * len(rec))
    assert kb.file_block(str(path), 3, 4)[:2] == (3 * len(rec), 3 * len(rec))       # nothing starts after 3/4 of the file
    # FASTA: the byte ranges as they are; format by extension or explicit
    fa = tmp_path / "g.fasta"
    fa.write_bytes(b">a\n" + b"ACGT" * 100 + b"\n")
    n = fa.stat().st_size
    assert [kb.file_block(str(fa), r, 3)[:2] for r in range(3)] == [(0, n // 3), (n // 3, 2 * n // 3), (2 * n // 3, n)]
    assert kb.file_block(str(path), 0, 1, format=2)[:2] == (0, 3 * len(rec))
    empty = tmp_path / "e.fastq"
    empty.write_bytes(b"")
    assert kb.file_block(str(empty), 1, 2) == (0, 0, 0)
    for bad in ((str(tmp_path / "missing.fastq"), 0, 1), (str(tmp_path / "x.txt"), 0, 1), (str(path), 2, 2), (str(path), -1, 2), (str(path), 0, 0)):
        with pytest.raises(kb.KmiError):
            kb.file_block(*bad)


# ---- pinned to the REFERENCE's own find_first_record, executed (tests/golden/make_golden_first_record.py) -------------------
def _golden_first_record():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_first_record.npz"))


def _golden_text(name):
    if name.startswith("synthetic."):
        import importlib.util
        import os
        spec = importlib.util.spec_from_file_location("mgfr", os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "make_golden_first_record.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return np.frombuffer(mod.synthetic()[name], dtype=np.uint8)
    return ou.read_fixture(name)


@pytest.mark.parametrize("name", sorted(_golden_first_record().files))
def test_find_first_record_equals_the_reference_function(name, tmp_path):
    """every row is one call of the reference's FASTQParser::find_first_record (fastq_loader.hpp:269-357; compiled from the
    reference, oracle tier T1): search range [b, e) of the file, result or 'throws'. The C ABI function, the C oracle's
    restatement and — for the rows that are p-rank block boundaries — kmi_file_block on the file on disk must all agree."""
    rows = _golden_first_record()[name]
    buf = _golden_text(name)
    n = buf.size
    throws = np.uint64(2**64 - 1)
    for b, e, want in rows.tolist():
        if want == int(throws):
            with pytest.raises(kb.KmiFormatError):
                kb.fastq_find_first_record(buf, b, e, parent=(0, n))
            with pytest.raises(ValueError):
                ou.fastq_find_first_record(buf, b, e, 0, (0, n))
        else:
            assert kb.fastq_find_first_record(buf, b, e, parent=(0, n)) == want, (name, b, e)
            assert ou.fastq_find_first_record(buf, b, e, 0, (0, n)) == want, (name, b, e)
    # partitioned_file::read_file: rank r's block begins at find_first_record over [n r / p, n)
    path = tmp_path / name
    path.write_bytes(buf.tobytes())
    whole = {int(b): int(w) for b, e, w in rows.tolist() if e == n and w != int(throws)}
    for p in (2, 3, 8, 64):
        # (rank 0 begins at byte 0 by definition — where the reference's search from 0 lands in any file that starts with a
        #  record; a text without records, like the junk one, is left to the parser to reject instead of being skipped)
        cuts = [0] + [whole[n * r // p] for r in range(1, p)] + [n]
        for r in range(p):
            b, e, size = kb.file_block(str(path), r, p, format=1)
            assert (b, e) == (cuts[r], max(cuts[r], cuts[r + 1])), (name, p, r)
