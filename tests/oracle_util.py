"""ctypes access to the CPU checkers under oracle/ (TEST INFRASTRUCTURE ONLY)."""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "tests", "golden", "data")
_P = ctypes.c_void_p
_SZ = ctypes.c_size_t
_U64 = ctypes.c_uint64


def _ptr(a):
    return a.ctypes.data_as(_P) if a is not None else None


def nwords(k):
    return 1 if k <= 32 else 2


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "libkmi_oracle.so"))
        lib.kmio_fastq_kmers.restype = ctypes.c_long
        lib.kmio_fastq_kmers.argtypes = [_P, _SZ, _U64, ctypes.c_int, ctypes.c_int, ctypes.c_int, _P, _P, _SZ, _P]
        lib.kmio_fasta_kmers.restype = ctypes.c_long
        lib.kmio_fasta_kmers.argtypes = [_P, _SZ, ctypes.c_int, ctypes.c_int, ctypes.c_int, _P, _P, _SZ, _P]
        lib.kmio_revcomp.argtypes = [ctypes.c_int, _P, _P, _SZ]
        lib.kmio_canonical.argtypes = [ctypes.c_int, _P, _P, _SZ]
        lib.kmio_hash.argtypes = [ctypes.c_int, ctypes.c_int, _P, _P, _SZ]
        lib.kmio_owner.argtypes = [_P, _SZ, ctypes.c_uint32, _P]
        lib.kmio_hash_prefix.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint, _P, _P, _SZ]
        lib.kmio_quality_lut.argtypes = [_P]
        lib.kmio_fastq_quality.restype = ctypes.c_long
        lib.kmio_fastq_quality.argtypes = [_P, _SZ, ctypes.c_int, ctypes.c_int, _P, _SZ]
        lib.kmio_fastq_find_first_record.argtypes = [_P, _U64, _SZ, _U64, _U64, _U64, _U64, _P]
        lib.kmio_count_sorted.restype = _SZ
        lib.kmio_count_sorted.argtypes = [ctypes.c_int, _P, _SZ, _P]
        lib.kmio_sort_tuples.argtypes = [ctypes.c_int, _P, _P, _SZ]
        lib.kmio_synth_record_bytes.restype = _SZ
        lib.kmio_synth_record_bytes.argtypes = [ctypes.c_uint32]
        lib.kmio_synth_fastq.argtypes = [_P, _U64, _U64, ctypes.c_uint32, _U64, _U64, ctypes.c_uint32, ctypes.c_uint32]
        lib.kmio_synth_fasta.restype = _SZ
        lib.kmio_synth_fasta.argtypes = [_P, ctypes.c_uint32, _U64, ctypes.c_uint32, _U64]
        lib.kmio_count_build_port.restype = _SZ
        lib.kmio_count_build_port.argtypes = [_P, _SZ, ctypes.c_int, ctypes.c_int, _P, _P, _SZ, _P]
        _oracle = lib
    return _oracle


def ref_lib(flavour="ndebug"):
    """The reference's own headers compiled into oracle/_ref (None when not built)."""
    p = os.path.join(ROOT, "oracle", "_ref", "libkmi_ref_%s.so" % flavour)
    if not os.path.exists(p):
        return None
    lib = ctypes.CDLL(p)
    lib.kmiref_windows.restype = _SZ
    lib.kmiref_windows.argtypes = [ctypes.c_int, _P, _SZ, _P]
    lib.kmiref_revcomp.argtypes = [ctypes.c_int, _P, _P, _SZ]
    lib.kmiref_canonical.argtypes = [ctypes.c_int, _P, _P, _SZ]
    lib.kmiref_hash.argtypes = [ctypes.c_int, ctypes.c_int, _P, _P, _SZ]
    lib.kmiref_sentinels.argtypes = [ctypes.c_int, _P]
    if hasattr(lib, "kmiref_quality"):
        lib.kmiref_quality.restype = _SZ
        lib.kmiref_quality.argtypes = [ctypes.c_int, _P, _SZ, _P]
        lib.kmiref_quality_lut.argtypes = [_P]
    if hasattr(lib, "kmiref_hash_prefix"):
        lib.kmiref_hash_prefix.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint, _P, _P, _SZ]
    lib.kmiref_count_build.restype = _SZ
    lib.kmiref_count_build.argtypes = [ctypes.c_int, _P, _P, _SZ, ctypes.c_int, ctypes.c_int, _P, _P, _SZ, _P]
    lib.kmiref_count_build_fastq.restype = _SZ
    lib.kmiref_count_build_fastq.argtypes = [ctypes.c_int, _P, _SZ, ctypes.c_int, ctypes.c_int, _P, _P, _P, _SZ, _P, _P]
    lib.kmiref_owner_checksums_fastq.restype = ctypes.c_int
    lib.kmiref_owner_checksums_fastq.argtypes = [ctypes.c_int, _P, _SZ, ctypes.c_int, ctypes.c_int, ctypes.c_int, _P]
    return lib


def ref_count_build_fastq(text, k, p, hash_kind=1, cuts=None, want_dump=False, flavour="ndebug"):
    """The reference's p-rank CountIndex build over FASTQ text (oracle/_ref: reference Kmer / lex_less / hashes /
    dense_hash_map, record walk restated). Returns dict(distinct, sizes[p], ck[p][2]) (+ keys, counts unsorted)."""
    lib = ref_lib(flavour)
    if lib is None:
        return None
    nw = nwords(k)
    sizes = np.zeros(p, dtype=np.uint64)
    ck = np.zeros(2 * p, dtype=np.uint64)
    c = np.ascontiguousarray(cuts, dtype=np.uint64) if cuts is not None else None
    keys = counts = None
    cap = 0
    if want_dump:
        cap = int(text.size) // 2 + 1
        keys = np.zeros(cap * nw, dtype=np.uint64)
        counts = np.zeros(cap, dtype=np.uint32)
    d = lib.kmiref_count_build_fastq(k, _ptr(text), int(text.size), p, hash_kind, _ptr(c), _ptr(keys), _ptr(counts), cap,
                                     _ptr(sizes), _ptr(ck))
    if d == (1 << 64) - 1:
        raise ValueError("reference walk: malformed FASTQ")
    out = {"distinct": int(d), "sizes": sizes, "ck": ck.reshape(p, 2)}
    if want_dump:
        out["keys"] = shape_keys(keys[:int(d) * nw], k)
        out["counts"] = counts[:int(d)]
    return out


def ref_owner_checksums_fastq(text, k, p_owners, hash_kind=1, threads=1, flavour="ndebug"):
    """per owner rank [sum of mix(key) over k-mer occurrences, number of occurrences] of this text (additive over texts)"""
    lib = ref_lib(flavour)
    if lib is None:
        return None
    out = np.zeros(2 * p_owners, dtype=np.uint64)
    rc = lib.kmiref_owner_checksums_fastq(k, _ptr(text), int(text.size), threads, p_owners, hash_kind, _ptr(out))
    if rc != 0:
        raise ValueError("reference walk: malformed FASTQ (%d)" % rc)
    return out.reshape(p_owners, 2)


def read_fixture(name):
    return np.fromfile(os.path.join(DATA, name), dtype=np.uint8)


def shape_keys(keys, k):
    return keys.reshape(-1, 2) if nwords(k) == 2 else keys.reshape(-1)


def fastq_kmers(buf, k, canonical=True, nsplit=False, want_ids=False, file_offset=0):
    """(keys, ids, n_reads) of a whole FASTQ buffer, reference order. Raises ValueError on malformed input."""
    n = int(buf.size)
    nw = nwords(k)
    keys = np.zeros(max(n, 1) * nw, dtype=np.uint64)
    ids = np.zeros(max(n, 1), dtype=np.uint64) if want_ids else None
    reads = _U64(0)
    m = oracle().kmio_fastq_kmers(_ptr(buf), n, file_offset, k, int(canonical), int(nsplit), _ptr(keys), _ptr(ids), max(n, 1),
                                  ctypes.byref(reads))
    if m < 0:
        raise ValueError("oracle: malformed FASTQ (%d)" % m)
    return shape_keys(keys[:m * nw].copy(), k), (ids[:m].copy() if want_ids else None), reads.value


def fastq_quality(buf, k, nsplit=False):
    """one float per k-mer, in the order fastq_kmers emits its tuples (PositionQualityIndex scores)"""
    n = int(buf.size)
    out = np.zeros(max(n, 1), dtype=np.float32)
    m = oracle().kmio_fastq_quality(_ptr(buf), n, k, int(nsplit), _ptr(out), max(n, 1))
    if m < 0:
        raise ValueError("oracle: malformed FASTQ (%d)" % m)
    return out[:m].copy()


def quality_lut():
    out = np.zeros(96, dtype=np.float32)
    oracle().kmio_quality_lut(_ptr(out))
    return out


def fastq_find_first_record(buf, s, e=None, mem=0, parent=None):
    n = int(buf.size)
    e = mem + n if e is None else e
    pb, pe = parent if parent is not None else (mem, mem + n)
    out = _U64(0)
    rc = oracle().kmio_fastq_find_first_record(_ptr(buf), mem, n, s, e, pb, pe, ctypes.byref(out))
    if rc != 0:
        raise ValueError("oracle: no FASTQ markers in range")
    return out.value


def fasta_kmers(buf, k, canonical=True, nsplit=False):
    n = int(buf.size)
    nw = nwords(k)
    keys = np.zeros(max(n, 1) * nw, dtype=np.uint64)
    ids = np.zeros(max(n, 1), dtype=np.uint64)
    seqs = _U64(0)
    m = oracle().kmio_fasta_kmers(_ptr(buf), n, k, int(canonical), int(nsplit), _ptr(keys), _ptr(ids), max(n, 1),
                                  ctypes.byref(seqs))
    if m < 0:
        raise ValueError("oracle: FASTA failure (%d)" % m)
    return shape_keys(keys[:m * nw].copy(), k), ids[:m].copy(), seqs.value


def fasta_block_kmers(buf, k, valid_begin, valid_end, canonical=True, nsplit=False):
    """What rank r of a partitioned FASTA file emits: the file's (k-mer, id) tuples whose first base lies in the rank's valid
    byte range — KmerParser::get_valid_iterator_range clips a sequence to [valid.start, valid.end) and lets the last windows
    run K-1 non-EOL characters past the end (kmer_parser.hpp:114-157, kmer_file_helper.hpp:153-173); sequence ids are
    global (FASTAParser::init_parser, fasta_loader.hpp:202-483)."""
    keys, ids, _ = fasta_kmers(buf, k, canonical=canonical, nsplit=nsplit)
    pos = ids & np.uint64((1 << 40) - 1)
    sel = (pos >= np.uint64(valid_begin)) & (pos < np.uint64(valid_end))
    return keys[sel], ids[sel]


def fasta_prefix_state(buf, off):
    """(header runs that start before byte off, kind of the last line start before off, off starts a line): the three
    facts a block beginning at off inherits. Plain restatement for the tests of kmerind_b200.fasta_chain."""
    b = bytes(np.asarray(buf[:off]).tobytes())
    runs, kind, have, pos = 0, 0, False, 0
    while pos < len(b):
        is_h = b[pos] in (0x3E, 0x3B)
        if is_h and not (have and kind):
            runs += 1
        kind, have = (1 if is_h else 0), True
        nl = b.find(b"\n", pos)
        if nl < 0:
            break
        pos = nl + 1
    bol = off == 0 or buf[off - 1] == 0x0A
    return runs, (kind if have else 0), bool(bol)


def revcomp(keys, k):
    a = np.ascontiguousarray(keys, dtype=np.uint64)
    out = np.zeros_like(a)
    oracle().kmio_revcomp(k, _ptr(a), _ptr(out), a.size // nwords(k))
    return out


def canonical(keys, k):
    a = np.ascontiguousarray(keys, dtype=np.uint64)
    out = np.zeros_like(a)
    oracle().kmio_canonical(k, _ptr(a), _ptr(out), a.size // nwords(k))
    return out


def hashes(keys, k, which):
    """which: 0 murmur store, 1 murmur dist, 2 farm store, 3 farm dist, 4/5 farm with DebugTweak."""
    a = np.ascontiguousarray(keys, dtype=np.uint64)
    n = a.size // nwords(k)
    out = np.zeros(n, dtype=np.uint64)
    oracle().kmio_hash(k, which, _ptr(a), _ptr(out), n)
    return out


def ceil_log2(p):
    b = 0
    while (1 << b) < p:
        b += 1
    return b


def hashes_prefix(keys, k, hash_kind, dist, nranks=1):
    """identity (hash_kind 3) / cpp_std (4) flavours; the dist hash is built with ceilLog2(nranks) prefix bits, as KeyToRank does."""
    a = np.ascontiguousarray(keys, dtype=np.uint64)
    n = a.size // nwords(k)
    out = np.zeros(n, dtype=np.uint64)
    oracle().kmio_hash_prefix(k, (hash_kind - 3) * 2 + (1 if dist else 0), ceil_log2(nranks), _ptr(a), _ptr(out), n)
    return out


def hash_which(hash_kind, dist):
    """map (kmi_hash_kind, dist?) to the oracle's `which`."""
    return {0: 0, 1: 2, 2: 4}[hash_kind] + (1 if dist else 0)


def count_sorted(keys, k):
    """sorted dump of a counting map: (distinct keys ascending, uint32 counts)."""
    a = np.ascontiguousarray(keys, dtype=np.uint64).reshape(-1).copy()
    nw = nwords(k)
    n = a.size // nw
    counts = np.zeros(max(n, 1), dtype=np.uint32)
    m = oracle().kmio_count_sorted(k, _ptr(a), n, _ptr(counts))
    return shape_keys(a[:m * nw].copy(), k), counts[:m].astype(np.uint64)


def sort_tuples(keys, ids, k):
    a = np.ascontiguousarray(keys, dtype=np.uint64).reshape(-1).copy()
    b = np.ascontiguousarray(ids, dtype=np.uint64).copy()
    oracle().kmio_sort_tuples(k, _ptr(a), _ptr(b), b.size)
    return shape_keys(a, k), b


def synth_fastq(first_read, n_reads, read_len=150, genome_len=100000, seed=1, err_ppm=5000, n_ppm=0):
    rec = oracle().kmio_synth_record_bytes(read_len)
    out = np.zeros(n_reads * rec, dtype=np.uint8)
    oracle().kmio_synth_fastq(_ptr(out), first_read, n_reads, read_len, genome_len, seed, err_ppm, n_ppm)
    return out


def synth_fasta(n_seq, seq_len, width=80, seed=1):
    n = oracle().kmio_synth_fasta(None, n_seq, seq_len, width, seed)
    out = np.zeros(n, dtype=np.uint8)
    oracle().kmio_synth_fasta(_ptr(out), n_seq, seq_len, width, seed)
    return out


def random_keys(k, n, seed=0):
    rng = np.random.default_rng(seed)
    nw = nwords(k)
    a = rng.integers(0, 2**64, size=(n, nw), dtype=np.uint64)
    if nw == 1:
        a[:, 0] &= np.uint64((1 << (2 * k)) - 1)
        return a.reshape(-1)
    a[:, 1] &= np.uint64((1 << (2 * (k - 32))) - 1)
    return a
