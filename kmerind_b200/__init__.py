"""kmerind_b200 — B200-native k-mer index construction path behind ParBLiSS/kmerind's Index API.

The product is the C-ABI CUDA library ``libkmi.so`` (include/kmi.h; C++ adapter in
include/kmi_index.hpp). This module is the thin ctypes binding the tests and bench.py use: it
mirrors ``bliss::index::kmer::Index`` member for member (reference src/index/kmer_index.hpp:98-394)
and never computes anything itself. There is no CPU fallback: if the library is missing or no CUDA
device is present, calls raise.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkmi.so")

MODE_COUNT, MODE_POSITION = 0, 1
HASH_MURMUR, HASH_FARM, HASH_FARM_DEBUG, HASH_IDENTITY, HASH_STD = 0, 1, 2, 3, 4
SUPPORTED_K = (15, 21, 31, 35, 47, 63)

KMI_OK = 0
ERR_ARG, ERR_CUDA, ERR_FORMAT, ERR_CAPACITY, ERR_UNSUPPORTED, ERR_NOMEM = -1, -2, -3, -4, -5, -6


class KmiError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("kmi error %d: %s" % (code, msg))
        self.code = code


class KmiFormatError(KmiError, ValueError):
    """std::logic_error in the reference (malformed FASTQ, fastq_loader.hpp:393,419,435)."""


class KmiConfig(ctypes.Structure):
    _fields_ = [("k", ctypes.c_int32), ("mode", ctypes.c_int32), ("canonical", ctypes.c_int32),
                ("hash", ctypes.c_int32), ("nsplit", ctypes.c_int32), ("device", ctypes.c_int32),
                ("rank", ctypes.c_int32), ("nranks", ctypes.c_int32), ("reserve", ctypes.c_uint64)]


_lib = None
_P = ctypes.c_void_p
_SZ = ctypes.c_size_t
_U64 = ctypes.c_uint64


def load_library(build_if_missing=False):
    """dlopen libkmi.so (in-tree). Raises if it has not been built — no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("KMI_LIB", LIB_PATH)   # tuning experiments: another build of the same library
    if not os.path.exists(path):
        if build_if_missing:
            from . import build as _b
            _b.build()
        else:
            raise ImportError("libkmi.so is not built (run python -m kmerind_b200.build); there is no CPU path")
    lib = ctypes.CDLL(path)
    lib.kmi_last_error.restype = ctypes.c_char_p
    lib.kmi_synth_record_bytes.restype = _SZ
    lib.kmi_synth_record_bytes.argtypes = [ctypes.c_uint32]
    lib.kmi_create.argtypes = [ctypes.POINTER(KmiConfig), ctypes.POINTER(_P)]
    lib.kmi_destroy.argtypes = [_P]
    lib.kmi_destroy.restype = None
    sig = {
        "kmi_device_count": [_P], "kmi_set_stream": [_P, _P], "kmi_set_param": [_P, ctypes.c_int, ctypes.c_int64], "kmi_flush": [_P], "kmi_bulk_stats": [_P, _P, _P], "kmi_reserve": [_P, _U64], "kmi_clear": [_P],
        "kmi_build_fastq": [_P, _P, _SZ, _U64, _P], "kmi_build_fasta": [_P, _P, _SZ, _P],
        "kmi_build_fastq_dev": [_P, _P, _SZ, _U64, _P], "kmi_build_fasta_dev": [_P, _P, _SZ, _P],
        "kmi_parse_fastq_dev": [_P, _P, _SZ, _U64, _P, _P, _SZ, _P],
        "kmi_parse_fasta_dev": [_P, _P, _SZ, _P, _P, _SZ, _P],
        "kmi_parse_fasta_block_dev": [_P, _P, _SZ, _P, _P, _P, _SZ, _P],
        "kmi_build_fasta_block": [_P, _P, _SZ, _P, _P], "kmi_build_fasta_block_dev": [_P, _P, _SZ, _P, _P],
        "kmi_fasta_block_summary_dev": [_P, _P, _SZ, ctypes.c_int, _P],
        "kmi_fasta_halo_end": [_P, _U64, _SZ, _U64, ctypes.c_int, _P],
        "kmi_build_file": [_P, ctypes.c_char_p, ctypes.c_int, _P],
        "kmi_file_block": [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, _P, _P, _P],
        "kmi_dist_build_file": [_P, ctypes.c_char_p, ctypes.c_int, _P],
        "kmi_save": [_P, ctypes.c_char_p], "kmi_load": [_P, ctypes.c_char_p, _P],
        "kmi_comm_unique_id": [_P], "kmi_comm_create": [_P, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(_P)],
        "kmi_comm_rank": [_P], "kmi_comm_size": [_P], "kmi_comm_barrier": [_P],
        "kmi_comm_allreduce_u64": [_P, _P, ctypes.c_int, ctypes.c_int], "kmi_comm_allgather_u64": [_P, _P, ctypes.c_int, _P],
        "kmi_dist_attach": [_P, _P],
        "kmi_dist_build_fastq": [_P, _P, _SZ, _U64, _P], "kmi_dist_build_fastq_dev": [_P, _P, _SZ, _U64, _P],
        "kmi_dist_build_fasta": [_P, _P, _SZ, _P, _P], "kmi_dist_build_fasta_dev": [_P, _P, _SZ, _P, _P],
        "kmi_dist_insert": [_P, _P, _P, _SZ], "kmi_dist_insert_dev": [_P, _P, _P, _SZ],
        "kmi_dist_count": [_P, _P, _SZ, _P], "kmi_dist_find": [_P, _P, _SZ, _P, _P],
        "kmi_dist_count_dev": [_P, _P, _SZ, _P], "kmi_dist_find_dev": [_P, _P, _SZ, _P, _P],
        "kmi_dist_find_positions": [_P, _P, _SZ, _P, _P, _SZ], "kmi_dist_erase": [_P, _P, _SZ],
        "kmi_dist_size": [_P, _P], "kmi_dist_last_stats": [_P, _P, _P, _P],
        "kmi_fastq_find_first_record": [_P, _U64, _SZ, _U64, _U64, _U64, _U64, _P],
        "kmi_canonicalize_dev": [_P, _P, _P, _SZ], "kmi_hash_dev": [_P, ctypes.c_int, _P, _P, _SZ],
        "kmi_partition_dev": [_P, _P, _P, _SZ, _P, _P, _P], "kmi_partition_noalloc_dev": [_P, _P, _P, _SZ, _P, _P, _P],
        "kmi_insert": [_P, _P, _P, _SZ], "kmi_insert_dev": [_P, _P, _P, _SZ],
        "kmi_count": [_P, _P, _SZ, _P], "kmi_find": [_P, _P, _SZ, _P, _P],
        "kmi_find_positions": [_P, _P, _SZ, _P, _P, _SZ], "kmi_erase": [_P, _P, _SZ],
        "kmi_count_dev": [_P, _P, _SZ, _P], "kmi_find_dev": [_P, _P, _SZ, _P, _P],
        "kmi_local_size": [_P, _P], "kmi_num_keys": [_P, _P], "kmi_dump": [_P, _P, _P, _SZ, _P],
        "kmi_checksum": [_P, _P, _P], "kmi_count_if": [_P, ctypes.c_int, _U64, _P], "kmi_dump_if": [_P, ctypes.c_int, _U64, _P, _P, _SZ, _P],
        "kmi_erase_if": [_P, ctypes.c_int, _U64, _P], "kmi_table_info": [_P, _P, _P, _P], "kmi_launch_count": [_P, _P],
        "kmi_last_stage_ms": [_P, _P], "kmi_last_kernel_ms": [_P, _P],
        "kmi_dump_sorted": [_P, _P, _P, _SZ, _P],
        "kmi_quality_lut": [_P], "kmi_fastq_quality_dev": [_P, _P, _SZ, _U64, _P, _SZ, _P],
        "kmi_posqual_create": [ctypes.POINTER(KmiConfig), ctypes.POINTER(_P)],
        "kmi_posqual_build_fastq": [_P, _P, _SZ, _U64, _P], "kmi_posqual_build_fastq_dev": [_P, _P, _SZ, _U64, _P],
        "kmi_posqual_local_size": [_P, _P], "kmi_posqual_count": [_P, _P, _SZ, _P],
        "kmi_posqual_find": [_P, _P, _SZ, _P, _P, _P, _SZ], "kmi_posqual_erase": [_P, _P, _SZ],
        "kmi_posqual_dump": [_P, _P, _P, _P, _SZ, _P], "kmi_posqual_launch_count": [_P, _P],
        "kmi_sorted_create": [ctypes.POINTER(KmiConfig), ctypes.POINTER(_P)],
        "kmi_sorted_build_fastq": [_P, _P, _SZ, _U64, _P], "kmi_sorted_build_fasta": [_P, _P, _SZ, _P],
        "kmi_sorted_build_fastq_dev": [_P, _P, _SZ, _U64, _P], "kmi_sorted_build_fasta_dev": [_P, _P, _SZ, _P],
        "kmi_sorted_insert": [_P, _P, _P, _SZ], "kmi_sorted_insert_dev": [_P, _P, _P, _SZ], "kmi_sorted_rehash": [_P],
        "kmi_sorted_local_size": [_P, _P], "kmi_sorted_count": [_P, _P, _SZ, _P], "kmi_sorted_find": [_P, _P, _SZ, _P, _P],
        "kmi_sorted_find_positions": [_P, _P, _SZ, _P, _P, _SZ], "kmi_sorted_dump": [_P, _P, _P, _SZ, _P],
        "kmi_sorted_stats": [_P, _P, _P], "kmi_sort_pairs_dev": [_P, _P, _P, _P, _P, _SZ, ctypes.c_int, _P],
        "kmi_synth_fasta_dev": [ctypes.c_int, _P, _U64, _U64, ctypes.c_uint32, _U64, ctypes.c_uint32, _U64],
        "kmi_synth_queries_dev": [ctypes.c_int, _P, _U64, _U64, ctypes.c_int, ctypes.c_uint32, _U64, _U64, _U64],
        "kmi_synth_fastq_dev": [ctypes.c_int, _P, _U64, _U64, ctypes.c_uint32, _U64, _U64, ctypes.c_uint32,
                                ctypes.c_uint32],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = ctypes.c_int
    lib.kmi_partition_noalloc_max_keys.restype = _SZ
    lib.kmi_partition_noalloc_max_keys.argtypes = [ctypes.c_int]
    lib.kmi_synth_fasta_bytes.restype = _SZ
    lib.kmi_synth_fasta_bytes.argtypes = [ctypes.c_uint32, _U64, ctypes.c_uint32]
    lib.kmi_comm_destroy.argtypes = [_P]
    lib.kmi_comm_destroy.restype = None
    lib.kmi_sorted_destroy.argtypes = [_P]
    lib.kmi_sorted_destroy.restype = None
    lib.kmi_posqual_destroy.argtypes = [_P]
    lib.kmi_posqual_destroy.restype = None
    _lib = lib
    return lib


EXPORTED_SYMBOLS = (
    "kmi_create", "kmi_destroy", "kmi_last_error", "kmi_device_count", "kmi_set_stream", "kmi_set_param", "kmi_flush", "kmi_bulk_stats", "kmi_reserve", "kmi_clear",
    "kmi_build_fastq", "kmi_build_fasta", "kmi_build_fastq_dev", "kmi_build_fasta_dev",
    "kmi_parse_fastq_dev", "kmi_parse_fasta_dev", "kmi_parse_fasta_block_dev", "kmi_build_fasta_block", "kmi_build_fasta_block_dev",
    "kmi_fasta_block_summary_dev", "kmi_fasta_halo_end", "kmi_build_file", "kmi_file_block", "kmi_dist_build_file", "kmi_save", "kmi_load", "kmi_fastq_find_first_record",
    "kmi_comm_unique_id", "kmi_comm_create", "kmi_comm_destroy", "kmi_comm_rank", "kmi_comm_size", "kmi_comm_barrier",
    "kmi_comm_allreduce_u64", "kmi_comm_allgather_u64", "kmi_dist_attach", "kmi_dist_build_fastq", "kmi_dist_build_fastq_dev",
    "kmi_dist_build_fasta", "kmi_dist_build_fasta_dev", "kmi_dist_insert", "kmi_dist_insert_dev", "kmi_dist_count", "kmi_dist_find",
    "kmi_dist_find_positions", "kmi_dist_erase", "kmi_dist_count_dev", "kmi_dist_find_dev", "kmi_dist_size", "kmi_dist_last_stats", "kmi_canonicalize_dev", "kmi_hash_dev",
    "kmi_partition_dev", "kmi_partition_noalloc_dev", "kmi_partition_noalloc_max_keys", "kmi_insert", "kmi_insert_dev", "kmi_count", "kmi_find", "kmi_find_positions",
    "kmi_erase", "kmi_count_dev", "kmi_find_dev", "kmi_local_size", "kmi_num_keys", "kmi_dump",
    "kmi_checksum", "kmi_count_if", "kmi_dump_if", "kmi_erase_if", "kmi_table_info", "kmi_launch_count", "kmi_last_stage_ms", "kmi_last_kernel_ms",
    "kmi_synth_record_bytes", "kmi_synth_fastq_dev", "kmi_synth_fasta_bytes", "kmi_synth_fasta_dev", "kmi_synth_queries_dev",
    "kmi_dump_sorted", "kmi_sorted_create", "kmi_sorted_destroy", "kmi_sorted_build_fastq", "kmi_sorted_build_fasta",
    "kmi_sorted_build_fastq_dev", "kmi_sorted_build_fasta_dev", "kmi_sorted_insert", "kmi_sorted_insert_dev", "kmi_sorted_rehash",
    "kmi_sorted_local_size", "kmi_sorted_count", "kmi_sorted_find", "kmi_sorted_find_positions", "kmi_sorted_dump", "kmi_sorted_stats",
    "kmi_sort_pairs_dev", "kmi_quality_lut", "kmi_fastq_quality_dev", "kmi_posqual_create", "kmi_posqual_destroy",
    "kmi_posqual_build_fastq", "kmi_posqual_build_fastq_dev", "kmi_posqual_local_size", "kmi_posqual_count", "kmi_posqual_find",
    "kmi_posqual_erase", "kmi_posqual_dump", "kmi_posqual_launch_count",
)


def _check(code):
    if code != KMI_OK:
        msg = load_library().kmi_last_error().decode("utf-8", "replace")
        if code == ERR_FORMAT:
            raise KmiFormatError(code, msg)
        raise KmiError(code, msg)


def _hptr(a):
    return a.ctypes.data_as(_P) if a is not None else None


def nwords(k):
    return 1 if k <= 32 else 2


def _as_keys(keys, nw):
    a = np.ascontiguousarray(keys, dtype=np.uint64)
    if nw == 2:
        a = a.reshape(-1, 2)
    else:
        a = a.reshape(-1)
    return a


def fastq_find_first_record(buf, search_begin, search_end=None, mem_offset=0, parent=None):
    """FASTQParser::find_first_record on host bytes (kmi_fastq_find_first_record): file offset of the first whole record
    at or after search_begin; buf holds file bytes [mem_offset, mem_offset + len(buf))."""
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    end = mem_offset + buf.size
    search_end = end if search_end is None else search_end
    pb, pe = parent if parent is not None else (mem_offset, end)
    out = _U64(0)
    _check(load_library().kmi_fastq_find_first_record(_hptr(buf), mem_offset, buf.size, search_begin, search_end, pb, pe, ctypes.byref(out)))
    return out.value


class KmiFastaBlock(ctypes.Structure):
    """mirror of kmi_fasta_block (include/kmi.h)"""
    _fields_ = [("file_offset", ctypes.c_uint64), ("n_valid", ctypes.c_uint64), ("header_runs_before", ctypes.c_uint32),
                ("flags", ctypes.c_uint32)]


FASTA_BOL, FASTA_PREV_HEADER, FASTA_FILE_STARTS_WITH_HEADER = 1, 2, 4


def fasta_halo_end(buf, valid_end, k, mem_offset=0):
    """kmi_fasta_halo_end: file offset up to which a block ending at valid_end needs bytes (K-1 non-EOL bytes)."""
    a = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf
    out = _U64(0)
    _check(load_library().kmi_fasta_halo_end(_hptr(a), mem_offset, a.size, valid_end, k, ctypes.byref(out)))
    return out.value


def fasta_chain(summaries, bol, file_starts_with_header):
    """Combine the per-block summaries of kmi_fasta_block_summary_dev (blocks in file order) into what each block needs
    from the bytes before it: an exclusive scan carrying (header runs so far, kind of the last line start) — the
    cross-rank part of FASTAParser::init_parser (fasta_loader.hpp:202-483). summaries[r] = (has_line_start,
    first_is_header, last_is_header, runs); bol[r] = block r's first byte starts a line. Returns [(header_runs_before,
    flags)] for kmi_fasta_block."""
    out, runs, kind = [], 0, 0
    for (has, first_h, last_h, r), b in zip(summaries, bol):
        flags = (FASTA_BOL if b else 0) | (FASTA_PREV_HEADER if kind else 0) | (FASTA_FILE_STARTS_WITH_HEADER if file_starts_with_header else 0)
        out.append((runs, flags))
        if has:
            runs += int(r) - (1 if (first_h and kind) else 0)
            kind = int(last_h)
    return out


def file_block(path, rank, nranks, format=0):
    """(begin, end, file size) of rank's block of ONE file shared by nranks ranks (kmi_file_block): equal byte ranges, FASTQ
    boundaries moved forward to whole records — partitioned_file::read_file + BlockPartitioner (file.hpp:1239-1251,1342-1431)"""
    b, e, sz = _U64(0), _U64(0), _U64(0)
    _check(load_library().kmi_file_block(os.fsencode(path), format, rank, nranks, ctypes.byref(b), ctypes.byref(e), ctypes.byref(sz)))
    return b.value, e.value, sz.value


def fastq_cuts(buf, parts):
    """Record-aligned block boundaries [0, c1, ..., len(buf)] for `parts` ranks, as partitioned_file::read_file cuts a file
    (equal byte ranges, each moved forward to its first whole record: file.hpp:1342-1431)."""
    n = int(np.asarray(buf).size)
    cuts = [0]
    for r in range(1, parts):
        cuts.append(max(cuts[-1], fastq_find_first_record(buf, n * r // parts)))
    cuts.append(n)
    return cuts


class Index:
    """CountIndex / PositionIndex over one GPU (one shard).

    Mirrors ``Index<MapType, KmerParser>``: insert, find, count, erase, build_*, size, local_size
    (reference src/index/kmer_index.hpp:115-392). Keys are numpy uint64 arrays of shape (n,) for
    K <= 31 or (n, 2) [low word, high word] for 33 <= K <= 63 — the memory layout of
    ``std::vector<Kmer<K, DNA, uint64_t>>``.
    """

    def __init__(self, k=31, mode=MODE_COUNT, canonical=True, hash=HASH_MURMUR, nsplit=False, device=0,
                 rank=0, nranks=1, reserve=0):
        self._lib = load_library()
        self.k, self.mode, self.nw = k, mode, nwords(k)
        cfg = KmiConfig(k, mode, 1 if canonical else 0, hash, 1 if nsplit else 0, device, rank, nranks, reserve)
        self._h = _P()
        _check(self._lib.kmi_create(ctypes.byref(cfg), ctypes.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.kmi_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- build ------------------------------------------------------------------------------
    def build_fastq(self, data, file_offset=0):
        """build_posix<FASTQParser>: host bytes -> index. Returns k-mer occurrences generated."""
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        n = _U64(0)
        _check(self._lib.kmi_build_fastq(self._h, _hptr(buf), buf.size, file_offset, ctypes.byref(n)))
        return n.value

    def build_file(self, path, format=0):
        """Index::build_posix: FASTQ / FASTA file by extension (format 1 / 2 to force), streamed from disk."""
        n = _U64(0)
        _check(self._lib.kmi_build_file(self._h, os.fsencode(path), format, ctypes.byref(n)))
        return n.value

    def save(self, path):
        """write this shard's entries, sorted by (key, value), to `path` (format: include/kmi.h, kmi_save)"""
        _check(self._lib.kmi_save(self._h, os.fsencode(path)))

    def load(self, path):
        """add the entries of a saved shard to this index; returns how many entries the file held"""
        n = _U64(0)
        _check(self._lib.kmi_load(self._h, os.fsencode(path), ctypes.byref(n)))
        return n.value

    def build_fasta(self, data, block=None):
        buf = np.frombuffer(data, dtype=np.uint8) if not isinstance(data, np.ndarray) else data
        n = _U64(0)
        blk = ctypes.byref(KmiFastaBlock(*block)) if block is not None else None
        _check(self._lib.kmi_build_fasta_block(self._h, _hptr(buf), buf.size, blk, ctypes.byref(n)))
        return n.value

    def build_fastq_dev(self, dptr, nbytes, file_offset=0):
        n = _U64(0)
        _check(self._lib.kmi_build_fastq_dev(self._h, dptr, nbytes, file_offset, ctypes.byref(n)))
        return n.value

    def build_fasta_dev(self, dptr, nbytes, block=None):
        n = _U64(0)
        blk = ctypes.byref(KmiFastaBlock(*block)) if block is not None else None
        _check(self._lib.kmi_build_fasta_block_dev(self._h, dptr, nbytes, blk, ctypes.byref(n)))
        return n.value

    # -- stages on device pointers ------------------------------------------------------------
    def parse_fastq_dev(self, dptr, nbytes, d_keys, cap, d_ids=None, file_offset=0):
        n = _U64(0)
        _check(self._lib.kmi_parse_fastq_dev(self._h, dptr, nbytes, file_offset, d_keys, d_ids, cap, ctypes.byref(n)))
        return n.value

    def parse_fasta_dev(self, dptr, nbytes, d_keys, cap, d_ids=None, block=None):
        """block = None (whole file) or (file_offset, n_valid, header_runs_before, flags): kmi_fasta_block"""
        n = _U64(0)
        blk = ctypes.byref(KmiFastaBlock(*block)) if block is not None else None
        _check(self._lib.kmi_parse_fasta_block_dev(self._h, dptr, nbytes, blk, d_keys, d_ids, cap, ctypes.byref(n)))
        return n.value

    def fasta_block_summary_dev(self, dptr, n_valid, bol):
        out = (ctypes.c_uint32 * 4)()
        _check(self._lib.kmi_fasta_block_summary_dev(self._h, dptr, n_valid, 1 if bol else 0, out))
        return tuple(int(x) for x in out)

    def canonicalize_dev(self, d_in, d_out, n):
        _check(self._lib.kmi_canonicalize_dev(self._h, d_in, d_out, n))

    def fastq_quality_dev(self, d_bytes, nbytes, file_offset, d_ids, n_ids, d_quals):
        _check(self._lib.kmi_fastq_quality_dev(self._h, d_bytes, nbytes, file_offset, d_ids, n_ids, d_quals))

    def hash_dev(self, which, d_keys, d_out, n):
        _check(self._lib.kmi_hash_dev(self._h, which, d_keys, d_out, n))

    def partition_dev(self, d_keys, d_vals, n, d_keys_out, d_vals_out, nranks):
        counts = np.zeros(nranks, dtype=np.uint64)
        _check(self._lib.kmi_partition_dev(self._h, d_keys, d_vals, n, d_keys_out, d_vals_out, _hptr(counts)))
        return counts

    def partition_noalloc_dev(self, d_keys, d_vals, n, d_keys_out, d_vals_out, d_counts):
        """kmi_partition_noalloc_dev: scratch in the library's static device array, counts to device memory, no sync"""
        _check(self._lib.kmi_partition_noalloc_dev(self._h, d_keys, d_vals, n, d_keys_out, d_vals_out, d_counts))

    def insert_dev(self, d_keys, n, d_vals=None):
        _check(self._lib.kmi_insert_dev(self._h, d_keys, d_vals, n))

    def count_dev(self, d_keys, n, d_counts):
        _check(self._lib.kmi_count_dev(self._h, d_keys, n, d_counts))

    def find_dev(self, d_keys, n, d_vals, d_found):
        _check(self._lib.kmi_find_dev(self._h, d_keys, n, d_vals, d_found))

    # -- Index API on host arrays ---------------------------------------------------------------
    def insert(self, keys, vals=None):
        a = _as_keys(keys, self.nw)
        v = None if vals is None else np.ascontiguousarray(vals, dtype=np.uint64)
        _check(self._lib.kmi_insert(self._h, _hptr(a), _hptr(v), a.shape[0]))

    def count(self, keys):
        a = _as_keys(keys, self.nw)
        out = np.zeros(a.shape[0], dtype=np.uint64)
        _check(self._lib.kmi_count(self._h, _hptr(a), a.shape[0], _hptr(out)))
        return out

    def find(self, keys):
        """COUNT index: (values, found) in input order; the reference returns only the hits."""
        a = _as_keys(keys, self.nw)
        vals = np.zeros(a.shape[0], dtype=np.uint64)
        found = np.zeros(a.shape[0], dtype=np.uint8)
        _check(self._lib.kmi_find(self._h, _hptr(a), a.shape[0], _hptr(vals), _hptr(found)))
        return vals, found

    def find_positions(self, keys, cap=None):
        a = _as_keys(keys, self.nw)
        n = a.shape[0]
        if cap is None:
            cap = int(self.count(a).sum())
        offsets = np.zeros(n + 1, dtype=np.uint64)
        vals = np.zeros(max(cap, 1), dtype=np.uint64)
        _check(self._lib.kmi_find_positions(self._h, _hptr(a), n, _hptr(offsets), _hptr(vals), cap))
        return offsets, vals[:int(offsets[-1])]

    def erase(self, keys):
        a = _as_keys(keys, self.nw)
        _check(self._lib.kmi_erase(self._h, _hptr(a), a.shape[0]))

    def reserve(self, n):
        _check(self._lib.kmi_reserve(self._h, n))

    def clear(self):
        _check(self._lib.kmi_clear(self._h))

    def local_size(self):
        n = _U64(0)
        _check(self._lib.kmi_local_size(self._h, ctypes.byref(n)))
        return n.value

    size = local_size  # single shard; the sharded size() is a sum over ranks (kmerind_b200.dist)

    def num_keys(self):
        n = _U64(0)
        _check(self._lib.kmi_num_keys(self._h, ctypes.byref(n)))
        return n.value

    def to_vector(self):
        """(keys, values), unordered — Index::get_map().to_vector()."""
        n = self.local_size()
        keys = np.zeros((n, 2) if self.nw == 2 else (n,), dtype=np.uint64)
        vals = np.zeros(n, dtype=np.uint64)
        got = _SZ(0)
        _check(self._lib.kmi_dump(self._h, _hptr(keys), _hptr(vals), n, ctypes.byref(got)))
        return keys[:got.value], vals[:got.value]

    def dump_sorted(self):
        """to_vector() ordered by (key, value) ON THE DEVICE (kmi_dump_sorted: the radix sort of the sorted back end)."""
        n = self.local_size()
        keys = np.zeros((n, 2) if self.nw == 2 else (n,), dtype=np.uint64)
        vals = np.zeros(n, dtype=np.uint64)
        got = _SZ(0)
        _check(self._lib.kmi_dump_sorted(self._h, _hptr(keys), _hptr(vals), n, ctypes.byref(got)))
        return keys[:got.value], vals[:got.value]

    def sort_pairs_dev(self, d_keys, d_vals, d_keys_tmp, d_vals_tmp, n, val_bits=0):
        """kmi_sort_pairs_dev; returns True when the result sits in the *_tmp pair"""
        w = ctypes.c_int(0)
        _check(self._lib.kmi_sort_pairs_dev(self._h, d_keys, d_vals, d_keys_tmp, d_vals_tmp, n, val_bits, ctypes.byref(w)))
        return bool(w.value)

    def sorted_dump(self):
        """Sorted by (key as multi-word unsigned integer, value): the parity dump."""
        keys, vals = self.to_vector()
        if self.nw == 2:
            order = np.lexsort((vals, keys[:, 0], keys[:, 1]))
        else:
            order = np.lexsort((vals, keys))
        return keys[order], vals[order]

    def count_if(self, op, value):
        """entries whose value is >= / <= / == value (op "ge" / "le" / "eq"), counted on the device"""
        n = _U64(0)
        _check(self._lib.kmi_count_if(self._h, {"ge": 0, "le": 1, "eq": 2}[op], value, ctypes.byref(n)))
        return n.value

    def find_if(self, op, value):
        n = self.count_if(op, value)
        keys = np.zeros(max(n, 1) * self.nw, dtype=np.uint64)
        vals = np.zeros(max(n, 1), dtype=np.uint64)
        got = _SZ(0)
        _check(self._lib.kmi_dump_if(self._h, {"ge": 0, "le": 1, "eq": 2}[op], value, _hptr(keys), _hptr(vals), n, ctypes.byref(got)))
        k = keys[:got.value * self.nw]
        return (k.reshape(-1, 2) if self.nw == 2 else k), vals[:got.value]

    def erase_if(self, op, value):
        n = _U64(0)
        _check(self._lib.kmi_erase_if(self._h, {"ge": 0, "le": 1, "eq": 2}[op], value, ctypes.byref(n)))
        return n.value

    def checksum(self):
        a, b = _U64(0), _U64(0)
        _check(self._lib.kmi_checksum(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def table_info(self):
        a, b, c = _U64(0), _U64(0), _U64(0)
        _check(self._lib.kmi_table_info(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return {"slots": a.value, "slot_bytes": b.value, "device_bytes": c.value}

    def launch_count(self):
        n = _U64(0)
        _check(self._lib.kmi_launch_count(self._h, ctypes.byref(n)))
        return n.value

    def last_stage_ms(self):
        ms = (ctypes.c_float * 3)()
        _check(self._lib.kmi_last_stage_ms(self._h, ms))
        return [float(x) for x in ms]

    def last_kernel_ms(self):
        ms = (ctypes.c_float * 5)()
        _check(self._lib.kmi_last_kernel_ms(self._h, ms))
        return [float(x) for x in ms]

    def set_build_path(self, path):
        """0 auto, 1 direct atomics, 2 hash-prefix buckets (KMI_PARAM_BUILD_PATH)."""
        _check(self._lib.kmi_set_param(self._h, 1, path))

    def set_defer_insert(self, budget_keys):
        """KMI_PARAM_DEFER_INSERT (bulk load): stage up to budget_keys k-mers in fine buckets, one table sweep per flush; 0 = off"""
        _check(self._lib.kmi_set_param(self._h, 5, int(budget_keys)))

    def flush(self):
        _check(self._lib.kmi_flush(self._h))

    def bulk_stats(self):
        """(keys staged right now, deferred segment inserts run so far)"""
        a, b = _U64(0), _U64(0)
        _check(self._lib.kmi_bulk_stats(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def set_max_load(self, percent):
        """KMI_PARAM_MAX_LOAD: fill limit of the table in per cent (10..70); set before reserve / the first build"""
        _check(self._lib.kmi_set_param(self._h, 6, int(percent)))

    def set_dist_trans(self, t):
        """KMI_PARAM_DIST_TRANS: 0 identity, 1 lex_less, 2 xor_rev_comp under the dist hash (non-canonical index)."""
        _check(self._lib.kmi_set_param(self._h, 4, int(t)))

    def set_saturate(self, on):
        """counts stick at 2^32 - 1 instead of wrapping (saturating_counting_densehash_map)"""
        _check(self._lib.kmi_set_param(self._h, 3, 1 if on else 0))

    def set_segment_insert(self, on):
        """bucketed path: shared-memory segment insert (default) or L2 atomics (KMI_PARAM_SEGMENT_INSERT)."""
        _check(self._lib.kmi_set_param(self._h, 2, 1 if on else 0))

    def set_stream(self, cuda_stream_ptr):
        _check(self._lib.kmi_set_stream(self._h, cuda_stream_ptr))


class SortedMap:
    """The sort-based back end (kmi_sorted_*): dsc::counting_sorted_map (MODE_COUNT) / dsc::sorted_multimap (MODE_POSITION)
    of reference src/containers/distributed_sorted_map.hpp, one shard. Entries live in HBM as sorted arrays; insert / build
    append, rehash() = device radix sort + run-length reduction, queries = binary searches."""

    def __init__(self, k=31, mode=MODE_COUNT, canonical=True, hash=HASH_MURMUR, nsplit=False, device=0):
        self._lib = load_library()
        self.k, self.mode, self.nw = k, mode, nwords(k)
        cfg = KmiConfig(k, mode, 1 if canonical else 0, hash, 1 if nsplit else 0, device, 0, 1, 0)
        self._h = _P()
        _check(self._lib.kmi_sorted_create(ctypes.byref(cfg), ctypes.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.kmi_sorted_destroy(self._h)
            self._h = _P()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def build_fastq(self, buf, file_offset=0):
        a = np.ascontiguousarray(np.frombuffer(buf, dtype=np.uint8))
        n = _U64(0)
        _check(self._lib.kmi_sorted_build_fastq(self._h, _hptr(a), a.size, file_offset, ctypes.byref(n)))
        return n.value

    def build_fasta(self, buf):
        a = np.ascontiguousarray(np.frombuffer(buf, dtype=np.uint8))
        n = _U64(0)
        _check(self._lib.kmi_sorted_build_fasta(self._h, _hptr(a), a.size, ctypes.byref(n)))
        return n.value

    def build_fastq_dev(self, dptr, nbytes, file_offset=0):
        n = _U64(0)
        _check(self._lib.kmi_sorted_build_fastq_dev(self._h, dptr, nbytes, file_offset, ctypes.byref(n)))
        return n.value

    def insert(self, keys, vals=None):
        a = _as_keys(keys, self.nw)
        v = np.ascontiguousarray(vals, dtype=np.uint64) if vals is not None else None
        _check(self._lib.kmi_sorted_insert(self._h, _hptr(a), _hptr(v), a.shape[0]))

    def rehash(self):
        _check(self._lib.kmi_sorted_rehash(self._h))

    def local_size(self):
        n = _U64(0)
        _check(self._lib.kmi_sorted_local_size(self._h, ctypes.byref(n)))
        return n.value

    def count(self, keys):
        a = _as_keys(keys, self.nw)
        out = np.zeros(a.shape[0], dtype=np.uint64)
        _check(self._lib.kmi_sorted_count(self._h, _hptr(a), a.shape[0], _hptr(out)))
        return out

    def find(self, keys):
        a = _as_keys(keys, self.nw)
        vals = np.zeros(a.shape[0], dtype=np.uint64)
        found = np.zeros(a.shape[0], dtype=np.uint8)
        _check(self._lib.kmi_sorted_find(self._h, _hptr(a), a.shape[0], _hptr(vals), _hptr(found)))
        return vals, found.astype(bool)

    def find_positions(self, keys):
        a = _as_keys(keys, self.nw)
        n = a.shape[0]
        cnt = self.count(keys)
        total = int(cnt.sum())
        off = np.zeros(n + 1, dtype=np.uint64)
        vals = np.zeros(max(total, 1), dtype=np.uint64)
        _check(self._lib.kmi_sorted_find_positions(self._h, _hptr(a), n, _hptr(off), _hptr(vals), max(total, 1)))
        return off, vals[:total]

    def dump(self):
        """entries in key order (POSITION: (key, id) order)"""
        n = self.local_size()
        keys = np.zeros((n, 2) if self.nw == 2 else (n,), dtype=np.uint64)
        vals = np.zeros(n, dtype=np.uint64)
        got = _SZ(0)
        _check(self._lib.kmi_sorted_dump(self._h, _hptr(keys), _hptr(vals), n, ctypes.byref(got)))
        return keys[:got.value], vals[:got.value]

    def stats(self):
        ms = (ctypes.c_float * 3)()
        n = _U64(0)
        _check(self._lib.kmi_sorted_stats(self._h, ms, ctypes.byref(n)))
        return [float(x) for x in ms], n.value


def quality_lut():
    out = np.zeros(96, dtype=np.float32)
    _check(load_library().kmi_quality_lut(_hptr(out)))
    return out


class PositionQualityIndex:
    """kmi_posqual_*: k-mer -> (ShortSequenceKmerId, float quality score) from FASTQ (PositionQualityIndex, kmer_index.hpp:406)."""

    def __init__(self, k=21, canonical=True, hash=HASH_MURMUR, nsplit=False, device=0):
        self._lib = load_library()
        self.k, self.nw = k, nwords(k)
        cfg = KmiConfig(k, MODE_POSITION, 1 if canonical else 0, hash, 1 if nsplit else 0, device, 0, 1, 0)
        self._h = _P()
        _check(self._lib.kmi_posqual_create(ctypes.byref(cfg), ctypes.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.kmi_posqual_destroy(self._h)
            self._h = _P()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def build_fastq(self, buf, file_offset=0):
        a = np.ascontiguousarray(np.frombuffer(buf, dtype=np.uint8))
        n = _U64(0)
        _check(self._lib.kmi_posqual_build_fastq(self._h, _hptr(a), a.size, file_offset, ctypes.byref(n)))
        return n.value

    def local_size(self):
        n = _U64(0)
        _check(self._lib.kmi_posqual_local_size(self._h, ctypes.byref(n)))
        return n.value

    def count(self, keys):
        a = _as_keys(keys, self.nw)
        out = np.zeros(a.shape[0], dtype=np.uint64)
        _check(self._lib.kmi_posqual_count(self._h, _hptr(a), a.shape[0], _hptr(out)))
        return out

    def erase(self, keys):
        a = _as_keys(keys, self.nw)
        _check(self._lib.kmi_posqual_erase(self._h, _hptr(a), a.shape[0]))

    def find(self, keys):
        """(offsets[n+1], ids, quals): CSR answer"""
        a = _as_keys(keys, self.nw)
        n = a.shape[0]
        total = int(self.count(keys).sum())
        off = np.zeros(n + 1, dtype=np.uint64)
        ids = np.zeros(max(total, 1), dtype=np.uint64)
        quals = np.zeros(max(total, 1), dtype=np.float32)
        _check(self._lib.kmi_posqual_find(self._h, _hptr(a), n, _hptr(off), _hptr(ids), _hptr(quals), max(total, 1)))
        return off, ids[:total], quals[:total]

    def to_vector(self):
        n = self.local_size()
        keys = np.zeros((n, 2) if self.nw == 2 else (n,), dtype=np.uint64)
        ids = np.zeros(max(n, 1), dtype=np.uint64)
        quals = np.zeros(max(n, 1), dtype=np.float32)
        got = _SZ(0)
        _check(self._lib.kmi_posqual_dump(self._h, _hptr(keys), _hptr(ids), _hptr(quals), n, ctypes.byref(got)))
        return keys[:got.value], ids[:got.value], quals[:got.value]

    def launch_count(self):
        n = _U64(0)
        _check(self._lib.kmi_posqual_launch_count(self._h, ctypes.byref(n)))
        return n.value


def CountIndex(k=31, **kw):
    """CountIndex<MapType> alias, reference src/index/kmer_index.hpp:408-409."""
    return Index(k=k, mode=MODE_COUNT, **kw)


def PositionIndex(k=21, **kw):
    """PositionIndex<MapType> alias, reference src/index/kmer_index.hpp:402-403."""
    return Index(k=k, mode=MODE_POSITION, **kw)


def synth_record_bytes(read_len=150):
    return load_library().kmi_synth_record_bytes(read_len)


def synth_fastq_dev(device, dptr, first_read, n_reads, read_len, genome_len, seed, err_ppm=5000, n_ppm=0):
    _check(load_library().kmi_synth_fastq_dev(device, dptr, first_read, n_reads, read_len, genome_len, seed, err_ppm,
                                              n_ppm))


def synth_fasta_bytes(n_seq, seq_len, width=80):
    return load_library().kmi_synth_fasta_bytes(n_seq, seq_len, width)


def synth_fasta_dev(device, dptr, file_offset, n_bytes, n_seq, seq_len, width=80, seed=1):
    _check(load_library().kmi_synth_fasta_dev(device, dptr, file_offset, n_bytes, n_seq, seq_len, width, seed))


def synth_queries_dev(device, dptr, first, n, k, n_seq, seq_len, genome_seed=1, seed=4):
    _check(load_library().kmi_synth_queries_dev(device, dptr, first, n, k, n_seq, seq_len, genome_seed, seed))
