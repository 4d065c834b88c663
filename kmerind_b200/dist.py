"""Sharded (multi-GPU) index: ctypes binding of the C sharded layer (kmi_comm_* / kmi_dist_*, csrc/kmi_dist.cu).

Everything collective — the owner exchange of imxx::distribute (reference src/io/incremental_mxx.hpp:1039-1109), the
NVLink push of the count-index build, the query round trip, the size all-reduce — runs inside libkmi.so over NCCL and
CUDA IPC. This module only (1) carries NCCL's unique id from rank 0 to the other ranks through the caller's
torch.distributed group (any backend: nccl on the GPU box, gloo in the CPU tests), the way an MPI program would use
MPI_Bcast, and (2) forwards calls. Test / bench harness; C++ callers use include/kmi.h directly (tests/cpp/dist_check.cpp).
"""
import ctypes
import os

import numpy as np

from . import (Index, KmiError, MODE_COUNT, MODE_POSITION, HASH_MURMUR, _check, _hptr, _as_keys, _U64, load_library, nwords,
               fasta_chain)

COMM_ID_BYTES = 128


def bootstrap_id(group=None):
    """rank 0: kmi_comm_unique_id; everyone: the same 128 bytes (broadcast over `group`)."""
    import torch.distributed as dist
    box = [None]
    if dist.get_rank(group) == 0:
        buf = np.zeros(COMM_ID_BYTES, dtype=np.uint8)
        try:
            _check(load_library().kmi_comm_unique_id(_hptr(buf)))
            box[0] = buf.tobytes()
        except KmiError as e:              # e.g. no libnccl: every rank must learn it, none may wait for an id
            box[0] = (e.code, str(e))
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    if isinstance(box[0], tuple):
        raise KmiError(box[0][0], box[0][1])
    return box[0]


class Comm:
    """kmi_comm: one NCCL communicator inside libkmi.so."""

    def __init__(self, id_bytes, nranks, rank, device):
        self._lib = load_library()
        self._h = ctypes.c_void_p()
        buf = np.frombuffer(id_bytes, dtype=np.uint8).copy()
        _check(self._lib.kmi_comm_create(_hptr(buf), nranks, rank, device, ctypes.byref(self._h)))
        self.rank, self.nranks, self.device = rank, nranks, device

    def close(self):
        if self._h:
            self._lib.kmi_comm_destroy(self._h)
            self._h = ctypes.c_void_p()

    def barrier(self):
        _check(self._lib.kmi_comm_barrier(self._h))

    def allreduce(self, vals, op="sum"):
        a = np.ascontiguousarray(vals, dtype=np.uint64).copy()
        _check(self._lib.kmi_comm_allreduce_u64(self._h, _hptr(a), a.size, {"sum": 0, "max": 1, "min": 2}[op]))
        return a

    def allgather(self, vals):
        a = np.ascontiguousarray(vals, dtype=np.uint64)
        out = np.zeros(a.size * self.nranks, dtype=np.uint64)
        _check(self._lib.kmi_comm_allgather_u64(self._h, _hptr(a), a.size, _hptr(out)))
        return out.reshape(self.nranks, a.size)


class KmiFastaPart(ctypes.Structure):
    _fields_ = [("file_offset", ctypes.c_uint64), ("n_valid", ctypes.c_uint64), ("bol", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class ShardedIndex:
    """CountIndex / PositionIndex sharded over the ranks of a communicator.

    Every rank must call every batch operation (also with empty input), exactly like the reference's distributed maps
    (distributed_densehash_map.hpp:2567-2573)."""

    def __init__(self, k=31, mode=MODE_COUNT, canonical=True, hash=HASH_MURMUR, nsplit=False, device=0, reserve=0, group=None,
                 comm=None):
        if comm is None:
            import torch.distributed as dist
            comm = Comm(bootstrap_id(group), dist.get_world_size(group), dist.get_rank(group), device)
            self._own_comm = True
        else:
            self._own_comm = False
        self.comm = comm
        self.rank, self.nranks = comm.rank, comm.nranks
        self.k, self.nw, self.mode = k, nwords(k), mode
        self.local = Index(k=k, mode=mode, canonical=canonical, hash=hash, nsplit=nsplit, device=device, rank=self.rank,
                           nranks=self.nranks, reserve=reserve)
        self._lib = self.local._lib
        _check(self._lib.kmi_dist_attach(self.local._h, comm._h))
        self.stage_ms, self.n_received, self.path = {}, 0, 0

    def close(self):
        self.local.close()
        if self._own_comm:
            self.comm.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _stats(self):
        ms = (ctypes.c_float * 4)()
        nr, path = _U64(0), ctypes.c_int32(0)
        _check(self._lib.kmi_dist_last_stats(self.local._h, ms, ctypes.byref(nr), ctypes.byref(path)))
        self.n_received, self.path = nr.value, path.value
        self.stage_ms = {"whole": ms[0], "parse+push" if path.value == 1 else "parse+exchange": ms[1], "insert": ms[2]}

    # -- build -----------------------------------------------------------------------------------
    def build_fastq_dev(self, dptr, nbytes, file_offset=0):
        """dptr: device address of this rank's whole FASTQ records. Returns the k-mers generated here."""
        n = _U64(0)
        _check(self._lib.kmi_dist_build_fastq_dev(self.local._h, dptr, nbytes, file_offset, ctypes.byref(n)))
        self._stats()
        return n.value

    def build_fastq(self, data, file_offset=0):
        """host bytes (numpy uint8, ideally pinned) of this rank's whole records; the copy is pipelined with the build"""
        a = np.ascontiguousarray(data, dtype=np.uint8)
        n = _U64(0)
        _check(self._lib.kmi_dist_build_fastq(self.local._h, _hptr(a), a.size, file_offset, ctypes.byref(n)))
        self._stats()
        return n.value

    def build_file(self, path, format=0):
        """Index::build_posix(filename, comm): every rank passes the same path and builds from ITS block of the file
        (kmi_dist_build_file); returns the k-mers generated from this rank's block"""
        n = _U64(0)
        _check(self._lib.kmi_dist_build_file(self.local._h, os.fsencode(path), format, ctypes.byref(n)))
        self._stats()
        return n.value

    def build_fasta(self, data, n_valid=None, file_offset=0, bol=True):
        """With the defaults every rank holds a whole file of its own; with n_valid rank r holds block r of ONE file:
        data = the block's n_valid bytes + halo (kmerind_b200.fasta_halo_end), file_offset = file byte of data[0]."""
        a = np.ascontiguousarray(data, dtype=np.uint8)
        part = ctypes.byref(KmiFastaPart(int(file_offset), int(n_valid), 1 if bol else 0, 0)) if n_valid is not None else None
        n = _U64(0)
        _check(self._lib.kmi_dist_build_fasta(self.local._h, _hptr(a), a.size, part, ctypes.byref(n)))
        self._stats()
        return n.value

    def build_fasta_dev(self, dptr, nbytes, n_valid=None, file_offset=0, bol=True):
        """as build_fasta, the bytes (block + halo) already on the device"""
        part = ctypes.byref(KmiFastaPart(int(file_offset), int(n_valid), 1 if bol else 0, 0)) if n_valid is not None else None
        n = _U64(0)
        _check(self._lib.kmi_dist_build_fasta_dev(self.local._h, dptr, nbytes, part, ctypes.byref(n)))
        self._stats()
        return n.value

    def insert(self, keys, vals=None):
        a = _as_keys(keys, self.nw)
        v = None if vals is None else np.ascontiguousarray(vals, dtype=np.uint64)
        _check(self._lib.kmi_dist_insert(self.local._h, _hptr(a), _hptr(v), a.shape[0]))

    # -- query -------------------------------------------------------------------------------------
    def count(self, keys):
        a = _as_keys(keys, self.nw)
        out = np.zeros(a.shape[0], dtype=np.uint64)
        _check(self._lib.kmi_dist_count(self.local._h, _hptr(a), a.shape[0], _hptr(out)))
        return out

    def count_dev(self, d_keys, n, d_counts):
        """queries resident in HBM, answers left there (kmi_dist_count_dev); collective"""
        _check(self._lib.kmi_dist_count_dev(self.local._h, d_keys, n, d_counts))

    def find_dev(self, d_keys, n, d_vals, d_found):
        _check(self._lib.kmi_dist_find_dev(self.local._h, d_keys, n, d_vals, d_found))

    def find(self, keys):
        a = _as_keys(keys, self.nw)
        vals = np.zeros(a.shape[0], dtype=np.uint64)
        found = np.zeros(a.shape[0], dtype=np.uint8)
        _check(self._lib.kmi_dist_find(self.local._h, _hptr(a), a.shape[0], _hptr(vals), _hptr(found)))
        return vals, found

    def find_positions(self, keys, cap=None):
        a = _as_keys(keys, self.nw)
        n = a.shape[0]
        if cap is None:
            cap = int(self.count(a).sum())
        offsets = np.zeros(n + 1, dtype=np.uint64)
        vals = np.zeros(max(cap, 1), dtype=np.uint64)
        _check(self._lib.kmi_dist_find_positions(self.local._h, _hptr(a), n, _hptr(offsets), _hptr(vals), cap))
        return offsets, vals[:int(offsets[-1])]

    def erase(self, keys):
        a = _as_keys(keys, self.nw)
        _check(self._lib.kmi_dist_erase(self.local._h, _hptr(a), a.shape[0]))

    # -- sizes -------------------------------------------------------------------------------------
    def local_size(self):
        return self.local.local_size()

    def size(self):
        """Index::size(): allreduce of local sizes (distributed_map_base.hpp:243)."""
        n = _U64(0)
        _check(self._lib.kmi_dist_size(self.local._h, ctypes.byref(n)))
        return n.value
