"""Sharded (multi-GPU) index: one process per GPU, keys owned by ``dist_hash(key) % nranks``.

This is the plumbing half of imxx::distribute (reference src/io/incremental_mxx.hpp:1039-1109):
the local half — owner computation, histogram, stable scatter — is the CUDA partition stage of
libkmi.so (kmi_partition_dev); the exchange itself (mxx::all2all of the counts :1087 and
mxx::all2allv of the keys :1098) is torch.distributed all_to_all_single, i.e. NCCL over NVLink /
NVSwitch on GPUs (gloo on CPU tensors in the world_size-2 tests). Nothing here computes on keys.
"""
import numpy as np
import torch
import torch.distributed as dist

from . import Index, MODE_COUNT, MODE_POSITION, HASH_MURMUR, nwords


def exchange_counts(send_counts, group=None):
    """mxx::all2all(send_counts): returns recv_counts (int64 tensor on the same device)."""
    recv = torch.empty_like(send_counts)
    dist.all_to_all_single(recv, send_counts, group=group)
    return recv


def exchange_rows(rows, send_counts, recv_counts, group=None):
    """mxx::all2allv: `rows` is (n, w) int64 grouped by destination rank in rank order; returns the
    rows received, grouped by source rank, each group in the sender's order."""
    w = rows.shape[1] if rows.dim() == 2 else 1
    out = torch.empty((int(recv_counts.sum().item()), w) if rows.dim() == 2 else (int(recv_counts.sum().item()),),
                      dtype=rows.dtype, device=rows.device)
    dist.all_to_all_single(out, rows.contiguous(), output_split_sizes=[int(x) for x in recv_counts.tolist()],
                           input_split_sizes=[int(x) for x in send_counts.tolist()], group=group)
    return out


def chain_fasta_blocks(summary, bol, file_offset, group=None, file_starts_with_header=None):
    """What THIS rank's block of a partitioned FASTA file inherits from the ranks before it: one all-gather of the block
    summaries (kmi_fasta_block_summary_dev: has a line start, first / last line start is a header, header runs) and the
    exclusive scan of kmerind_b200.fasta_chain over them — the cross-rank part of FASTAParser::init_parser
    (fasta_loader.hpp:202-483). Blocks are in rank order. Returns (header_runs_before, flags) for kmi_fasta_block."""
    from . import fasta_chain
    world = dist.get_world_size(group)
    mine = tuple(int(x) for x in summary) + (1 if bol else 0, int(file_offset))
    every = [None] * world
    dist.all_gather_object(every, mine, group=group)
    if file_starts_with_header is None:   # the first non-empty block at file offset 0 opens the file
        file_starts_with_header = False
        for e in every:
            if e[0] and e[4] and e[5] == 0:
                file_starts_with_header = bool(e[1])
                break
    return fasta_chain([e[:4] for e in every], [e[4] for e in every], file_starts_with_header)[dist.get_rank(group)]


class ShardedIndex:
    """CountIndex / PositionIndex sharded over the ranks of a torch.distributed group.

    Every rank must call every batch operation (also with empty input), exactly like the
    reference's distributed maps (distributed_densehash_map.hpp:2567-2573).
    """

    def __init__(self, k=31, mode=MODE_COUNT, canonical=True, hash=HASH_MURMUR, nsplit=False, device=0, reserve=0,
                 group=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.nranks = dist.get_world_size(group)
        self.k, self.nw, self.mode = k, nwords(k), mode
        self.device = torch.device("cuda", device)
        # the parser side canonicalises; keys arriving after the exchange are stored as they are
        self.local = Index(k=k, mode=mode, canonical=canonical, hash=hash, nsplit=nsplit, device=device, rank=self.rank,
                           nranks=self.nranks, reserve=reserve)
        self.canonical = canonical
        self.stage_ms = {}

    def close(self):
        self.local.close()

    # -- build -----------------------------------------------------------------------------------
    def _distribute(self, d_keys, d_vals, n):
        """partition (CUDA) -> count exchange -> key exchange (NCCL). Returns (recv_keys, recv_vals, n_recv)."""
        p, nw = self.nranks, self.nw
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        ev[0].record()
        part_k = torch.empty(max(n, 1) * nw, dtype=torch.int64, device=self.device)
        part_v = torch.empty(max(n, 1), dtype=torch.int64, device=self.device) if d_vals is not None else None
        counts = self.local.partition_dev(d_keys.data_ptr(), d_vals.data_ptr() if d_vals is not None else None, n,
                                          part_k.data_ptr(), part_v.data_ptr() if part_v is not None else None, p)
        ev[1].record()
        send = torch.from_numpy(counts.astype(np.int64)).to(self.device)
        recv = exchange_counts(send, self.group)
        rk = exchange_rows(part_k[:n * nw].view(n, nw), send, recv, self.group)
        rv = exchange_rows(part_v[:n], send, recv, self.group) if part_v is not None else None
        ev[2].record()
        self._ev = ev
        return rk.reshape(-1), rv, int(recv.sum().item())

    def _insert_received(self, rk, rv, n_recv):
        # received keys are already canonical (or intentionally not): store verbatim
        import ctypes
        from . import _check
        lib = self.local._lib
        # temporarily bypass the input transform: kmi_insert_dev canonicalises when cfg.canonical is set,
        # and canonical(canonical(x)) == canonical(x), so inserting as-is is exact either way.
        _check(lib.kmi_insert_dev(self.local._h, rk.data_ptr(), rv.data_ptr() if rv is not None else None, n_recv))

    def _buffer(self, name, n_words):
        """persistent device scratch (int64 words), grown geometrically"""
        t = getattr(self, name, None)
        if t is None or t.numel() < n_words:
            t = None
            setattr(self, name, None)
            t = torch.empty(int(n_words * 1.1) + 1024, dtype=torch.int64, device=self.device)
            setattr(self, name, t)
        return t

    def enable_peer_exchange(self, region_cap):
        """Map every rank's receive buffer into this process (CUDA IPC over NVLink) so that the parse
        kernel writes keys straight into their owners' memory. Collective. Returns False (and keeps the
        NCCL all-to-all path) when IPC mapping is not possible."""
        from . import KmiError
        ok = 1
        try:
            handle = self.local.exchange_setup(int(region_cap))
        except KmiError:
            handle, ok = b"\0" * 64, 0
        handles = [None] * self.nranks
        dist.all_gather_object(handles, (ok, handle), group=self.group)
        if all(h[0] for h in handles):
            try:
                self.local.exchange_connect([h[1] for h in handles])
            except KmiError:
                ok = 0
        else:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int64, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        self._peer_cap = int(region_cap) if int(flag.item()) == 1 else 0
        return self._peer_cap > 0

    def _build_fastq_dev_peer(self, text, nbytes):
        """parse + owner + scatter + exchange as one kernel (stores land in the owners' HBM over NVLink)."""
        p = self.nranks
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        dist.barrier(group=self.group)          # nobody is still reading its receive buffer of the last round
        from . import KmiError, ERR_CAPACITY
        ev[0].record()
        try:
            counts, n = self.local.exchange_parse_fastq_dev(text.data_ptr(), nbytes, p)
            send_counts = torch.from_numpy(counts.astype(np.int64)).to(self.device)
        except KmiError as e:
            if e.code != ERR_CAPACITY:
                raise
            n, send_counts = None, torch.full((p * self.local.exchange_bins(),), -1, dtype=torch.int64, device=self.device)  # skewed owners here
        ev[1].record()
        recv_counts = exchange_counts(send_counts, self.group).cpu().numpy()   # also: "the writes are complete"
        bad = torch.tensor([1 if (n is None or (recv_counts < 0).any()) else 0], dtype=torch.int64, device=self.device)
        dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=self.group)
        if int(bad.item()):
            return None                        # every rank leaves the peer path together
        ev[2].record()
        self.local.exchange_insert(recv_counts.astype(np.uint64))
        ev[3].record()
        torch.cuda.synchronize(self.device)
        self.stage_ms = {"parse+scatter+exchange": ev[0].elapsed_time(ev[1]), "counts": ev[1].elapsed_time(ev[2]),
                         "insert": ev[2].elapsed_time(ev[3])}
        self.n_received = int(recv_counts.sum())
        return n

    def build_fastq_dev(self, text, nbytes, max_kmers=None):
        """text: CUDA uint8 tensor holding this rank's whole records. Returns k-mers generated here.

        parse + owner + scatter run as ONE kernel straight into the per-owner send regions
        (kmi_parse_fastq_by_owner_dev); the regions are exchanged with one grouped NCCL all-to-all and the
        received keys go through the bucketed insert. Falls back to parse -> partition when an owner
        region overflows (skewed owners)."""
        from . import KmiError, ERR_CAPACITY
        nw, p = self.nw, self.nranks
        if getattr(self, "_peer_cap", 0):
            n = self._build_fastq_dev_peer(text, nbytes)
            if n is not None:
                return n
        upper = max_kmers if max_kmers is not None else max(nbytes // 2 + 1, 1)
        region_cap = upper // p + upper // (8 * p) + 4096
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        send = self._buffer("_send", p * region_cap * nw)
        try:
            counts, n = self.local.parse_fastq_by_owner_dev(text.data_ptr(), nbytes, send.data_ptr(), region_cap, p)
        except KmiError as e:
            if e.code != ERR_CAPACITY:
                raise
            return self._build_fastq_dev_unfused(text, nbytes, max_kmers)
        ev[1].record()
        send_counts = torch.from_numpy(counts.astype(np.int64)).to(self.device)
        recv_counts = exchange_counts(send_counts, self.group).tolist()
        n_recv = int(sum(recv_counts))
        recv = self._buffer("_recv", max(n_recv, 1) * nw)
        ins, outs, off = [], [], 0
        for r in range(p):
            ins.append(send[r * region_cap * nw:(r * region_cap + int(counts[r])) * nw])
            outs.append(recv[off * nw:(off + int(recv_counts[r])) * nw])
            off += int(recv_counts[r])
        dist.all_to_all(outs, ins, group=self.group)
        ev[2].record()
        self._insert_received(recv, None, n_recv)
        ev[3].record()
        torch.cuda.synchronize(self.device)
        self.stage_ms = {"parse+scatter": ev[0].elapsed_time(ev[1]), "a2a": ev[1].elapsed_time(ev[2]),
                         "insert": ev[2].elapsed_time(ev[3])}
        self.n_received = n_recv
        return n

    def build_fastq_host(self, host, cuts, max_kmers_per_chunk=None):
        """Host-buffer build of this rank's share (the sharded counterpart of kmi_build_fastq).

        host: pinned CPU uint8 tensor with whole FASTQ records; cuts: record-aligned offsets [0, c1, ..., nbytes] (every
        rank must pass the same NUMBER of chunks: each chunk is one collective exchange round). Chunk i+1 crosses PCIe on
        a copy stream while chunk i is parsed, exchanged and inserted, so the transfer hides all but the last round."""
        n_chunks = len(cuts) - 1
        big = max(int(cuts[i + 1] - cuts[i]) for i in range(n_chunks)) if n_chunks else 0
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream(device=self.device)
        bufs = [self._buffer("_text%d" % j, (big + 64 + 7) // 8).view(torch.uint8) for j in range(2)]
        done = [torch.cuda.Event() for _ in range(n_chunks)]

        def copy(i):
            a, b = int(cuts[i]), int(cuts[i + 1])
            with torch.cuda.stream(self._copy_stream):
                bufs[i % 2][:b - a].copy_(host[a:b], non_blocking=True)
                done[i].record(self._copy_stream)

        self._copy_stream.wait_stream(torch.cuda.current_stream(self.device))   # the staging buffers may still be in use
        total = 0
        if n_chunks:
            copy(0)
        for i in range(n_chunks):
            if i + 1 < n_chunks:
                copy(i + 1)       # its buffer held chunk i-1, whose parse has completed (the call below is synchronous)
            torch.cuda.current_stream(self.device).wait_event(done[i])
            total += self.build_fastq_dev(bufs[i % 2], int(cuts[i + 1] - cuts[i]), max_kmers=max_kmers_per_chunk)
        return total

    def _build_fastq_dev_unfused(self, text, nbytes, max_kmers=None):
        nw = self.nw
        cap = max_kmers if max_kmers is not None else max(nbytes // 2 + 1, 1)
        keys = torch.empty(cap * nw, dtype=torch.int64, device=self.device)
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t2 = torch.cuda.Event(enable_timing=True)
        t0.record()
        n = self.local.parse_fastq_dev(text.data_ptr(), nbytes, keys.data_ptr(), cap)
        t1.record()
        rk, rv, n_recv = self._distribute(keys, None, n)
        del keys
        self._insert_received(rk, rv, n_recv)
        t2.record()
        torch.cuda.synchronize(self.device)
        self.stage_ms = {"parse": t0.elapsed_time(t1), "partition": self._ev[0].elapsed_time(self._ev[1]),
                         "a2a": self._ev[1].elapsed_time(self._ev[2]), "insert": self._ev[2].elapsed_time(t2)}
        self.n_received = n_recv
        return n

    def build_fasta_dev(self, text, nbytes, n_valid=None, file_offset=0, bol=True, file_starts_with_header=None):
        """Sharded FASTA build. With the defaults every rank holds a whole file of its own. With n_valid given, rank r
        holds block r of ONE file (blocks in rank order, cut at any byte): text = the block's n_valid bytes + halo
        (kmerind_b200.fasta_halo_end), file_offset = file byte of text[0], bol = text[0] starts a line. The ranks combine
        their block summaries (header runs, kind of the last line) with one all-gather — the cross-rank part of
        FASTAParser::init_parser, fasta_loader.hpp:202-483 — and each parses its block with global sequence ids and the
        K-1 tail rule (kmer_parser.hpp:114-157)."""
        nw = self.nw
        block = None
        if n_valid is not None:
            summary = self.local.fasta_block_summary_dev(text.data_ptr(), n_valid, bol)
            runs0, flags = chain_fasta_blocks(summary, bol, file_offset, self.group, file_starts_with_header)
            block = (int(file_offset), int(n_valid), runs0, flags)
        cap = max(nbytes if n_valid is None else n_valid, 1)
        keys = torch.empty(cap * nw, dtype=torch.int64, device=self.device)
        ids = torch.empty(cap, dtype=torch.int64, device=self.device) if self.mode == MODE_POSITION else None
        n = self.local.parse_fasta_dev(text.data_ptr(), nbytes, keys.data_ptr(), cap, ids.data_ptr() if ids is not None else None,
                                       block=block)
        rk, rv, n_recv = self._distribute(keys, ids, n)
        self._insert_received(rk, rv, n_recv)
        torch.cuda.synchronize(self.device)
        return n

    def insert(self, keys, vals=None):
        """Index::insert on host arrays (any rank may hold any keys)."""
        a = np.ascontiguousarray(keys, dtype=np.uint64).reshape(-1)
        n = a.size // self.nw
        dk = torch.from_numpy(a.view(np.int64)).to(self.device)
        if self.canonical and n:
            self.local.canonicalize_dev(dk.data_ptr(), dk.data_ptr(), n)
        dv = torch.from_numpy(np.ascontiguousarray(vals, dtype=np.uint64).view(np.int64)).to(self.device) if vals is not None else None
        rk, rv, n_recv = self._distribute(dk, dv, n)
        self._insert_received(rk, rv, n_recv)
        torch.cuda.synchronize(self.device)

    # -- query -------------------------------------------------------------------------------------
    def _query(self, keys, want_vals):
        """canonicalise -> distribute -> local probe -> answers back along the transposed counts
        (distributed_densehash_map.hpp:1122-1216). Returns answers in INPUT order."""
        a = np.ascontiguousarray(keys, dtype=np.uint64).reshape(-1)
        nw, p = self.nw, self.nranks
        n = a.size // nw
        dk = torch.from_numpy(a.view(np.int64)).to(self.device)
        if self.canonical and n:
            self.local.canonicalize_dev(dk.data_ptr(), dk.data_ptr(), n)
        tag = torch.arange(n, dtype=torch.int64, device=self.device)  # original index travels as the value
        part_k = torch.empty(max(n, 1) * nw, dtype=torch.int64, device=self.device)
        part_t = torch.empty(max(n, 1), dtype=torch.int64, device=self.device)
        counts = self.local.partition_dev(dk.data_ptr(), tag.data_ptr(), n, part_k.data_ptr(), part_t.data_ptr(), p)
        send = torch.from_numpy(counts.astype(np.int64)).to(self.device)
        recv = exchange_counts(send, self.group)
        rk = exchange_rows(part_k[:n * nw].view(n, nw), send, recv, self.group).reshape(-1)
        m = int(recv.sum().item())
        vals = torch.zeros(max(m, 1), dtype=torch.int64, device=self.device)
        found = torch.zeros(max(m, 1), dtype=torch.uint8, device=self.device)
        if want_vals:
            self.local.find_dev(rk.data_ptr(), m, vals.data_ptr(), found.data_ptr())
        else:
            self.local.count_dev(rk.data_ptr(), m, vals.data_ptr())
        back_v = exchange_rows(vals[:m], recv, send, self.group)
        back_f = exchange_rows(found[:m].to(torch.int64), recv, send, self.group)
        out_v = torch.zeros(n, dtype=torch.int64, device=self.device)
        out_f = torch.zeros(n, dtype=torch.int64, device=self.device)
        out_v[part_t[:n]] = back_v
        out_f[part_t[:n]] = back_f
        return out_v.cpu().numpy().view(np.uint64), out_f.cpu().numpy().astype(np.uint8)

    def count(self, keys):
        return self._query(keys, False)[0]

    def find(self, keys):
        return self._query(keys, True)

    # -- sizes -------------------------------------------------------------------------------------
    def local_size(self):
        return self.local.local_size()

    def size(self):
        """Index::size(): allreduce of local sizes (distributed_map_base.hpp:243)."""
        t = torch.tensor([self.local.local_size()], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, group=self.group)
        return int(t.item())
