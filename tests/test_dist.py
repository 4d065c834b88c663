"""Exchange plumbing of the sharded index (kmerind_b200/dist.py).

CPU (gloo, world_size 2): the count exchange and the all-to-all of key rows deliver what
mxx::all2all / mxx::all2allv deliver in imxx::distribute (reference src/io/incremental_mxx.hpp:1087,1098):
received rows grouped by source rank, each group in the sender's order. The owner rule is checked
against the oracle's dist hash. GPU (nccl, 2 GPUs when present): a sharded CountIndex equals the
single-shard one.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_util as ou


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _gloo_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from kmerind_b200.dist import exchange_counts, exchange_rows
        rng = np.random.default_rng(100 + rank)
        n = 1000 + 37 * rank
        keys = rng.integers(0, 1 << 62, size=n, dtype=np.int64)  # any 62-bit value is a 31-mer
        # owner by the oracle's dist hash (farm seed 83) % world — KeyToRank
        owner = (ou.hashes(keys.view(np.uint64), 31, 3) % np.uint64(world)).astype(np.int64)
        order = np.argsort(owner, kind="stable")           # stable: sender order kept inside a group
        send_rows = torch.from_numpy(keys[order]).view(-1, 1)
        send_counts = torch.from_numpy(np.bincount(owner, minlength=world).astype(np.int64))
        recv_counts = exchange_counts(send_counts)
        got = exchange_rows(send_rows, send_counts, recv_counts).view(-1).numpy()
        np.save(os.path.join(out_dir, "keys%d.npy" % rank), keys)
        np.save(os.path.join(out_dir, "owner%d.npy" % rank), owner)
        np.save(os.path.join(out_dir, "got%d.npy" % rank), got)
        np.save(os.path.join(out_dir, "rc%d.npy" % rank), recv_counts.numpy())
        # empty contribution from one rank must still work (every rank calls every collective)
        empty = torch.zeros((0, 1), dtype=torch.int64)
        zc = torch.zeros(world, dtype=torch.int64)
        if rank == 0:
            r2 = exchange_counts(zc)
            g2 = exchange_rows(empty, zc, r2)
        else:
            r2 = exchange_counts(send_counts)
            g2 = exchange_rows(send_rows, send_counts, r2)
        np.save(os.path.join(out_dir, "got2_%d.npy" % rank), g2.view(-1).numpy())
    finally:
        dist.destroy_process_group()


def _gloo_fasta_chain_worker(rank, world, port, out_dir, cuts):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from kmerind_b200.dist import chain_fasta_blocks
        from test_partition import block_summary
        buf = ou.read_fixture("test.fasta")
        b, e = cuts[rank], cuts[rank + 1]
        bol = b == 0 or buf[b - 1] == 0x0A
        runs0, flags = chain_fasta_blocks(block_summary(buf[b:e], bol), bol, b)
        np.save(os.path.join(out_dir, "chain%d.npy" % rank), np.array([runs0, flags]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("cuts", ([0, 200, 512], [0, 1, 512], [0, 0, 512], [0, 333, 512]))
def test_fasta_block_chain_gloo_world2(tmp_path, cuts):
    """the host side of the sharded FASTA build: every rank learns, from one all-gather of block summaries, how many header
    runs began before its block and what kind of line it starts in"""
    import kmerind_b200 as kb
    world = 2
    mp.spawn(_gloo_fasta_chain_worker, args=(world, _free_port(), str(tmp_path), cuts), nprocs=world, join=True)
    buf = ou.read_fixture("test.fasta")
    fsh = buf[0] in (0x3E, 0x3B)
    for r in range(world):
        runs0, flags = np.load(tmp_path / ("chain%d.npy" % r)).tolist()
        want = ou.fasta_prefix_state(buf, cuts[r])
        assert (runs0, (flags >> 1) & 1, bool(flags & 1)) == want, (r, cuts)
        assert bool(flags & kb.FASTA_FILE_STARTS_WITH_HEADER) == bool(fsh)


def test_exchange_gloo_world2(tmp_path):
    world = 2
    mp.spawn(_gloo_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    keys = [np.load(tmp_path / ("keys%d.npy" % r)) for r in range(world)]
    owner = [np.load(tmp_path / ("owner%d.npy" % r)) for r in range(world)]
    for dst in range(world):
        got = np.load(tmp_path / ("got%d.npy" % dst))
        rc = np.load(tmp_path / ("rc%d.npy" % dst))
        want = np.concatenate([keys[src][owner[src] == dst] for src in range(world)])  # by source, sender order
        assert rc.tolist() == [int((owner[src] == dst).sum()) for src in range(world)]
        assert np.array_equal(got, want)
    # second exchange: rank 0 sent nothing
    for dst in range(world):
        got2 = np.load(tmp_path / ("got2_%d.npy" % dst))
        assert np.array_equal(got2, keys[1][owner[1] == dst])


def _nccl_worker(rank, world, port, out_dir, peer=False, k=31):
    if peer == "prefix":          # sub-bin layout of the receive buffer (kmi_exchange_bins > 1)
        os.environ["KMI_EXCHANGE_PREFIX"] = "1"
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import kmerind_b200 as kb
        from kmerind_b200.dist import ShardedIndex
        n_reads = 30000
        buf = ou.synth_fastq(rank * n_reads, n_reads, 150, 400000, 11, 5000, 0)
        text = torch.from_numpy(np.concatenate([buf, np.zeros(64, np.uint8)])).cuda()
        idx = ShardedIndex(k, hash=kb.HASH_FARM, device=rank)
        if peer:
            idx.local.set_build_path(2)   # received keys take the bucketed insert even at this size
            assert idx.enable_peer_exchange(n_reads * 120), "CUDA IPC peer mapping failed"
            if k > 32:
                assert idx.local.exchange_bins() == 1
            if k <= 32:
                assert idx.local.exchange_bins() == (128 if peer == "prefix" else 1)
        n = idx.build_fastq_dev(text, buf.size)
        assert n == n_reads * (150 - k + 1)
        if peer:
            assert "parse+scatter+exchange" in idx.stage_ms
            assert idx.build_fastq_dev(text, buf.size) == n   # a second round through the same buffers
            # third time from a pinned host buffer in 3 record-aligned chunks (H2D copies pipelined with the rounds)
            host = torch.from_numpy(buf.copy()).pin_memory()
            rec = buf.size // n_reads
            assert idx.build_fastq_host(host, [0, 10000 * rec, 20000 * rec, buf.size]) == n

        dk, dc = idx.local.sorted_dump()
        np.save(os.path.join(out_dir, "k%d.npy" % rank), dk)
        np.save(os.path.join(out_dir, "c%d.npy" % rank), dc)
        total = idx.size()
        # batched query through the exchange: every rank asks for its own first 1000 k-mers + 1000 absent keys
        keys, _, _ = ou.fastq_kmers(buf, k)
        absent = np.arange(1000, dtype=np.uint64) * np.uint64(7919) + np.uint64(1 << 40)
        if k > 32:
            absent = np.stack([absent, absent >> np.uint64(3)], axis=1)
        q = np.concatenate([keys[:1000], absent])
        vals, found = idx.find(q)
        np.save(os.path.join(out_dir, "q%d.npy" % rank), q)
        np.save(os.path.join(out_dir, "v%d.npy" % rank), vals)
        np.save(os.path.join(out_dir, "f%d.npy" % rank), found)
        np.save(os.path.join(out_dir, "size%d.npy" % rank), np.array([total]))
        idx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("peer,K", ((False, 31), (True, 31), ("prefix", 31), (True, 63)))
def test_sharded_count_index_nccl_world2(tmp_path, peer, K):
    """peer=False: keys travel by NCCL all-to-all; peer=True: the parse kernel stores them straight into
    the owners' HBM over NVLink (kmi_exchange_*), and the build is run three times, the last one from pinned host memory in
    chunks (counts triple); K=63: two-word keys through the same exchange and the two-word segment insert."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    world = 2
    mp.spawn(_nccl_worker, args=(world, _free_port(), str(tmp_path), peer, K), nprocs=world, join=True)
    n_reads = 30000
    nw = ou.nwords(K)
    bufs = [ou.synth_fastq(r * n_reads, n_reads, 150, 400000, 11, 5000, 0) for r in range(world)]
    keys = np.concatenate([ou.fastq_kmers(b, K)[0] for b in bufs])
    wk, wc = ou.count_sorted(keys, K)
    if peer:
        wc = 3 * wc
    owner = (ou.hashes(wk, K, 3) % np.uint64(world)).astype(np.int64)
    rows = (lambda a: [tuple(np.atleast_1d(r).tolist()) for r in a])
    for r in range(world):
        k = np.load(tmp_path / ("k%d.npy" % r))
        c = np.load(tmp_path / ("c%d.npy" % r))
        assert np.array_equal(k, wk[owner == r]) and np.array_equal(c, wc[owner == r])   # per-rank local contents
        assert int(np.load(tmp_path / ("size%d.npy" % r))[0]) == len(wk)
        q = np.load(tmp_path / ("q%d.npy" % r))
        v = np.load(tmp_path / ("v%d.npy" % r))
        f = np.load(tmp_path / ("f%d.npy" % r))
        cq = ou.shape_keys(np.ascontiguousarray(ou.canonical(q, K)).reshape(-1), K)
        want = dict(zip(rows(wk), wc.tolist()))
        got_hit = [kk in want for kk in rows(cq)]
        assert f.astype(bool).tolist() == got_hit
        assert [int(x) for x, h in zip(v, got_hit) if h] == [want[kk] for kk in rows(cq) if kk in want]


def _nccl_fasta_worker(rank, world, port, out_dir, cut):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import kmerind_b200 as kb
        from kmerind_b200.dist import ShardedIndex
        k = 21
        buf = ou.read_fixture("test.medium.fasta")
        cuts = [0, cut, int(buf.size)]
        b, e = cuts[rank], cuts[rank + 1]
        he = kb.fasta_halo_end(buf, e, k)
        text = torch.from_numpy(np.concatenate([buf[b:he], np.zeros(64, np.uint8)])).cuda()
        idx = ShardedIndex(k, mode=kb.MODE_POSITION, hash=kb.HASH_FARM, device=rank)
        n = idx.build_fasta_dev(text, he - b, n_valid=e - b, file_offset=b, bol=(b == 0 or buf[b - 1] == 0x0A))
        dk, dv = idx.local.sorted_dump()
        np.save(os.path.join(out_dir, "fk%d.npy" % rank), dk)
        np.save(os.path.join(out_dir, "fv%d.npy" % rank), dv)
        np.save(os.path.join(out_dir, "fn%d.npy" % rank), np.array([n]))
        idx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_sharded_position_index_from_fasta_blocks_nccl_world2(tmp_path):
    """One FASTA file, two ranks, cut in the middle of a line: every rank parses its byte range (+ K-1 halo) with global
    sequence ids, the tuples travel to their owners, and each rank ends up with exactly the reference's tuples it owns."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    world, k = 2, 21
    buf = ou.read_fixture("test.medium.fasta")
    cut = int(buf.size) // 2 + 17
    mp.spawn(_nccl_fasta_worker, args=(world, _free_port(), str(tmp_path), cut), nprocs=world, join=True)
    wk, wi, _ = ou.fasta_kmers(buf, k)
    owner = (ou.hashes(wk, k, 3) % np.uint64(world)).astype(np.int64)
    cuts = [0, cut, int(buf.size)]
    for r in range(world):
        assert int(np.load(tmp_path / ("fn%d.npy" % r))[0]) == len(ou.fasta_block_kmers(buf, k, cuts[r], cuts[r + 1])[0])
        sk, si = ou.sort_tuples(wk[owner == r], wi[owner == r], k)
        assert np.array_equal(np.load(tmp_path / ("fk%d.npy" % r)), sk)
        assert np.array_equal(np.load(tmp_path / ("fv%d.npy" % r)), si)
