"""The sharded index (csrc/kmi_dist.cu behind kmi_comm_* / kmi_dist_*; kmerind_b200/dist.py is its ctypes binding).

CPU (gloo, world_size 2): what the host side does — the NCCL id travels from rank 0 to every rank through the caller's
group; without a CUDA device kmi_comm_create fails on every rank (no CPU fallback, nobody hangs); the FASTA block chain
(the exclusive scan the C layer runs over its all-gather) equals a plain prefix walk.
GPU: tests/cpp/dist_check.cpp — C++ main, one forked process per visible GPU, no Python in the loop — and a Python
world-N build whose per-rank local_size / checksum must equal the REFERENCE code's per-rank values at p = N.
On a 1-GPU box both run with one rank (the NVLink push path degenerates to local copies but runs the same kernels).
"""
import os
import socket
import subprocess

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_util as ou

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _init(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)


# ------------------------------------------------------------------------------------------------------------------
# CPU
# ------------------------------------------------------------------------------------------------------------------
def _gloo_bootstrap_worker(rank, world, port, out_dir):
    _init(rank, world, port)
    try:
        import kmerind_b200 as kb
        from kmerind_b200.dist import bootstrap_id, Comm
        try:
            ident = bootstrap_id()
        except kb.KmiError as e:            # libnccl missing on this machine: the failure must be the same on every rank
            ident = b"unsupported:%d" % e.code
        np.save(os.path.join(out_dir, "id%d.npy" % rank), np.frombuffer(ident, dtype=np.uint8))
        outcome = "created"
        if not ident.startswith(b"unsupported"):
            try:
                Comm(ident, world, rank, 0).close()
            except kb.KmiError as e:
                outcome = "error %d" % e.code
        with open(os.path.join(out_dir, "outcome%d.txt" % rank), "w") as f:
            f.write(outcome)
    finally:
        dist.destroy_process_group()


def test_bootstrap_id_reaches_every_rank_and_no_cpu_fallback(tmp_path):
    world = 2
    mp.spawn(_gloo_bootstrap_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    ids = [np.load(tmp_path / ("id%d.npy" % r)) for r in range(world)]
    assert np.array_equal(ids[0], ids[1])
    if not bytes(ids[0]).startswith(b"unsupported"):
        assert ids[0].size == 128 and ids[0].any()
    if not torch.cuda.is_available():
        import kmerind_b200 as kb
        for r in range(world):
            out = open(tmp_path / ("outcome%d.txt" % r)).read()
            assert out in ("error %d" % kb.ERR_CUDA, "created") and out != "created", out


def _gloo_fasta_chain_worker(rank, world, port, out_dir, cuts):
    _init(rank, world, port)
    try:
        import kmerind_b200 as kb
        from test_partition import block_summary
        buf = ou.read_fixture("test.fasta")
        b, e = cuts[rank], cuts[rank + 1]
        bol = b == 0 or buf[b - 1] == 0x0A
        # what kmi_dist_build_fasta_dev all-gathers per rank: its block summary, bol, file offset, first byte is a header
        mine = tuple(int(x) for x in block_summary(buf[b:e], bol)) + (1 if bol else 0, int(b), int(e > b and b == 0 and buf[0] in (0x3E, 0x3B)))
        every = [None] * world
        dist.all_gather_object(every, mine)
        fsh = False
        for s in every:
            if s[0] and s[5] == 0:
                fsh = bool(s[6])
                break
        runs0, flags = kb.fasta_chain([s[:4] for s in every], [s[4] for s in every], fsh)[rank]
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


# ------------------------------------------------------------------------------------------------------------------
# GPU
# ------------------------------------------------------------------------------------------------------------------
def _build_dist_check(tmp_path):
    import kmerind_b200 as kb
    exe = str(tmp_path / "dist_check")
    subprocess.check_call(["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "dist_check.cpp"), kb.LIB_PATH, os.path.join(ROOT, "oracle", "libkmi_oracle.so"),
                           "-o", exe, "-Wl,-rpath," + os.path.dirname(kb.LIB_PATH), "-Wl,-rpath," + os.path.join(ROOT, "oracle")])
    return exe


@pytest.mark.gpu
@pytest.mark.parametrize("k,reads,push", [(31, 240000, None), (31, 60000, "0"), (31, 3000, "1"), (63, 30000, "1"), (21, 40000, "1")])
def test_sharded_index_from_cpp_main(tmp_path, k, reads, push):
    """C++ main(), one process per visible GPU (1 on a single-GPU box): sharded build from host blocks, size, per-shard dump,
    find / count / erase from every rank, sharded PositionIndex from FASTA blocks + find_positions — all against the oracle.
    push None = the library's choice (NVLink push when every rank's block is >= 8 MiB: 240000 reads = 76 MB over <= 8 ranks),
    "0" = the NCCL all-to-all path, "1" = push forced on a small input (small tables also send the received regions down
    the key-stream fallback of the table build)."""
    exe = _build_dist_check(tmp_path)
    env = dict(os.environ)
    if push is not None:
        env["KMI_DIST_PUSH"] = push
    n = max(torch.cuda.device_count(), 1)
    r = subprocess.run([exe, str(n), str(reads), str(k)], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "dist ok: %d ranks" % n in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    assert ("path 2" if push == "0" else "path 1") in r.stdout, r.stdout


def _world_worker(rank, world, port, out_dir, k, n_reads, rounds_env, defer=0):
    _init(rank, world, port)
    if rounds_env:
        os.environ["KMI_DIST_ROUND_BYTES"] = rounds_env
    try:
        import kmerind_b200 as kb
        from kmerind_b200.dist import ShardedIndex
        torch.cuda.set_device(rank)
        buf = ou.synth_fastq(rank * n_reads, n_reads, 150, 5 * n_reads * world, 11, 5000, 0)
        with ShardedIndex(k, hash=kb.HASH_FARM, device=rank) as idx:
            if defer:
                idx.local.set_defer_insert(defer)       # bulk load: received keys are staged, one table sweep per flush
            text = torch.from_numpy(np.concatenate([buf, np.zeros(64, np.uint8)])).cuda()
            n = idx.build_fastq_dev(text.data_ptr(), buf.size)
            assert n == n_reads * (150 - k + 1)
            assert idx.path == 1, "expected the NVLink push path"
            ck = idx.local.checksum()
            np.save(os.path.join(out_dir, "r%d.npy" % rank), np.array([idx.local_size(), ck[0], ck[1], idx.size()], dtype=np.uint64))
            # a second build from pinned host memory into the same index doubles every count
            host = torch.from_numpy(buf.copy()).pin_memory()
            assert idx.build_fastq(host.numpy()) == n
            ck2 = idx.local.checksum()
            with np.errstate(over="ignore"):
                assert ck2[1] == 2 * ck[1] and np.uint64(ck2[0]) == np.uint64(ck[0]) * np.uint64(2)
            keys, _, _ = ou.fastq_kmers(buf[:317 * 50], k)
            vals, found = idx.find(keys)
            np.save(os.path.join(out_dir, "q%d.npy" % rank), keys)
            np.save(os.path.join(out_dir, "v%d.npy" % rank), vals)
            assert found.all()
            # the same batch with the queries resident in HBM and the answers left there
            dk = torch.from_numpy(np.ascontiguousarray(keys).view(np.int64).reshape(-1)).cuda()
            nq = len(keys)
            dv = torch.zeros(nq, dtype=torch.int64, device="cuda")
            df = torch.zeros(nq, dtype=torch.uint8, device="cuda")
            dc = torch.zeros(nq, dtype=torch.int64, device="cuda")
            idx.find_dev(dk.data_ptr(), nq, dv.data_ptr(), df.data_ptr())
            idx.count_dev(dk.data_ptr(), nq, dc.data_ptr())
            torch.cuda.synchronize()
            assert np.array_equal(dv.cpu().numpy().view(np.uint64), vals) and bool(df.cpu().numpy().all())
            assert np.array_equal(dc.cpu().numpy().view(np.uint64), idx.count(keys))
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("k,rounds,defer", [(31, None, 0), (31, str(8 << 20), 0), (63, None, 0), (31, None, 40000000), (63, str(8 << 20), 40000000)])
def test_sharded_build_equals_reference_per_rank(tmp_path, k, rounds, defer):
    """Every rank's shard must be what the REFERENCE's rank holds at p = world: local_size and the content checksum against
    oracle/_ref (reference Kmer / lex_less / farm KeyToRank / dense_hash_map, p ranks over the same blocks), where that was
    built; else against the C oracle's owner split. rounds: several parse/push rounds per build."""
    import kmerind_b200 as kb
    world = max(torch.cuda.device_count(), 1)
    n_reads = 90000
    mp.spawn(_world_worker, args=(world, _free_port(), str(tmp_path), k, n_reads, rounds, defer), nprocs=world, join=True)
    texts = [ou.synth_fastq(r * n_reads, n_reads, 150, 5 * n_reads * world, 11, 5000, 0) for r in range(world)]
    whole = np.concatenate(texts)
    cuts = [r * texts[0].size for r in range(world + 1)]
    ref = ou.ref_count_build_fastq(whole, k, world, 1, cuts=cuts, want_dump=True)
    if ref is None:
        keys, _, _ = ou.fastq_kmers(whole, k)
        owner = (ou.hashes(keys, k, 3) % np.uint64(world)).astype(np.int64)
        sizes = [len(ou.count_sorted(keys[owner == r], k)[0]) for r in range(world)]
        cks = None
    else:
        sizes, cks = ref["sizes"].tolist(), ref["ck"]
    got = [np.load(tmp_path / ("r%d.npy" % r)) for r in range(world)]
    for r in range(world):
        assert int(got[r][0]) == int(sizes[r]), "rank %d local_size %d, reference %d" % (r, got[r][0], sizes[r])
        assert int(got[r][3]) == int(sum(sizes))
        if cks is not None:
            assert int(got[r][1]) == int(cks[r, 0]) and int(got[r][2]) == int(cks[r, 1]), "rank %d checksum differs from the reference's rank" % r
    if ref is not None:   # sharded find(): stored frequency of keys other ranks own, doubled by the second build
        order = np.lexsort(ref["keys"].reshape(len(ref["counts"]), -1).T[::-1]) if k > 32 else np.argsort(ref["keys"])
        for r in range(world):
            q = ou.canonical(np.load(tmp_path / ("q%d.npy" % r)), k)
            v = np.load(tmp_path / ("v%d.npy" % r))
            if k <= 32:
                sk = ref["keys"][order]
                pos = np.searchsorted(sk, q)
                assert np.array_equal(sk[pos], q) and np.array_equal(v, 2 * ref["counts"][order][pos].astype(np.uint64))


# ---- Index::build_posix(filename, comm): ONE file, every rank builds from its block (kmi_dist_build_file) -------------------
def _file_worker(rank, world, port, out_dir, path, k, fasta, stage_threads):
    _init(rank, world, port)
    os.environ["KMI_STAGE_THREADS"] = stage_threads
    try:
        import kmerind_b200 as kb
        from kmerind_b200.dist import ShardedIndex
        torch.cuda.set_device(rank)
        mode = kb.MODE_POSITION if fasta else kb.MODE_COUNT
        with ShardedIndex(k, mode=mode, hash=kb.HASH_FARM, device=rank) as idx:
            n = idx.build_file(path)
            if fasta:
                gk, gi = idx.local.sorted_dump()
                np.save(os.path.join(out_dir, "fk%d.npy" % rank), gk)
                np.save(os.path.join(out_dir, "fi%d.npy" % rank), gi)
                np.save(os.path.join(out_dir, "r%d.npy" % rank), np.array([n, idx.size()], dtype=np.uint64))
                # sharded find_positions (answered on the device, ids exchanged back, CSR in input order) from every rank:
                # 20000 k-mers of the file on the other strand + absent keys, against the oracle's tuples
                fa = np.fromfile(path, dtype=np.uint8)
                keys, ids, _ = ou.fasta_kmers(fa, k)
                sk, si = ou.sort_tuples(keys, ids, k)
                q = np.concatenate([ou.revcomp(keys[rank::37][:20000], k), ou.random_keys(k, 3000, 5 + rank)])
                off, vals = idx.find_positions(q)
                cq = ou.canonical(q, k)
                lo, hi = np.searchsorted(sk, cq, "left"), np.searchsorted(sk, cq, "right")
                assert np.array_equal(np.diff(off.astype(np.int64)), hi - lo) and int(off[-1]) == len(vals)
                want = np.concatenate([si[a:b] for a, b in zip(lo, hi)])
                grp = np.repeat(np.arange(len(q)), hi - lo)
                assert np.array_equal(vals[np.lexsort((vals, grp))], want)       # (ids of one key come in any order)
            else:
                ck = idx.local.checksum()
                np.save(os.path.join(out_dir, "r%d.npy" % rank), np.array([idx.local_size(), ck[0], ck[1], idx.size(), n, idx.path], dtype=np.uint64))
            with pytest.raises(kb.KmiError):                      # every rank gets the error, nobody waits in a collective
                idx.build_file(path + ".missing.fastq")
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("k,stage_threads", [(31, "4"), (63, "4"), (31, "0")])
def test_sharded_build_from_one_fastq_file(tmp_path, k, stage_threads):
    """every rank reads ITS record-aligned block of the one file (streamed through the pinned slice ring, or read whole when
    the ring is off) and the shards must be the REFERENCE's ranks at p = world over the same blocks"""
    import kmerind_b200 as kb
    world = max(torch.cuda.device_count(), 1)
    n_reads = 90000 * world
    whole = ou.synth_fastq(0, n_reads, 150, 5 * n_reads, 17, 5000, 0)
    path = tmp_path / "reads.fastq"
    path.write_bytes(whole.tobytes())
    mp.spawn(_file_worker, args=(world, _free_port(), str(tmp_path), str(path), k, False, stage_threads), nprocs=world, join=True)
    cuts = [kb.file_block(str(path), r, world)[0] for r in range(world)] + [whole.size]
    assert cuts == kb.fastq_cuts(whole, world)
    got = [np.load(tmp_path / ("r%d.npy" % r)) for r in range(world)]
    for r in range(world):
        assert int(got[r][4]) == (cuts[r + 1] - cuts[r]) // 317 * (150 - k + 1) and int(got[r][5]) == 1      # its block's k-mers, NVLink push path
    ref = ou.ref_count_build_fastq(whole, k, world, 1, cuts=cuts)
    if ref is None:
        keys, _, _ = ou.fastq_kmers(whole, k)
        owner = (ou.hashes(keys, k, 3) % np.uint64(world)).astype(np.int64)
        sizes, cks = [len(ou.count_sorted(shape_keys(keys, k)[owner == r], k)[0]) for r in range(world)], None
    else:
        sizes, cks = ref["sizes"].tolist(), ref["ck"]
    for r in range(world):
        assert int(got[r][0]) == int(sizes[r]) and int(got[r][3]) == int(sum(sizes))
        if cks is not None:
            assert int(got[r][1]) == int(cks[r, 0]) and int(got[r][2]) == int(cks[r, 1]), "rank %d differs from the reference's rank" % r


def shape_keys(keys, k):
    return ou.shape_keys(np.ascontiguousarray(keys).reshape(-1), k)


@pytest.mark.gpu
def test_sharded_position_index_from_one_fasta_file(tmp_path):
    """FASTA: rank r takes the byte range [size r / p, size (r + 1) / p) + K-1 halo of the one file; the union of the shards is
    the oracle's (k-mer, LongSequenceKmerId) list of the whole file"""
    world = max(torch.cuda.device_count(), 1)
    k = 21
    fa = ou.synth_fasta(5, 150000, 70, 3)
    path = tmp_path / "genome.fasta"
    path.write_bytes(fa.tobytes())
    mp.spawn(_file_worker, args=(world, _free_port(), str(tmp_path), str(path), k, True, "4"), nprocs=world, join=True)
    keys, ids, _ = ou.fasta_kmers(fa, k)
    sk, si = ou.sort_tuples(keys, ids, k)
    gk = np.concatenate([np.load(tmp_path / ("fk%d.npy" % r)) for r in range(world)])
    gi = np.concatenate([np.load(tmp_path / ("fi%d.npy" % r)) for r in range(world)])
    uk, ui = ou.sort_tuples(gk, gi, k)
    assert np.array_equal(uk, sk) and np.array_equal(ui, si)
    tot = sum(int(np.load(tmp_path / ("r%d.npy" % r))[0]) for r in range(world))
    assert tot == len(keys) and int(np.load(tmp_path / "r0.npy")[1]) == len(keys)


# ---- CPU, world_size 2: the blocks two ranks take from ONE file (kmi_file_block: host code, no GPU) tile it at record starts ---
def _gloo_file_block_worker(rank, world, port, out_dir, path):
    _init(rank, world, port)
    try:
        import kmerind_b200 as kb
        b, e, size = kb.file_block(path, rank, world)
        every = [None] * world
        dist.all_gather_object(every, (b, e, size))
        np.save(os.path.join(out_dir, "blocks%d.npy" % rank), np.array(every, dtype=np.int64))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ("natural.fastq", "test.medium.fastq"))
def test_file_blocks_gloo_world2(tmp_path, name):
    """what kmi_dist_build_file does first on every rank: rank r's block of the shared file. Both ranks must see the same
    partition, it must tile the file, and the inner boundary must be the record start the in-memory cut finds"""
    import kmerind_b200 as kb
    world = 2
    buf = ou.read_fixture(name)
    path = tmp_path / name
    path.write_bytes(buf.tobytes())
    mp.spawn(_gloo_file_block_worker, args=(world, _free_port(), str(tmp_path), str(path)), nprocs=world, join=True)
    a, b = (np.load(tmp_path / ("blocks%d.npy" % r)) for r in range(world))
    assert np.array_equal(a, b)
    cuts = kb.fastq_cuts(buf, world)
    assert a[:, :2].tolist() == [[cuts[0], cuts[1]], [cuts[1], cuts[2]]] and (a[:, 2] == buf.size).all()
