"""kmi_hoststage.hpp: the slice ring that moves PAGEABLE host memory (the reference's std::vector<unsigned char> file data,
std::vector<Kmer> queries and answer vectors: kmer_index.hpp:132-225, file.hpp:160-221) to and from the device.

CPU part: tests/cpp/stage_check.cpp drives the template over a fake device whose copy stream runs LATER on a worker thread —
ring reuse, marks, empty ranges, a failing copy, an abandoned job — plain and under ThreadSanitizer.
GPU part: every host-pointer entry point with the ring forced on for small inputs (tiny slices, few threads, so slots are
reused many times) must give exactly what the device-resident / direct path gives."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "stage_check.cpp")


@pytest.mark.parametrize("flags", ([], ["-fsanitize=thread", "-g"]), ids=("plain", "tsan"))
def test_stage_threads_on_fake_device(tmp_path, flags):
    exe = str(tmp_path / "stage_check")
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-pthread"] + flags + [SRC, "-o", exe], capture_output=True, text=True)
    if r.returncode != 0 and flags:
        pytest.skip("no ThreadSanitizer runtime here: " + r.stderr[-200:])
    assert r.returncode == 0, r.stderr
    for _ in range(3):
        out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
        if flags and "FATAL: ThreadSanitizer" in out.stderr and "memory mapping" in out.stderr:
            pytest.skip("ThreadSanitizer cannot run under this kernel's address-space layout")
        assert out.returncode == 0 and "stage_check ok" in out.stdout, out.stdout + out.stderr


# ---------------------------------------------------------------------------------------------------------------- GPU
@pytest.fixture
def tiny_ring(monkeypatch):
    """64 KiB slices, 3 threads (9 slots), no size threshold: a few MB already wrap the ring many times"""
    monkeypatch.setenv("KMI_STAGE_SLICE_BYTES", "65536")
    monkeypatch.setenv("KMI_STAGE_THREADS", "3")
    monkeypatch.setenv("KMI_STAGE_MIN_BYTES", "0")


def _dev_build(kb, torch, text, k, **kw):
    with kb.CountIndex(k, hash=kb.HASH_FARM, **kw) as idx:
        t = torch.from_numpy(text).cuda()
        n = idx.build_fastq_dev(t.data_ptr(), text.size)
        return n, idx.local_size(), idx.checksum()


@pytest.mark.gpu
@pytest.mark.parametrize("k", (31, 63))
def test_pageable_fastq_build_equals_oracle(tiny_ring, k):
    torch = pytest.importorskip("torch")
    import kmerind_b200 as kb
    import oracle_util as ou
    text = ou.synth_fastq(0, 30000, 150, 700000, 4, 5000, 0)          # 9.5 MB: the bucketed build, one round
    keys, _, _ = ou.fastq_kmers(text, k)
    wk, wc = ou.count_sorted(keys, k)
    with kb.CountIndex(k, hash=kb.HASH_FARM) as idx:
        assert idx.build_fastq(text) == len(keys)                      # numpy memory is pageable
        gk, gc = idx.sorted_dump()                                     # (staged device -> host as well)
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
        ms = idx.last_kernel_ms()
        assert ms[2] > 0 and ms[3] > 0, "not the bucketed build"


@pytest.mark.gpu
@pytest.mark.parametrize("threads", ("0", "1", "5"))
def test_pageable_multi_round_build_equals_resident(monkeypatch, threads):
    """120 MB of reads: the host-pointer build runs in rounds (one mark per round); pageable (staged, or KMI_STAGE_THREADS=0:
    the driver's own path) and pinned sources must leave exactly the index the device-resident build leaves"""
    torch = pytest.importorskip("torch")
    import kmerind_b200 as kb
    monkeypatch.setenv("KMI_STAGE_THREADS", threads)
    monkeypatch.setenv("KMI_STAGE_SLICE_BYTES", str(1 << 20))
    monkeypatch.setenv("KMI_STAGE_MIN_BYTES", "0")
    reads, k = 380000, 31
    rec = kb.synth_record_bytes(150)
    dev = torch.empty(reads * rec, dtype=torch.uint8, device="cuda")
    kb.synth_fastq_dev(0, dev.data_ptr(), 0, reads, 150, 3_000_000, 11, 5000, 0)
    torch.cuda.synchronize()
    text = dev.cpu().numpy()
    want = _dev_build(kb, torch, text, k)
    assert want[0] == reads * (150 - k + 1)
    pinned = torch.empty(text.size, dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[:] = text
    for src in (text, pinned.numpy(), text):
        with kb.CountIndex(k, hash=kb.HASH_FARM) as idx:
            n = idx.build_fastq(src)
            assert (n, idx.local_size(), idx.checksum()) == want
            idx.clear()                                               # the ring is kept with the index: a second job on it
            n = idx.build_fastq(src)
            assert (n, idx.local_size(), idx.checksum()) == want


@pytest.mark.gpu
def test_pageable_queries_inserts_and_dumps(tiny_ring):
    torch = pytest.importorskip("torch")
    import kmerind_b200 as kb
    import oracle_util as ou
    k = 31
    rng = np.random.default_rng(3)
    keys = ou.random_keys(k, 400000, 21)
    keys = np.concatenate([keys, keys[:150000]])                      # some twice
    canon = ou.canonical(keys, k)
    uk, cnt = np.unique(canon, return_counts=True)
    with kb.CountIndex(k, hash=kb.HASH_MURMUR) as idx:
        idx.insert(keys)                                              # 4.4 MB of keys through the ring
        assert idx.local_size() == len(uk)
        q = np.concatenate([keys[rng.integers(0, len(keys), 300000)], ou.random_keys(k, 300000, 22)])
        cq = ou.canonical(q, k)
        pos = np.minimum(np.searchsorted(uk, cq), len(uk) - 1)
        hit = uk[pos] == cq
        assert np.array_equal(idx.count(q), hit.astype(np.uint64))
        vals, found = idx.find(q)
        assert np.array_equal(found.astype(bool), hit)
        assert np.array_equal(vals[hit], cnt[pos[hit]].astype(np.uint64))
        dk, dv = idx.to_vector()
        o = np.argsort(dk, kind="stable")
        assert np.array_equal(dk[o], uk) and np.array_equal(dv[o], cnt.astype(np.uint64))
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gk, uk) and np.array_equal(gc, cnt.astype(np.uint64))
        # (key, count) tuples: the counts travel through the ring too
        idx.insert(uk, np.full(len(uk), 5, dtype=np.uint64))
        gk, gc = idx.sorted_dump()
        assert np.array_equal(gc, cnt.astype(np.uint64) + 5)


@pytest.mark.gpu
def test_pageable_fasta_positions(tiny_ring):
    torch = pytest.importorskip("torch")
    import kmerind_b200 as kb
    import oracle_util as ou
    k = 21
    fa = ou.synth_fasta(3, 400000, 80, 5)
    keys, ids, _ = ou.fasta_kmers(fa, k)
    sk, si = ou.sort_tuples(keys, ids, k)
    with kb.PositionIndex(k, hash=kb.HASH_MURMUR) as idx:
        assert idx.build_fasta(fa) == len(keys)
        gk, gi = idx.sorted_dump()
        assert np.array_equal(gk, sk) and np.array_equal(gi, si)
        q = keys[::7]
        off, vals = idx.find_positions(q)
        uk, first, cnt = np.unique(sk, return_index=True, return_counts=True)
        p = np.searchsorted(uk, ou.canonical(q, k))
        assert np.array_equal(np.diff(off), cnt[p].astype(np.uint64))
        assert np.array_equal(np.sort(vals[off[0]:off[1]]), si[first[p[0]]:first[p[0]] + cnt[p[0]]])


@pytest.mark.gpu
@pytest.mark.parametrize("chunk", (12 << 20, 1 << 32))
def test_file_streamed_through_the_ring_equals_resident(monkeypatch, tmp_path, chunk):
    """kmi_build_file on a 48 MB FASTQ file: the ring's threads pread the file's slices (no host copy of the file); with
    12 MiB chunks the file is cut at record boundaries found in each chunk's tail and every chunk is a bucketed build"""
    torch = pytest.importorskip("torch")
    import kmerind_b200 as kb
    monkeypatch.setenv("KMI_FILE_CHUNK_BYTES", str(chunk))
    monkeypatch.setenv("KMI_STAGE_SLICE_BYTES", str(1 << 20))
    reads, k = 150000, 31
    rec = kb.synth_record_bytes(150)
    dev = torch.empty(reads * rec, dtype=torch.uint8, device="cuda")
    kb.synth_fastq_dev(0, dev.data_ptr(), 0, reads, 150, 2_000_000, 13, 5000, 0)
    torch.cuda.synchronize()
    text = dev.cpu().numpy()
    path = tmp_path / "reads.fastq"
    path.write_bytes(text.tobytes())
    want = _dev_build(kb, torch, text, k)
    for threads in ("4", "0"):
        monkeypatch.setenv("KMI_STAGE_THREADS", threads)
        with kb.CountIndex(k, hash=kb.HASH_FARM) as idx:
            n = idx.build_file(str(path))
            assert (n, idx.local_size(), idx.checksum()) == want
