"""GPU parity of the PositionQualityIndex path: kmi_fastq_quality_dev / kmi_posqual_* against the oracle's
restatement of QualityScoreSlidingWindow, which tests/test_oracle.py pins bit for bit to the reference's own class
(tests/golden/ref_quality.json). This is the path's only floating-point kernel: the running float sum is replayed in the
reference's operation order, so zeros (windows with an uncallable base) must match exactly and every other score within
RTOL — the device's exp2f against glibc's."""
import numpy as np
import pytest

import kmerind_b200 as kb
import oracle_util as ou

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

RTOL = 1e-6


def check_scores(got, want):
    assert got.shape == want.shape
    assert np.array_equal(got == 0, want == 0)                   # bad windows are exactly 0 on both sides
    np.testing.assert_allclose(got, want, rtol=RTOL, atol=0)


def with_qualities(buf, seed):
    """fixture FASTQ with its quality lines replaced by seeded random Phred+33 bytes (some '!' = Phred 0 and '~' + 1 = out
    of range mixed in), same line lengths"""
    rng = np.random.default_rng(seed)
    out = buf.copy()
    i, n, line = 0, buf.size, 0
    while i < n:
        while i < n and buf[i] in (10, 13):
            i += 1
        j = i
        while j < n and buf[j] not in (10, 13):
            j += 1
        if j > i and line % 4 == 3:
            q = rng.integers(34, 75, j - i, dtype=np.uint8)
            bad = rng.random(j - i) < 0.01
            q[bad] = rng.choice(np.array([33, 127], dtype=np.uint8), int(bad.sum()))
            out[i:j] = q
        if j > i:
            line += 1
        i = j
    return out


def test_lut_equals_oracle_table():
    assert np.array_equal(kb.quality_lut().view(np.uint32), ou.quality_lut().view(np.uint32))


@pytest.mark.parametrize("k", (15, 21, 31, 63))
@pytest.mark.parametrize("name", ("natural.fastq", "natural.withN.fastq", "test.small.fastq", "test.unitiq1.short.fastq"))
def test_quality_stage_matches_oracle(name, k):
    base = ou.read_fixture(name)
    for variant, buf in (("file", base), ("random", with_qualities(base, k))):
        for nsplit in (False, True):
            keys, ids, _ = ou.fastq_kmers(buf, k, nsplit=nsplit, want_ids=True, file_offset=1 << 20)
            want = ou.fastq_quality(buf, k, nsplit=nsplit)
            assert len(want) == len(ids)
            if len(ids) == 0:
                continue
            by_id = dict(zip(ids.tolist(), want.tolist()))
            nw = ou.nwords(k)
            with kb.Index(k=k, mode=kb.MODE_POSITION, nsplit=nsplit) as idx:
                text = torch.from_numpy(buf).cuda()
                cap = buf.size
                d_keys = torch.zeros(cap * nw, dtype=torch.int64, device="cuda")
                d_ids = torch.zeros(cap, dtype=torch.int64, device="cuda")
                n = idx.parse_fastq_dev(text.data_ptr(), buf.size, d_keys.data_ptr(), cap, d_ids.data_ptr(), 1 << 20)
                assert n == len(ids)
                d_q = torch.zeros(n, dtype=torch.float32, device="cuda")
                idx.fastq_quality_dev(text.data_ptr(), buf.size, 1 << 20, d_ids.data_ptr(), n, d_q.data_ptr())
                torch.cuda.synchronize()
                g_ids = d_ids[:n].cpu().numpy().view(np.uint64)
                got = d_q.cpu().numpy()
                want_in_gpu_order = np.array([by_id[int(i)] for i in g_ids], dtype=np.float32)
                check_scores(got, want_in_gpu_order)


@pytest.mark.parametrize("k", (21, 35))
def test_position_quality_index(k):
    buf = with_qualities(ou.read_fixture("natural.fastq"), 5)
    keys, ids, _ = ou.fastq_kmers(buf, k, want_ids=True)
    want = ou.fastq_quality(buf, k)
    by_id = dict(zip(ids.tolist(), want.tolist()))
    key_of = dict(zip(ids.tolist(), [tuple(np.atleast_1d(x).tolist()) for x in keys]))
    with kb.PositionQualityIndex(k) as pq:
        assert pq.build_fastq(buf) == len(keys) and pq.local_size() == len(keys)
        gk, gi, gq = pq.to_vector()
        assert sorted(gi.tolist()) == sorted(ids.tolist())
        assert all(tuple(np.atleast_1d(a).tolist()) == key_of[int(i)] for a, i in zip(gk, gi))
        check_scores(gq, np.array([by_id[int(i)] for i in gi], dtype=np.float32))
        # find(): both strands of present k-mers and some absent ones; every hit comes with its score
        sel = keys[::37]
        q = np.concatenate([np.ascontiguousarray(ou.revcomp(np.ascontiguousarray(sel).reshape(-1), k)).reshape(-1),
                            np.ascontiguousarray(ou.random_keys(k, 50, seed=2)).reshape(-1)])
        off, fi, fq = pq.find(q)
        rows = {}
        for a, i in zip(keys, ids):
            rows.setdefault(tuple(np.atleast_1d(a).tolist()), []).append(int(i))
        for j, a in enumerate(sel):
            got_ids = sorted(fi[int(off[j]):int(off[j + 1])].tolist())
            assert got_ids == sorted(rows[tuple(np.atleast_1d(a).tolist())])
        assert int(off[-1]) == int(off[len(sel)])                    # the random keys are absent
        check_scores(fq, np.array([by_id[int(i)] for i in fi], dtype=np.float32))
        assert pq.launch_count() > 0
        pq.erase(sel[:5])
        assert np.all(pq.count(sel[:5]) == 0)
