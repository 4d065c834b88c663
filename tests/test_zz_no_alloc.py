"""kmi_partition_noalloc_dev — the entry point that allocates nothing (scratch in a file-scope __device__ array of the
library, include/kmi.h) — against kmi_partition_dev, through the C ABI. No counterpart in the reference; the parity target
is this repository's own allocating entry point (bit-exact: integer work), with the oracle's owner arithmetic as a second
witness. Named test_zz_* so that pytest collects it last.

Two calls in a row on DIFFERENT inputs: the second call finds the first call's histogram in the shared scratch and must not
be affected by it. A torch-side check around the calls shows that the CUDA caching allocator saw no request and — through
cudaMemGetInfo — that the device's free memory did not move.
"""
import numpy as np
import pytest
import torch

import kmerind_b200 as kb
import oracle_util as ou

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()


def host_u64(t):
    return t.cpu().numpy().view(np.uint64)


def _both(idx, keys, vals, n, nw, p, bufs, check=True):
    """run the allocating entry point, then the static-scratch one, on the same device inputs; returns host copies"""
    dk, dv = dev(keys), dev(vals)
    ok_a, ov_a = torch.zeros_like(dk), torch.zeros_like(dv)
    counts_a = idx.partition_dev(dk.data_ptr(), dv.data_ptr(), n, ok_a.data_ptr(), ov_a.data_ptr(), p)
    torch.cuda.synchronize()
    ok_b, ov_b, cnt_b = bufs["k"][:n * nw], bufs["v"][:n], bufs["c"][:p]
    ok_b.zero_(); ov_b.zero_(); cnt_b.fill_(-1)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = torch.cuda.memory_stats()["allocation.all.allocated"]
    idx.partition_noalloc_dev(dk.data_ptr(), dv.data_ptr(), n, ok_b.data_ptr(), ov_b.data_ptr(), cnt_b.data_ptr())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    if check:   # (the warm-up call is exempt: the first launch of a kernel may make the driver load its module)
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == stats0
        assert free1 == free0, "device free memory moved across the no-allocation call"
    return counts_a, host_u64(ok_a), host_u64(ov_a), host_u64(cnt_b), host_u64(ok_b), host_u64(ov_b)


@pytest.mark.parametrize("k,p", [(31, 8), (63, 3), (21, 1)])
def test_no_alloc_partition_equals_partition_twice_in_a_row(k, p):
    nw = ou.nwords(k)
    sizes = (100003, 41017)                      # second call smaller: stale histogram words of the first stay behind it
    n_max = max(sizes)
    bufs = {"k": torch.zeros(n_max * nw, dtype=torch.int64, device="cuda"), "v": torch.zeros(n_max, dtype=torch.int64, device="cuda"),
            "c": torch.zeros(p, dtype=torch.int64, device="cuda")}
    with kb.Index(k=k, hash=kb.HASH_FARM, nranks=p) as idx:
        # warm both paths once so that first-use set-up (address look-up, the index's own scratch) is behind us
        w = np.ascontiguousarray(ou.canonical(ou.random_keys(k, 4099, seed=5), k)).reshape(-1)
        _both(idx, w, np.arange(4099, dtype=np.uint64), 4099, nw, p, bufs, check=False)
        for call, n in enumerate(sizes):
            keys = np.ascontiguousarray(ou.canonical(ou.random_keys(k, n, seed=101 + 7 * call + k), k)).reshape(-1)
            vals = np.arange(n, dtype=np.uint64) * np.uint64(3) + np.uint64(call)
            counts_a, ok_a, ov_a, counts_b, ok_b, ov_b = _both(idx, keys, vals, n, nw, p, bufs)
            assert counts_b.tolist() == counts_a.tolist(), "call %d" % call
            assert np.array_equal(ok_b, ok_a) and np.array_equal(ov_b, ov_a), "call %d" % call
            owner = (ou.hashes(keys, k, ou.hash_which(kb.HASH_FARM, True)) % np.uint64(p)).astype(np.int64)
            assert counts_b.tolist() == np.bincount(owner, minlength=p).tolist()
            assert np.array_equal(ov_b, vals[np.argsort(owner, kind="stable")])


def test_no_alloc_partition_limits_and_empty():
    p = 8
    lib = kb.load_library()
    assert lib.kmi_partition_noalloc_max_keys(p) == (1 << 20) // p * 2048
    assert lib.kmi_partition_noalloc_max_keys(0) == 0
    with kb.Index(k=31, hash=kb.HASH_MURMUR, nranks=p) as idx:
        cnt = torch.full((p,), 7, dtype=torch.int64, device="cuda")
        one = torch.zeros(16, dtype=torch.int64, device="cuda")
        idx.partition_noalloc_dev(None, None, 0, None, None, cnt.data_ptr())
        torch.cuda.synchronize()
        assert cnt.cpu().tolist() == [0] * p
        # one key beyond the limit: rejected with a clear message before anything is launched (pointers are never read)
        too_many = lib.kmi_partition_noalloc_max_keys(p) + 1
        cnt.fill_(7)
        with pytest.raises(kb.KmiError) as e:
            idx.partition_noalloc_dev(one.data_ptr(), None, too_many, one.data_ptr(), None, cnt.data_ptr())
        assert e.value.code == kb.ERR_ARG and "exceed the static scratch" in str(e.value)
        torch.cuda.synchronize()
        assert cnt.cpu().tolist() == [7] * p
