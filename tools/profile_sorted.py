"""The alternative back ends on the bench workload's shape (device-resident text): the sort-based counting map
(kmi_sorted_*: parse -> radix sort -> run-length reduction) and the PositionQualityIndex build (parse with ids -> quality
scores -> position insert), next to the hash CountIndex. python tools/profile_sorted.py [n_reads]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmerind_b200 as kb  # noqa: E402

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 1666667          # 0.25 Gbp by default
REC, KPR = 317, 120
nbytes = n_reads * REC
genome = int(33333333 * n_reads / 6666667)
text = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
kb.synth_fastq_dev(0, text.data_ptr(), 0, n_reads, 150, genome, 1, 5000, 0)
torch.cuda.synchronize()
with kb.CountIndex(31, hash=kb.HASH_FARM, reserve=int(0.2 * n_reads * KPR)) as idx:
    for _ in range(2):
        idx.clear()
        idx.build_fastq_dev(text.data_ptr(), nbytes)
    print("hash CountIndex build ms", idx.last_kernel_ms()[0], "distinct", idx.size())
    want = idx.size()
for _ in range(2):
    with kb.SortedMap(31, hash=kb.HASH_FARM) as sm:
        n = sm.build_fastq_dev(text.data_ptr(), nbytes)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sm.rehash()
        t1 = time.perf_counter()
        ms, launches = sm.stats()
        size = sm.local_size()
print("sorted counting map: %d k-mers, parse %.2f ms, sort %.2f ms, reduce %.2f ms (rehash wall %.2f ms), distinct %d (hash index: %d), %d launches"
      % (n, ms[0], ms[1], ms[2], (t1 - t0) * 1e3, size, want, launches))
assert size == want
import numpy as np
host = text[:nbytes].cpu().numpy()
with kb.PositionQualityIndex(21) as pq:
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = pq.build_fastq(host[: min(nbytes, 200000 * REC)])
    t1 = time.perf_counter()
    print("PositionQualityIndex K=21: %d tuples from %d reads in %.1f ms (host bytes in)" % (n, min(n_reads, 200000), (t1 - t0) * 1e3))
