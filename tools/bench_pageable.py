"""What the reference's caller sees: input and output in ORDINARY (pageable) host memory. The 1 Gbp bench text is built into a
CountIndex through kmi_build_fastq from a numpy array, and a batch of count() queries is answered from / into numpy arrays,
with the pinned slice ring of kmi_hoststage.hpp at several thread counts / slice sizes and without it (threads 0 = the
driver's own pageable copies, what every call did before). Wall clock around the blocking calls, best of 3 after a warm-up.
    python tools/bench_pageable.py [n_reads] [n_queries]
One JSON line per configuration."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmerind_b200 as kb  # noqa: E402


def best_of(f, reps=3):
    f()
    best = None
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        f()
        dt = (time.perf_counter() - t0) * 1e3
        best = dt if best is None else min(best, dt)
    return best


def main():
    n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 6666667
    n_q = int(sys.argv[2]) if len(sys.argv) > 2 else 125_000_000
    k, rec = 31, kb.synth_record_bytes(150)
    nbytes = n_reads * rec
    dev = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
    kb.synth_fastq_dev(0, dev.data_ptr(), 0, n_reads, 150, max(n_reads * 5, 1000), 1, 5000, 0)
    torch.cuda.synchronize()
    text = dev[:nbytes].cpu().numpy()          # pageable
    pinned = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[:] = text
    del dev
    n_kmers = n_reads * (150 - k + 1)
    rng = np.random.default_rng(1)
    queries = rng.integers(0, 1 << 62, size=n_q, dtype=np.uint64)
    configs = [("pinned", None, None), ("pageable", 0, 4 << 20)]
    for thr in (2, 4, 8, 12, 16):
        configs.append(("pageable", thr, 4 << 20))
    for sl in (1 << 20, 16 << 20):
        configs.append(("pageable", 8, sl))
    want = None
    for kind, thr, sl in configs:
        if thr is not None:
            os.environ["KMI_STAGE_THREADS"] = str(thr)
            os.environ["KMI_STAGE_SLICE_BYTES"] = str(sl)
        src = pinned.numpy() if kind == "pinned" else text
        with kb.CountIndex(k, hash=kb.HASH_FARM, reserve=int(0.25 * n_kmers)) as idx:
            def build():
                idx.clear()
                assert idx.build_fastq(src) == n_kmers
                idx.local_size()
            ms = best_of(build)
            got = (idx.local_size(), idx.checksum())
            want = want or got
            assert got == want, (got, want)
            out = {"source": kind, "stage_threads": thr, "slice_bytes": sl, "build_ms": ms, "build_gkmers_s": n_kmers / ms / 1e6,
                   "text_gb_s": nbytes / ms / 1e6}
            if kind == "pageable" and thr in (0, 8) and sl == 4 << 20 and n_q:
                counts = np.ones(n_q, dtype=np.uint64)      # touched: the answer vector of a caller that reuses its buffers

                def count():
                    kb._check(idx._lib.kmi_count(idx._h, kb._hptr(queries), n_q, kb._hptr(counts)))
                qms = best_of(count, reps=2)
                out.update({"count_ms": qms, "count_gq_s": n_q / qms / 1e6, "count_bytes_gb_s": 16 * n_q / qms / 1e6})
            print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
