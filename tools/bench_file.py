"""End-to-end build from a FILE (kmi_build_file): the synthetic 1 Gbp FASTQ of the bench workload is written to a scratch
directory (so it sits in the page cache) and the CountIndex is built from its path: pread into two pinned buffers by reader
threads, H2D copy and build pipelined behind them. Wall clock around the call (it is blocking), 3 repetitions.
    python tools/bench_file.py [n_reads] [dir]
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmerind_b200 as kb  # noqa: E402


def main():
    n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 6666667
    d = sys.argv[2] if len(sys.argv) > 2 else "/tmp"
    k, rec = 31, kb.synth_record_bytes(150)
    nbytes = n_reads * rec
    text = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
    kb.synth_fastq_dev(0, text.data_ptr(), 0, n_reads, 150, max(n_reads * 5, 1000), 1, 5000, 0)
    torch.cuda.synchronize()
    path = os.path.join(d, "kmi_bench_%d.fastq" % n_reads)
    with open(path, "wb") as f:
        f.write(text[:nbytes].cpu().numpy().tobytes())
    del text
    n_kmers = n_reads * (150 - k + 1)
    try:
        # stage threads 0 = the chunked reader (whole chunk into a pinned buffer, then its build); > 0 = slices pread straight
        # into the pinned ring by that many threads while the build runs
        for threads, chunk in ((0, 1 << 30), (4, 4 << 30), (8, 4 << 30), (8, 1 << 30), (12, 4 << 30), (16, 4 << 30)):
            os.environ["KMI_STAGE_THREADS"] = str(threads)
            os.environ["KMI_FILE_CHUNK_BYTES"] = str(chunk)
            with kb.CountIndex(k, hash=kb.HASH_FARM, reserve=int(0.25 * n_kmers)) as idx:
                best = None
                for rep in range(4):
                    idx.clear()
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    n = idx.build_file(path)
                    dt = (time.perf_counter() - t0) * 1e3
                    assert n == n_kmers and idx.checksum()[1] == n
                    if rep:
                        best = dt if best is None else min(best, dt)
                print("build_file, %.2f GB FASTQ, stage threads %d, chunk %d MiB: %.1f ms = %.2f G k-mers/s (%.1f GB/s of text)"
                      % (nbytes / 1e9, threads, chunk >> 20, best, n_kmers / best / 1e6, nbytes / best / 1e6), flush=True)
    finally:
        os.remove(path)


if __name__ == "__main__":
    main()
