"""Single-GPU measurements of the two BASELINE.json configurations that are parity cases, not bench lines:
  cfg 4 shape: CountIndex K=63 (two-word k-mers) on synthetic 150 bp reads, device-resident text
  cfg 5 shape: PositionIndex K=21 on a synthetic FASTA genome (80 bases per line) + batched count()/find() of
               query k-mers (half sampled from the genome, half random), device-resident keys
CUDA events around the calls, inputs resident in HBM; every result is checked by a size-independent property
(sum of counts == k-mers generated; every sampled query is found, its count >= 1).
    python tools/bench_configs.py [reads_k63] [genome_bp] [n_queries]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmerind_b200 as kb  # noqa: E402


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def k63(n_reads):
    rec, kpr = 317, 150 - 63 + 1
    nbytes = n_reads * rec
    text = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
    kb.synth_fastq_dev(0, text.data_ptr(), 0, n_reads, 150, max(n_reads * 5, 1000), 1, 5000, 0)
    torch.cuda.synchronize()
    idx = kb.CountIndex(63, hash=kb.HASH_FARM, reserve=int(0.5 * n_reads * kpr))
    n = [0]

    def build():
        idx.clear()
        n[0] = idx.build_fastq_dev(text.data_ptr(), nbytes)
    ms = timed(build)
    assert n[0] == n_reads * kpr
    assert idx.checksum()[1] == n[0]
    print("cfg4 shape: CountIndex K=63, %d reads (%.2f Gbp), %d k-mers, %d distinct: build %.2f ms = %.2f G k-mers/s"
          % (n_reads, n_reads * 150 / 1e9, n[0], idx.local_size(), ms, n[0] / ms / 1e6))
    idx.close()


def fasta_positions(genome_bp, n_q):
    rng = np.random.default_rng(4)
    n_seq = 4
    per = genome_bp // n_seq
    parts = []
    for s in range(n_seq):
        bases = rng.integers(0, 4, size=per, dtype=np.uint8)
        seq = np.frombuffer(b"ACGT", dtype=np.uint8)[bases]
        lines = np.full((per // 80, 81), ord("\n"), dtype=np.uint8)
        lines[:, :80] = seq[:per // 80 * 80].reshape(-1, 80)
        parts.append(np.frombuffer(b">chr%02d\n" % s, dtype=np.uint8))
        parts.append(lines.reshape(-1))
    buf = np.concatenate(parts)
    text = torch.empty(buf.size + 64, dtype=torch.uint8, device="cuda")
    text[:buf.size] = torch.from_numpy(buf).cuda()
    k = 21
    n_kmers = n_seq * (per // 80 * 80 - k + 1)
    idx = kb.PositionIndex(k, hash=kb.HASH_FARM)
    n = [0]

    def build():
        idx.clear()
        n[0] = idx.build_fasta_dev(text.data_ptr(), buf.size)
    ms = timed(build, reps=2)
    assert n[0] == n_kmers, (n[0], n_kmers)
    print("cfg5 shape: PositionIndex K=21, FASTA genome %.0f Mbp in %d sequences (%d k-mers, %d distinct keys): build %.2f ms = %.2f G k-mers/s"
          % (genome_bp / 1e6, n_seq, n[0], idx.num_keys(), ms, n[0] / ms / 1e6))
    host = torch.from_numpy(buf).pin_memory()
    hbuf = host.numpy()

    def build_host():
        idx.clear()
        n[0] = idx.build_fasta(hbuf)
    ms_h = timed(build_host, reps=2)
    assert n[0] == n_kmers
    print("            from pinned host memory (chunks streamed under the build): %.2f ms = %.2f G k-mers/s" % (ms_h, n[0] / ms_h / 1e6))
    idx.clear()
    idx.build_fasta_dev(text.data_ptr(), buf.size)
    # queries: half = k-mers that exist (parsed from the text itself), half random
    cap = n_kmers + 1024
    keys = torch.empty(cap, dtype=torch.int64, device="cuda")
    with kb.Index(k=k, hash=kb.HASH_FARM) as p:
        m = p.parse_fasta_dev(text.data_ptr(), buf.size, keys.data_ptr(), cap)
    half = n_q // 2
    sel = torch.randint(0, m, (half,), device="cuda")
    rnd = torch.randint(0, 1 << 42, (n_q - half,), dtype=torch.int64, device="cuda")
    q = torch.cat([keys[:m][sel], rnd]).contiguous()
    counts = torch.empty(n_q, dtype=torch.int64, device="cuda")
    vals = torch.empty(n_q, dtype=torch.int64, device="cuda")
    found = torch.empty(n_q, dtype=torch.uint8, device="cuda")
    ms_c = timed(lambda: idx.count_dev(q.data_ptr(), n_q, counts.data_ptr()))
    ms_f = timed(lambda: idx.find_dev(q.data_ptr(), n_q, vals.data_ptr(), found.data_ptr()))
    assert int((counts[:half] >= 1).all().item()) == 1 and int(found[:half].all().item()) == 1
    print("            %d queries (half present): count() %.2f ms = %.2f G queries/s (36 B/query -> %.0f GB/s), find() %.2f ms = %.2f G queries/s"
          % (n_q, ms_c, n_q / ms_c / 1e6, 36.0 * n_q / ms_c / 1e6, ms_f, n_q / ms_f / 1e6))
    idx.close()


if __name__ == "__main__":
    k63(int(sys.argv[1]) if len(sys.argv) > 1 else 2000000)
    fasta_positions(int(sys.argv[2]) if len(sys.argv) > 2 else 200000000, int(sys.argv[3]) if len(sys.argv) > 3 else 100000000)
