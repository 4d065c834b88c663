"""One warm-up + one measured bucketed CountIndex build of the bench workload, for ncu:
    ncu --set full --clock-control none --import-source on -k regex:'k_fastq_kmers|k_scatter_fine|k_insert_segments' \
        --launch-skip 3 --launch-count 3 -o gpurun_out/build python tools/profile_build.py [n_reads]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmerind_b200 as kb  # noqa: E402

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 6666667
REC, KPR = 317, 120
nbytes = n_reads * REC
genome = int(33333333 * n_reads / 6666667)
text = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
kb.synth_fastq_dev(0, text.data_ptr(), 0, n_reads, 150, genome, 1, 5000, 0)
torch.cuda.synchronize()
idx = kb.CountIndex(31, hash=kb.HASH_FARM, reserve=int(0.2 * n_reads * KPR))
idx.set_build_path(2)
for _ in range(2):
    idx.clear()
    n = idx.build_fastq_dev(text.data_ptr(), nbytes)
    assert n == n_reads * KPR
print("kernel ms [build, parse+scatter, fine scatter, insert, overflow]:", idx.last_kernel_ms(), "distinct", idx.size())
