"""Sharded count build with ONE rank (k_owner_push degenerates to local stores but runs the same instructions), for ncu:
    ncu --set full --import-source on -k regex:k_owner_push --launch-skip 8 --launch-count 1 -o gpurun_out/push python tools/profile_dist1.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import kmerind_b200 as kb
from kmerind_b200.dist import ShardedIndex, Comm
from kmerind_b200 import load_library, _hptr, _check
buf = np.zeros(128, dtype=np.uint8)
_check(load_library().kmi_comm_unique_id(_hptr(buf)))
comm = Comm(buf.tobytes(), 1, 0, 0)
n_reads = 6666667
text = torch.empty(n_reads * 317 + 64, dtype=torch.uint8, device="cuda")
kb.synth_fastq_dev(0, text.data_ptr(), 0, n_reads, 150, 33333333, 1, 5000, 0)
idx = ShardedIndex(31, hash=kb.HASH_FARM, device=0, reserve=int(0.2 * n_reads * 120), comm=comm)
for _ in range(2):
    idx.local.clear()
    n = idx.build_fastq_dev(text.data_ptr(), n_reads * 317)
print("stage ms", idx.stage_ms, "kernel ms", idx.local.last_kernel_ms(), "path", idx.path)
