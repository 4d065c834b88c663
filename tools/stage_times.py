"""Times the stages of the count-index build separately on one GPU (CUDA events, resident inputs):
scan, parse->keys (sink 0), insert of the key stream, fused text->table. Usage:
    python tools/stage_times.py [n_reads] [reserve_frac]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmerind_b200 as kb  # noqa: E402

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 6666667
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
REC, KPR = 317, 120
nbytes = n_reads * REC
n_kmers = n_reads * KPR
genome = int(33333333 * n_reads / 6666667)
text = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
kb.synth_fastq_dev(0, text.data_ptr(), 0, n_reads, 150, genome, 1, 5000, 0)
keys = torch.empty(n_kmers, dtype=torch.int64, device="cuda")
torch.cuda.synchronize()


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


idx = kb.CountIndex(31, hash=kb.HASH_FARM, reserve=int(frac * n_kmers))
t_parse = timed(lambda: idx.parse_fastq_dev(text.data_ptr(), nbytes, keys.data_ptr(), n_kmers))


def ins():
    idx.clear()
    idx.insert_dev(keys.data_ptr(), n_kmers)


t_clear = timed(lambda: idx.clear())


def fused():
    idx.clear()
    idx.build_fastq_dev(text.data_ptr(), nbytes)


print("reads %d kmers %d reserve %d" % (n_reads, n_kmers, int(frac * n_kmers)))
print("parse->keys (incl. scan) %.3f ms | clear %.3f ms" % (t_parse, t_clear))
for path, name in ((1, "direct"), (3, "bucketed-L2"), (2, "bucketed-smem")):
    idx.set_build_path(min(path, 2))
    idx.set_segment_insert(path != 3)
    t_ins = timed(ins) - t_clear
    t_fused = timed(fused) - t_clear
    st = idx.last_stage_ms()
    info = idx.table_info()
    print("[%s] distinct %d slots %d | insert(keys) %.3f ms (%.2f G/s) | build %.3f ms (%.2f G/s): scan %.3f, parse(+scatter) %.3f, insert %.3f"
          % (name, idx.size(), info["slots"], t_ins, n_kmers / t_ins / 1e6, t_fused, n_kmers / t_fused / 1e6, st[0], st[1], st[2]))
