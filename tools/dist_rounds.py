"""Sharded count build at N GPUs, timed per stage for several round sizes (KMI_DIST_ROUND_BYTES): one huge round shows
parse and push back to back (no overlap), smaller rounds show how much of the push hides under the next parse.
torchrun --nproc-per-node N tools/dist_rounds.py"""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import kmerind_b200 as kb
from kmerind_b200.dist import ShardedIndex

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("gloo")
n_reads = 6666667
nbytes = n_reads * 317
text = torch.empty(nbytes + 64, dtype=torch.uint8, device="cuda")
kb.synth_fastq_dev(lr, text.data_ptr(), rank * n_reads, n_reads, 150, 33333333 * world, 1, 5000, 0)
torch.cuda.synchronize()
idx = ShardedIndex(31, hash=kb.HASH_FARM, device=lr, reserve=int(0.2 * n_reads * 120))
for rb in [1 << 40, 1 << 30, 512 << 20, 256 << 20, 128 << 20, 64 << 20]:
    os.environ["KMI_DIST_ROUND_BYTES"] = str(rb)
    rows = []
    for it in range(4):
        idx.local.clear()
        dist.barrier()
        n = idx.build_fastq_dev(text.data_ptr(), nbytes)
        if it:
            rows.append([idx.stage_ms["whole"], idx.stage_ms.get("parse+push", 0), idx.stage_ms["insert"]] + list(idx.local.last_kernel_ms()))
    if rank == 0:
        m = np.mean(np.array(rows), axis=0)
        print(json.dumps({"round_bytes": rb, "rounds": -(-nbytes // rb), "whole_ms": m[0], "parse_push_ms": m[1], "insert_ms": m[2],
                          "fine_scatter_ms": m[5], "segment_insert_ms": m[6], "overflow_ms": m[7]}), flush=True)
idx.close()
dist.destroy_process_group()
