"""cfg 5's query batch on one GPU (PositionIndex K=21 over a synthetic FASTA genome, 1.25e8 count() queries resident in HBM),
for a launch list:  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/q.csv python tools/profile_query.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmerind_b200 as kb  # noqa: E402

n_seq, seq_len, k = 3, 125_000_000, 21
n_q = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000_000
fbytes = kb.synth_fasta_bytes(n_seq, seq_len, 80)
text = torch.empty(fbytes + 64, dtype=torch.uint8, device="cuda")
kb.synth_fasta_dev(0, text.data_ptr(), 0, fbytes, n_seq, seq_len, 80, 1)
idx = kb.PositionIndex(k, hash=kb.HASH_MURMUR)
idx.set_max_load(int(os.environ.get("KMI_MAX_LOAD", "70")))
idx.reserve(n_seq * seq_len)
print("table", idx.table_info())
idx.build_fasta_dev(text.data_ptr(), fbytes)
q = torch.empty(n_q, dtype=torch.int64, device="cuda")
c = torch.zeros(n_q, dtype=torch.int64, device="cuda")
kb.synth_queries_dev(0, q.data_ptr(), 0, n_q, k, n_seq, seq_len, 1, 4)
torch.cuda.synchronize()
for _ in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    idx.count_dev(q.data_ptr(), n_q, c.data_ptr())
    e1.record()
    torch.cuda.synchronize()
    print("count_dev ms", e0.elapsed_time(e1), "hits", int((c > 0).sum()))
