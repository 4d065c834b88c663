import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
import kmerind_b200 as kb
rng=np.random.default_rng(4); per=50000000
parts=[]
for s in range(4):
    seq=np.frombuffer(b"ACGT",dtype=np.uint8)[rng.integers(0,4,size=per,dtype=np.uint8)]
    lines=np.full((per//80,81),10,dtype=np.uint8); lines[:,:80]=seq.reshape(-1,80)
    parts += [np.frombuffer(b">chr%02d\n"%s,dtype=np.uint8), lines.reshape(-1)]
buf=np.concatenate(parts)
text=torch.empty(buf.size+64,dtype=torch.uint8,device="cuda"); text[:buf.size]=torch.from_numpy(buf).cuda()
idx=kb.PositionIndex(21,hash=kb.HASH_FARM)
for _ in range(3):
    idx.clear(); torch.cuda.synchronize()
    e=[torch.cuda.Event(enable_timing=True) for _ in range(2)]
    e[0].record(); n=idx.build_fasta_dev(text.data_ptr(),buf.size); e[1].record(); torch.cuda.synchronize()
    print("build ms", e[0].elapsed_time(e[1]), "stage ms [scan, parse, -]", idx.last_stage_ms(), "kernel", idx.last_kernel_ms())
