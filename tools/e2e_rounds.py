import os, sys, time, torch, numpy as np
sys.path.insert(0, os.getcwd())
import kmerind_b200 as kb
n_reads=6666667; REC=317; nbytes=n_reads*REC
text=torch.empty(nbytes+64,dtype=torch.uint8,device="cuda")
kb.synth_fastq_dev(0,text.data_ptr(),0,n_reads,150,33333333,1,5000,0)
host=torch.empty(nbytes,dtype=torch.uint8,pin_memory=True); host.copy_(text[:nbytes]); torch.cuda.synchronize()
h=host.numpy()
idx=kb.CountIndex(31,hash=kb.HASH_FARM,reserve=int(0.2*n_reads*120))
idx.clear(); idx.build_fastq(h); torch.cuda.synchronize()
ts=[]
for _ in range(4):
    t0=time.perf_counter(); idx.clear(); n=idx.build_fastq(h); sz=idx.size(); torch.cuda.synchronize(); ts.append((time.perf_counter()-t0)*1e3)
# pure copy time
d=torch.empty(nbytes,dtype=torch.uint8,device="cuda")
torch.cuda.synchronize(); t0=time.perf_counter(); d.copy_(host,non_blocking=True); torch.cuda.synchronize(); tc=(time.perf_counter()-t0)*1e3
print(os.environ.get("KMI_E2E_ROUNDS"), "e2e ms", ["%.2f"%t for t in ts], "copy ms %.2f"%tc)
