import os, sys, torch
sys.path.insert(0, os.getcwd())
import kmerind_b200 as kb
n_reads=4000000; nbytes=n_reads*317
text=torch.empty(nbytes+64,dtype=torch.uint8,device="cuda")
kb.synth_fastq_dev(0,text.data_ptr(),0,n_reads,150,n_reads*5,1,5000,0)
idx=kb.CountIndex(63,hash=kb.HASH_FARM,reserve=int(0.5*n_reads*88))
for _ in range(3):
    idx.clear(); n=idx.build_fastq_dev(text.data_ptr(),nbytes)
print("K=63 kmers",n,"kernel ms [build, parse+scatter, fine, insert, overflow]", idx.last_kernel_ms(), "distinct", idx.local_size(), idx.table_info())
