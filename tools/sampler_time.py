import sys,time,random,torch
sys.path.insert(0,'/root/repo/pytorch-dppo_b200')
import b200_native as nv
rng=random.Random(1); out=torch.empty(10,4096,dtype=torch.int32).pin_memory()
for _ in range(3): nv.sample_indices(65536,4096,10,rng=rng,out=out)
t=time.perf_counter()
for _ in range(50): nv.sample_indices(65536,4096,10,rng=rng,out=out)
print('sampler 10x4096 of 65536: %.3f ms per call'%((time.perf_counter()-t)/50*1e3))
