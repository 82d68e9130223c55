import sys,time,random,torch
sys.path.insert(0,'/root/repo/pytorch-dppo_b200'); sys.path.insert(0,'/root/repo')
import b200_native as nv
from b200_engine import UpdateEngine, RolloutBuffer
from model import Model, Shared_grad_buffers, Shared_obs_stats
from synthetic import make_rollout
torch.manual_seed(1)
m=Model(17,6); stats=Shared_obs_stats(17); sgb=Shared_grad_buffers(m)
eng=UpdateEngine(m, nv.FLAVOUR_DPPO, 4096, 10, 3e-4)
host,dev=make_rollout(m,stats,65536,seed=2,flavour=nv.FLAVOUR_DPPO)
dev2=RolloutBuffer(65536,17,6,dev.obs.device)
rng=random.Random(1)
eng.update_stream([host],[dev,dev2],3,obs_stats=stats,rng=rng,gshared=sgb.flat)
for t in (stats.n,stats.mean,stats.mean_diff,stats.var): t.zero_()
evs=[]
orig=eng.run_epochs
def wrapped(*a,**k):
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    # graph replay only: scalars staged before e0
    eng._stage_scalars_orig()
    e0.record(); r=orig(*a,**k); e1.record(); evs.append((e0,e1)); return r
eng._stage_scalars_orig=eng._stage_scalars
eng._stage_scalars=lambda: None
eng.run_epochs=wrapped
host_t=[]
def on_step(i): host_t.append(time.perf_counter())
t0=time.perf_counter()
eng.update_stream([host],[dev,dev2],40,obs_stats=stats,rng=rng,gshared=sgb.flat,on_step=on_step)
torch.cuda.synchronize()
dur=[a.elapsed_time(b) for a,b in evs]
gaps=[evs[i][1].elapsed_time(evs[i+1][0]) for i in range(len(evs)-1)]
print('graph dur avg %.3f ms; gap avg %.3f ms max %.3f; host per-iteration avg %.3f ms'%(sum(dur)/len(dur), sum(gaps)/len(gaps), max(gaps), (host_t[-1]-host_t[0])/(len(host_t)-1)*1e3))
print('gaps', ['%.2f'%g for g in gaps[:12]])
