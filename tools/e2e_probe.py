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
def run(n, label):
    eng.update_stream([host],[dev,dev2],3,obs_stats=stats,rng=rng,gshared=sgb.flat)
    for t in (stats.n,stats.mean,stats.mean_diff,stats.var): t.zero_()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    t0=time.perf_counter(); e0.record()
    eng.update_stream([host],[dev,dev2],n,obs_stats=stats,rng=rng,gshared=sgb.flat)
    e1.record(); torch.cuda.synchronize()
    print('%-28s %.3f ms/step (events)  %.3f ms/step (wall)'%(label, e0.elapsed_time(e1)/n, (time.perf_counter()-t0)/n*1e3))
run(50,'full')
orig=RolloutBuffer.copy_from
RolloutBuffer.copy_from=lambda self,other,fields,non_blocking=True: 0
run(50,'no rollout upload')
RolloutBuffer.copy_from=orig
orig_s=nv.sample_indices
nv.sample_indices=lambda *a,**k: None
run(50,'no sampling')
nv.sample_indices=orig_s
# pure graph replays back to back
for t in (stats.n,stats.mean,stats.mean_diff,stats.var): t.zero_()
eng.idx_dev.copy_(eng.idx_host)
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(50):
    eng.run_epochs(dev,gshared=sgb.flat,use_graph=True,indices_staged=True,obs_stats=stats,with_gae=True)
e1.record(); torch.cuda.synchronize()
print('%-28s %.3f ms/step'%('graph replays only', e0.elapsed_time(e1)/50))
# interference test: graph replays on the main stream while a side stream uploads the rollout continuously
cs=torch.cuda.Stream()
stop=False
fields=["act","rew","val","boot","seg_end","raw_obs"]
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(50):
    with torch.cuda.stream(cs):
        dev2.copy_from(host, fields)
    eng.run_epochs(dev,gshared=sgb.flat,use_graph=True,indices_staged=True,obs_stats=stats,with_gae=True)
e1.record(); torch.cuda.synchronize()
print('%-28s %.3f ms/step'%('replays + free-running upload', e0.elapsed_time(e1)/50))
e0.record()
for i in range(50):
    with torch.cuda.stream(cs):
        dev2.raw_obs.copy_(host.raw_obs, non_blocking=True)
    eng.run_epochs(dev,gshared=sgb.flat,use_graph=True,indices_staged=True,obs_stats=stats,with_gae=True)
e1.record(); torch.cuda.synchronize()
print('%-28s %.3f ms/step'%('replays + raw_obs upload only', e0.elapsed_time(e1)/50))
