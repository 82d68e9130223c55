import sys,time,random,torch,collections
sys.path.insert(0,'/root/repo/pytorch-dppo_b200'); sys.path.insert(0,'/root/repo')
import b200_native as nv
import b200_engine
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
acc=collections.defaultdict(float)
def timed(name, fn):
    def w(*a,**k):
        t=time.perf_counter(); r=fn(*a,**k); acc[name]+=time.perf_counter()-t; return r
    return w
eng._stage_scalars=timed('stage_scalars', eng._stage_scalars)
torch.cuda.Event.synchronize=timed('event.synchronize', torch.cuda.Event.synchronize)
RolloutBuffer.copy_from=timed('copy_from', RolloutBuffer.copy_from)
import concurrent.futures
concurrent.futures.Future.result=timed('fut.result', concurrent.futures.Future.result)
torch.cuda.CUDAGraph.replay=timed('graph.replay', torch.cuda.CUDAGraph.replay)
torch.Tensor.copy_=timed('tensor.copy_', torch.Tensor.copy_)
n=40
t0=time.perf_counter()
eng.update_stream([host],[dev,dev2],n,obs_stats=stats,rng=rng,gshared=sgb.flat)
tot=time.perf_counter()-t0
print('total per step %.3f ms'%(tot/n*1e3))
for k,v in sorted(acc.items(), key=lambda x:-x[1]): print('  %-20s %.3f ms/step'%(k, v/n*1e3))
