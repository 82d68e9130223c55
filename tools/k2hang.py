import sys,torch
sys.path.insert(0,'/root/repo/pytorch-dppo_b200'); sys.path.insert(0,'/root/repo')
import b200_native as nv
from model import Model
O,A,B,T=int(sys.argv[1]),int(sys.argv[2]),int(sys.argv[3]),int(sys.argv[4])
torch.manual_seed(1)
m=Model(O,A)
obs=torch.randn(T,O,device='cuda');act=torch.randn(T,A,device='cuda');adv=torch.randn(T,device='cuda');ret=torch.randn(T,device='cuda')
idx=torch.randperm(T,device='cuda')[:B].int()
npart=nv.update_nparts(m.dims,B)
parts=torch.empty(npart,m.P_pad,device='cuda');lp=torch.empty(npart,4,device='cuda')
old=m.flat.clone()
nv.update_fused(0,m.dims,obs,act,adv,ret,None,idx,m.flat,old,0.2,0.0,parts,lp)
torch.cuda.synchronize()
print('ok', O,A,B,T, float(parts.sum()))
