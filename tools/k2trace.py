"""Timeline of the wide-layer pipeline (V2 forward of the last K2 launch, CTA 0); needs a -DDPPO_K2_TRACE build."""
import sys,ctypes,torch
sys.path.insert(0,'/root/repo/pytorch-dppo_b200'); sys.path.insert(0,'/root/repo')
import b200_native as nv
from model import Model
torch.manual_seed(1)
m=Model(17,6); T=65536;B=4096
obs=torch.randn(T,17,device='cuda');act=torch.randn(T,6,device='cuda');adv=torch.randn(T,device='cuda');ret=torch.randn(T,device='cuda')
idx=torch.randperm(T,device='cuda')[:B].int()
npart=nv.update_nparts(m.dims,B)
parts=torch.empty(npart,m.P_pad,device='cuda');lp=torch.empty(npart,4,device='cuda')
old=m.flat.clone()
import os
if os.environ.get('K2_TRACE_PARAM'):          # name of the bias whose forward layer is traced, e.g. p_fc2.bias / v_fc2.bias
    off=m.offsets[nv.PARAM_NAMES.index(os.environ['K2_TRACE_PARAM'])]
    nv.lib.dppo_debug_k2_trace_bias(ctypes.c_void_p(m.flat.data_ptr()+4*off))
for _ in range(3):
    nv.update_fused(0,m.dims,obs,act,adv,ret,None,idx,m.flat,old,0.2,0.0,parts,lp)
torch.cuda.synchronize()
buf=(ctypes.c_longlong*(4*64*8))()
nv.lib.dppo_debug_k2_trace(buf)
def ev(role,gp,e): return buf[((role*64+gp)<<3)+e]
print('pair | A: start raw-seen split0 md-seen st-done arrived | B: start md-seen arrived | TMA: start free-seen issued | I: start af-seen committed [tile-done]')
for gp in range(int(os.environ.get('K2_TRACE_PAIRS','8'))):
    print('%4d | %6d %6d %6d %6d %6d %6d | %6d %6d %6d | %6d %6d %6d | %6d %6d %6d %6d'%((gp,)+tuple(ev(0,gp,e) for e in range(6))+tuple(ev(1,gp,e) for e in range(3))+tuple(ev(2,gp,e) for e in range(3))+tuple(ev(3,gp,e) for e in range(4))))
print('A detail (relative to start): raw-seen split0 md-seen st0-done split1-done st1-done arrived')
for gp in range(int(os.environ.get('K2_TRACE_PAIRS','8'))):
    b=ev(0,gp,0); print('%4d | %s'%(gp,' '.join('%6d'%(ev(0,gp,e)-b) for e in (1,2,3,6,7,4,5))))
print('epilogue: start %d end %d'%(ev(0,63,0),ev(0,63,1)))
