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
for _ in range(3):
    nv.update_fused(0,m.dims,obs,act,adv,ret,None,idx,m.flat,old,0.2,0.0,parts,lp)
torch.cuda.synchronize()
buf=(ctypes.c_longlong*64)()
nv.lib.dppo_debug_k2_clocks(buf)
names={0:'heads loaded',1:'tile start',2:'zero+gather',3:'old W1',4:'old V1',5:'old W2',6:'old V2',7:'old heads',8:'new W1',9:'new V1',10:'new W2',11:'new V2',12:'new heads',13:'loss',14:'dls',15:'dW heads',16:'hdx mu',17:'hdx v',18:'dW2',19:'dV2',20:'dx W2',21:'dx V2',22:'dW1',23:'dV1',24:'-',25:'end tiles'}
prev=buf[0]
for i in range(1,26):
    print('%-14s %8d cycles'%(names[i],buf[i]-prev)); prev=buf[i]
print('total',buf[25]-buf[0])
for i,nm in ((40,'dW ts: B stage'),(41,'dW ts: A transform'),(42,'dW ts: block barrier'),(43,'dW ts: MMAs (wait)'),(44,'dW ts: epilogue'),(45,'dW ts: end-of-tile barrier'),(46,'dW ts: db')):
    print('acc %-26s per tile (3 launches x 5 tiles): thread 0 %7.1f   thread 224 (issuer warp) %7.1f'%(nm,buf[i]/15.0,buf[i+8]/15.0))
