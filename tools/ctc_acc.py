import os, sys
sys.path.insert(0, "/root/repo" if os.path.exists("/root/repo/tests") else ".")
sys.path.insert(0, ".")
import numpy as np, torch
from deepspeech_b200 import ops
from deepspeech_b200._lib import lib
def case(T,B,L,ragged,seed):
    rng=np.random.default_rng(seed)
    x=rng.standard_normal((T,B,29)).astype(np.float32)*2
    labs=[]
    for b in range(B):
        r=rng.integers(0,28,size=int(rng.integers(L//2,L+1)) if ragged else L)
        labs.append([int(v) for i,v in enumerate(r) if i==0 or v!=r[i-1]])
    Lm=max(len(r) for r in labs); lab=np.zeros((B,Lm),np.int32)
    for b,r in enumerate(labs): lab[b,:len(r)]=r
    ll=np.asarray([len(r) for r in labs],np.int32)
    sl=(rng.integers(T//2+Lm,T+1,size=B) if ragged else np.full(B,T)).astype(np.int32)
    return x,lab,ll,sl,labs
for (T,B,L,rg) in [(200,16,50,False),(800,16,200,True),(1600,8,400,True)]:
    x,lab,ll,sl,labs=case(T,B,L,rg,5)
    x64=torch.tensor(x,dtype=torch.float64,requires_grad=True)
    ref=torch.nn.functional.ctc_loss(torch.log_softmax(x64,2),torch.tensor(sum(labs,[]),dtype=torch.long),torch.tensor(sl,dtype=torch.long),torch.tensor(ll,dtype=torch.long),blank=28,reduction="none")
    ref.sum().backward()
    for v1 in (0,1):
        if v1: os.environ["DS2_CTC_V1"]="1"
        else: os.environ.pop("DS2_CTC_V1",None)
        lib.ds2_debug_reload_env()
        xt=torch.tensor(x,device="cuda",requires_grad=True)
        loss=ops.ctc_loss_dense(xt,lab,ll,np.zeros(B,np.int32),sl); loss.sum().backward()
        g=xt.grad.cpu().numpy()
        el=float(((loss.detach().cpu().double()-ref.detach()).abs()/ref.detach().abs()).max())
        eg=np.abs(g-x64.grad.numpy()).max()/np.abs(x64.grad.numpy()).max()
        print("T=%d L=%d %s: loss rel %.2e grad %.2e rowsum %.2e"%(T,L,"v1" if v1 else "v2",el,eg,np.abs(g.sum(2)).max()))
