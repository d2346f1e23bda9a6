import sys, os, statistics
ROOT="/root/repo"
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")): sys.path.insert(0,p)
import torch
from b200attn import ops
D,PAGE=128,256
Hq,Hkv=8,1; branches=64; prefix=8192; tail=64
nblk=prefix//PAGE+branches
g=torch.Generator(device="cuda").manual_seed(1)
kc=torch.randn(nblk,PAGE,Hkv,D,generator=g,device="cuda",dtype=torch.bfloat16); vc=torch.randn(nblk,PAGE,Hkv,D,generator=g,device="cuda",dtype=torch.bfloat16)
bt=torch.empty(branches,prefix//PAGE+1,dtype=torch.int32); bt[:,:prefix//PAGE]=torch.arange(prefix//PAGE,dtype=torch.int32); bt[:,-1]=torch.arange(prefix//PAGE,nblk,dtype=torch.int32); bt=bt.cuda()
lq=[tail]*branches; lk=[prefix+tail]*branches
T=sum(lq)
q=torch.randn(T,Hq,D,generator=g,device="cuda",dtype=torch.bfloat16)
cu_q=torch.tensor([0]+torch.tensor(lq).cumsum(0).tolist(),dtype=torch.int32,device="cuda"); cu_k=torch.tensor([0]+torch.tensor(lk).cumsum(0).tolist(),dtype=torch.int32,device="cuda")
plan=ops.attn_plan(cu_q,cu_k,T,Hq,Hkv,route_decode=True); out=torch.empty_like(q)
run=lambda: ops.attn_varlen_planned(q,kc,vc,cu_q,cu_k,tail,prefix+tail,D**-0.5,block_table=bt,plan=plan,out=out)
for _ in range(5): run()
torch.cuda.synchronize(); ts=[]
for _ in range(10):
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True); e0.record(); run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
print("NOPACK" if os.environ.get("B200ATTN_PREFILL_NOPACK") else "PACK", round(statistics.median(ts)*1000,1),"us", "checksum", out.float().abs().mean().item())
