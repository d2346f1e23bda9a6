import sys, os, statistics
ROOT="/root/repo"
for p in (ROOT, os.path.join(ROOT, "nano-vllm-v1_b200")): sys.path.insert(0,p)
import torch
from b200attn import ops
D=128; B,L,Hq,Hkv=4,4096,32,8; T=B*L
g=torch.Generator(device="cuda").manual_seed(7)
q=torch.randn(T,Hq,D,generator=g,device="cuda",dtype=torch.bfloat16); k=torch.randn(T,Hkv,D,generator=g,device="cuda",dtype=torch.bfloat16); v=torch.randn(T,Hkv,D,generator=g,device="cuda",dtype=torch.bfloat16)
cu=torch.arange(0,T+1,L,dtype=torch.int32,device="cuda")
plan=ops.attn_plan(cu,cu,T,Hq,Hkv,route_decode=False); out=torch.empty_like(q)
run=lambda: ops.attn_varlen_planned(q,k,v,cu,cu,L,L,D**-0.5,plan=plan,out=out)
flush=torch.zeros(256*1024*1024//4,dtype=torch.int32,device="cuda")
small=torch.zeros(1024,device="cuda")
def t(pre, n=15):
    ts=[]
    for _ in range(n):
        pre()
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return round(statistics.median(ts)*1000,1), round(min(ts)*1000,1)
for _ in range(5): run()
torch.cuda.synchronize()
print("no flush, idle before      ", t(lambda: None))
print("read sweep 256MB (max)     ", t(lambda: flush.max()))
print("write sweep 256MB (zero_)  ", t(lambda: flush.zero_()))
print("spin kernel 100us no L2    ", t(lambda: torch.cuda._sleep(200000)))
import time
def sl(ms):
    def f():
        time.sleep(ms / 1e3); flush.max()
    return f
print("20 ms sleep + read sweep   ", t(sl(20)))
print("2 ms sleep + read sweep    ", t(sl(2)))
print("read sweep only            ", t(lambda: flush.max()))
print("200 ms sleep + read sweep  ", t(sl(200), 8))
print("no flush again             ", t(lambda: None))
# two launches back to back, time the second
def t2(n=15):
    ts=[]
    for _ in range(n):
        run(); e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return round(statistics.median(ts)*1000,1), round(min(ts)*1000,1)
print("second of two back to back ", t2())
