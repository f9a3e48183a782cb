import torch, sys, os, json
sys.path.insert(0, os.getcwd())
from introspective_awareness_b200 import kernels
dev="cuda:0"
flush=torch.empty(256*1024*1024,dtype=torch.uint8,device=dev)
for rows,width in ((25600,14336),(2400,14336),(1200,14336),(200,14336),(25600,27648)):
    gu=(torch.randn(rows,2*width,device=dev)*2).to(torch.bfloat16)
    g,u=gu.split(width,dim=-1)
    for _ in range(3): kernels.silu_mul(g,u)
    ts=[]
    for _ in range(10):
        flush.zero_(); a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
        a.record(); o=kernels.silu_mul(g,u); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    want=torch.nn.functional.silu(g)*u
    ms=sum(ts)/len(ts); bytes_=rows*width*2*3
    print(json.dumps({"rows":rows,"width":width,"us":ms*1e3,"gbs":bytes_/ms/1e6,"equal":bool(torch.equal(o,want))}))
