import torch, json
a = torch.empty(1<<30, dtype=torch.bfloat16, device='cuda'); b = torch.empty_like(a)
ev = lambda: torch.cuda.Event(enable_timing=True)
for _ in range(3): b.copy_(a)
torch.cuda.synchronize()
ts=[]
for _ in range(10):
    s,e=ev(),ev(); s.record(); b.copy_(a); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
gb = 2*a.numel()*2/1e9
s,e=ev(),ev(); s.record()
for _ in range(200): b.copy_(a)
e.record(); torch.cuda.synchronize()
print(json.dumps({"copy_best_GBps": gb/min(ts)*1e3, "copy_mean_GBps": gb/(sum(ts)/len(ts))*1e3, "copy_sustained200_GBps": gb*200/s.elapsed_time(e)*1e3}))
