import torch
dev="cuda"
flush=torch.empty(256*1024*1024,dtype=torch.uint8,device=dev)
for nbytes in (84_000_000, 96_000_000, 512_000_000):
    out=torch.empty(nbytes//8,dtype=torch.float64,device=dev)
    src=torch.randn(nbytes//16,dtype=torch.float64,device=dev)
    ts=[];tc=[]
    for rep in range(30):
        flush.zero_()
        e0=torch.cuda.Event(enable_timing=True);e1=torch.cuda.Event(enable_timing=True)
        e0.record(); out.fill_(1.5); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        flush.zero_()
        e0=torch.cuda.Event(enable_timing=True);e1=torch.cuda.Event(enable_timing=True)
        e0.record(); out[:nbytes//16].copy_(src); e1.record(); torch.cuda.synchronize(); tc.append(e0.elapsed_time(e1))
    ts.sort();tc.sort()
    print("bytes %d: fill median %.1f us (%.0f GB/s); copy (half in, half out) median %.1f us (%.0f GB/s)"%(nbytes, ts[15]*1e3, nbytes/ts[15]/1e6, tc[15]*1e3, nbytes/tc[15]/1e6))
