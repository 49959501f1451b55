import sys, numpy as np, torch
sys.path.insert(0,'.')
from ydorbslam_b200 import OrbMatcher, synth
dev=torch.device('cuda:0')
m=OrbMatcher(0.6)
q=torch.from_numpy(synth.synth_descriptors(2000,7)).to(dev)
out=torch.zeros(2000,4,dtype=torch.int32,device=dev)
s=torch.cuda.Stream(dev); m.set_stream(s.cuda_stream)
for rows in (1000000, 500000, 250000, 125000, 62500):
    db=torch.from_numpy(synth.synth_descriptors(rows,11)).to(dev)
    with torch.cuda.stream(s):
        for _ in range(5): m.top2_device(q.data_ptr(),2000,db.data_ptr(),rows,0,out.data_ptr())
        e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(50): m.top2_device(q.data_ptr(),2000,db.data_ptr(),rows,0,out.data_ptr())
        e1.record(s)
    torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/50
    print(rows, round(ms,4), "ms", round(2000*rows/ms/1e6,1), "Gpairs/s", "ideal", round(2.64*rows/1e6,4))
