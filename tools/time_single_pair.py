import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import time, numpy as np, torch, racing_slam_b200 as pkg
syn=pkg.synthetic
with pkg.Context(0) as ctx:
    st = torch.cuda.ExternalStream(ctx.stream)
    rng=np.random.default_rng(8)
    for n in (8000, 4000, 2000):
        dq=torch.from_numpy(rng.integers(0,256,(n,32),dtype=np.uint8)).cuda(); dt=torch.from_numpy(rng.integers(0,256,(n,32),dtype=np.uint8)).cuda()
        idx=torch.empty((n,2),dtype=torch.int32,device="cuda"); dist=torch.empty((n,2),dtype=torch.int32,device="cuda")
        cnt=torch.empty(1,dtype=torch.int32,device="cuda"); ot=torch.empty(n,dtype=torch.int32,device="cuda")
        for name,fn in (("knn2",lambda: ctx.knn2_dev(dq.data_ptr(),n,dt.data_ptr(),n,idx.data_ptr(),dist.data_ptr())),
                        ("match_descriptors",lambda: ctx.match_descriptors_dev(dq.data_ptr(),n,dt.data_ptr(),n,64.0,0.75,ot.data_ptr(),cnt.data_ptr())),
                        ("crosscheck",lambda: ctx.crosscheck_dev(dq.data_ptr(),n,dt.data_ptr(),n,idx.data_ptr(),dist.data_ptr()))):
            for _ in range(5): fn()
            ctx.sync()
            a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
            a.record(st)
            for _ in range(50): fn()
            b.record(st); torch.cuda.synchronize()
            us=a.elapsed_time(b)/50*1e3
            print(f"{name} {n}x{n} device: {us:.1f} us  {n*n/us/1e3:.0f} G/s")
    q,t=syn.bf_correlated(2000,2000,1)
    for name,fn in (("match_descriptors",lambda: ctx.match_descriptors(q,t,64.0,0.75)),("match_features",lambda: ctx.match_features(t,q,64.0)),("knn2",lambda: ctx.knn2(q,t))):
        for _ in range(5): fn()
        ts=[]
        for _ in range(200):
            a=time.perf_counter(); fn(); ts.append(time.perf_counter()-a)
        print(name, "2000x2000 host us", np.median(ts)*1e6)
