import sys,time; sys.path.insert(0,'/root/repo')
import numpy as np, racing_slam_b200 as pkg
syn=pkg.synthetic
v=syn.random_vocabulary(10,5,6); q=syn.vocabulary_queries(v,8000,7)
def wall(fn,reps=100):
    for _ in range(5): fn()
    ts=[]
    for _ in range(reps):
        t=time.perf_counter(); fn(); ts.append(time.perf_counter()-t)
    return np.median(ts)*1e6
with pkg.Context(0) as ctx:
    ctx.vocab_upload(1,v["child_offset"],v["child_ids"],v["desc"],v["weight"],v["word_id"])
    print('transform host us',wall(lambda: ctx.vocab_transform(1,q)))
    print('bow host us',wall(lambda: ctx.vocab_bow(1,q)))
    print('bow 1000 feats us',wall(lambda: ctx.vocab_bow(1,q[:1000])))
    print('sync only us',wall(lambda: ctx.sync()))
    print('knn2 64x64 us',wall(lambda: ctx.knn2(q[:64],q[:64])))
