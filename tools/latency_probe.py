import time, numpy as np, sys
sys.path.insert(0,'/root/repo')
import orb_slam2_2021_b200 as pkg
from orb_slam2_2021_b200 import synth
W,H,nf,nl,sf,ini,mn,fx,bf=synth.CONFIGS["kitti"]
L,R=synth.stereo_pair(W,H,5)
e=pkg.ORBextractor(nf,sf,nl,ini,mn); er=pkg.ORBextractor(nf,sf,nl,ini,mn)
def t(fn,n=60):
    for _ in range(5): fn()
    ts=[]
    for _ in range(n):
        t0=time.perf_counter(); fn(); ts.append(time.perf_counter()-t0)
    return np.median(ts)*1e3
print("extract one image", t(lambda: e(L)))
kl,dl=e(L); kr,dr=er(R)
print("stereo match", t(lambda: pkg.compute_stereo_matches(e,er,kl,dl,kr,dr,bf,bf/fx)))
b=pkg.StereoFrameBatch(nf,sf,nl,ini,mn,H,W,1,stereo=True)
pl=pkg.PinnedArray((1,H,W),np.uint8); pr=pkg.PinnedArray((1,H,W),np.uint8); pl.array[0]=L; pr.array[0]=R
out=b.alloc_outputs(1,pinned=True)
print("batch engine, 1 stereo frame, process()", t(lambda: b.process(pl.array,pr.array,bf,bf/fx,out=out)))
b.upload(pl.array,pr.array)
print("batch engine, 1 stereo frame, run() only", t(lambda: b.run(1,bf,bf/fx)))
b.set_profiling(True)
for _ in range(20): b.run(1,bf,bf/fx)
st,runs=b.stage_times(); print({k:round(v/runs*1e3,1) for k,v in st.items()}, "us per stage")
