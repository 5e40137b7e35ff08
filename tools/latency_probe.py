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
# the three single-frame flows through the C-ABI (see bench.py: latency)
from bench import single_frame_latency
Ls, Rs = synth.stereo_batch(W, H, 16, seed0=40)
lat = single_frame_latency(pkg, synth.CONFIGS["kitti"], Ls, Rs, 0, n=1000)
print({k: (v if not isinstance(v, dict) else {a: round(b, 4) if isinstance(b, float) else b for a, b in v.items()}) for k, v in lat.items() if k != "api"})
# the reference's OWN Frame::Frame(stereo) (src/Frame.cc:80-152 compiled with INTEGRATION.md's patch, oracle/_ref/libref_frame_gpu.so):
# its two std::threads -> operator() on the GPU extractors, ComputeStereoMatchesGPU (resident match), AssignFeaturesToGrid, ...
try:
    from oracle import ref as RR
    gl, gr = RR.GpuShimExtractor(nf, sf, nl, ini, mn), RR.GpuShimExtractor(nf, sf, nl, ini, mn)
    K4 = np.array([fx, fx, W / 2, H / 2], np.float32)
    ts = []
    for i in range(120):
        t0 = time.perf_counter()
        F = RR.frame_stereo_gpu(gl, gr, Ls[i % 16], Rs[i % 16], K4, bf, 35 * bf / fx, bf / fx, capacity=4096)
        ts.append(time.perf_counter() - t0)
    print("reference Frame::Frame(stereo) on the GPU extractor (two threads, whole constructor incl. grid, + copy-out of the Frame): p50 %.3f ms"
          % (np.median(ts[20:]) * 1e3), "N", F["N"])
    cl, cr = RR.RefExtractor(nf, sf, nl, ini, mn), RR.RefExtractor(nf, sf, nl, ini, mn)
    ts = []
    for i in range(6):
        t0 = time.perf_counter()
        RR.frame_stereo(cl, cr, Ls[i], Rs[i], K4, bf, 35 * bf / fx, bf / fx, capacity=4096)
        ts.append(time.perf_counter() - t0)
    print("the same constructor on the reference's CPU extractor (scalar OpenCV stand-ins): p50 %.1f ms" % (np.median(ts[1:]) * 1e3))
except Exception as e:
    print("drop-in Frame timing unavailable:", e)
