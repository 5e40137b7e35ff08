"""GPU debug aid: stage-by-stage parity of the CUDA path against the oracle on a few configs."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import orb_slam2_2021_b200 as pkg
from oracle import oracle
from orb_slam2_2021_b200 import synth


def cmp_frame(name, seed):
    W, H, nf, nl, sf, ini, mn, fx, bf = synth.CONFIGS[name]
    left, right = synth.stereo_pair(W, H, seed)
    b = pkg.StereoFrameBatch(nf, sf, nl, ini, mn, H, W, 2, stereo=True)
    L = np.stack([left, left[::-1].copy()])
    R = np.stack([right, right[::-1].copy()])
    out = b.process(L, R, mbf=bf, mb=bf / fx)
    ok = True
    for f in range(2):
        ol, orr = oracle.OracleExtractor(nf, sf, nl, ini, mn), oracle.OracleExtractor(nf, sf, nl, ini, mn)
        okl, odl = ol(L[f])
        okr, odr = orr(R[f])
        for l in range(nl):
            for s, o in ((0, ol), (1, orr)):
                if not np.array_equal(b.level(f, s, l), o.level(l)):
                    print(name, f, "LEVEL MISMATCH", l, s); ok = False
                if not np.array_equal(b.level(f, s, l, True), o.level(l, True)):
                    print(name, f, "BLUR MISMATCH", l, s); ok = False
                gc, oc = b.candidates(f, s, l), o.candidates(l)
                if len(gc) != len(oc) or not all(np.array_equal(gc[k], oc[k]) for k in ("x", "y", "response")):
                    print(name, f, "CAND MISMATCH", l, s, len(gc), len(oc)); ok = False
        nL, nR = out["n_left"][f], out["n_right"][f]
        kl, dl = out["kps_left"][f][:nL], out["desc_left"][f][:nL]
        kr, dr = out["kps_right"][f][:nR], out["desc_right"][f][:nR]
        print(name, "frame", f, "n", nL, len(okl), nR, len(okr))
        if nL != len(okl) or nR != len(okr):
            ok = False
            continue
        for k in okl.dtype.names:
            if not np.array_equal(kl[k], okl[k]):
                print("  left field mismatch", k, (kl[k] != okl[k]).sum()); ok = False
            if not np.array_equal(kr[k], okr[k]):
                print("  right field mismatch", k, (kr[k] != okr[k]).sum()); ok = False
        if not np.array_equal(dl, odl):
            print("  desc L mismatch", (dl != odl).any(1).sum()); ok = False
        if not np.array_equal(dr, odr):
            print("  desc R mismatch", (dr != odr).any(1).sum()); ok = False
        our, odp = oracle.stereo_match(ol, orr, okl, odl, okr, odr, bf, bf / fx)
        ur, dp = out["u_right"][f][:nL], out["depth"][f][:nL]
        print("  stereo valid", (ur >= 0).sum(), (our >= 0).sum(), "equal", np.array_equal(ur, our), np.array_equal(dp, odp))
        if not (np.array_equal(ur, our) and np.array_equal(dp, odp)):
            ok = False
    return ok


if __name__ == "__main__":
    allok = True
    for name in sys.argv[1:] or ["arducam", "kitti", "euroc"]:
        t = time.time()
        allok &= cmp_frame(name, 3)
        print(name, "done %.1fs" % (time.time() - t))
    print("ALL OK" if allok else "MISMATCHES")
    import __graft_entry__ as g

    g.smoke()
