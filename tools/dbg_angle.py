import sys, numpy as np
sys.path.insert(0,'.')
import orb_slam2_2021_b200 as pkg
from oracle import oracle as O
from tests.test_reference_pin import images
nf,sf,nl,ini,mn = 800,2.0,3,20,7
img = images(701,397,23)["left"]
ex = pkg.ORBextractor(nf,sf,nl,ini,mn); o = O.OracleExtractor(nf,sf,nl,ini,mn)
k,d = ex(img); ko,do = o(img)
bad = np.nonzero(k["angle"]!=ko["angle"])[0]
print(len(k), len(bad))
for l in range(nl): print("level", l, o.level(l).shape)
for i in bad[:30]:
    s = sf**k["octave"][i]
    print(i, k["octave"][i], k["x"][i]/s, k["y"][i]/s, k["angle"][i], ko["angle"][i], (d[i]!=do[i]).sum())
