import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0)
n = 8192
hA = la.pinned(n*n, "f32"); hB = la.pinned(n*n, "f32"); hC = la.pinned(n*n, "f32")
rng = np.random.default_rng(0)
hA[:] = rng.random(n*n, dtype=np.float32); hB[:] = rng.random(n*n, dtype=np.float32)
for i in range(4):
    w = time.perf_counter()
    la.mul_host(hA, hB, n, n, n, out=hC)
    print("call %d: %.2f ms" % (i, (time.perf_counter()-w)*1e3), flush=True)
