"""One i32 8192^2 transpose per variant given on the command line (run under ncu: tools/ncu_summary.py condenses the report)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0)
L = la._lib
R = C = 8192
X = la.Vector.synth(0x5EED, R * C, "i32"); Z = la.Vector.uninitialized(R * C, "i32")
for v in [int(x) for x in sys.argv[1:]] or [0]:
    L.lib.la_transpose_set_variant(v)
    for _ in range(2):
        L.call("la_transpose", 5, Z.ptr, X.ptr, R, C, None)
    la.synchronize()
