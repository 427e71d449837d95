#!/usr/bin/env python3
"""GB/s of la_transpose for every element size (GPU only)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la  # noqa: E402

la.init(0)
lib = la._lib
variants = [int(x) for x in sys.argv[1:]] or [0]
for variant in variants:
  lib.lib.la_transpose_set_variant(variant)
  print("--- variant", variant, flush=True)
  for name, code in ((("i32", 5),) if variant else (("u8", 0), ("u16", 2), ("i32", 5), ("u64", 6))):
      for R, C in ((8192, 8192), (16384, 4096), (32768, 32768) if name == "i32" else (4096, 16384)):
          sz = lib.lib.la_dtype_size(code)
          X = la.Vector.uninitialized(R * C, name); Z = la.Vector.uninitialized(R * C, name)
          v = np.array([3], la.parameter(name).np)
          lib.call("la_fill", code, X.ptr, v.ctypes.data, R * C, None)
          for _ in range(3):
              lib.call("la_transpose", code, Z.ptr, X.ptr, R, C, None)
          la.synchronize()
          e0, e1 = la.Event(), la.Event()
          e0.record()
          for _ in range(10):
              lib.call("la_transpose", code, Z.ptr, X.ptr, R, C, None)
          e1.record()
          ms = e0.elapsed_ms(e1) / 10
          print("%s %dx%d: %.4f ms  %.0f GB/s" % (name, R, C, ms, 2.0 * R * C * sz / ms / 1e6), flush=True)
          del X, Z
