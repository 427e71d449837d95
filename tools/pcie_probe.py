#!/usr/bin/env python3
"""PCIe probe: H2D alone, D2H alone, both at once on two streams (pinned host memory)."""
import ctypes
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la  # noqa: E402

la.init(0)
lib = la._lib
n = 64 << 20      # 256 MB of f32
h1 = la.pinned(n, "f32"); h2 = la.pinned(n, "f32"); h1[:] = 1; h2[:] = 2
d1 = la.Vector.uninitialized(n, "f32"); d2 = la.Vector.uninitialized(n, "f32")
s1 = ctypes.c_void_p(); s2 = ctypes.c_void_p()
lib.call("la_stream_create", ctypes.byref(s1)); lib.call("la_stream_create", ctypes.byref(s2))


def t(fn, reps=5):
    fn()
    lib.call("la_stream_sync", s1); lib.call("la_stream_sync", s2)
    w = time.perf_counter()
    for _ in range(reps):
        fn()
    lib.call("la_stream_sync", s1); lib.call("la_stream_sync", s2)
    return (time.perf_counter() - w) / reps


cuda = ctypes.CDLL("libcudart.so.12") if False else None
up = lambda: lib.call("la_upload", d1.ptr, h1.ctypes.data, n * 4, s1)
# la_download syncs; use la_copy-like async path through cudaMemcpyAsync is not exported: emulate with a thread
import threading


def down():
    lib.call("la_download", h2.ctypes.data, d2.ptr, n * 4, s2)


def both():
    th = threading.Thread(target=down)
    th.start()
    up()
    lib.call("la_stream_sync", s1)
    th.join()


tu = t(up); td = t(down); tb = t(both)
print("H2D 256MB: %.2f ms (%.1f GB/s)  D2H: %.2f ms (%.1f GB/s)  both at once: %.2f ms (%.1f GB/s combined)" % (tu * 1e3, n * 4 / tu / 1e9, td * 1e3, n * 4 / td / 1e9, tb * 1e3, 2 * n * 4 / tb / 1e9))
