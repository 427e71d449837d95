#!/usr/bin/env python3
"""oracle/build_ref.py -- TEST INFRASTRUCTURE ONLY.

Compiles the reference's own kernel source, where it lies under /root/reference, into
oracle/_ref/libref_kernels.so (git-ignored; it travels to the GPU box like our own .so).

Recipe (mirrors /root/reference/src/lib.rs:337-354 `get_src`):
  for ty in TYPES (lib.rs:84-92) + "double" (opt-in type, traits.rs:51):
      "#define IS_<TY>\n" + kernels.cl + "\n"   with every "{T}" replaced by ty
      "\n#undef IS_<TY>\n"
The text is piped to g++ on stdin between oracle/cl_shim.hpp and oracle/cl_driver.hpp, so
no reference source is ever written into this repository.  The runtime-loaded
src/extra_kernels.cl (lib.rs:173-180) is a test-only sample that nothing launches; it is
left out.

Usage: python oracle/build_ref.py [--reference /root/reference] [--contract fast|off]
"""
import argparse
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
# lib.rs:84-92, in order, plus the opt-in "double".
TYPES = ["uchar", "char", "ushort", "short", "uint", "int", "ulong", "long", "float", "double"]


def templated_source(kernels_cl: str) -> str:
    out = []
    for ty in TYPES:
        header = "#define IS_%s\n" % ty.upper()
        src = header + kernels_cl + "\n"
        out.append(src.replace("{T}", ty))
        out.append("\n#undef IS_%s\n" % ty.upper())
    return "".join(out)


def build(reference: str, contract: str = "off", out_name: str = "libref_kernels.so") -> str:
    path = os.path.join(reference, "src", "kernels.cl")
    with open(path, "r") as f:
        kernels_cl = f.read()
    tu = ['#include "cl_shim.hpp"\n', templated_source(kernels_cl), '\n#include "cl_driver.hpp"\n']
    tu += ["REF_WRAPPERS(%s)\n" % ty for ty in TYPES]
    out_dir = os.path.join(HERE, "_ref")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, out_name)
    cmd = ["g++", "-x", "c++", "-std=c++17", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-shared",
           "-fwrapv",                       # integer wrap-around is the defined semantics here
           "-ffp-contract=%s" % contract,   # OpenCL permits contraction; bracket with off/fast
           "-Wno-unused-function", "-I", HERE, "-o", out, "-"]
    subprocess.run(cmd, input="".join(tu).encode(), check=True)
    return out


def build_cl_source(reference: str) -> str:
    """oracle/_ref/libref_clsrc.so: the OpenCL program text exactly as lib.rs:337-354 `get_src` assembles it for
    the default TYPES (lib.rs:84-92), embedded in a binary as one C string (ref_cl_source()).  Used by
    oracle/cl_ref.py to run the reference's UNMODIFIED kernels through a real OpenCL driver where one exists
    (the GPU box ships NVIDIA's).  No reference text is written to the repository, only this binary."""
    path = os.path.join(reference, "src", "kernels.cl")
    with open(path, "r") as f:
        kernels_cl = f.read()
    text = []
    for ty in TYPES[:9]:                       # the reference's default list: no "double"
        text.append(("#define IS_%s\n" % ty.upper() + kernels_cl + "\n").replace("{T}", ty))
        text.append("\n#undef IS_%s\n" % ty.upper())
    src = "".join(text).encode()
    raw = kernels_cl.encode()                  # untemplated: cl_ref.py re-runs get_src with a user's extra_kernels.cl text appended
    c = "static const unsigned char text[] = {" + ",".join(str(b) for b in src) + ",0};\n" \
        "const char* ref_cl_source(void) { return (const char*)text; }\n" \
        "static const unsigned char raw[] = {" + ",".join(str(b) for b in raw) + ",0};\n" \
        "const char* ref_cl_raw_source(void) { return (const char*)raw; }\n"
    out_dir = os.path.join(HERE, "_ref")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "libref_clsrc.so")
    subprocess.run(["gcc", "-x", "c", "-O1", "-fPIC", "-shared", "-o", out, "-"], input=c.encode(), check=True)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--contract", default="off", choices=["off", "fast"])
    args = ap.parse_args()
    if not os.path.isdir(args.reference):
        print("reference tree %s not present (GPU box?): keeping prebuilt oracle/_ref" % args.reference)
        sys.exit(0)
    print(build(args.reference, "off"))
    print(build(args.reference, "fast", "libref_kernels_fma.so"))
    print(build_cl_source(args.reference))
