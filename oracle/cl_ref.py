"""oracle/cl_ref.py -- TEST INFRASTRUCTURE ONLY.  The reference's UNMODIFIED OpenCL kernels, run through a
real OpenCL driver (kind == "reference-opencl").

Where an OpenCL platform exists (the GPU box ships NVIDIA's libnvidia-opencl.so.1, though without an ICD
registration, so it is opened directly), this builds the program exactly as /root/reference/src/lib.rs:337-354
does -- the text comes from oracle/_ref/libref_clsrc.so, built by oracle/build_ref.py where /root/reference
exists -- creates the `{T}_...` kernels and launches them with the NDRanges of the reference's Rust host code
(cited per method).  Same method names as oracle.Port / oracle.Ref.  Used by tests/test_gpu_opencl_reference.py
(our CUDA results against the reference on the SAME B200) and by bench.py's `extra.reference_opencl` figures.
The product never imports it."""
import ctypes
import os

import numpy as np

from . import CL, HERE, WORK_GROUP_COUNT, name_of

CL_DEVICE_TYPE_GPU, CL_DEVICE_TYPE_ALL = 4, 0xFFFFFFFF
CL_MEM_READ_WRITE, CL_MEM_COPY_HOST_PTR = 1, 32
CL_QUEUE_PROFILING_ENABLE = 2
CL_PROGRAM_BUILD_LOG = 0x1183
CL_KERNEL_PREFERRED_WORK_GROUP_SIZE_MULTIPLE = 0x11B3
CL_PROFILING_COMMAND_START, CL_PROFILING_COMMAND_END = 0x1282, 0x1283
CL_DEVICE_NAME = 0x102B
_P, _U, _I, _Z, _UL = ctypes.c_void_p, ctypes.c_uint, ctypes.c_int, ctypes.c_size_t, ctypes.c_ulong


class ClUnavailable(RuntimeError):
    pass


def _open_library():
    """The ICD loader (libOpenCL.so.1, shipped with the CUDA toolkit) sees no platform on the GPU box because the
    driver's libnvidia-opencl.so.1 is installed without an /etc/OpenCL/vendors/*.icd registration.  The loader's own
    environment hooks (OCL_ICD_FILENAMES / OCL_ICD_VENDORS) point it at that vendor library; nothing is installed."""
    errs = []
    vendor = None
    for cand in ("/usr/lib/libnvidia-opencl.so.1", "/usr/local/nvidia/lib/libnvidia-opencl.so.1", "/usr/lib/x86_64-linux-gnu/libnvidia-opencl.so.1"):
        if os.path.exists(cand):
            vendor = cand
            break
    if vendor and not os.path.isdir("/etc/OpenCL/vendors"):
        icd_dir = os.path.join(HERE, "_ref", "icd")
        os.makedirs(icd_dir, exist_ok=True)
        with open(os.path.join(icd_dir, "nvidia.icd"), "w") as f:
            f.write(vendor + "\n")
        os.environ.setdefault("OCL_ICD_VENDORS", icd_dir)
        os.environ.setdefault("OPENCL_VENDOR_PATH", icd_dir)
        os.environ.setdefault("OCL_ICD_FILENAMES", vendor)
    for name in ("libOpenCL.so.1", "/usr/local/cuda/targets/x86_64-linux/lib/libOpenCL.so.1"):
        try:
            lib = ctypes.CDLL(name)
            n = _U(0)
            lib.clGetPlatformIDs.restype = _I
            r = lib.clGetPlatformIDs(0, None, ctypes.byref(n))
            if r == 0 and n.value > 0:
                return lib, name
            errs.append("%s: clGetPlatformIDs -> %d, %d platforms (vendor library: %s)" % (name, r, n.value, vendor))
        except (OSError, AttributeError) as e:
            errs.append("%s: %s" % (name, e))
    raise ClUnavailable("no OpenCL platform: " + "; ".join(errs))


class ClRef:
    kind = "reference-opencl"

    TYPES = ["uchar", "char", "ushort", "short", "uint", "int", "ulong", "long", "float"]      # lib.rs:84-92

    def __init__(self, extra_source=None):
        """`extra_source`: the text of a user's src/extra_kernels.cl; the program is then assembled as lib.rs:337-354
        `get_src` does when that file exists (core source + "\n" + extra, per type)."""
        src_so = os.path.join(HERE, "_ref", "libref_clsrc.so")
        if not os.path.exists(src_so):
            raise ClUnavailable(src_so + " missing (built by oracle/build_ref.py where /root/reference exists)")
        srclib = ctypes.CDLL(src_so)
        srclib.ref_cl_source.restype = ctypes.c_char_p
        source = srclib.ref_cl_source()
        if extra_source is not None:
            if not hasattr(srclib, "ref_cl_raw_source"):
                raise ClUnavailable(src_so + " predates user-kernel support: rebuild with oracle/build_ref.py")
            srclib.ref_cl_raw_source.restype = ctypes.c_char_p
            raw = srclib.ref_cl_raw_source().decode()
            parts = []
            for ty in self.TYPES:
                parts.append(("#define IS_%s\n" % ty.upper() + raw + "\n" + extra_source).replace("{T}", ty))
                parts.append("\n#undef IS_%s\n" % ty.upper())
            source = "".join(parts).encode()
        self.cl, self.libname = _open_library()
        cl = self.cl
        for f, res in (("clCreateContext", _P), ("clCreateCommandQueue", _P), ("clCreateProgramWithSource", _P), ("clCreateKernel", _P), ("clCreateBuffer", _P)):
            getattr(cl, f).restype = res
        n = _U(0)
        cl.clGetPlatformIDs(0, None, ctypes.byref(n))
        plats = (_P * n.value)()
        cl.clGetPlatformIDs(n.value, plats, None)
        # lib.rs:356-388: first GPU on any platform, else the first device of any type
        self.device = None
        for want in (CL_DEVICE_TYPE_GPU, CL_DEVICE_TYPE_ALL):
            for p in plats:
                nd = _U(0)
                if cl.clGetDeviceIDs(_P(p), _UL(want), 0, None, ctypes.byref(nd)) == 0 and nd.value:
                    devs = (_P * nd.value)()
                    cl.clGetDeviceIDs(_P(p), _UL(want), nd.value, devs, None)
                    self.device = _P(devs[0])
                    break
            if self.device:
                break
        if not self.device:
            raise ClUnavailable("no OpenCL device")
        buf = ctypes.create_string_buffer(256)
        cl.clGetDeviceInfo(self.device, CL_DEVICE_NAME, 256, buf, None)
        self.device_name = buf.value.decode()
        err = _I(0)
        dev_arr = (_P * 1)(self.device)
        self.ctx = _P(cl.clCreateContext(None, 1, dev_arr, None, None, ctypes.byref(err)))
        self._ok(err.value, "clCreateContext")
        self.queue = _P(cl.clCreateCommandQueue(self.ctx, self.device, _UL(CL_QUEUE_PROFILING_ENABLE), ctypes.byref(err)))
        self._ok(err.value, "clCreateCommandQueue")
        srcs = (ctypes.c_char_p * 1)(source)
        self.program = _P(cl.clCreateProgramWithSource(self.ctx, 1, srcs, None, ctypes.byref(err)))
        self._ok(err.value, "clCreateProgramWithSource")
        r = cl.clBuildProgram(self.program, 1, dev_arr, None, None, None)        # lib.rs:326-332
        if r != 0:
            sz = _Z(0)
            cl.clGetProgramBuildInfo(self.program, self.device, CL_PROGRAM_BUILD_LOG, 0, None, ctypes.byref(sz))
            log = ctypes.create_string_buffer(sz.value + 1)
            cl.clGetProgramBuildInfo(self.program, self.device, CL_PROGRAM_BUILD_LOG, sz.value, log, None)
            raise ClUnavailable("Failed to compile kernels with error: %d\n%s" % (r, log.value.decode(errors="replace")[-2000:]))
        self._kernels = {}
        self.last_kernel_ns = None

    @staticmethod
    def _ok(code, what):
        if code != 0:
            raise RuntimeError("%s failed: %d" % (what, code))

    # ---- plumbing
    def _kernel(self, ty, name):
        key = ty + "_" + name                                   # lib.rs:187-189
        if key not in self._kernels:
            err = _I(0)
            k = _P(self.cl.clCreateKernel(self.program, key.encode(), ctypes.byref(err)))
            self._ok(err.value, "clCreateKernel(%s)" % key)
            self._kernels[key] = k
        return self._kernels[key]

    def _buf(self, arr=None, nbytes=None):
        err = _I(0)
        if arr is not None and arr.size:                          # vector.rs:58-69 (COPY_HOST_PTR)
            m = self.cl.clCreateBuffer(self.ctx, _UL(CL_MEM_READ_WRITE | CL_MEM_COPY_HOST_PTR), _Z(arr.nbytes), arr.ctypes.data_as(_P), ctypes.byref(err))
        else:                                                     # vector.rs:36-46
            m = self.cl.clCreateBuffer(self.ctx, _UL(CL_MEM_READ_WRITE), _Z(max(1, nbytes if nbytes is not None else 1)), None, ctypes.byref(err))
        self._ok(err.value, "clCreateBuffer")
        return _P(m)

    def _read(self, mem, dtype, n):                               # vector.rs:71-78 (blocking read)
        out = np.empty(n, dtype=dtype)
        if n:
            self._ok(self.cl.clEnqueueReadBuffer(self.queue, mem, 1, _Z(0), _Z(out.nbytes), out.ctypes.data_as(_P), 0, None, None), "clEnqueueReadBuffer")
        return out

    def _free(self, *mems):
        for m in mems:
            self.cl.clReleaseMemObject(m)

    def _launch(self, kernel, args, gws, lws=None):
        """args: list of ('buf', cl_mem) | ('val', numpy scalar array) | ('local', nbytes).  Waits for the
        kernel's event like every reference op (event.wait_for(), vector.rs:108) and records its device time."""
        for i, (kind, v) in enumerate(args):
            if kind == "buf":
                self._ok(self.cl.clSetKernelArg(kernel, i, _Z(8), ctypes.byref(v)), "clSetKernelArg")
            elif kind == "val":
                self._ok(self.cl.clSetKernelArg(kernel, i, _Z(v.nbytes), v.ctypes.data_as(_P)), "clSetKernelArg")
            else:
                self._ok(self.cl.clSetKernelArg(kernel, i, _Z(v), None), "clSetKernelArg")
        g = (_Z * 1)(gws)
        l = (_Z * 1)(lws) if lws else None
        ev = _P()
        self._ok(self.cl.clEnqueueNDRangeKernel(self.queue, kernel, 1, None, g, l, 0, None, ctypes.byref(ev)), "clEnqueueNDRangeKernel")
        evs = (_P * 1)(ev)
        self._ok(self.cl.clWaitForEvents(1, evs), "clWaitForEvents")
        t0, t1 = _UL(0), _UL(0)
        self.cl.clGetEventProfilingInfo(ev, CL_PROFILING_COMMAND_START, 8, ctypes.byref(t0), None)
        self.cl.clGetEventProfilingInfo(ev, CL_PROFILING_COMMAND_END, 8, ctypes.byref(t1), None)
        self.last_kernel_ns = t1.value - t0.value
        self.cl.clReleaseEvent(ev)

    def run_user_kernel(self, full_name, args, gws):
        """create_kernel (lib.rs:165-171) + one launch with `gws` work-items and no local size (vector.rs:80-151).
        args: numpy arrays (buffers, returned with their contents after the launch) or ('val', numpy scalar array)."""
        err = _I(0)
        k = _P(self.cl.clCreateKernel(self.program, full_name.encode(), ctypes.byref(err)))
        self._ok(err.value, "clCreateKernel(%s)" % full_name)
        mems, largs = [], []
        for a in args:
            if isinstance(a, tuple):
                largs.append(a)
            else:
                m = self._buf(np.ascontiguousarray(a))
                mems.append((m, a))
                largs.append(("buf", m))
        self._launch(k, largs, gws)
        outs = [self._read(m, a.dtype, a.size) for m, a in mems]
        self._free(*[m for m, _ in mems])
        self.cl.clReleaseKernel(k)
        return outs

    def preferred_multiple(self, kernel):                         # lib.rs:101-113
        v = _Z(0)
        self._ok(self.cl.clGetKernelWorkGroupInfo(kernel, self.device, CL_KERNEL_PREFERRED_WORK_GROUP_SIZE_MULTIPLE, 8, ctypes.byref(v), None), "clGetKernelWorkGroupInfo")
        return v.value

    # ---- elementwise (vector.rs:96-125, 430-470, 598-643)
    _OPS = {0: "add", 1: "sub", 2: "mul", 3: "div"}

    def ew_vec_vec(self, op, a, b):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        A, B, C = self._buf(a), self._buf(b), self._buf(nbytes=a.nbytes)
        self._launch(self._kernel(CL[name_of(a)], self._OPS[op] + "_vec_vec"), [("buf", C), ("buf", A), ("buf", B)], a.size)
        out = self._read(C, a.dtype, a.size); self._free(A, B, C)
        return out

    def ew_assign_vec_vec(self, op, c, b):
        c = np.ascontiguousarray(c); b = np.ascontiguousarray(b, dtype=c.dtype)
        C, B = self._buf(c), self._buf(b)
        self._launch(self._kernel(CL[name_of(c)], self._OPS[op] + "_assign_vec_vec"), [("buf", C), ("buf", B)], c.size)
        out = self._read(C, c.dtype, c.size); self._free(C, B)
        return out

    def ew_vec_scl(self, op, a, s):
        a = np.ascontiguousarray(a); sv = np.array([s], dtype=a.dtype)
        A, C = self._buf(a), self._buf(nbytes=a.nbytes)
        self._launch(self._kernel(CL[name_of(a)], self._OPS[op] + "_vec_scl"), [("buf", C), ("buf", A), ("val", sv)], a.size)
        out = self._read(C, a.dtype, a.size); self._free(A, C)
        return out

    def eq_vec(self, a, b):                                       # vector.rs:645-675
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        if a.size != b.size:
            return False
        A, B, F = self._buf(a), self._buf(b), self._buf(np.array([1], np.uint8))
        self._launch(self._kernel(CL[name_of(a)], "eq_vec"), [("buf", F), ("buf", A), ("buf", B)], a.size)
        flag = self._read(F, np.uint8, 1); self._free(A, B, F)
        return bool(flag[0] != 0)

    # ---- sum / dot (vector.rs:256-293, 326-363; lib.rs:101-122)
    def _reduce(self, a, b):
        ty = CL[name_of(a)]
        kernel = self._kernel(ty, "dot_vec_vec" if b is not None else "sum_vec")
        W = self.preferred_multiple(kernel)
        if a.size <= WORK_GROUP_COUNT:                            # host path, vector.rs:268-270 / 338-340
            with np.errstate(over="ignore"):
                vals = a if b is None else (a * b).astype(a.dtype)
                acc = a.dtype.type(0)
                for x in vals:
                    acc = a.dtype.type(acc + x)
            return acc, W
        A = self._buf(a); T = self._buf(nbytes=WORK_GROUP_COUNT * a.itemsize)
        count = np.array([a.size], np.int32)
        if b is None:
            args = [("buf", A), ("buf", T), ("val", count), ("local", W * a.itemsize)]
            B = None
        else:
            B = self._buf(b)
            args = [("buf", A), ("buf", B), ("buf", T), ("val", count), ("local", W * a.itemsize)]
        self._launch(kernel, args, WORK_GROUP_COUNT * W, W)
        parts = self._read(T, a.dtype, WORK_GROUP_COUNT)
        self._free(A, T, *([B] if B is not None else []))
        with np.errstate(over="ignore"):
            acc = a.dtype.type(0)
            for x in parts:                                       # host iter::sum(), vector.rs:290 / 361
                acc = a.dtype.type(acc + x)
        return acc, W

    def sum_vec(self, a):
        return self._reduce(np.ascontiguousarray(a), None)[0]

    def dot_vec_vec(self, a, b):
        a = np.ascontiguousarray(a)
        return self._reduce(a, np.ascontiguousarray(b, dtype=a.dtype))[0]

    # ---- matrix-vector / matrix-matrix (vector.rs:508-569, matrix.rs:240-265)
    def mul_vec_mat(self, v, m, rows, cols):
        v = np.ascontiguousarray(v); m = np.ascontiguousarray(m, dtype=v.dtype)
        A, B, C = self._buf(v), self._buf(m), self._buf(nbytes=cols * v.itemsize)
        self._launch(self._kernel(CL[name_of(v)], "mul_vec_mat"), [("buf", C), ("buf", A), ("buf", B), ("val", np.array([cols], np.int32)), ("val", np.array([rows], np.int32))], cols)
        out = self._read(C, v.dtype, cols); self._free(A, B, C)
        return out

    def mul_vec_transpose_mat(self, v, m, rows, cols):
        v = np.ascontiguousarray(v); m = np.ascontiguousarray(m, dtype=v.dtype)
        A, B, C = self._buf(v), self._buf(m), self._buf(nbytes=rows * v.itemsize)
        self._launch(self._kernel(CL[name_of(v)], "mul_vec_transpose_mat"), [("buf", C), ("buf", A), ("buf", B), ("val", np.array([cols], np.int32)), ("val", np.array([cols], np.int32))], rows)
        out = self._read(C, v.dtype, rows); self._free(A, B, C)
        return out

    def mul_mat_mat(self, a, b, M, K, N):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        A, B, C = self._buf(a), self._buf(b), self._buf(nbytes=M * N * a.itemsize)
        self._launch(self._kernel(CL[name_of(a)], "mul_mat_mat"), [("buf", C), ("buf", A), ("buf", B), ("val", np.array([N], np.int32)), ("val", np.array([K], np.int32))], M * N)
        out = self._read(C, a.dtype, M * N); self._free(A, B, C)
        return out


_instance = None


def cl_ref():
    """The shared ClRef, or raises ClUnavailable."""
    global _instance
    if _instance is None:
        _instance = ClRef()
    return _instance
