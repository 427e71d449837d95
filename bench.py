#!/usr/bin/env python3
"""bench.py -- the hot path's headline metric on B200.

Headline (BASELINE.json `metric`): f32 GEMM GFLOP/s at 8192^2.  A "step" is one
C = A x B of 8192 x 8192 f32 matrices through la_mul_mat_mat with the library's default engine
selection (LA_GEMM_AUTO: the tcgen05 3xTF32 engine, which meets the same f32 tolerance gate as
the FP32 SIMT engine; `--engine simt` times the SIMT engine instead, and `extra.gemm_f32_sweep`
always carries both), operands resident in HBM.  With N GPUs the 8192^2 problem is row-block sharded (A and C split by
rows, B replicated, no data-path collective): strong scaling, value = total flop / max-over-ranks
time.  `e2e` is the same GEMM on pinned HOST buffers through the public streamed call mul_host /
la_mul_mat_mat_host (upload A and B, multiply, download C, all inside the timed region);
`e2e.three_call` is the reference-style from_vec, from_vec, `*`, to_vec sequence.  `extra` carries the vector-op HBM GB/s figures
the metric also names (elementwise 2^28 f32, dot/sum, Matrix x Vector at 32768^2, i32 8192^2).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-extra]

`--impl reference` times the reference's own kernels (src/kernels.cl compiled as C++ into
oracle/_ref, all host cores) on a bounded sample of the same GEMM.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_GEMM = 8192
METRIC = "f32_gemm_gflops_8192"
UNIT = "GFLOP/s"


def bind_near_gpu(local):
    """Run this rank on the CPU cores next to its GPU (NVML's ideal affinity), BEFORE any pinned buffer is allocated, so that
    the page-locked host buffers land on the GPU's own NUMA node.  What `numactl` / a job launcher would do; with 8 ranks
    streaming host buffers at once the far socket otherwise carries half of the traffic.  Returns a note for the bench line."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local)
        bus = "%08X:%02X:%02X.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return "rank bound to the %d cores next to its GPU (NVML affinity)" % len(cpus)
    except Exception as e:  # noqa: BLE001
        return "not bound (%s)" % str(e)[:80]
    return "not bound"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("sm_max_mhz", 1965.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.samples, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def window(self, t0, t1):
        rows = [s.split(", ") for t, s in self.samples if t0 - 0.05 <= t <= t1 + 0.15] or [s.split(", ") for _, s in self.samples[-3:]]
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        reasons = set()
        for r in rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.strip().lower() == "active":
                    reasons.add(name)
        mx = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(rows)}

    def stop(self):
        if self.proc:
            self.proc.terminate()


# ----------------------------------------------------------------------------- reference arm / cpu baseline
def omp_use_all_cores():
    """The reference kernels run as OpenMP loops over work-items (oracle/cl_driver.hpp).  torch.distributed.run exports
    OMP_NUM_THREADS=1 to its workers, which libgomp reads when it loads: set the team size explicitly instead, and report the
    number libgomp will really use."""
    want = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    os.environ["OMP_NUM_THREADS"] = str(want)             # for a libgomp that is not loaded yet
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(want)                     # for one that is
        return int(gomp.omp_get_max_threads())
    except OSError:
        return want


def time_reference_c0():
    """BASELINE.json configs[0] on the host cores through the reference's own kernels (oracle/_ref; the C port where that
    was not built): 1024^2 f32 mul_mat_mat (gws = M*N), 2^20-element add_vec_vec (gws = N) and dot_vec_vec (64 work-groups
    of W = 32 work-items + host sum of the 64 partials, src/lib.rs:101-122, src/vector.rs:326-363)."""
    import oracle
    try:
        R = oracle.ref(); kind = "reference"
    except Exception:
        R = oracle.port(); kind = "port"
    rng = np.random.default_rng(2)
    n0 = 1024; v0 = 1 << 20
    a = rng.uniform(-1, 1, n0 * n0).astype(np.float32); b = rng.uniform(-1, 1, n0 * n0).astype(np.float32)
    x = rng.uniform(-1, 1, v0).astype(np.float32); y = rng.uniform(-1, 1, v0).astype(np.float32)

    def best_of(fn, reps):
        fn()
        ts = []
        for _ in range(reps):
            t = time.perf_counter(); fn(); ts.append(time.perf_counter() - t)
        return min(ts) * 1e3, statistics.median(ts) * 1e3

    out = {"kind": kind}
    b_ms, m_ms = best_of(lambda: R.mul_mat_mat(a, b, n0, n0, n0), 3)
    out["gemm_f32_1024^2"] = {"ms": round(m_ms, 2), "best_ms": round(b_ms, 2), "GFLOP/s": round(2.0 * n0 ** 3 / m_ms / 1e6, 2)}
    b_ms, m_ms = best_of(lambda: R.ew_vec_vec(oracle.ADD, x, y), 10)
    out["add_f32_2^20"] = {"ms": round(m_ms, 4), "best_ms": round(b_ms, 4), "GB/s": round(12.0 * v0 / m_ms / 1e6, 2)}
    b_ms, m_ms = best_of(lambda: R.dot_vec_vec(x, y), 10)
    out["dot_f32_2^20"] = {"ms": round(m_ms, 4), "best_ms": round(b_ms, 4), "GB/s": round(8.0 * v0 / m_ms / 1e6, 2)}
    return out


def time_reference_gemm(budget_s):
    """The reference's {T}_mul_mat_mat (kernels.cl:161-169), compiled verbatim as C++ (oracle/_ref),
    all host cores, on the first `rows` rows of the 8192^2 product (gws = rows*8192 work-items)."""
    import ctypes
    import oracle
    kind = "reference"
    try:
        R = oracle.ref()
        fn = R.lib.ref_float_mul_mat_mat
    except Exception:
        R, fn, kind = oracle.port(), None, "port"
    n = N_GEMM
    rng = np.random.default_rng(1)
    a = rng.uniform(-1, 1, n * n).astype(np.float32); b = rng.uniform(-1, 1, n * n).astype(np.float32)

    def run(rows):
        c = np.empty(rows * n, np.float32)
        t = time.perf_counter()
        if fn is not None:
            fn(c.ctypes.data_as(ctypes.c_void_p), a.ctypes.data_as(ctypes.c_void_p), b.ctypes.data_as(ctypes.c_void_p),
               ctypes.c_int(n), ctypes.c_int(n), ctypes.c_long(rows))
        else:
            R.lib.oracle_mul_mat_mat(8, c.ctypes.data_as(ctypes.c_void_p), a.ctypes.data_as(ctypes.c_void_p), b.ctypes.data_as(ctypes.c_void_p), rows, n, n, 0)
        return time.perf_counter() - t

    cores = omp_use_all_cores()
    t1 = run(cores)                                       # one row per core: what a row costs with every core busy
    rows = int(max(cores, min(512, cores * budget_s / max(t1, 1e-6))))
    t = run(rows)
    return {"value": 2.0 * n * n * rows / t / 1e9, "unit": UNIT, "cores": cores, "kind": kind, "seconds": t,
            "sample": "first %d of %d rows of the 8192^2 f32 product (%d work-items of kernels.cl mul_mat_mat, OpenMP over work-items)" % (rows, n, rows * n)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    budget = float(os.environ.get("LA_BENCH_REF_BUDGET_S", "30"))      # CPU seconds over all timed steps (tests shrink it)
    for _ in range(args.warmup):
        time_reference_gemm(min(1.0, budget))
    for _ in range(args.steps):
        vals.append(time_reference_gemm(max(min(2.0, budget), budget / max(1, args.steps))))
    v = statistics.mean(x["value"] for x in vals)
    cb = dict(vals[-1]); cb["value"] = v
    ms = statistics.mean(x["seconds"] for x in vals) * 1e3
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": "f32 GEMM 8192x8192x8192 (bounded row sample per step)", "M": N_GEMM, "K": N_GEMM, "N": N_GEMM},
                      "cpu_baseline": cb, "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ----------------------------------------------------------------------------- our arm
def timed_stats(la, fn, iters=10, warmup=3):
    """Device times from events on the launching stream.  SURVEY.md section 8(d): >= 3 warm-ups, best and median of >= 10.
    A sample is one launch for kernels of a millisecond or more; shorter kernels are timed in back-to-back batches of k
    launches (k chosen so that a batch lasts ~1 ms) and a sample is the batch average, so that the few microseconds an event
    record costs between two launches do not count as kernel time."""
    for _ in range(warmup):
        fn()
    la.synchronize()
    p0, p1 = la.Event(), la.Event()
    p0.record(); fn(); p1.record()
    once = max(p0.elapsed_ms(p1), 1e-3)
    batch = int(max(1, min(64, round(1.0 / once))))
    ev = [la.Event() for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        for _ in range(batch):
            fn()
        ev[i + 1].record()
    ms = [ev[i].elapsed_ms(ev[i + 1]) / batch for i in range(iters)]
    return {"best": min(ms), "median": statistics.median(ms), "mean": sum(ms) / len(ms), "n": iters * batch, "launches_per_sample": batch}


def timed(la, fn, iters, warmup=3):
    return timed_stats(la, fn, iters, warmup)["median"]


def extras(la, hbm_peak, peak_src):
    """Vector-op HBM numbers (BASELINE.json configs[1], [2], [4]); operands >> L2 (126 MB)."""
    lib = la._lib
    out = {}
    n = 1 << 28
    A = la.Vector.uninitialized(n, "f32"); B = la.Vector.uninitialized(n, "f32"); C = la.Vector.uninitialized(n, "f32")
    one = np.array([1.0], np.float32); two = np.array([1.0000001], np.float32)
    lib.call("la_synth", 8, A.ptr, 0x5EED0001, n, None); lib.call("la_synth", 8, B.ptr, 0x5EED1001, n, None)

    def rec(name, st, nbytes):
        """st = timed_stats(...): GB/s from the MEDIAN launch, the best beside it."""
        gbs = nbytes / st["median"] / 1e6
        out[name] = {"ms": round(st["median"], 4), "best_ms": round(st["best"], 4), "GB/s": round(gbs, 1), "best_GB/s": round(nbytes / st["best"] / 1e6, 1),
                     "frac_of_measured_hbm": round(gbs / hbm_peak, 3), "frac_of_nominal_8TBs": round(gbs / 8000.0, 3), "launches": st["n"],
                     "launches_per_sample": st["launches_per_sample"]}

    for opname, op in (("add", 0), ("sub", 1), ("mul", 2)):
        rec("ew_%s_f32_2^28" % opname, timed_stats(la, lambda: lib.call("la_ew_vec_vec", 8, op, C.ptr, A.ptr, B.ptr, n, None), 10), 12 * n)
    rec("ew_add_assign_f32_2^28", timed_stats(la, lambda: lib.call("la_ew_assign_vec_vec", 8, 0, C.ptr, B.ptr, n, None), 10), 12 * n)
    rec("scale_f32_2^28", timed_stats(la, lambda: lib.call("la_ew_vec_scl", 8, 2, C.ptr, A.ptr, two.ctypes.data, n, None), 10), 8 * n)
    rec("clone_f32_2^28", timed_stats(la, lambda: lib.call("la_copy", C.ptr, A.ptr, 4 * n, None), 10), 8 * n)        # Vector::clone (vector.rs:700-708)
    rec("fill_f32_2^28", timed_stats(la, lambda: lib.call("la_fill", 8, C.ptr, one.ctypes.data, n, None), 10), 4 * n)   # Vector::new (vector.rs:30-34): writes only
    eqflag = ctypes.c_int(0)
    lib.call("la_copy", C.ptr, A.ptr, 4 * n, None)                 # equal operands: the whole of both vectors is read, the answer is "equal"
    rec("eq_f32_2^28", timed_stats(la, lambda: lib.call("la_eq_vec", 8, A.ptr, C.ptr, n, ctypes.byref(eqflag), None), 10), 8 * n)
    assert eqflag.value == 1
    res = la.Vector.uninitialized(4, "f32")
    rec("dot_f32_2^28", timed_stats(la, lambda: lib.call("la_dot_vec_vec_dev", 8, A.ptr, B.ptr, n, res.ptr, None), 10), 8 * n)
    rec("sum_f32_2^28", timed_stats(la, lambda: lib.call("la_sum_vec_dev", 8, A.ptr, n, res.ptr, None), 10), 4 * n)
    del B, C
    # C2: dot / sum over 2^30 elements (4 GiB per operand)
    n30 = 1 << 30
    del A
    A = la.Vector.uninitialized(n30, "f32"); B = la.Vector.uninitialized(n30, "f32")
    lib.call("la_synth", 8, A.ptr, 0x5EED0005, n30, None); lib.call("la_synth", 8, B.ptr, 0x5EED1005, n30, None)
    rec("dot_f32_2^30", timed_stats(la, lambda: lib.call("la_dot_vec_vec_dev", 8, A.ptr, B.ptr, n30, res.ptr, None), 10), 8 * n30)
    rec("sum_f32_2^30", timed_stats(la, lambda: lib.call("la_sum_vec_dev", 8, A.ptr, n30, res.ptr, None), 10), 4 * n30)
    del B
    # Matrix x Vector at 32768^2 (4 GiB): reuse A's storage? allocate a fresh matrix
    r = 32768
    M = la.Vector.uninitialized(r * r, "f32"); v = la.Vector.uninitialized(r, "f32"); y = la.Vector.uninitialized(r, "f32")
    lib.call("la_synth", 8, M.ptr, 0x5EED0002, r * r, None); lib.call("la_synth", 8, v.ptr, 0x5EED1002, r, None)
    rec("mat_x_vec_f32_32768^2", timed_stats(la, lambda: lib.call("la_mul_vec_transpose_mat", 8, y.ptr, v.ptr, M.ptr, r, r, None), 10, 3), 4 * (r * r + 2 * r))
    rec("vec_x_mat_f32_32768^2", timed_stats(la, lambda: lib.call("la_mul_vec_mat", 8, y.ptr, v.ptr, M.ptr, r, r, None), 10, 3), 4 * (r * r + 2 * r))
    del M, v, y, A
    # i32 at 8192^2: transpose + GEMM
    k = 8192
    X = la.Vector.uninitialized(k * k, "i32"); Y = la.Vector.uninitialized(k * k, "i32"); Z = la.Vector.uninitialized(k * k, "i32")
    three = np.array([3], np.int32)
    lib.call("la_synth", 5, X.ptr, 0x5EED0004, k * k, None); lib.call("la_synth", 5, Y.ptr, 0x5EED1004, k * k, None)
    rec("transpose_i32_8192^2", timed_stats(la, lambda: lib.call("la_transpose", 5, Z.ptr, X.ptr, k, k, None), 20), 8 * k * k)
    # C4: i32 elementwise at 8192^2 (wrap-around arithmetic; 768 MB of traffic per launch)
    for opname, op in (("add", 0), ("sub", 1), ("mul", 2)):
        rec("ew_%s_i32_8192^2" % opname, timed_stats(la, lambda: lib.call("la_ew_vec_vec", 5, op, Z.ptr, X.ptr, Y.ptr, k * k, None), 10), 12 * k * k)
    sm = __import__("ctypes").c_int(); lib.call("la_sm_count", __import__("ctypes").byref(sm))
    imad_peak = sm.value * 64 * 2 * 1965e6 / 1e12
    peaks0 = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    int8_peak = 2.0 * peaks0.get("bf16_tflops", 1590.0)            # int8 runs at twice the bf16 rate
    lib.call("la_set_gemm_mode", 0)                                 # IMAD engine (mode SIMT)
    ms = timed(la, lambda: lib.call("la_mul_mat_mat", 5, Z.ptr, X.ptr, Y.ptr, k, k, k, None), 3, 1)
    out["gemm_i32_8192^2_imad"] = {"ms": round(ms, 3), "Top/s": round(2.0 * k ** 3 / ms / 1e9, 2),
                                   "frac_of_nominal_int32_mad_peak": round(2.0 * k ** 3 / ms / 1e9 / imad_peak, 3),
                                   "peak_note": "nominal: %d SMs x 64 IMAD lanes x 2 x 1965 MHz = %.1f Top/s" % (sm.value, imad_peak)}
    lib.call("la_set_gemm_mode", 2)                                 # default: exact byte-limb products on the INT8 tensor cores
    ms = timed(la, lambda: lib.call("la_mul_mat_mat", 5, Z.ptr, X.ptr, Y.ptr, k, k, k, None), 5, 2)
    tops = 2.0 * k ** 3 / ms / 1e9
    out["gemm_i32_8192^2"] = {"ms": round(ms, 3), "Top/s": round(tops, 2), "engine": "int8-limb tcgen05 (10 u8 x u8 MMAs per 32-bit product, bit-exact mod 2^32)",
                              "roofline": {"bound": "tensor", "achieved": round(10 * tops, 1), "peak": 4500.0, "unit": "Top/s",
                                           "frac": round(10 * tops / 4500.0, 3), "frac_of_2x_measured_bf16": round(10 * tops / int8_peak, 3),
                                           "note": "achieved = int8 MMA work issued (10 u8 x u8 MMAs per 32-bit product, pre-pass in the time); peak = NOMINAL dense int8 rate (4.5 Pop/s: "
                                                   "MEASURED_PEAKS.json has no int8 figure); twice the measured bf16 burst (%.0f) is what the int8 pipe would give if it throttled like bf16 -- it does not quite" % int8_peak}}
    del X, Y, Z
    # f32 GEMM sweep (BASELINE.json configs[3]) on both engines (LA_GEMM_SIMT: FP32 FMA pipe; LA_GEMM_3XTF32: tf32
    # hi/lo split, 3 tcgen05 MMAs per product, chunked FP32 accumulation, same f32 bar; LA_GEMM_AUTO picks the latter)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    tf32_peak = peaks.get("bf16_tflops", 1590.0) / 2.0          # tf32 runs at half the bf16 rate
    sweep = {}
    for nn in (2048, 4096, 8192, 16384):
        P = la.Vector.uninitialized(nn * nn, "f32"); Q = la.Vector.uninitialized(nn * nn, "f32"); R = la.Vector.uninitialized(nn * nn, "f32")
        lib.call("la_synth", 8, P.ptr, 0x5EED0003, nn * nn, None); lib.call("la_synth", 8, Q.ptr, 0x5EED1003, nn * nn, None)   # U[-1,1), generated on the device
        row = {}
        for mode, label in ((0, "simt"), (1, "3xtf32")):
            lib.call("la_set_gemm_mode", mode)
            st = timed_stats(la, lambda: lib.call("la_mul_mat_mat", 8, R.ptr, P.ptr, Q.ptr, nn, nn, nn, None), 10, 3)
            ms = st["median"]
            row[label] = {"ms": round(ms, 4), "best_ms": round(st["best"], 4), "TFLOP/s": round(2.0 * nn ** 3 / ms / 1e9, 2), "launches": st["n"]}
        # steady state: B prepared once (la_gemm_prepare_b), bit-identical products
        lib.call("la_set_gemm_mode", 1)
        hnd = ctypes.c_void_p()
        lib.call("la_gemm_prepare_b", 8, Q.ptr, nn, nn, None, ctypes.byref(hnd))
        st = timed_stats(la, lambda: lib.call("la_mul_mat_mat_prepared", R.ptr, P.ptr, hnd, nn, None), 10, 3)
        lib.call("la_gemm_prepared_free", hnd)
        row["3xtf32_prepared_b"] = {"ms": round(st["median"], 4), "best_ms": round(st["best"], 4), "TFLOP/s": round(2.0 * nn ** 3 / st["median"] / 1e9, 2)}
        lib.call("la_set_gemm_mode", 2)
        units = 1.5 if nn >= 1024 else 3       # K >= 1024: three fp16 MMAs (half a tf32 MMA's time each) per product; below: 3 tf32 MMAs
        row["3xtf32"]["roofline"] = {"bound": "tensor", "achieved": round(units * row["3xtf32"]["TFLOP/s"], 1), "peak": round(tf32_peak, 1), "unit": "TFLOP/s",
                                      "frac": round(units * row["3xtf32"]["TFLOP/s"] / tf32_peak, 3),
                                      "note": "achieved = tensor-core work issued in tf32-equivalent flops (%.1f per useful f32 flop; pre-pass included in the time); peak = measured bf16 / 2" % units}
        sweep[str(nn)] = row
        del P, Q, R
    out["gemm_f32_sweep"] = sweep
    # C0 (BASELINE.json configs[0], the reference's CPU-runnable case): 1024^2 f32 GEMM, 2^20-element add and dot on the GPU;
    # the same three through the reference's kernels on the host cores are in cpu_baseline.c0
    c0 = {}
    n0 = 1024; v0 = 1 << 20
    P = la.Vector.synth(0x5EED0010, n0 * n0, "f32"); Q = la.Vector.synth(0x5EED1010, n0 * n0, "f32"); R = la.Vector.uninitialized(n0 * n0, "f32")
    for mode, label in ((2, "auto"), (0, "simt")):
        lib.call("la_set_gemm_mode", mode)
        st = timed_stats(la, lambda: lib.call("la_mul_mat_mat", 8, R.ptr, P.ptr, Q.ptr, n0, n0, n0, None), 20, 3)
        c0["gemm_f32_1024^2_" + label] = {"ms": round(st["median"], 4), "best_ms": round(st["best"], 4), "GFLOP/s": round(2.0 * n0 ** 3 / st["median"] / 1e6, 1)}
    lib.call("la_set_gemm_mode", 2)
    st = timed_stats(la, lambda: lib.call("la_ew_vec_vec", 8, 0, R.ptr, P.ptr, Q.ptr, v0, None), 20, 3)
    c0["add_f32_2^20"] = {"ms": round(st["median"], 5), "best_ms": round(st["best"], 5), "GB/s": round(12.0 * v0 / st["median"] / 1e6, 1), "note": "12 MB: L2-resident, launch-bound"}
    res0 = la.Vector.uninitialized(4, "f32")
    st = timed_stats(la, lambda: lib.call("la_dot_vec_vec_dev", 8, P.ptr, Q.ptr, v0, res0.ptr, None), 20, 3)
    c0["dot_f32_2^20"] = {"ms": round(st["median"], 5), "best_ms": round(st["best"], 5), "GB/s": round(8.0 * v0 / st["median"] / 1e6, 1)}
    hostres = np.zeros(1, np.float32)
    w0 = time.perf_counter()
    for _ in range(50):
        lib.call("la_dot_vec_vec", 8, P.ptr, Q.ptr, v0, hostres.ctypes.data, None)
    c0["dot_f32_2^20"]["blocking_call_ms"] = round((time.perf_counter() - w0) * 1e3 / 50, 5)      # launch + scalar back on the host, as the reference's dot() returns it
    out["c0"] = c0
    del P, Q, R
    # the reference's own OpenCL kernels on this very GPU (separate process: its OpenCL context stays apart from ours)
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "cl_bench.py")], capture_output=True, text=True, timeout=240)
        js = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        out["reference_opencl_on_this_gpu"] = json.loads(js[-1]) if js else {"unavailable": (r.stderr or "no output")[-300:]}
    except Exception as e:  # noqa: BLE001
        out["reference_opencl_on_this_gpu"] = {"unavailable": str(e)[:300]}
    out["_hbm_peak_GB/s"] = hbm_peak
    out["_hbm_peak_source"] = peak_src
    return out


def extras_multi(la, comm, hbm_peak):
    """BASELINE.json configs[2] on N GPUs: dot/sum over 2^28 f32 and both matrix-vector products at
    32768^2, row-block sharded (linear_algebra_b200/dist.py).  Times include the exchange and the host
    round trip of the scalar results; max over ranks; GB/s = whole-job algorithmic bytes / time."""
    import torch
    import torch.distributed as td
    from linear_algebra_b200 import dist as ladist
    out = {}
    world, rank = comm.world, comm.rank

    def wall(fn, iters):
        for _ in range(3):
            fn()
        la.synchronize(); comm.barrier()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        la.synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / iters
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        return float(t[0])

    def device_ms(fn, iters):
        """CUDA events around each call on the launching stream: the fused kernel alone (it includes the wait for the peers'
        partials), without the host's share of the round trip.  Max over ranks of the median."""
        la.synchronize(); comm.barrier()
        ms = []
        for _ in range(iters):
            e0, e1 = la.Event(), la.Event()
            e0.record(); fn(); e1.record()
            ms.append(e0.elapsed_ms(e1))
        t = torch.tensor([statistics.median(ms)], device="cuda", dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        return float(t[0])

    def rec(name, ms, nbytes, kernel_ms=None):
        gbs = nbytes / ms / 1e6
        out[name] = {"ms": round(ms, 4), "GB/s": round(gbs, 1), "frac_of_measured_hbm_x_n": round(gbs / (hbm_peak * world), 3)}
        if kernel_ms:
            out[name].update({"kernel_ms": round(kernel_ms, 4), "kernel_GB/s": round(nbytes / kernel_ms / 1e6, 1),
                              "kernel_frac_of_measured_hbm_x_n": round(nbytes / kernel_ms / 1e6 / (hbm_peak * world), 3)})

    n = 1 << 28
    lo, hi = ladist.partition(n, world, rank)
    A = ladist.ShardedVector(comm, la.Vector.synth(0x5EED0001 + rank, hi - lo, "f32"), n)
    B = ladist.ShardedVector(comm, la.Vector.synth(0x5EED1001 + rank, hi - lo, "f32"), n)
    rec("dot_f32_2^28_sharded", wall(lambda: ladist.dot(A, B), 20), 8 * n, device_ms(lambda: ladist.dot(A, B), 20))
    rec("sum_f32_2^28_sharded", wall(lambda: A.sum(), 20), 4 * n, device_ms(lambda: A.sum(), 20))
    rec("add_f32_2^28_sharded", wall(lambda: A + B, 20), 12 * n)
    del A, B
    # BASELINE.json configs[2]: dot / sum over the 2^30 elements of a 32768 x 32768 f32 operand, sharded
    n30 = 1 << 30
    lo, hi = ladist.partition(n30, world, rank)
    A = ladist.ShardedVector(comm, la.Vector.synth(0x5EED0005 + rank, hi - lo, "f32"), n30)
    B = ladist.ShardedVector(comm, la.Vector.synth(0x5EED1005 + rank, hi - lo, "f32"), n30)
    rec("dot_f32_2^30_sharded", wall(lambda: ladist.dot(A, B), 20), 8 * n30, device_ms(lambda: ladist.dot(A, B), 20))
    rec("sum_f32_2^30_sharded", wall(lambda: A.sum(), 20), 4 * n30, device_ms(lambda: A.sum(), 20))
    del A, B
    r = 32768
    rlo, rhi = ladist.partition(r, world, rank)
    M = ladist.ShardedMatrix(comm, la.Matrix(la.Vector.synth(0x5EED0002 + rank, (rhi - rlo) * r, "f32"), rhi - rlo, r), r)
    v = la.Vector.synth(0x5EED1002, r, "f32")
    sv = ladist.ShardedVector(comm, la.Vector.synth(0x5EED2002 + rank, rhi - rlo, "f32"), r)
    rec("mat_x_vec_f32_32768^2_sharded", wall(lambda: ladist.mul_transpose_mat(v, M), 10), 4 * (r * r + 2 * r))
    rec("vec_x_mat_f32_32768^2_sharded", wall(lambda: sv * M, 10), 4 * (r * r + 2 * r), device_ms(lambda: sv * M, 10))
    # BASELINE.json configs[3] on N GPUs: the square f32 GEMM sweep, A / C row-block sharded, B replicated (no exchange on the
    # data path): FP32 SIMT engine (the north star's), tensor engine cold and with B prepared.  Whole-job TFLOP/s, max over ranks.
    lib = la._lib

    def gemm_ms(fn, iters):
        for _ in range(3):
            fn()
        la.synchronize(); comm.barrier()
        e0, e1 = la.Event(), la.Event()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record(); la.synchronize()
        t = torch.tensor([e0.elapsed_ms(e1) / iters], device="cuda", dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        return float(t[0])

    sweep = {}
    for nn in (2048, 4096, 8192, 16384):
        rlo, rhi = ladist.partition(nn, world, rank)
        rows_nn = rhi - rlo
        P = la.Vector.synth(0x5EED0003 + rank, rows_nn * nn, "f32"); Q = la.Vector.synth(0x5EED1003, nn * nn, "f32"); R = la.Vector.uninitialized(rows_nn * nn, "f32")
        row = {"rows_per_gpu": rows_nn}
        for mode, label, iters in ((0, "simt", 5), (1, "tensor_cold", 10)):
            lib.call("la_set_gemm_mode", mode)
            ms = gemm_ms(lambda: lib.call("la_mul_mat_mat", 8, R.ptr, P.ptr, Q.ptr, rows_nn, nn, nn, None), iters)
            row[label] = {"ms": round(ms, 4), "TFLOP/s": round(2.0 * nn ** 3 / ms / 1e9, 1)}
        hnd = ctypes.c_void_p()
        lib.call("la_gemm_prepare_b", 8, Q.ptr, nn, nn, None, ctypes.byref(hnd))
        ms = gemm_ms(lambda: lib.call("la_mul_mat_mat_prepared", R.ptr, P.ptr, hnd, rows_nn, None), 10)
        lib.call("la_gemm_prepared_free", hnd)
        row["tensor_b_prepared"] = {"ms": round(ms, 4), "TFLOP/s": round(2.0 * nn ** 3 / ms / 1e9, 1)}
        lib.call("la_set_gemm_mode", 2)
        sweep[str(nn)] = row
        del P, Q, R
    out["gemm_f32_sweep_sharded"] = sweep
    out["_note"] = ("la_*_sharded entry points (csrc/comm.cu): exchange fused into the kernels -- partials stored into / read from the peers' cudaIpc-mapped buffers over NVLink, "
                    "rank-ordered (deterministic); `ms` = wall time of the blocking call with the scalar on the host (zero-copy word, no D2H copy), `kernel_ms` = CUDA events around the launch")
    return out


def parity_check_multi(la, comm):
    """Before anything is timed on N > 1 GPUs: the sharded paths (fused NVLink reductions, v x M all-reduce, sharded transpose,
    row-block GEMM) against the CPU oracle on small seeded inputs -- integers bit-exact, f32 inside the gate of
    tests/test_gpu_parity.py.  The oracle is the checker here, never the thing measured.  Returns {check: bool}, each the AND
    over all ranks."""
    import oracle
    from linear_algebra_b200 import dist as ladist
    P = oracle.port()
    world, rank = comm.world, comm.rank
    rng = np.random.default_rng(4242)                       # same stream on every rank: every rank holds the global operands
    res = {}
    n = (1 << 20) + 37
    a = rng.integers(-2 ** 31, 2 ** 31 - 1, n, dtype=np.int32); b = rng.integers(-2 ** 31, 2 ** 31 - 1, n, dtype=np.int32)
    A = ladist.ShardedVector.from_global(comm, a); B = ladist.ShardedVector.from_global(comm, b)
    res["sum_i32_bit_exact"] = int(A.sum()) == int(P.sum_vec(a))
    res["dot_i32_bit_exact"] = int(ladist.dot(A, B)) == int(P.dot_vec_vec(a, b))
    x = rng.uniform(-1, 1, n).astype(np.float32); y = rng.uniform(-1, 1, n).astype(np.float32)
    X = ladist.ShardedVector.from_global(comm, x); Y = ladist.ShardedVector.from_global(comm, y)
    d64, ad = P.f32_dot_truth(x, y)
    bar = max(abs(float(P.dot_vec_vec(x, y)) - d64), abs(float(P.dot_vec_vec(x, y, fma=True)) - d64), 2.0 ** -22 * ad)
    res["dot_f32_in_gate"] = abs(float(ladist.dot(X, Y)) - d64) <= bar
    rows, cols = 512 * world + 12, 2048
    m = rng.integers(-2 ** 31, 2 ** 31 - 1, rows * cols, dtype=np.int32); v = rng.integers(-2 ** 31, 2 ** 31 - 1, rows, dtype=np.int32)
    SM = ladist.ShardedMatrix.from_global(comm, m, rows, cols); SV = ladist.ShardedVector.from_global(comm, v)
    res["vec_x_mat_i32_bit_exact"] = (SV * SM).to_vec().tobytes() == P.mul_vec_mat(v, m, rows, cols).tobytes()
    vc = rng.integers(-2 ** 31, 2 ** 31 - 1, cols, dtype=np.int32)
    mv = ladist.mul_transpose_mat(la.Vector.from_vec(vc), SM)
    res["mat_x_vec_i32_bit_exact"] = mv.gather().tobytes() == P.mul_vec_transpose_mat(vc, m, rows, cols).tobytes()
    res["transpose_i32_bit_exact"] = ladist.transpose(SM).gather().tobytes() == P.transpose(m, rows, cols).tobytes()
    Mg, K, N = 256 * world, 1024, 512
    ga = rng.uniform(-1, 1, Mg * K).astype(np.float32); gb = rng.uniform(-1, 1, K * N).astype(np.float32)
    lo, hi = ladist.partition(Mg, world, rank)
    mine = ga.reshape(Mg, K)[lo:hi].reshape(-1)
    Cs = (la.Matrix.from_vec(mine, hi - lo, K) * la.Matrix.from_vec(gb, K, N)).to_vec().astype(np.float64)
    truth, absab = P.f32_mul_mat_mat_truth(mine, gb, hi - lo, K, N)
    e_ref = max((np.abs(P.mul_mat_mat(mine, gb, hi - lo, K, N).astype(np.float64) - truth) / absab).max(),
                (np.abs(P.mul_mat_mat(mine, gb, hi - lo, K, N, fma=True).astype(np.float64) - truth) / absab).max())
    res["gemm_f32_row_block_in_gate"] = float((np.abs(Cs - truth) / absab).max()) <= max(e_ref, 2.0 ** -22)
    hc = np.empty((hi - lo) * N, np.float32)
    ladist.mul_host(comm, mine, gb, Mg, K, N, out=hc)
    res["mul_host_f32_row_block_in_gate"] = float((np.abs(hc.astype(np.float64) - truth) / absab).max()) <= max(e_ref, 2.0 ** -22)
    res["nvlink_peer_path_active"] = bool(getattr(comm, "_p2p", {}).get("ok"))
    keys = sorted(res)
    flags = comm.all_gather_host(bytes(1 if res[k] else 0 for k in keys))
    return {k: all(f[i] == 1 for f in flags) for i, k in enumerate(keys)}


def ncu_traffic(path, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of `kernel` per launch, from the committed ncu summary."""
    try:
        rd = wr = None
        hit = False
        for ln in open(os.path.join(ROOT, path)):
            if ln.startswith("== "):
                if hit:
                    break
                hit = kernel in ln
            elif hit and "dram__bytes_read.sum " in ln:
                v, u = ln.split()[1:3]; rd = float(v) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[u]
            elif hit and "dram__bytes_write.sum " in ln:
                v, u = ln.split()[1:3]; wr = float(v) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[u]
        return None if rd is None or wr is None else {"dram_bytes_per_launch": int(rd + wr), "source": path}
    except OSError:
        return None


def run_ours(args):
    import linear_algebra_b200 as la
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sampler = ClockSampler(local); sampler.start()
    numa_note = bind_near_gpu(local) if os.environ.get("LA_BENCH_BIND", "1") != "0" else "not bound (LA_BENCH_BIND=0)"
    la.init(local)
    lib = la._lib
    parity = None
    comm = None
    if dist is not None:
        from linear_algebra_b200 import dist as ladist
        comm = ladist.Comm()
        parity = parity_check_multi(la, comm)              # outside every timed region
        if not all(parity.values()):
            if rank == 0:
                print(json.dumps({"metric": METRIC, "error": "multi-GPU parity check failed; nothing was timed", "parity_checked": parity, "n_gpus": world}))
            sampler.stop()
            dist.barrier(); dist.destroy_process_group()
            sys.exit(1)
    hbm_peak, sm_max_mhz, peak_src = measured_peaks()
    sm_count = __import__("ctypes").c_int(); lib.call("la_sm_count", __import__("ctypes").byref(sm_count))

    n = N_GEMM
    rows = n // world                              # row-block shard of A and C; B replicated
    rng = np.random.default_rng(1000 + rank)
    hA = la.pinned(rows * n, "f32"); hB = la.pinned(n * n, "f32"); hC = la.pinned(rows * n, "f32")
    hA[:] = rng.uniform(-1, 1, rows * n).astype(np.float32)
    hB[:] = np.random.default_rng(999).uniform(-1, 1, n * n).astype(np.float32)
    A = la.Matrix.from_vec(hA, rows, n); B = la.Matrix.from_vec(hB, n, n)
    C = la.Matrix.uninitialized(rows, n, "f32")

    mode = {"auto": lib.GEMM_AUTO, "simt": lib.GEMM_SIMT, "3xtf32": lib.GEMM_3XTF32}[args.engine]
    lib.call("la_set_gemm_mode", mode)

    def step():
        lib.call("la_mul_mat_mat", 8, C.data.ptr, A.data.ptr, B.data.ptr, rows, n, n, None)

    def barrier():
        la.synchronize()
        if dist is not None:
            dist.barrier()
        la.synchronize()

    wait_until = time.time() + 3.0                  # nvidia-smi (started at the top of this function) needs a moment before its first line;
    while not sampler.samples and time.time() < wait_until:      # wait for it with the GPU IDLE, so that the W warm-up steps below are all
        time.sleep(0.02)                                          # the load the GPU has seen when the timed region starts
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    import ctypes as _ct
    # Steady state of the sharded product ("B broadcast once; report steady-state and cold separately"):
    # B is replicated and does not change between products, so its share of the pre-pass is done once (la_gemm_prepare_b)
    # and every step is la_mul_mat_mat_prepared -- bit-identical results.  `value` stays the COLD figure and is timed first.
    hnd = _ct.c_void_p()
    lib.call("la_gemm_prepare_b", 8, B.data.ptr, n, n, None, _ct.byref(hnd))

    def steady_step():
        lib.call("la_mul_mat_mat_prepared", C.data.ptr, A.data.ptr, hnd, rows, None)

    def time_steady():
        for _ in range(3):
            steady_step()
        barrier()
        lib.lib.la_gemm_kernel_timing(1)
        s0, s1 = la.Event(), la.Event()
        s0.record()
        for _ in range(args.steps):
            steady_step()
        s1.record()
        total = s0.elapsed_ms(s1)
        kms, kn = _ct.c_float(0), _ct.c_int(0)
        lib.lib.la_gemm_kernel_time_ms(_ct.byref(kms), _ct.byref(kn))
        lib.lib.la_gemm_kernel_timing(0)
        barrier()
        return total, ((kms.value / kn.value) if kn.value else None)

    t0 = time.time()
    e0, e1 = la.Event(), la.Event()
    lib.lib.la_gemm_kernel_timing(1)                # brackets every launch of the dominant (tensor GEMM) kernel with events on its stream
    step_ev = [la.Event() for _ in range(args.steps)]
    e0.record()
    for i in range(args.steps):
        step()
        step_ev[i].record()
    e1.record()
    ms_total = e0.elapsed_ms(e1)
    per_step = [(e0 if i == 0 else step_ev[i - 1]).elapsed_ms(step_ev[i]) for i in range(args.steps)]
    _kms, _kn = _ct.c_float(0), _ct.c_int(0)
    lib.lib.la_gemm_kernel_time_ms(_ct.byref(_kms), _ct.byref(_kn))
    lib.lib.la_gemm_kernel_timing(0)
    kernel_ms = (_kms.value / _kn.value) if _kn.value else None      # average duration of the dominant kernel inside the timed region
    barrier()
    t1 = time.time()

    engine_used = lib.lib.la_gemm_last_engine()     # 1 = TMA SIMT, 2 = tcgen05 split-f32
    tc_form = lib.lib.la_gemm_tc_last_form()        # of the cold step just timed -- 1: f16-split form (three fp16 MMAs per product), 0: tf32 form (three tf32 MMAs)
    tc_launches = lib.lib.la_gemm_tc_last_launches()

    time.sleep(1.0)                                     # same protocol as the headline region: GPU idle, then W warm-up steps, then K timed steps
    steady_total, steady_kernel_ms = time_steady()
    t2 = time.time()
    lib.call("la_gemm_prepared_free", hnd)
    clocks = sampler.window(t0, t1)                  # the headline (cold) timed region
    if clocks["samples"] and t1 - t0 < 0.2:
        clocks["note"] = "the timed region lasts %.0f ms, sampling period 100 ms: samples are from -50 / +150 ms around it" % ((t1 - t0) * 1e3)

    # e2e: public API on HOST buffers; H2D of A and B and D2H of C inside the timed region.
    def e2e_step():                                 # one streamed call (N > 1: each rank uploads 1/N of B, NVLink all_gather)
        if comm is None:
            la.mul_host(hA, hB, rows, n, n, out=hC)
        else:
            ladist.mul_host(comm, hA, hB, n, n, n, out=hC)

    def three_call_step():                          # the reference user's sequence
        a = la.Matrix.from_vec(hA, rows, n); b = la.Matrix.from_vec(hB, n, n)
        (a * b).to_vec(out=hC)

    def wall(fn, iters):
        for _ in range(2):
            fn()
        barrier()
        w0 = time.perf_counter()
        for _ in range(iters):
            fn()
        la.synchronize()
        return (time.perf_counter() - w0) * 1e3    # host-blocking calls: the wall clock is the honest timer

    e2e_steps = max(3, min(args.steps, 10))
    e2e_ms = wall(e2e_step, e2e_steps)
    three_ms = wall(three_call_step, e2e_steps)
    barrier()
    sampler.stop()

    if dist is not None:
        import torch
        t = torch.tensor([ms_total, e2e_ms, three_ms, steady_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, e2e_ms, three_ms, steady_total = (float(x) for x in t)
    ms_step = ms_total / args.steps
    steady_step_ms = steady_total / args.steps
    flop = 2.0 * n * n * n                         # whole job: all ranks' row blocks together
    value = flop / ms_step / 1e6
    e2e_value = flop / (e2e_ms / e2e_steps) / 1e6
    three_value = flop / (three_ms / e2e_steps) / 1e6

    if rank == 0:
        simt_peak = sm_count.value * 128 * 2 * sm_max_mhz * 1e6 / 1e12        # TFLOP/s at max clock
        per_gpu_tflops = value / world / 1e3
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        if engine_used == 2:
            mma16_peak = peaks.get("bf16_tflops", 1590.0)       # 16-bit MMA rate (fp16 = bf16); tf32 runs at half of it
            mma16_sustained = peaks.get("bf16_tflops_sustained", 1400.0)
            tf32_peak = mma16_peak / 2.0
            # tensor-core work per useful f32 flop, in 16-bit MMA flops: f16-split form 3 (hi*hi, hs*ls, ls*hs, all fp16); tf32 form 3 tf32 MMAs = 6
            work = 3.0 if tc_form == 1 else 6.0
            roofline = {"bound": "tensor", "achieved": round(work * per_gpu_tflops, 1), "peak": round(mma16_peak, 1), "unit": "TFLOP/s",
                        "frac": round(work * per_gpu_tflops / mma16_peak, 4), "frac_of_sustained_peak": round(work * per_gpu_tflops / mma16_sustained, 4),
                        "peak_source": "measured (MEASURED_PEAKS.json): bf16 burst %.1f (fp16 MMAs run at the bf16 rate, tf32 at half of it)" % mma16_peak,
                        "achieved_note": ("tensor-core work per second in 16-bit MMA flops: per useful f32 product three fp16 x fp16 MMAs (hi*hi, hs*ls, ls*hs) = 3 units; the pre-pass is inside the time"
                                          if tc_form == 1 else "tensor-core work per second in 16-bit-equivalent MMA flops: per useful f32 product three tf32 MMAs at half the 16-bit rate = 6 units; the pre-pass is inside the time"),
                        "useful_f32_frac_of_tf32_peak": round(per_gpu_tflops / tf32_peak, 4),
                        "kernel_only": (None if not kernel_ms else {
                            "kernel": "gemm_f16split_2cta_kernel" if tc_form == 1 else "gemm_3xtf32_2cta_kernel", "avg_launch_ms": round(kernel_ms, 4), "launches_timed": _kn.value,
                            "share_of_step": round(kernel_ms / ms_step, 4),
                            "achieved": round(work * 2.0 * (n / world) * n * n / kernel_ms / 1e9, 1), "frac": round(work * 2.0 * (n / world) * n * n / kernel_ms / 1e9 / mma16_peak, 4),
                            "algorithmic_TFLOP/s": round(2.0 * (n / world) * n * n / kernel_ms / 1e9, 1),
                            "note": "CUDA events around each launch of the kernel on its own stream, inside the timed region (rank 0)"}),
                        "traffic": ncu_traffic("profiles/r2_gemm_f16split_8192_ncu_full.txt", "gemm_f16split_2cta_kernel")}
            engine_name = ("tcgen05 split-f32, f16 form (rows of A / columns of B scaled by powers of two; hi = fp16, residual lifted by 2^6 into fp16; three fp16 MMAs per product; cta_group::2 CTA pairs, TMA, chunked FP32 accumulation; same f32 gate as SIMT)"
                           if tc_form == 1 else "tcgen05 split-f32, tf32 form (three tf32 MMAs per product)")
            launches = tc_launches                              # pre-pass (2 launches), product, fix-ups of left-out elements (0 - 2) per cold step
        else:
            roofline = {"bound": "fp32-fma (SIMT; MEASURED_PEAKS.json has no FP32 figure)", "achieved": round(per_gpu_tflops, 2),
                        "peak": round(simt_peak, 2), "unit": "TFLOP/s", "frac": round(per_gpu_tflops / simt_peak, 4),
                        "peak_source": "nominal: %d SMs x 128 lanes x 2 x %.0f MHz" % (sm_count.value, sm_max_mhz),
                        "frac_at_observed_clock": (round(per_gpu_tflops / (sm_count.value * 256 * clocks["sm_mhz"] * 1e6 / 1e12), 4) if clocks.get("sm_mhz") else None),
                        "traffic": ncu_traffic("profiles/r1_gemm_f32_8192_ncu_full.txt", "gemm_tma_kernel")}
            engine_name = "TMA-fed FP32 SIMT (FFMA2), A^T staged per call"
            launches = 2
        roofline["timing"] = {"per_step_ms_best": round(min(per_step), 4), "per_step_ms_median": round(statistics.median(per_step), 4), "steps": args.steps,
                              "note": "rank 0, an event after every step; ms_per_step / value use the whole K-step region (max over ranks)"}
        steady_value = flop / steady_step_ms / 1e6
        roofline["steady_prepared_b"] = {
            "value": round(steady_value, 1), "unit": UNIT, "ms_per_step": round(steady_step_ms, 4),
            "timed": "after the headline region and one idle second, with the same protocol (3 warm-up steps, then K timed steps): within ~50 ms of tensor load the GPU settles into a lower power state, so two regions are only comparable when both start from idle",
            "frac": (round(work * steady_value / world / 1e3 / mma16_peak, 4) if engine_used == 2 else round(steady_value / world / 1e3 / simt_peak, 4)),
            "kernel_share_of_step": (round(steady_kernel_ms / steady_step_ms, 4) if steady_kernel_ms else None),
            "note": "same products through la_gemm_prepare_b (once, outside) + la_mul_mat_mat_prepared: the replicated B's split/transpose planes are reused, results bit-identical; `value` of this line is the COLD figure with B's pre-pass inside every step"}
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32" if engine_used != 2 else "f32 (each product as hi*hi + hi*lo + lo*hi of an fp16 hi / lifted-residual split on the tensor cores, FP32 accumulate; gate: every output's error <= max(worst error of the reference's own k-ascending order on the same problem, 2^-22 sum|a||b|))",
            "data": "synthetic",
            "config": {"workload": "f32 GEMM 8192x8192x8192, C = A x B, A/C row-block sharded over %d GPU(s), B replicated" % world,
                       "M": n, "K": n, "N": n, "rows_per_gpu": rows, "l2": "operands 768 MB >> 126 MB L2 (no flush needed)",
                       "engine": engine_name, "engine_flag": args.engine},
            "clocks": clocks,
            "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "h2d_bytes_per_step": int((rows * n + n * n // world) * 4), "d2h_bytes_per_step": int(rows * n * 4),
                    "bytes_note": "per rank: its rows of A and its 1/N of B up (the other slices of B arrive over NVLink), its rows of C down", "host": numa_note,
                    "steps": e2e_steps, "api": "mul_host(pinned A, pinned B) -> pinned C  (la_mul_mat_mat_host: row-block streaming, copies overlap the multiply)",
                    "three_call": {"value": round(three_value, 1), "api": "Matrix.from_vec(pinned) x2 -> A * B -> to_vec(pinned)"}},
            "gpu_launches": launches * args.steps,
            "roofline": roofline,
        }
        if world == 1:
            if not args.no_extra:
                ex = extras(la, hbm_peak, peak_src)
                line["extra"] = ex
                # the whole metric in keys the driver's record keeps: vector-op HBM figures, SIMT engine, sweep, C0
                short = {"ew_add_f32_2^28": "add", "ew_sub_f32_2^28": "sub", "ew_mul_f32_2^28": "mul", "ew_add_assign_f32_2^28": "add_assign", "scale_f32_2^28": "scale", "eq_f32_2^28": "eq", "clone_f32_2^28": "clone", "fill_f32_2^28": "fill",
                         "dot_f32_2^28": "dot", "sum_f32_2^28": "sum", "dot_f32_2^30": "dot_2^30", "sum_f32_2^30": "sum_2^30",
                         "mat_x_vec_f32_32768^2": "matvec", "vec_x_mat_f32_32768^2": "vecmat", "transpose_i32_8192^2": "transpose_i32",
                         "ew_add_i32_8192^2": "i32_add", "ew_sub_i32_8192^2": "i32_sub", "ew_mul_i32_8192^2": "i32_mul"}
                roofline["hbm"] = {"_columns": ["GB/s (median sample)", "frac of measured copy peak %.1f" % hbm_peak, "GB/s (best sample)"],
                                   **{v: [ex[k]["GB/s"], ex[k]["frac_of_measured_hbm"], ex[k]["best_GB/s"]] for k, v in short.items() if k in ex}}
                sw = ex.get("gemm_f32_sweep", {})
                if "8192" in sw:
                    st_ = sw["8192"]["simt"]
                    roofline["simt"] = {"ms": st_["ms"], "TFLOP/s": st_["TFLOP/s"], "frac_of_nominal": round(st_["TFLOP/s"] / simt_peak, 3),
                                        "frac_of_measured_ffma_loop": round(st_["TFLOP/s"] / 72.3, 3),
                                        "note": "north-star engine (TMA-fed FP32 SIMT, FFMA2) at 8192^3; nominal = SMs x 128 x 2 x max clock; 72.3 TFLOP/s = register-resident FFMA loop on this GPU (tools/pipe_peaks.cu)"}
                roofline["sweep_f32"] = {"_columns": ["simt TFLOP/s", "tensor TFLOP/s (cold)", "tensor TFLOP/s (B prepared)", "tensor frac of measured peak (cold)"],
                                         **{k: [v["simt"]["TFLOP/s"], v["3xtf32"]["TFLOP/s"], v["3xtf32_prepared_b"]["TFLOP/s"], v["3xtf32"]["roofline"]["frac"]] for k, v in sw.items()}}
                if "gemm_i32_8192^2" in ex:
                    roofline["i32_gemm_8192"] = {"Top/s": ex["gemm_i32_8192^2"]["Top/s"], "frac_of_nominal_int8": ex["gemm_i32_8192^2"]["roofline"]["frac"], "frac_of_2x_measured_bf16": ex["gemm_i32_8192^2"]["roofline"]["frac_of_2x_measured_bf16"],
                                                 "imad_Top/s": ex["gemm_i32_8192^2_imad"]["Top/s"], "imad_frac_of_nominal": ex["gemm_i32_8192^2_imad"]["frac_of_nominal_int32_mad_peak"]}
                roofline["c0_gpu"] = ex.get("c0")
            cb = time_reference_gemm(12.0)
            cb.pop("seconds", None)
            cb["c0"] = time_reference_c0()
            line["cpu_baseline"] = cb
    if world > 1 and not args.no_extra:
        em = extras_multi(la, comm, hbm_peak)
        if rank == 0:
            line["extra"] = em
            line["roofline"]["hbm_sharded"] = {"_columns": ["GB/s whole job (wall, scalar on the host)", "frac of N x measured copy peak", "GB/s kernel only (events)", "kernel frac"],
                                               **{k: [v["GB/s"], v["frac_of_measured_hbm_x_n"]] + ([v["kernel_GB/s"], v["kernel_frac_of_measured_hbm_x_n"]] if "kernel_GB/s" in v else []) for k, v in em.items() if isinstance(v, dict) and "GB/s" in v}}
            sw = em.get("gemm_f32_sweep_sharded", {})
            line["roofline"]["sweep_f32_sharded"] = {"_columns": ["simt TFLOP/s (whole job)", "simt frac of N x nominal FP32 peak", "tensor TFLOP/s (cold)", "tensor TFLOP/s (B prepared)"],
                                                     "_note": "BASELINE.json configs[3] on %d GPUs (A / C row-block sharded, B replicated); taken after seconds of load: a GPU settles 5-10 %% below its first-50-ms rate" % world,
                                                     **{k: [v["simt"]["TFLOP/s"], round(v["simt"]["TFLOP/s"] / (simt_peak * world), 3), v["tensor_cold"]["TFLOP/s"], v["tensor_b_prepared"]["TFLOP/s"]] for k, v in sw.items()}}
    if rank == 0 and parity is not None:
        line["parity_checked"] = parity
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--engine", default="auto", choices=["auto", "simt", "3xtf32"], help="f32 GEMM engine for the timed steps (auto = library default)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
