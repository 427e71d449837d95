// oracle/cl_shim.hpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// A ~60-line OpenCL-C -> C++ shim that lets the reference's own kernel source
// (/root/reference/src/kernels.cl, text-templated on {T} exactly the way
// /root/reference/src/lib.rs:337-354 `get_src` does it) compile with g++ and run on host
// cores.  oracle/build_ref.py pastes the templated text between `cl_shim.hpp` and
// `cl_driver.hpp`; nothing from the reference is stored in this repository.
//
// Execution model: one OpenCL work-item == one call of the kernel function with the
// thread-local `cl_wi` describing its ids.  Barrier-free kernels run in an OpenMP loop
// over work-items.  The two barrier kernels (sum_vec / dot_vec_vec, kernels.cl:83-124) run
// VERBATIM with one OpenMP team per work-group and `barrier()` mapped to `omp barrier`.
#pragma once
#include <cstdint>
#include <cstddef>
#include <omp.h>

typedef unsigned char  uchar;
typedef unsigned short ushort;
typedef unsigned int   uint;
typedef unsigned long  ulong;   // LP64: 64 bit, as OpenCL's ulong
static_assert(sizeof(long) == 8 && sizeof(ulong) == 8, "OpenCL long/ulong are 64-bit");

struct cl_work_item { size_t gid, lid, wgid, lsz, gsz; };
static thread_local cl_work_item cl_wi;

#define kernel static
#define global
#define local
#define CLK_LOCAL_MEM_FENCE 0
#define get_global_id(d)   ((int)cl_wi.gid)      /* kernels.cl does its index math in int */
#define get_local_id(d)    ((int)cl_wi.lid)
#define get_group_id(d)    ((int)cl_wi.wgid)
#define get_local_size(d)  ((int)cl_wi.lsz)
#define get_global_size(d) ((int)cl_wi.gsz)
#define barrier(flags)     _Pragma("omp barrier")
