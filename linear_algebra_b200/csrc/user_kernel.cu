// csrc/user_kernel.cu -- user-supplied elementwise kernels, compiled at run time.
//
// Replaces the reference's extension point: `create_kernel` (src/lib.rs:156-171), the text of
// src/extra_kernels.cl appended to the program source and templated on {T} (`load_extra_src`
// src/lib.rs:173-180, `get_src` :337-354), and the Vector helpers that launch such a kernel with
// one work-item per element (src/vector.rs:80-151: generate, from_for_each2, for_each_mut,
// map_mut, map).  There the OpenCL driver compiles the text; here NVRTC compiles it for sm_100a.
// The text may stay in the OpenCL-C dialect the reference documents: `kernel`, `global`, `local`,
// `constant` are rewritten to their CUDA spellings, a prelude maps the work-item functions,
// `barrier`, the unsigned type names and the reference's `i` (= get_global_id(0),
// src/kernels.cl:14) onto CUDA, and NVRTC's
// "unannotated functions are device functions" mode takes care of plain helper functions.
//
// NVRTC is opened with dlopen at first use, so the rest of the library loads on a machine
// without it; asking for a user kernel there is an error, never a fallback.
#include "common.cuh"
#include <nvrtc.h>
#include <dlfcn.h>
#include <cctype>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace {

struct Nvrtc {
    decltype(&nvrtcCreateProgram) create = nullptr;
    decltype(&nvrtcCompileProgram) compile = nullptr;
    decltype(&nvrtcGetCUBINSize) cubin_size = nullptr;
    decltype(&nvrtcGetCUBIN) cubin = nullptr;
    decltype(&nvrtcGetProgramLogSize) log_size = nullptr;
    decltype(&nvrtcGetProgramLog) log = nullptr;
    decltype(&nvrtcDestroyProgram) destroy = nullptr;
    decltype(&nvrtcGetErrorString) error_string = nullptr;
};

la_status get_nvrtc(const Nvrtc** out) {
    static Nvrtc api;
    static bool ready = false;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (!ready) {
        const char* names[] = {"libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so"};
        void* h = nullptr;
        for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_LOCAL); if (h) break; }
        if (!h) return la_fail(LA_ERR_UNSUPPORTED, "user kernels need NVRTC (libnvrtc.so.12), which could not be loaded: %s", dlerror());
#define LA_NVRTC_SYM(field, sym)                                                                          \
        api.field = (decltype(api.field))dlsym(h, #sym);                                                   \
        if (!api.field) return la_fail(LA_ERR_UNSUPPORTED, "libnvrtc lacks %s", #sym);
        LA_NVRTC_SYM(create, nvrtcCreateProgram)
        LA_NVRTC_SYM(compile, nvrtcCompileProgram)
        LA_NVRTC_SYM(cubin_size, nvrtcGetCUBINSize)
        LA_NVRTC_SYM(cubin, nvrtcGetCUBIN)
        LA_NVRTC_SYM(log_size, nvrtcGetProgramLogSize)
        LA_NVRTC_SYM(log, nvrtcGetProgramLog)
        LA_NVRTC_SYM(destroy, nvrtcDestroyProgram)
        LA_NVRTC_SYM(error_string, nvrtcGetErrorString)
#undef LA_NVRTC_SYM
        ready = true;
    }
    *out = &api;
    return LA_OK;
}

// The {T} spellings of the reference (src/traits.rs:30-57), with i8 as a true signed char
// (the reference's "uchar" for i8 is reachable by passing LA_U8, as the Python mirror does).
const char* type_name(int dtype) {
    static const char* names[] = {"uchar", "char", "ushort", "short", "uint", "int", "ulong", "long", "float", "double"};
    return la_valid_dtype(dtype) ? names[dtype] : nullptr;
}

// What stands where the reference pastes its own kernel source: the dialect mapping, the launch
// state, `i`, and the one helper of the core source a user kernel may call ({T}_dot_product,
// src/kernels.cl:134-140).  {T} is substituted in here as well.
const char* const PRELUDE = R"LA(
typedef unsigned char  uchar;
typedef unsigned short ushort;
typedef unsigned int   uint;
typedef unsigned long  ulong;
static_assert(sizeof(long) == 8 && sizeof(ulong) == 8, "OpenCL long and ulong are 64 bits wide");

__device__ unsigned long la_uk_first;     // global id of thread 0 of block 0 of the running launch
__device__ unsigned long la_uk_count;     // the NDRange's global size
extern "C" __global__ void la_uk_set_state(unsigned long first, unsigned long count) { la_uk_first = first; la_uk_count = count; }

#define CLK_LOCAL_MEM_FENCE  1
#define CLK_GLOBAL_MEM_FENCE 2
#define barrier(flags)       __syncthreads()
#define get_work_dim()       1u
#define get_global_id(d)     ((d) == 0 ? (size_t)(la_uk_first + (unsigned long)blockIdx.x * blockDim.x + threadIdx.x) : (size_t)0)
#define get_global_size(d)   ((d) == 0 ? (size_t)la_uk_count : (size_t)1)
#define get_local_id(d)      ((d) == 0 ? (size_t)threadIdx.x : (size_t)0)
#define get_local_size(d)    ((d) == 0 ? (size_t)blockDim.x : (size_t)1)
#define get_group_id(d)      ((d) == 0 ? (size_t)(la_uk_first / 256u + blockIdx.x) : (size_t)0)
#define get_num_groups(d)    ((d) == 0 ? (size_t)((la_uk_count + 255u) / 256u) : (size_t)1)
#define mad(a, b, c)         ((a) * (b) + (c))
#define native_exp(x)  __expf(x)
#define native_log(x)  __logf(x)
#define native_sin(x)  __sinf(x)
#define native_cos(x)  __cosf(x)
#define native_sqrt(x) sqrtf(x)
#define native_recip(x) __frcp_rn(x)
template <typename X> __device__ inline X clamp(X x, X lo, X hi) { return x < lo ? lo : (x > hi ? hi : x); }

inline {T} {T}_dot_product(const {T}* a, int stride_a, const {T}* b, int stride_b, int count) {
    {T} total = 0;
    for (int k = 0; k < count; ++k) total += a[k * stride_a] * b[k * stride_b];
    return total;
}

#define i get_global_id(0)
#line 1 "extra_kernels.cl"
)LA";

// The address-space and kernel qualifiers of OpenCL C cannot be macros here: `__global__` and `__constant__` are
// themselves spelled through the identifiers `global` and `constant` inside NVRTC's front end, so a `#define global`
// silently turns every kernel into a plain function.  They are rewritten as whole identifiers instead, outside
// comments and string literals.
std::string translate_qualifiers(const char* src) {
    static const struct { const char* from; const char* to; } words[] = {
        {"kernel", "extern \"C\" __global__"}, {"__kernel", "extern \"C\" __global__"},
        {"global", ""}, {"__global", ""},
        {"local", "__shared__"}, {"__local", "__shared__"},
        {"constant", "__constant__"}, {"__constant", "__constant__"},
        {"restrict", "__restrict__"},
    };
    std::string out;
    const size_t n = strlen(src);
    out.reserve(n + 256);
    size_t p = 0;
    while (p < n) {
        const char c = src[p];
        if (c == '/' && p + 1 < n && src[p + 1] == '/') {                    // line comment
            const size_t e = std::min(n, (size_t)(strchr(src + p, '\n') ? strchr(src + p, '\n') - src : n));
            out.append(src + p, e - p); p = e;
        } else if (c == '/' && p + 1 < n && src[p + 1] == '*') {             // block comment
            const char* e = strstr(src + p + 2, "*/");
            const size_t end = e ? (size_t)(e - src) + 2 : n;
            out.append(src + p, end - p); p = end;
        } else if (c == '"' || c == '\'') {                                   // string / character literal
            size_t e = p + 1;
            while (e < n && src[e] != c) e += (src[e] == '\\' && e + 1 < n) ? 2 : 1;
            e = std::min(n, e + 1);
            out.append(src + p, e - p); p = e;
        } else if (isalpha((unsigned char)c) || c == '_') {                  // identifier
            size_t e = p + 1;
            while (e < n && (isalnum((unsigned char)src[e]) || src[e] == '_')) ++e;
            const std::string word(src + p, e - p);
            const char* to = nullptr;
            for (const auto& w : words) if (word == w.from) { to = w.to; break; }
            if (to) out += to; else out += word;
            p = e;
        } else if (isdigit((unsigned char)c)) {                              // number (keeps suffixes such as 1.0f intact)
            size_t e = p + 1;
            while (e < n && (isalnum((unsigned char)src[e]) || src[e] == '.' || src[e] == '_')) ++e;
            out.append(src + p, e - p); p = e;
        } else {
            out += c; ++p;
        }
    }
    return out;
}

void replace_all(std::string& s, const std::string& from, const std::string& to) {
    size_t pos = 0;
    while ((pos = s.find(from, pos)) != std::string::npos) { s.replace(pos, from.size(), to); pos += to.size(); }
}

thread_local std::string g_build_log;

// NVRTC: source text -> sm_100a cubin.  No device needed.
la_status compile_source(int dtype, const char* source, std::vector<char>* cubin) {
    const char* ty = type_name(dtype);
    if (!ty) return la_fail(LA_ERR_ARG, "user kernel: unknown dtype %d", dtype);
    if (!source) return la_fail(LA_ERR_ARG, "user kernel: source is NULL");
    const Nvrtc* rt; LA_TRY(get_nvrtc(&rt));

    std::string upper(ty);
    for (char& c : upper) c = (char)toupper((unsigned char)c);
    std::string text = "#define IS_" + upper + "\n" + PRELUDE + translate_qualifiers(source) + "\n";     // get_src, src/lib.rs:342-350
    replace_all(text, "{T}", ty);

    nvrtcProgram prog = nullptr;
    nvrtcResult r = rt->create(&prog, text.c_str(), "linalg_b200_user_kernels.cu", 0, nullptr, nullptr);
    if (r != NVRTC_SUCCESS) return la_fail(LA_ERR_CUDA, "nvrtcCreateProgram: %s", rt->error_string(r));
    const char* opts[] = {"--gpu-architecture=sm_100a", "--std=c++17", "-default-device", "--fmad=true"};
    r = rt->compile(prog, (int)(sizeof(opts) / sizeof(opts[0])), opts);
    size_t log_bytes = 0;
    g_build_log.clear();
    if (rt->log_size(prog, &log_bytes) == NVRTC_SUCCESS && log_bytes > 1) {
        g_build_log.resize(log_bytes);
        rt->log(prog, &g_build_log[0]);
        g_build_log.resize(strlen(g_build_log.c_str()));
    }
    if (r != NVRTC_SUCCESS) {
        rt->destroy(&prog);
        // the reference panics with the driver's build log (src/lib.rs:329-331); the whole log is in la_user_kernel_build_log()
        return la_fail(LA_ERR_ARG, "user kernel source for {T}=%s failed to compile (%s): %.380s", ty, rt->error_string(r), g_build_log.c_str());
    }
    size_t bytes = 0;
    r = rt->cubin_size(prog, &bytes);
    if (r == NVRTC_SUCCESS) { cubin->resize(bytes); r = rt->cubin(prog, cubin->data()); }
    rt->destroy(&prog);
    if (r != NVRTC_SUCCESS || bytes == 0) return la_fail(LA_ERR_CUDA, "nvrtcGetCUBIN: %s", rt->error_string(r));
    return LA_OK;
}

constexpr unsigned UK_BLOCK = 256;     // get_group_id / get_num_groups in the prelude assume it

}  // namespace

struct la_user_kernel {
    cudaLibrary_t lib = nullptr;
    cudaKernel_t fn = nullptr;
    cudaKernel_t set_state = nullptr;
    int device = -1;
};

extern "C" const char* la_user_kernel_type_name(int dtype) { return type_name(dtype); }
extern "C" const char* la_user_kernel_build_log(void) { return g_build_log.c_str(); }

extern "C" la_status la_user_kernel_compile(int dtype, const char* source, size_t* cubin_bytes) {
    std::vector<char> cubin;
    LA_TRY(compile_source(dtype, source, &cubin));
    if (cubin_bytes) *cubin_bytes = cubin.size();
    return LA_OK;
}

extern "C" la_status la_user_kernel_create(int dtype, const char* source, const char* kernel_name, la_user_kernel** out) {
    if (!out || !kernel_name) return la_fail(LA_ERR_ARG, "la_user_kernel_create: NULL argument");
    *out = nullptr;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    std::vector<char> cubin;
    LA_TRY(compile_source(dtype, source, &cubin));
    la_user_kernel* k = new la_user_kernel;
    k->device = info->device;
    cudaError_t e = cudaLibraryLoadData(&k->lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e != cudaSuccess) { delete k; return la_cuda_fail(e, "cudaLibraryLoadData(user kernel cubin)", __FILE__, __LINE__); }
    e = cudaLibraryGetKernel(&k->fn, k->lib, kernel_name);
    if (e != cudaSuccess) {
        cudaLibraryUnload(k->lib); delete k; (void)cudaGetLastError();
        // src/lib.rs:168-169: "Failed to create kernel. Forgot to add kernel source to kernel.cl?"
        return la_fail(LA_ERR_ARG, "failed to create kernel '%s': no such `kernel void` in the source compiled for {T}=%s (forgot the {T}_ prefix?)", kernel_name, type_name(dtype));
    }
    e = cudaLibraryGetKernel(&k->set_state, k->lib, "la_uk_set_state");
    if (e != cudaSuccess) { cudaLibraryUnload(k->lib); delete k; return la_cuda_fail(e, "cudaLibraryGetKernel(la_uk_set_state)", __FILE__, __LINE__); }
    *out = k;
    return LA_OK;
}

extern "C" la_status la_user_kernel_destroy(la_user_kernel* k) {
    if (!k) return LA_OK;
    cudaError_t e = cudaLibraryUnload(k->lib);
    delete k;
    if (e != cudaSuccess) return la_cuda_fail(e, "cudaLibraryUnload", __FILE__, __LINE__);
    return LA_OK;
}

extern "C" la_status la_user_kernel_launch(la_user_kernel* k, size_t global_size, void** arg_values, void* stream) {
    if (!k) return la_fail(LA_ERR_ARG, "la_user_kernel_launch: kernel is NULL");
    if (global_size == 0) return LA_OK;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    if (info->device != k->device) return la_fail(LA_ERR_ARG, "la_user_kernel_launch: kernel was created on device %d, current device is %d", k->device, info->device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t blocks = global_size / UK_BLOCK;
    const unsigned tail = (unsigned)(global_size % UK_BLOCK);
    if (blocks > 0x7fffffffull) return la_fail(LA_ERR_ARG, "la_user_kernel_launch: global size %zu is beyond one grid", global_size);
    unsigned long first = 0, count = (unsigned long)global_size;
    void* state_args[] = {&first, &count};
    if (blocks) {
        LA_CUDA(cudaLaunchKernel((const void*)k->set_state, dim3(1), dim3(1), state_args, 0, st));
        LA_CUDA(cudaLaunchKernel((const void*)k->fn, dim3((unsigned)blocks), dim3(UK_BLOCK), arg_values, 0, st));
    }
    if (tail) {                         // the last, partial block: its own launch, so no work-item beyond global_size ever exists
        first = (unsigned long)blocks * UK_BLOCK;
        LA_CUDA(cudaLaunchKernel((const void*)k->set_state, dim3(1), dim3(1), state_args, 0, st));
        LA_CUDA(cudaLaunchKernel((const void*)k->fn, dim3(1), dim3(tail), arg_values, 0, st));
    }
    return LA_OK;
}
