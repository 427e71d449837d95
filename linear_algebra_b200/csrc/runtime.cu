// csrc/runtime.cu -- device binding, error reporting, buffers, streams, scratch.
//
// Replaces the OpenCL context / queue / buffer plumbing of the reference (src/lib.rs:24-27,
// 125-149, 310-335; ocl::Buffer use in src/vector.rs:36-78, 700-708).  Kernels are compiled
// ahead of time for sm_100a, so "init" is only: pick the device, size the allocator pool.
#include "common.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>

#define LA_VERSION_STR "linalg_b200 0.1 (sm_100a)"

// ---------------------------------------------------------------- errors
static thread_local char g_err[512] = "";

void la_set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
la_status la_fail(la_status code, const char* fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
la_status la_cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) in %s at %s:%d", (int)e, cudaGetErrorString(e), what, file, line);
    (void)cudaGetLastError();   // clear the sticky-less error so later calls report their own
    return e == cudaErrorMemoryAllocation ? LA_ERR_OOM : LA_ERR_CUDA;
}

extern "C" const char* la_last_error(void) { return g_err; }
extern "C" const char* la_version(void) { return LA_VERSION_STR; }
extern "C" size_t la_dtype_size(int dtype) {
    static const size_t sz[] = {1, 1, 2, 2, 4, 4, 8, 8, 4, 8};
    return la_valid_dtype(dtype) ? sz[dtype] : 0;
}

// ---------------------------------------------------------------- device state
namespace {
struct DeviceState {
    LaDeviceInfo info;
    std::map<cudaStream_t, std::pair<void*, size_t>> scratch;   // per stream: block, capacity
    std::map<cudaStream_t, unsigned*> host_flags;                // per stream: one mapped, page-locked word
    std::map<cudaStream_t, cudaEvent_t> events;                  // per stream: one timing-less event
    std::map<cudaStream_t, cudaEvent_t> user_streams;            // streams made by la_stream_create, each with the event la_malloc / la_free order it by
    cudaEvent_t legacy_event = nullptr;                          // recorded on the legacy stream by la_malloc
};
std::mutex g_mu;
std::map<int, DeviceState*> g_devices;
}  // namespace

static la_status device_state(DeviceState** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return la_cuda_fail(e, "cudaGetDevice (is a CUDA device present? there is no CPU fallback)", __FILE__, __LINE__);
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_devices.find(dev);
    if (it != g_devices.end()) { *out = it->second; return LA_OK; }
    cudaDeviceProp prop;
    LA_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (prop.major != 10) {
        return la_fail(LA_ERR_CUDA, "device %d (%s) is sm_%d%d; liblinalg_b200 carries sm_100a code only",
                       dev, prop.name, prop.major, prop.minor);
    }
    DeviceState* st = new DeviceState();
    st->info.device = dev;
    st->info.sm_count = prop.multiProcessorCount;
    st->info.max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    // Result buffers are allocated per out-of-place op (as the reference does, src/vector.rs:99):
    // keep freed blocks in the stream-ordered pool instead of returning them to the driver.
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        uint64_t keep = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    (void)cudaGetLastError();
    g_devices[dev] = st;
    *out = st;
    return LA_OK;
}

la_status la_ensure_init(LaDeviceInfo** info) {
    DeviceState* st = nullptr;
    LA_TRY(device_state(&st));
    *info = &st->info;
    return LA_OK;
}

la_status la_scratch(cudaStream_t stream, size_t bytes, void** ptr) {
    DeviceState* st = nullptr;
    LA_TRY(device_state(&st));
    bytes += LA_SCRATCH_ZEROED;
    std::lock_guard<std::mutex> lk(g_mu);
    auto& slot = st->scratch[stream];
    if (slot.second < bytes) {
        size_t cap = bytes < (size_t(1) << 20) ? (size_t(1) << 20) : bytes;
        void* p = nullptr;
        // A plain cudaMalloc (not stream-ordered): the block outlives any one stream op.
        if (slot.first) { LA_CUDA(cudaStreamSynchronize(stream)); LA_CUDA(cudaFree(slot.first)); slot = {nullptr, 0}; }
        LA_CUDA(cudaMalloc(&p, cap));
        LA_CUDA(cudaMemsetAsync(p, 0, LA_SCRATCH_ZEROED, stream));
        slot = {p, cap};
    }
    *ptr = slot.first;
    return LA_OK;
}

// A 4-byte flag in page-locked host memory that kernels on `stream` can set directly (zero-copy),
// so the host can read it after an event wait without queueing a device-to-host copy behind
// whatever bulk transfer a copy engine is busy with.
la_status la_host_flag(cudaStream_t stream, volatile unsigned** host_ptr, unsigned** dev_ptr) {
    DeviceState* st = nullptr;
    LA_TRY(device_state(&st));
    std::lock_guard<std::mutex> lk(g_mu);
    unsigned*& slot = st->host_flags[stream];
    if (!slot) {
        void* p = nullptr;
        LA_CUDA(cudaHostAlloc(&p, 64, cudaHostAllocMapped));
        *(volatile unsigned*)p = 0u;
        slot = (unsigned*)p;
    }
    void* d = nullptr;
    LA_CUDA(cudaHostGetDevicePointer(&d, slot, 0));
    *host_ptr = slot; *dev_ptr = (unsigned*)d;
    return LA_OK;
}

la_status la_stream_event(cudaStream_t stream, cudaEvent_t* event) {
    DeviceState* st = nullptr;
    LA_TRY(device_state(&st));
    std::lock_guard<std::mutex> lk(g_mu);
    cudaEvent_t& slot = st->events[stream];
    if (!slot) LA_CUDA(cudaEventCreateWithFlags(&slot, cudaEventDisableTiming));
    *event = slot;
    return LA_OK;
}

extern "C" la_status la_init(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        (void)cudaGetLastError();
        return la_fail(LA_ERR_CUDA, "no CUDA device visible (%s); liblinalg_b200 has no CPU fallback",
                       e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= n) return la_fail(LA_ERR_ARG, "device %d out of range [0,%d)", device, n);
    LA_CUDA(cudaSetDevice(device));
    LaDeviceInfo* info;
    return la_ensure_init(&info);
}
extern "C" la_status la_device_count(int* count) {
    if (!count) return la_fail(LA_ERR_ARG, "count is NULL");
    cudaError_t e = cudaGetDeviceCount(count);
    if (e != cudaSuccess) { *count = 0; return la_cuda_fail(e, "cudaGetDeviceCount", __FILE__, __LINE__); }
    return LA_OK;
}
extern "C" la_status la_device_name(char* buf, size_t buf_len) {
    if (!buf || !buf_len) return la_fail(LA_ERR_ARG, "buf is NULL");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    cudaDeviceProp prop; LA_CUDA(cudaGetDeviceProperties(&prop, info->device));
    snprintf(buf, buf_len, "%s", prop.name);
    return LA_OK;
}
extern "C" la_status la_sm_count(int* count) {
    if (!count) return la_fail(LA_ERR_ARG, "count is NULL");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    *count = info->sm_count;
    return LA_OK;
}

// ---------------------------------------------------------------- buffers
// la_malloc / la_free are stream-ordered pool operations on the legacy default stream.  The streams la_stream_create hands
// out are cudaStreamNonBlocking, which the legacy stream does NOT order itself against, so the two calls do that explicitly:
//   la_free    the legacy stream first waits for everything queued so far on every la_stream_create stream (an event
//              each), so the pool cannot hand the block to anyone while a kernel on such a stream still uses it;
//   la_malloc  every la_stream_create stream then waits for the legacy stream (one event), so work that the new owner queues
//              on such a stream starts after legacy-stream work on the block's previous owner has drained.
// With no la_stream_create stream alive (the reference-style synchronous use) both are a single pool call.  Streams the
// caller made itself with cudaStreamNonBlocking are outside this: synchronise those before la_free (header).
static la_status order_legacy_after_user_streams(DeviceState* st) {
    for (auto& kv : st->user_streams) {
        LA_CUDA(cudaEventRecord(kv.second, kv.first));
        LA_CUDA(cudaStreamWaitEvent((cudaStream_t)0, kv.second, 0));
    }
    return LA_OK;
}
extern "C" la_status la_malloc(void** dev_ptr, size_t bytes) {
    if (!dev_ptr) return la_fail(LA_ERR_ARG, "dev_ptr is NULL");
    DeviceState* st = nullptr; LA_TRY(device_state(&st));
    *dev_ptr = nullptr;
    // bytes == 0: Vector::from_vec(vec![]) is legal in the reference; hand out a real pointer.
    cudaError_t e = cudaMallocAsync(dev_ptr, bytes ? bytes : 16, (cudaStream_t)0);
    if (e != cudaSuccess) return la_cuda_fail(e, "cudaMallocAsync", __FILE__, __LINE__);
    std::lock_guard<std::mutex> lk(g_mu);
    if (!st->user_streams.empty()) {
        if (!st->legacy_event) LA_CUDA(cudaEventCreateWithFlags(&st->legacy_event, cudaEventDisableTiming));
        LA_CUDA(cudaEventRecord(st->legacy_event, (cudaStream_t)0));
        for (auto& kv : st->user_streams) LA_CUDA(cudaStreamWaitEvent(kv.first, st->legacy_event, 0));
    }
    return LA_OK;
}
extern "C" la_status la_free(void* dev_ptr) {
    if (!dev_ptr) return LA_OK;
    DeviceState* st = nullptr; LA_TRY(device_state(&st));
    {
        std::lock_guard<std::mutex> lk(g_mu);
        LA_TRY(order_legacy_after_user_streams(st));
    }
    LA_CUDA(cudaFreeAsync(dev_ptr, (cudaStream_t)0));
    return LA_OK;
}
extern "C" la_status la_upload(void* dst_dev, const void* src_host, size_t bytes, void* stream) {
    if (bytes == 0) return LA_OK;
    if (!dst_dev || !src_host) return la_fail(LA_ERR_ARG, "la_upload: NULL pointer");
    LA_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return LA_OK;
}
extern "C" la_status la_download(void* dst_host, const void* src_dev, size_t bytes, void* stream) {
    if (bytes == 0) return LA_OK;
    if (!dst_host || !src_dev) return la_fail(LA_ERR_ARG, "la_download: NULL pointer");
    LA_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    LA_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return LA_OK;
}
extern "C" la_status la_copy(void* dst_dev, const void* src_dev, size_t bytes, void* stream) {
    if (bytes == 0) return LA_OK;
    if (!dst_dev || !src_dev) return la_fail(LA_ERR_ARG, "la_copy: NULL pointer");
    LA_CUDA(cudaMemcpyAsync(dst_dev, src_dev, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return LA_OK;
}
extern "C" la_status la_copy_2d(void* dst_dev, size_t dst_pitch, const void* src_dev, size_t src_pitch, size_t width_bytes, size_t height, void* stream) {
    if (width_bytes == 0 || height == 0) return LA_OK;
    if (!dst_dev || !src_dev) return la_fail(LA_ERR_ARG, "la_copy_2d: NULL pointer");
    if (dst_pitch < width_bytes || src_pitch < width_bytes) return la_fail(LA_ERR_ARG, "la_copy_2d: pitch smaller than the row width");
    LA_CUDA(cudaMemcpy2DAsync(dst_dev, dst_pitch, src_dev, src_pitch, width_bytes, height, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return LA_OK;
}
extern "C" la_status la_host_alloc(void** host_ptr, size_t bytes) {
    if (!host_ptr) return la_fail(LA_ERR_ARG, "host_ptr is NULL");
    LA_CUDA(cudaMallocHost(host_ptr, bytes ? bytes : 16));
    return LA_OK;
}
extern "C" la_status la_host_free(void* host_ptr) {
    if (!host_ptr) return LA_OK;
    LA_CUDA(cudaFreeHost(host_ptr));
    return LA_OK;
}
extern "C" la_status la_stream_create(void** stream) {
    if (!stream) return la_fail(LA_ERR_ARG, "stream is NULL");
    DeviceState* st = nullptr; LA_TRY(device_state(&st));
    cudaStream_t s; cudaEvent_t ev;
    LA_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    LA_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    {
        std::lock_guard<std::mutex> lk(g_mu);
        st->user_streams[s] = ev;          // la_malloc / la_free order themselves against it from now on
    }
    *stream = (void*)s;
    return LA_OK;
}
extern "C" la_status la_stream_destroy(void* stream) {
    if (!stream) return LA_OK;
    {
        DeviceState* st = nullptr;
        if (device_state(&st) == LA_OK) {
            std::lock_guard<std::mutex> lk(g_mu);
            auto it = st->scratch.find((cudaStream_t)stream);
            if (it != st->scratch.end()) {
                cudaStreamSynchronize((cudaStream_t)stream);
                cudaFree(it->second.first);
                st->scratch.erase(it);
            }
            auto us = st->user_streams.find((cudaStream_t)stream);
            if (us != st->user_streams.end()) {
                // frees issued later must still come after this stream's last work: let the legacy stream wait for it now
                cudaEventRecord(us->second, us->first);
                cudaStreamWaitEvent((cudaStream_t)0, us->second, 0);
                cudaEventDestroy(us->second);
                st->user_streams.erase(us);
            }
        }
    }
    LA_CUDA(cudaStreamDestroy((cudaStream_t)stream));
    return LA_OK;
}
extern "C" la_status la_stream_sync(void* stream) {
    LA_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return LA_OK;
}

// ---------------------------------------------------------------- events
extern "C" la_status la_event_create(void** event) {
    if (!event) return la_fail(LA_ERR_ARG, "event is NULL");
    cudaEvent_t e;
    LA_CUDA(cudaEventCreate(&e));
    *event = (void*)e;
    return LA_OK;
}
extern "C" la_status la_event_destroy(void* event) {
    if (!event) return LA_OK;
    LA_CUDA(cudaEventDestroy((cudaEvent_t)event));
    return LA_OK;
}
extern "C" la_status la_event_record(void* event, void* stream) {
    if (!event) return la_fail(LA_ERR_ARG, "event is NULL");
    LA_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
    return LA_OK;
}
extern "C" la_status la_event_elapsed_ms(void* start, void* stop, float* ms) {
    if (!start || !stop || !ms) return la_fail(LA_ERR_ARG, "la_event_elapsed_ms: NULL argument");
    LA_CUDA(cudaEventSynchronize((cudaEvent_t)stop));
    LA_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
    return LA_OK;
}
