// csrc/gemm_host.cu -- Matrix x Matrix on HOST buffers, streamed.
//
// The reference's user writes  Matrix::from_vec(a) * Matrix::from_vec(b)  and then to_vec():
// upload, upload, multiply, download, one after the other (src/matrix.rs:39-46, 221-265,
// src/vector.rs:53-78).  la_mul_mat_mat_host is that whole sequence as ONE call that overlaps its
// parts: B goes up once (and is pre-processed once for the tensor engine); A and C move in row
// blocks on their own copy streams, so block i+1 uploads and block i-1 downloads while block i
// multiplies (PCIe is full duplex).  Results are those of la_mul_mat_mat on the same operands.
// Any of the three pointers may also be a device pointer (the copies use cudaMemcpyDefault): the
// multi-GPU path passes a B that ranks have gathered over NVLink instead of each uploading all of it.
#include "common.cuh"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>

la_status la_mul_mat_mat_impl(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, void* stream, int mode, const la_gemm_prepared* prep = nullptr);
int la_gemm_current_mode();
la_status la_tc_split_a(float* a_hi, float* a_lo, const float* A, size_t M, size_t K, unsigned* flag, cudaStream_t st);
la_status la_tc_split_bt(float* bt_hi, float* bt_lo, const float* B, size_t K, size_t N, unsigned* flag, cudaStream_t st);
la_status la_tc_run(float* C, const float* a_hi, const float* a_lo, const float* bt_hi, const float* bt_lo, size_t M, size_t K, size_t N, cudaStream_t st);
bool la_tc_shape_ok(const void* A, const void* B, const void* C, size_t M, size_t K, size_t N);
bool la_tc_a_direct(size_t K);
bool la_f16_wanted(size_t K);                          // the f16-split form of the f32 tensor engine, in pieces (gemm_tc.cu)
size_t la_f16_b_bytes(size_t K, size_t N);
size_t la_f16_a_bytes(size_t rows, size_t K);
la_status la_f16_split_b(void* b_block, const float* B, size_t K, size_t N, unsigned* flag, cudaStream_t st);
la_status la_f16_block(float* C, const float* A, const float* B, size_t rows, size_t K, size_t N, void* a_block, void* b_block, unsigned* flag, cudaStream_t st);

namespace {

struct HostGemmStreams {
    std::mutex busy;      // one la_mul_mat_mat_host at a time per device: the streams, events and the range-flag word below are shared
    cudaStream_t in = nullptr, cmp = nullptr, out = nullptr;
    cudaEvent_t b_up = nullptr, in_done[2] = {nullptr, nullptr}, cmp_done[2] = {nullptr, nullptr}, out_done[2] = {nullptr, nullptr};
};
std::mutex g_hs_mu;
std::map<int, HostGemmStreams*> g_hs;

la_status host_streams(HostGemmStreams** out) {
    int dev = -1;
    LA_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_hs_mu);
    auto it = g_hs.find(dev);
    if (it != g_hs.end()) { *out = it->second; return LA_OK; }
    HostGemmStreams* s = new HostGemmStreams();
    LA_CUDA(cudaStreamCreateWithFlags(&s->in, cudaStreamNonBlocking));
    LA_CUDA(cudaStreamCreateWithFlags(&s->cmp, cudaStreamNonBlocking));
    LA_CUDA(cudaStreamCreateWithFlags(&s->out, cudaStreamNonBlocking));
    LA_CUDA(cudaEventCreateWithFlags(&s->b_up, cudaEventDisableTiming));
    for (int i = 0; i < 2; ++i) {
        LA_CUDA(cudaEventCreateWithFlags(&s->in_done[i], cudaEventDisableTiming));
        LA_CUDA(cudaEventCreateWithFlags(&s->cmp_done[i], cudaEventDisableTiming));
        LA_CUDA(cudaEventCreateWithFlags(&s->out_done[i], cudaEventDisableTiming));
    }
    g_hs[dev] = s;
    *out = s;
    return LA_OK;
}

struct DevBuf {       // stream-ordered pool allocation, freed on scope exit
    void* p = nullptr; cudaStream_t st;
    explicit DevBuf(cudaStream_t s) : st(s) {}
    la_status alloc(size_t bytes) { LA_CUDA(cudaMallocAsync(&p, bytes ? bytes : 16, st)); return LA_OK; }
    ~DevBuf() { if (p) cudaFreeAsync(p, st); }
};

// Declared after the DevBufs of a pass, so it runs before their destructors on EVERY way out: the frees are ordered on
// the compute stream only, while copies on the two copy streams may still be reading or writing the blocks when an
// error return cuts the pipeline short.
struct DrainStreams {
    HostGemmStreams* hs;
    ~DrainStreams() { cudaStreamSynchronize(hs->in); cudaStreamSynchronize(hs->out); cudaStreamSynchronize(hs->cmp); (void)cudaGetLastError(); }
};

}  // namespace

// `b_ready` != NULL: B_host is a DEVICE buffer of this GPU that is complete once that event has happened (la_mul_mat_mat_host_sharded
// gathers it over NVLink on another stream); it is used in place, and the uploads of A start at once, in the shadow of the gather.
static la_status host_gemm_pass(int dtype, void* C_host, const void* A_host, const void* B_host, size_t M, size_t K, size_t N, bool allow_tensor, bool* redo,
                                cudaEvent_t b_ready = nullptr) {
    *redo = false;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    HostGemmStreams* hs; LA_TRY(host_streams(&hs));
    std::lock_guard<std::mutex> one_at_a_time(hs->busy);
    const size_t es = la_dtype_size(dtype);
    const int mode = la_gemm_current_mode();
    static const bool dbg = getenv("LA_DEBUG_HOSTGEMM") != nullptr;
    auto t0 = std::chrono::steady_clock::now();
    auto ms_since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };

    // Row block.  Tensor engine: about a sixteenth of M (short pipeline tail) but at least 512 rows, so that
    // one block still fills most SM pairs (512 x N / (256 x 256) tiles).  Other engines are compute bound and
    // tile-quantised: a quarter of M, at least 1024 rows.  Multiples of 128 / 256 rows.
    const bool want_tensor = allow_tensor && dtype == LA_F32 && mode != LA_GEMM_SIMT && K > 0 &&
                             (mode == LA_GEMM_3XTF32 || ((double)M * (double)N * (double)K >= 134217728.0 && K >= LA_TC_AUTO_MIN_K));   // la_mul_mat_mat's policy
    size_t rb;
    if (want_tensor) { rb = ((M / 16 + 127) / 128) * 128; if (rb < 512) rb = 512; if (rb > 2048) rb = 2048; }
    else             { rb = ((M / 4 + 255) / 256) * 256;  if (rb < 1024) rb = 1024; if (rb > 4096) rb = 4096; }
    if (rb > M) rb = M;
    const size_t nblocks = (M + rb - 1) / rb;

    DevBuf B(hs->cmp), Abuf0(hs->cmp), Abuf1(hs->cmp), Cbuf0(hs->cmp), Cbuf1(hs->cmp), bt_hi(hs->cmp), bt_lo(hs->cmp), a_hi0(hs->cmp), a_lo0(hs->cmp), a_hi1(hs->cmp), a_lo1(hs->cmp);
    DrainStreams drain{hs};
    if (!b_ready) LA_TRY(B.alloc(K * N * es));
    const void* Bp = b_ready ? B_host : B.p;
    LA_TRY(Abuf0.alloc(rb * K * es)); LA_TRY(Abuf1.alloc(rb * K * es));
    LA_TRY(Cbuf0.alloc(rb * N * es)); LA_TRY(Cbuf1.alloc(rb * N * es));
    void* Ab[2] = {Abuf0.p, Abuf1.p}; void* Cb[2] = {Cbuf0.p, Cbuf1.p};

    // tensor engine for f32 when the mode and the shapes allow: B is split once, up front
    bool tensor = want_tensor && la_tc_shape_ok(Ab[0], Bp, Cb[0], rb, K, N);
    float* ah[2] = {nullptr, nullptr}; float* al[2] = {nullptr, nullptr};
    volatile unsigned* flag_host = nullptr; unsigned* flag = nullptr;
    const bool f16 = tensor && la_f16_wanted(K);           // the f16-split form (from K = 1024), else the tf32 form: la_mul_mat_mat's choice
    if (tensor) {
        LA_TRY(la_host_flag(hs->cmp, &flag_host, &flag));
        flag_host[0] = 0u; flag_host[1] = 0u;
        if (f16) {
            LA_TRY(bt_hi.alloc(la_f16_b_bytes(K, N)));
            LA_TRY(a_hi0.alloc(la_f16_a_bytes(rb, K))); LA_TRY(a_hi1.alloc(la_f16_a_bytes(rb, K)));
        } else {
            LA_TRY(bt_hi.alloc(K * N * 4)); LA_TRY(bt_lo.alloc(K * N * 4));
            LA_TRY(a_hi0.alloc(rb * K * 4)); LA_TRY(a_lo0.alloc(rb * K * 4)); LA_TRY(a_hi1.alloc(rb * K * 4)); LA_TRY(a_lo1.alloc(rb * K * 4));
        }
        ah[0] = (float*)a_hi0.p; al[0] = (float*)a_lo0.p; ah[1] = (float*)a_hi1.p; al[1] = (float*)a_lo1.p;
    }
    // the copy streams may touch the buffers only after the (stream-ordered) allocations exist
    LA_CUDA(cudaEventRecord(hs->b_up, hs->cmp));
    LA_CUDA(cudaStreamWaitEvent(hs->in, hs->b_up, 0));
    LA_CUDA(cudaStreamWaitEvent(hs->out, hs->b_up, 0));

    if (b_ready) {
        LA_CUDA(cudaStreamWaitEvent(hs->cmp, b_ready, 0));
    } else {
        LA_CUDA(cudaMemcpyAsync(B.p, B_host, K * N * es, cudaMemcpyDefault, hs->in));
        LA_CUDA(cudaEventRecord(hs->b_up, hs->in));
        LA_CUDA(cudaStreamWaitEvent(hs->cmp, hs->b_up, 0));
    }
    if (f16) LA_TRY(la_f16_split_b(bt_hi.p, (const float*)Bp, K, N, flag, hs->cmp));
    else if (tensor) LA_TRY(la_tc_split_bt((float*)bt_hi.p, (float*)bt_lo.p, (const float*)Bp, K, N, flag, hs->cmp));
    if (dbg) fprintf(stderr, "[hostgemm] setup enqueued at %.3f ms (tensor=%d rb=%zu nblocks=%zu)\n", ms_since(), (int)tensor, rb, nblocks);

    for (size_t i = 0; i < nblocks; ++i) {
        const int b = (int)(i & 1);
        const size_t r0 = i * rb, rows = (r0 + rb <= M) ? rb : M - r0;
        // upload block i once the multiply that last read this A buffer is done
        if (i >= 2) LA_CUDA(cudaStreamWaitEvent(hs->in, hs->cmp_done[b], 0));
        LA_CUDA(cudaMemcpyAsync(Ab[b], (const char*)A_host + r0 * K * es, rows * K * es, cudaMemcpyDefault, hs->in));
        LA_CUDA(cudaEventRecord(hs->in_done[b], hs->in));
        // multiply once the block is up and the download that last read this C buffer is done
        LA_CUDA(cudaStreamWaitEvent(hs->cmp, hs->in_done[b], 0));
        if (i >= 2) LA_CUDA(cudaStreamWaitEvent(hs->cmp, hs->out_done[b], 0));
        // The range check of the split is read back once, after everything is queued: no host
        // wait inside the pipeline.  (If it trips, the whole product is redone in plain f32.)
        if (f16 && rows >= 128 && rows % 4 == 0) {
            LA_TRY(la_f16_block((float*)Cb[b], (const float*)Ab[b], (const float*)Bp, rows, K, N, ah[b], bt_hi.p, flag, hs->cmp));
        } else if (tensor && rows >= 128 && rows % 4 == 0) {
            LA_TRY(la_tc_split_a(ah[b], al[b], (const float*)Ab[b], rows, K, flag, hs->cmp));
            LA_TRY(la_tc_run((float*)Cb[b], la_tc_a_direct(K) ? (const float*)Ab[b] : ah[b], al[b], (const float*)bt_hi.p, (const float*)bt_lo.p, rows, K, N, hs->cmp));
        } else {
            // (a ragged tail block of a tensor-engine product goes to the SIMT engine: la_gemm_3xtf32 would
            //  reset the range flag this pass is still collecting)
            LA_TRY(la_mul_mat_mat_impl(dtype, Cb[b], Ab[b], Bp, rows, K, N, hs->cmp, (tensor || !allow_tensor) ? LA_GEMM_SIMT : mode));
        }
        LA_CUDA(cudaEventRecord(hs->cmp_done[b], hs->cmp));
        // download
        LA_CUDA(cudaStreamWaitEvent(hs->out, hs->cmp_done[b], 0));
        LA_CUDA(cudaMemcpyAsync((char*)C_host + r0 * N * es, Cb[b], rows * N * es, cudaMemcpyDefault, hs->out));
        LA_CUDA(cudaEventRecord(hs->out_done[b], hs->out));
    }
    if (dbg) fprintf(stderr, "[hostgemm] all enqueued at %.3f ms\n", ms_since());
    LA_CUDA(cudaStreamSynchronize(hs->out));
    if (dbg) fprintf(stderr, "[hostgemm] out drained at %.3f ms\n", ms_since());
    LA_CUDA(cudaStreamSynchronize(hs->in));
    LA_CUDA(cudaStreamSynchronize(hs->cmp));
    // an operand the split cannot represent (inf, NaN, extreme exponents or scales) or, f16 form, one with one-product outputs:
    // the whole product again on the SIMT engine
    if (tensor && (flag_host[0] || (f16 && flag_host[1]))) *redo = true;
    return LA_OK;
}

extern "C" la_status la_mul_mat_mat_host(int dtype, void* C_host, const void* A_host, const void* B_host, size_t M, size_t K, size_t N) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_mul_mat_mat_host: unknown dtype %d", dtype);
    if (M == 0 || N == 0) return LA_OK;
    if (!C_host || !A_host || !B_host) return la_fail(LA_ERR_ARG, "la_mul_mat_mat_host: NULL operand");
    bool redo = false;
    LA_TRY(host_gemm_pass(dtype, C_host, A_host, B_host, M, K, N, true, &redo));
    if (redo) LA_TRY(host_gemm_pass(dtype, C_host, A_host, B_host, M, K, N, false, &redo));
    return LA_OK;
}

// comm.cu: the same streamed product against a B that is (or is about to be, `b_ready`) resident on this GPU.
la_status la_mul_mat_mat_host_dev_b(int dtype, void* C_host, const void* A_host, const void* B_dev, cudaEvent_t b_ready, size_t M, size_t K, size_t N) {
    if (M == 0 || N == 0) return LA_OK;
    bool redo = false;
    LA_TRY(host_gemm_pass(dtype, C_host, A_host, B_dev, M, K, N, true, &redo, b_ready));
    if (redo) LA_TRY(host_gemm_pass(dtype, C_host, A_host, B_dev, M, K, N, false, &redo, b_ready));
    return LA_OK;
}
