// csrc/comm.cu -- the multi-GPU layer below the C ABI: one process per GPU, row-block shards.
//
// The reference has one context and one in-order queue for one device (src/lib.rs:125-149, 310-335); what replaces that for
// N devices has to be reachable from the same boundary as everything else, not only from a Python layer.  An la_comm is one
// rank's view of a group of `world` processes, each driving its own B200.  The library needs exactly one thing from the
// caller: a way to all-gather a few bytes between the processes while the group is set up (and for the few host-side
// agreements later) -- the `exchange` callback, which the caller implements with whatever it already has (MPI, a torch
// process group, files, pipes).  Everything else is done here: the buffers that peers must reach are exported with cudaIpc and
// mapped into every process, and the data path runs over NVLink / NVSwitch peer memory with plain loads and stores:
//   sum / dot          ONE launch per rank: the reduction kernel stores its partial into every peer's mailbox and adds the
//                      `world` partials in rank order (p2p.cu); the scalar comes back through a zero-copy word the host polls
//   v x M              local partial straight into this rank's exported vector buffer, one kernel adds the peers' partials
//                      read over NVLink, rank order
//   transpose          one kernel per rank turns its row block and stores the pieces straight into the owners' exported
//                      staging buffers (the all-to-all as peer stores), then a local copy into the caller's output
//   M x M on host buffers   every rank uploads 1/world of B over its own PCIe link, fetches the other slices from the peers'
//                      staging buffers over NVLink, then streams its rows of A and C (gemm_host.cu)
//   == / fallbacks     one byte / one element per rank through the exchange callback
// Where ranks have to meet inside an operation (before and after the transpose's peer stores, around the gather of B) they do so
// in a small kernel ON the stream (peer_barrier_kernel, p2p.cu): no host synchronisation and no callback on the data path.
// All sharded calls of one communicator must be issued from one thread, and the stream-ordered ones on one stream.
// No collective library is involved and no TMA multicast: peers' memory is only touched by ordinary loads, stores and copies.
// Every rank must call the sharded entry points in the same order with the same global shapes (they are collectives).
// Results are deterministic and identical on every rank (rank-ordered combination with the element type's own arithmetic).
#include "common.cuh"
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

la_status la_fused_reduce(int dtype, bool dot, const void* a, const void* b, size_t n, void* out_dev, unsigned long long* done_flag,
                          void* const* peer_mailboxes, int world, int rank, unsigned long long seq, void* stream);      // p2p.cu
la_status la_peer_barrier(void* const* peer_mailboxes, int world, int rank, unsigned long long seq, cudaStream_t st);    // p2p.cu
la_status la_mul_mat_mat_host_dev_b(int dtype, void* C_host, const void* A_host, const void* B_dev, cudaEvent_t b_ready, size_t M, size_t K, size_t N);   // gemm_host.cu

namespace {

constexpr int COMM_MAX_WORLD = 16;

struct Exported {                  // a device buffer of this rank that every peer has mapped, and the peers' buffers as mapped here
    void* ptr[COMM_MAX_WORLD] = {};
    size_t bytes = 0;
};

}  // namespace

struct la_comm {
    int world = 1, rank = 0, device = -1;
    la_exchange_fn exchange = nullptr; void* user = nullptr;
    bool p2p = false;              // the peer-memory paths are usable on ALL ranks (decided together at creation)
    Exported box, vec, stage;
    size_t vec_half = 0;
    unsigned long long seq = 0;
    unsigned long long* res_host = nullptr; unsigned long long* res_dev = nullptr;     // zero-copy: [0] value, [1] sequence number of the round it belongs to
    cudaStream_t side = nullptr; cudaEvent_t side_done = nullptr;                       // the gather of B runs beside the uploads of A
    std::string note;
};

namespace {

la_status xchg(la_comm* c, const void* send, void* recv, size_t bytes) {
    if (c->world == 1) { memcpy(recv, send, bytes); return LA_OK; }
    const int r = c->exchange(c->user, send, recv, bytes);
    if (r != 0) return la_fail(LA_ERR_ARG, "la_comm: the exchange callback failed with %d", r);
    return LA_OK;
}

// Every rank says whether it is fine; the answer is the same everywhere.  Also serves as a barrier.
la_status agree(la_comm* c, bool mine, bool* all) {
    unsigned char s = mine ? 1 : 0, r[COMM_MAX_WORLD] = {};
    LA_TRY(xchg(c, &s, r, 1));
    *all = true;
    for (int p = 0; p < c->world; ++p) *all = *all && r[p] == 1;
    return LA_OK;
}
la_status barrier(la_comm* c) { bool all; return agree(c, true, &all); }

// Giving up an exported buffer has two halves: unmapping the peers' buffers here, and freeing this rank's own -- which the peers
// must have unmapped first.  The collective paths (a buffer being replaced, la_comm_destroy) put a barrier between the two;
// unexport() is the immediate form for the error paths, where no further exchange can be relied on.
void close_imports(la_comm* c, Exported& e) {
    for (int p = 0; p < c->world; ++p)
        if (p != c->rank && e.ptr[p]) { cudaIpcCloseMemHandle(e.ptr[p]); e.ptr[p] = nullptr; }
    (void)cudaGetLastError();
}
void free_own(la_comm* c, Exported& e) {
    if (e.ptr[c->rank]) { cudaFree(e.ptr[c->rank]); e.ptr[c->rank] = nullptr; }
    e.bytes = 0;
    (void)cudaGetLastError();
}
void unexport(la_comm* c, Exported& e) { close_imports(c, e); free_own(c, e); }

// Collective: every rank creates a zeroed buffer of `bytes`, the cudaIpc handles go round, every rank maps the others'.  Every
// rank issues the same sequence of exchanges whatever fails locally, and all ranks end with the same answer in *ok.
la_status export_collective(la_comm* c, Exported& e, size_t bytes, bool* ok) {
    LA_CUDA(cudaDeviceSynchronize());
    LA_TRY(barrier(c));                                   // nobody is still reading the buffers about to be replaced
    if (e.bytes) {
        close_imports(c, e);
        LA_TRY(barrier(c));                               // every peer has unmapped this rank's old buffer
        free_own(c, e);
    }
    void* mine = nullptr;
    cudaIpcMemHandle_t h; memset(&h, 0, sizeof(h));
    bool fine = cudaMalloc(&mine, bytes) == cudaSuccess && cudaMemset(mine, 0, bytes) == cudaSuccess && cudaIpcGetMemHandle(&h, mine) == cudaSuccess;
    if (!fine) { c->note = std::string("exportable buffer: ") + cudaGetErrorString(cudaGetLastError()); }
    bool all = false;
    LA_TRY(agree(c, fine, &all));
    if (!all) { if (mine) cudaFree(mine); (void)cudaGetLastError(); *ok = false; return LA_OK; }
    cudaIpcMemHandle_t hs[COMM_MAX_WORLD];
    LA_TRY(xchg(c, &h, hs, sizeof(h)));
    e.ptr[c->rank] = mine;
    fine = true;
    for (int p = 0; p < c->world && fine; ++p) {
        if (p == c->rank) continue;
        if (cudaIpcOpenMemHandle(&e.ptr[p], hs[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            c->note = std::string("mapping a peer's buffer: ") + cudaGetErrorString(cudaGetLastError());
            e.ptr[p] = nullptr; fine = false;
        }
    }
    LA_TRY(agree(c, fine, &all));
    if (!all) { unexport(c, e); *ok = false; return LA_OK; }
    e.bytes = bytes;
    *ok = true;
    return LA_OK;
}

// Stream-ordered barrier between the ranks (peer memory; p2p.cu) -- no host round trip, unlike barrier() above.
la_status stream_barrier(la_comm* c, cudaStream_t st) { return la_peer_barrier(c->box.ptr, c->world, c->rank, ++c->seq, st); }

la_status check_comm(const la_comm* c, const char* who) {
    if (!c) return la_fail(LA_ERR_ARG, "%s: comm is NULL", who);
    int dev = -1; LA_CUDA(cudaGetDevice(&dev));
    if (dev != c->device) return la_fail(LA_ERR_ARG, "%s: the communicator belongs to device %d, current device is %d", who, c->device, dev);
    return LA_OK;
}

// rank-ordered sum of `world` elements with the element type's own arithmetic (integers wrap, floats round after every add)
void combine_host(int dtype, const unsigned char* parts, size_t stride, int world, void* out) {
    switch (dtype) {
#define LA_CASE(code, T) case code: { T acc = T(0); for (int p = 0; p < world; ++p) { T x; memcpy(&x, parts + p * stride, sizeof(T)); acc = Arith<T>::add(acc, x); } memcpy(out, &acc, sizeof(T)); } break;
        LA_CASE(LA_U8, uint8_t) LA_CASE(LA_I8, int8_t) LA_CASE(LA_U16, uint16_t) LA_CASE(LA_I16, int16_t) LA_CASE(LA_U32, uint32_t) LA_CASE(LA_I32, int32_t)
        LA_CASE(LA_U64, uint64_t) LA_CASE(LA_I64, int64_t) LA_CASE(LA_F32, float) LA_CASE(LA_F64, double)
#undef LA_CASE
    default: break;
    }
}

}  // namespace

// ---------------------------------------------------------------- group set-up
extern "C" void la_partition(size_t n, int world, int rank, size_t* lo, size_t* hi) {
    // contiguous blocks; the first n % world ranks hold one element (row) more
    const size_t base = n / (size_t)world, extra = n % (size_t)world, r = (size_t)rank;
    const size_t b = r * base + (r < extra ? r : extra);
    if (lo) *lo = b;
    if (hi) *hi = b + base + (r < extra ? 1 : 0);
}

extern "C" la_status la_comm_create(int world, int rank, la_exchange_fn exchange, void* user, la_comm** comm) {
    if (!comm) return la_fail(LA_ERR_ARG, "la_comm_create: comm is NULL");
    *comm = nullptr;
    if (world < 1 || world > COMM_MAX_WORLD || rank < 0 || rank >= world) return la_fail(LA_ERR_ARG, "la_comm_create: world %d / rank %d out of range (at most %d ranks)", world, rank, COMM_MAX_WORLD);
    if (world > 1 && !exchange) return la_fail(LA_ERR_ARG, "la_comm_create: an exchange callback is needed for more than one rank");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    la_comm* c = new la_comm();
    c->world = world; c->rank = rank; c->device = info->device; c->exchange = exchange; c->user = user;
    void* rh = nullptr;
    if (cudaHostAlloc(&rh, 64, cudaHostAllocMapped) != cudaSuccess) { delete c; return la_cuda_fail(cudaGetLastError(), "cudaHostAlloc", __FILE__, __LINE__); }
    memset(rh, 0, 64);
    c->res_host = (unsigned long long*)rh;
    void* rd = nullptr; cudaHostGetDevicePointer(&rd, rh, 0); c->res_dev = (unsigned long long*)rd;
    if (world > 1 && getenv("LA_COMM_NO_PEER_MEMORY")) {      // test hook: behave as on a box without peer access (same on every rank)
        c->note = "peer memory switched off by LA_COMM_NO_PEER_MEMORY";
    } else if (world > 1) {
        bool ok = false;
        la_status s = export_collective(c, c->box, la_p2p_mailbox_bytes(), &ok);
        if (s == LA_OK && ok) { c->vec_half = size_t(4) << 20; s = export_collective(c, c->vec, 2 * c->vec_half, &ok); }
        if (s != LA_OK) { unexport(c, c->box); unexport(c, c->vec); cudaFreeHost(rh); delete c; return s; }
        c->p2p = ok;                                       // false: no peer access / IPC refused somewhere -- the callback paths remain
        if (!ok) { unexport(c, c->box); unexport(c, c->vec); }
    }
    *comm = c;
    return LA_OK;
}

extern "C" la_status la_comm_destroy(la_comm* c) {
    if (!c) return LA_OK;
    cudaDeviceSynchronize();
    if (c->world > 1) barrier(c);                          // peers are done with this rank's buffers
    close_imports(c, c->box); close_imports(c, c->vec); close_imports(c, c->stage);
    if (c->world > 1) barrier(c);                          // ... and have unmapped them
    free_own(c, c->box); free_own(c, c->vec); free_own(c, c->stage);
    if (c->res_host) cudaFreeHost(c->res_host);
    if (c->side) cudaStreamDestroy(c->side);
    if (c->side_done) cudaEventDestroy(c->side_done);
    (void)cudaGetLastError();
    delete c;
    return LA_OK;
}

extern "C" la_status la_comm_info(const la_comm* c, int* world, int* rank, int* peer_memory) {
    if (!c) return la_fail(LA_ERR_ARG, "la_comm_info: comm is NULL");
    if (world) *world = c->world;
    if (rank) *rank = c->rank;
    if (peer_memory) *peer_memory = c->p2p ? 1 : 0;
    if (!c->note.empty()) la_set_error("%s", c->note.c_str());     // why peer memory is off, if it is
    return LA_OK;
}

extern "C" la_status la_comm_barrier(la_comm* c) {
    LA_TRY(check_comm(c, "la_comm_barrier"));
    return barrier(c);
}

// ---------------------------------------------------------------- sum / dot / ==
static la_status reduce_sharded(la_comm* c, int dtype, bool dot, const void* a, const void* b, size_t n, void* out_host, void* stream) {
    const char* who = dot ? "la_dot_vec_vec_sharded" : "la_sum_vec_sharded";
    LA_TRY(check_comm(c, who));
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "%s: unknown dtype %d", who, dtype);
    if (!out_host) return la_fail(LA_ERR_ARG, "%s: out_host is NULL", who);
    if (c->world == 1) return dot ? la_dot_vec_vec(dtype, a, b, n, out_host, stream) : la_sum_vec(dtype, a, n, out_host, stream);
    const size_t es = la_dtype_size(dtype);
    if (c->p2p) {
        const unsigned long long seq = ++c->seq;
        LA_TRY(la_fused_reduce(dtype, dot, a, b, n, c->res_dev, c->res_dev + 1, c->box.ptr, c->world, c->rank, seq, stream));
        // the kernel stores the total and then the round's number into mapped host memory: no copy, no stream synchronisation
        volatile unsigned long long* done = c->res_host + 1;
        const auto t0 = std::chrono::steady_clock::now();
        unsigned spins = 0;
        while (*done != seq) {
            if ((++spins & 0xfffu) == 0 && std::chrono::steady_clock::now() - t0 > std::chrono::seconds(20)) {
                LA_CUDA(cudaStreamSynchronize((cudaStream_t)stream));          // a failed launch / trapped kernel reports here
                if (*done != seq) return la_fail(LA_ERR_CUDA, "%s: the reduction finished without delivering its result", who);
            }
        }
        std::atomic_thread_fence(std::memory_order_acquire);
        memcpy(out_host, (const void*)c->res_host, es);
        return LA_OK;
    }
    unsigned long long mine = 0, parts[COMM_MAX_WORLD] = {};
    LA_TRY(dot ? la_dot_vec_vec(dtype, a, b, n, &mine, stream) : la_sum_vec(dtype, a, n, &mine, stream));
    LA_TRY(xchg(c, &mine, parts, 8));
    combine_host(dtype, (const unsigned char*)parts, 8, c->world, out_host);
    return LA_OK;
}
extern "C" la_status la_sum_vec_sharded(la_comm* c, int dtype, const void* data, size_t n_local, void* out_host, void* stream) {
    return reduce_sharded(c, dtype, false, data, nullptr, n_local, out_host, stream);
}
extern "C" la_status la_dot_vec_vec_sharded(la_comm* c, int dtype, const void* a, const void* b, size_t n_local, void* out_host, void* stream) {
    return reduce_sharded(c, dtype, true, a, b, n_local, out_host, stream);
}
extern "C" la_status la_eq_vec_sharded(la_comm* c, int dtype, const void* A, const void* B, size_t n_local, int* equal_host, void* stream) {
    LA_TRY(check_comm(c, "la_eq_vec_sharded"));
    if (!equal_host) return la_fail(LA_ERR_ARG, "la_eq_vec_sharded: equal_host is NULL");
    int mine = 1;
    LA_TRY(la_eq_vec(dtype, A, B, n_local, &mine, stream));
    bool all = false;
    LA_TRY(agree(c, mine != 0, &all));
    *equal_host = all ? 1 : 0;
    return LA_OK;
}

// ---------------------------------------------------------------- v x M (v and M co-sharded by rows)
// C (cols elements, the full result, on every rank) = sum over ranks of  v_local x M_local.
extern "C" la_status la_mul_vec_mat_sharded(la_comm* c, int dtype, void* C, const void* v_local, const void* M_local, size_t rows_local, size_t cols, void* stream) {
    LA_TRY(check_comm(c, "la_mul_vec_mat_sharded"));
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_mul_vec_mat_sharded: unknown dtype %d", dtype);
    if (c->world == 1) return la_mul_vec_mat(dtype, C, v_local, M_local, rows_local, cols, stream);
    if (cols == 0) return LA_OK;
    const size_t es = la_dtype_size(dtype), bytes = cols * es;
    cudaStream_t st = (cudaStream_t)stream;
    if (c->p2p) {
        if (bytes > c->vec_half) {                          // same `cols` on every rank: all of them grow together
            LA_CUDA(cudaStreamSynchronize(st));
            const size_t half = (bytes + (size_t(1) << 20) - 1) >> 20 << 20;
            bool ok = false;
            LA_TRY(export_collective(c, c->vec, 2 * half, &ok));
            if (!ok) return la_fail(LA_ERR_OOM, "la_mul_vec_mat_sharded: could not grow the exported vector buffers to 2 x %zu bytes", half);
            c->vec_half = half;
        }
        const unsigned long long seq = ++c->seq;
        void* mine = (char*)c->vec.ptr[c->rank] + (size_t)(seq & 1ull) * c->vec_half;
        // rows_local == 0 (more ranks than rows): the partial is all zeros
        if (rows_local == 0) LA_CUDA(cudaMemsetAsync(mine, 0, bytes, st));
        else LA_TRY(la_mul_vec_mat(dtype, mine, v_local, M_local, rows_local, cols, stream));
        return la_vec_allreduce(dtype, C, cols, c->vec.ptr, c->vec_half, c->box.ptr, c->world, c->rank, seq, stream);
    }
    // no peer memory: partial through the host and the exchange callback, rank-ordered adds on the host
    void* part = nullptr; LA_TRY(la_malloc(&part, bytes));
    la_status s = rows_local ? la_mul_vec_mat(dtype, part, v_local, M_local, rows_local, cols, stream) : LA_OK;
    if (s == LA_OK && rows_local == 0 && cudaMemsetAsync(part, 0, bytes, st) != cudaSuccess) s = LA_ERR_CUDA;
    unsigned char* mine = (unsigned char*)malloc(bytes); unsigned char* all = (unsigned char*)malloc(bytes * c->world); unsigned char* sum = (unsigned char*)malloc(bytes);
    if (s == LA_OK && (!mine || !all || !sum)) s = la_fail(LA_ERR_OOM, "la_mul_vec_mat_sharded: out of host memory for %zu bytes x %d ranks", bytes, c->world);
    if (s == LA_OK) s = la_download(mine, part, bytes, stream);
    if (s == LA_OK) s = xchg(c, mine, all, bytes);
    if (s == LA_OK) {
        for (size_t i = 0; i < cols; ++i) combine_host(dtype, all + i * es, bytes, c->world, sum + i * es);
        s = la_upload(C, sum, bytes, stream);
        if (s == LA_OK && cudaStreamSynchronize(st) != cudaSuccess) s = LA_ERR_CUDA;
    }
    free(mine); free(all); free(sum); la_free(part);
    return s;
}

// ---------------------------------------------------------------- transpose of a row-sharded matrix
namespace {

struct PeerPtrs { void* p[COMM_MAX_WORLD]; };

__host__ __device__ inline int span_owner(size_t i, size_t base, size_t extra) {            // which rank's block holds index i (la_partition)
    const size_t cut = extra * (base + 1);
    return i < cut ? (int)(i / (base + 1)) : (int)(extra + (i - cut) / (base ? base : 1));
}
__host__ __device__ inline size_t span_lo(int r, size_t base, size_t extra) { return (size_t)r * base + ((size_t)r < extra ? (size_t)r : extra); }

// The local row block (rows_p x cols_g, rows lo_p.. of the global matrix) turned and scattered: column c of it is row c of the
// transpose, which belongs to the rank that owns index c of cols_g; it lands in that rank's staging buffer at
// [(c - first column of the owner) * rows_g + lo_p + r].  32 x 32 tiles through shared memory, both sides coalesced; the
// writes are plain stores into peer memory (NVLink) or local memory.
template <typename E>
__global__ void __launch_bounds__(256)
transpose_scatter_kernel(PeerPtrs stage, const E* __restrict__ in, size_t rows_p, size_t cols_g, size_t rows_g, size_t lo_p, size_t col_base, size_t col_extra, unsigned tiles_x) {
    __shared__ E tile[32][33];
    const size_t r0 = (size_t)(blockIdx.x / tiles_x) * 32, c0 = (size_t)(blockIdx.x % tiles_x) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
    for (int j = ty; j < 32; j += 8)
        if (r0 + j < rows_p && c0 + tx < cols_g) tile[j][tx] = in[(r0 + j) * cols_g + c0 + tx];
    __syncthreads();
#pragma unroll
    for (int j = ty; j < 32; j += 8) {
        const size_t cidx = c0 + j;
        if (cidx < cols_g && r0 + tx < rows_p) {
            const int q = span_owner(cidx, col_base, col_extra);
            E* dst = reinterpret_cast<E*>(stage.p[q]);
            dst[(cidx - span_lo(q, col_base, col_extra)) * rows_g + lo_p + r0 + tx] = tile[tx][j];
        }
    }
}

la_status ensure_stage(la_comm* c, size_t bytes, const char* who) {
    if (c->stage.bytes >= bytes) return LA_OK;
    bool ok = false;
    LA_TRY(export_collective(c, c->stage, bytes, &ok));
    if (!ok) return la_fail(LA_ERR_OOM, "%s: could not create the exported staging buffers (%zu bytes per rank): %s", who, bytes, c->note.c_str());
    return LA_OK;
}

}  // namespace

// in_local: this rank's rows [lo_p, hi_p) of the rows_g x cols_g matrix; out_local: this rank's rows of the cols_g x rows_g
// transpose (la_partition of cols_g), rows_g elements each.  Collective; returns with out_local's copy queued on `stream`.
extern "C" la_status la_transpose_sharded(la_comm* c, int dtype, void* out_local, const void* in_local, size_t rows_g, size_t cols_g, void* stream) {
    LA_TRY(check_comm(c, "la_transpose_sharded"));
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_transpose_sharded: unknown dtype %d", dtype);
    size_t lo_p, hi_p, clo, chi;
    la_partition(rows_g, c->world, c->rank, &lo_p, &hi_p);
    la_partition(cols_g, c->world, c->rank, &clo, &chi);
    if (c->world == 1) return la_transpose(dtype, out_local, in_local, rows_g, cols_g, stream);
    if (!c->p2p) return la_fail(LA_ERR_UNSUPPORTED, "la_transpose_sharded: needs peer memory access between the GPUs (%s)", c->note.c_str());
    if (rows_g == 0 || cols_g == 0) return LA_OK;
    const size_t es = la_dtype_size(dtype);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t max_cols = (cols_g + c->world - 1) / c->world;
    if (c->stage.bytes < max_cols * rows_g * es) {          // (same decision on every rank: it depends on the global shape only)
        LA_CUDA(cudaStreamSynchronize(st));
        LA_TRY(ensure_stage(c, max_cols * rows_g * es, "la_transpose_sharded"));
    }
    // Everything below is stream-ordered; the ranks meet in two barriers ON the stream (p2p.cu), not on the host.
    LA_TRY(stream_barrier(c, st));                          // every rank's earlier copy out of its staging buffer is done: they may be overwritten
    const size_t rows_p = hi_p - lo_p;
    if (rows_p) {
        PeerPtrs pp; for (int p = 0; p < COMM_MAX_WORLD; ++p) pp.p[p] = c->stage.ptr[p];
        const size_t tx = (cols_g + 31) / 32, ty = (rows_p + 31) / 32;
        if (tx * ty > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "la_transpose_sharded: the local block is too large");
        const size_t cb = cols_g / c->world, ce = cols_g % c->world;
        switch (es) {
        case 1: transpose_scatter_kernel<uint8_t><<<(unsigned)(tx * ty), 256, 0, st>>>(pp, (const uint8_t*)in_local, rows_p, cols_g, rows_g, lo_p, cb, ce, (unsigned)tx); break;
        case 2: transpose_scatter_kernel<uint16_t><<<(unsigned)(tx * ty), 256, 0, st>>>(pp, (const uint16_t*)in_local, rows_p, cols_g, rows_g, lo_p, cb, ce, (unsigned)tx); break;
        case 4: transpose_scatter_kernel<uint32_t><<<(unsigned)(tx * ty), 256, 0, st>>>(pp, (const uint32_t*)in_local, rows_p, cols_g, rows_g, lo_p, cb, ce, (unsigned)tx); break;
        default: transpose_scatter_kernel<uint64_t><<<(unsigned)(tx * ty), 256, 0, st>>>(pp, (const uint64_t*)in_local, rows_p, cols_g, rows_g, lo_p, cb, ce, (unsigned)tx); break;
        }
        LA_LAUNCH_CHECK();
    }
    LA_TRY(stream_barrier(c, st));                          // my stores have landed in the peers' memory and everybody's in mine
    if (chi > clo) LA_CUDA(cudaMemcpyAsync(out_local, c->stage.ptr[c->rank], (chi - clo) * rows_g * es, cudaMemcpyDeviceToDevice, st));
    return LA_OK;
}

// ---------------------------------------------------------------- M x M on host buffers, A and C row-sharded
// A_host_local / C_host_local: this rank's rows (la_partition of M_global) of A (x K) and C (x N); B_host: all of B (K x N) on
// every rank.  Each rank uploads only its K / world rows of B over its own PCIe link; the other slices arrive from the peers'
// staging buffers over NVLink; then la_mul_mat_mat_host streams the local rows against the device-resident B.
extern "C" la_status la_mul_mat_mat_host_sharded(la_comm* c, int dtype, void* C_host_local, const void* A_host_local, const void* B_host, size_t M_global, size_t K, size_t N) {
    LA_TRY(check_comm(c, "la_mul_mat_mat_host_sharded"));
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_mul_mat_mat_host_sharded: unknown dtype %d", dtype);
    size_t lo, hi; la_partition(M_global, c->world, c->rank, &lo, &hi);
    const size_t rows = hi - lo, es = la_dtype_size(dtype);
    if (c->world == 1 || !c->p2p || K == 0 || N == 0 || K % (size_t)c->world != 0)
        return la_mul_mat_mat_host(dtype, C_host_local, A_host_local, B_host, rows, K, N);       // every rank uploads all of its B
    static const bool dbg = getenv("LA_DEBUG_HOSTGEMM") != nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    auto ms_since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
    if (c->stage.bytes < K * N * es) LA_TRY(ensure_stage(c, K * N * es, "la_mul_mat_mat_host_sharded"));
    if (!c->side) {
        LA_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
        LA_CUDA(cudaEventCreateWithFlags(&c->side_done, cudaEventDisableTiming));
    }
    // The gather of B runs on a stream of its own, beside the uploads of A that la_mul_mat_mat_host_dev_b starts at once; its
    // compute stream waits for `side_done` before it touches B.  Barriers are kernels on that stream (no host round trip).
    cudaStream_t st = c->side;
    LA_TRY(stream_barrier(c, st));                          // nobody still reads a slice of an earlier B from my staging buffer
    const size_t slice = (K / (size_t)c->world) * N * es;
    char* mine = (char*)c->stage.ptr[c->rank];
    LA_CUDA(cudaMemcpyAsync(mine + c->rank * slice, (const char*)B_host + c->rank * slice, slice, cudaMemcpyHostToDevice, st));
    LA_TRY(stream_barrier(c, st));                          // every slice is on its owner's GPU
    for (int k = 1; k < c->world; ++k) {                    // start with the next rank, so the pulls spread over the peers
        const int p = (c->rank + k) % c->world;
        LA_CUDA(cudaMemcpyAsync(mine + p * slice, (const char*)c->stage.ptr[p] + p * slice, slice, cudaMemcpyDefault, st));
    }
    LA_CUDA(cudaEventRecord(c->side_done, st));
    if (dbg) {
        fprintf(stderr, "[hostgemm rank %d] gather of B enqueued at %.3f ms\n", c->rank, ms_since());
        cudaStreamSynchronize(st);
        fprintf(stderr, "[hostgemm rank %d] gather of B complete at %.3f ms (debug mode waits here)\n", c->rank, ms_since());
    }
    const la_status rs = la_mul_mat_mat_host_dev_b(dtype, C_host_local, A_host_local, mine, c->side_done, rows, K, N);
    if (dbg) fprintf(stderr, "[hostgemm rank %d] done at %.3f ms\n", c->rank, ms_since());
    return rs;
}
