// csrc/p2p.cu -- reductions fused with their cross-GPU exchange over NVLink peer memory.
//
// Multi-GPU sum / dot (no counterpart in the single-device reference) need one
// exchange: every rank's partial has to reach every other rank.  Instead of a separate NCCL collective after
// the reduction kernel, the reduction kernel itself finishes the job: the last block of rank r stores r's
// partial straight into slot r of every peer's "mailbox" (cudaIpc-mapped device memory, plain st.global over
// NVLink / NVSwitch), then waits until its own mailbox holds all `world` partials of this round and adds them
// in rank order.  One launch per rank, deterministic, identical result on every rank, any element type.
//
// Mailbox: 2 rounds (parity of the sequence number) x world slots x {u64 seq, 8-byte value}.  A rank can be at
// most one round ahead of the slowest peer (it needs everybody's value of round s to finish round s), so two
// parities never collide.  Every rank must call the collectives in the same order with the same sequence
// numbers; waits are bounded and trap rather than hang.
#include "common.cuh"

namespace {

struct alignas(16) MailSlot { unsigned long long seq; unsigned long long value; };
constexpr int P2P_MAX_WORLD = 16;
constexpr int P2P_THREADS = 256;
constexpr int P2P_UNROLL = 4;

struct PeerTable { MailSlot* box[P2P_MAX_WORLD]; };      // box[p] = rank p's mailbox as mapped in this process

template <typename T, bool DOT>
__global__ void __launch_bounds__(P2P_THREADS)
reduce_allreduce_kernel(const T* __restrict__ a, const T* __restrict__ b, size_t n, bool packed,
                        typename Arith<T>::Acc* partials, unsigned* ticket, T* out, unsigned long long* done_flag,
                        PeerTable peers, int world, int rank, unsigned long long seq) {
    using A = typename Arith<T>::Acc;
    using P = Pack<T>;
    constexpr int N = P::N;
    __shared__ A smem[32];
    __shared__ bool is_last;
    const size_t stride = (size_t)gridDim.x * P2P_THREADS;
    const size_t tid = (size_t)blockIdx.x * P2P_THREADS + threadIdx.x;
    A acc[P2P_UNROLL];
#pragma unroll
    for (int j = 0; j < P2P_UNROLL; ++j) acc[j] = A(0);
    size_t done = 0;
    if (packed) {
        const size_t n_pack = n / N;
        const P* ap = reinterpret_cast<const P*>(a);
        const P* bp = reinterpret_cast<const P*>(b);
        size_t p = tid;
        for (; p + (P2P_UNROLL - 1) * stride < n_pack; p += P2P_UNROLL * stride) {
            P x[P2P_UNROLL], y[P2P_UNROLL];
#pragma unroll
            for (int j = 0; j < P2P_UNROLL; ++j) { x[j] = ld_stream(ap + p + j * stride); if (DOT) y[j] = ld_stream(bp + p + j * stride); }
#pragma unroll
            for (int j = 0; j < P2P_UNROLL; ++j)
#pragma unroll
                for (int e = 0; e < N; ++e) acc[j] = DOT ? Arith<T>::fma(x[j].v[e], y[j].v[e], acc[j]) : acc[j] + Arith<T>::widen(x[j].v[e]);
        }
        for (; p < n_pack; p += stride) {
            P x = ld_stream(ap + p), y;
            if (DOT) y = ld_stream(bp + p);
#pragma unroll
            for (int e = 0; e < N; ++e) acc[0] = DOT ? Arith<T>::fma(x.v[e], y.v[e], acc[0]) : acc[0] + Arith<T>::widen(x.v[e]);
        }
        done = n_pack * N;
    }
    for (size_t i = done + tid; i < n; i += stride) acc[0] = DOT ? Arith<T>::fma(a[i], b[i], acc[0]) : acc[0] + Arith<T>::widen(a[i]);
    A v = block_sum((acc[0] + acc[1]) + (acc[2] + acc[3]), smem);

    if (threadIdx.x == 0) {
        partials[blockIdx.x] = v;
        __threadfence();
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    A t = A(0);
    for (unsigned k = threadIdx.x; k < gridDim.x; k += P2P_THREADS) t += __ldcg(partials + k);
    t = block_sum(t, smem);                               // this rank's partial, in thread 0
    __shared__ unsigned long long mine;
    if (threadIdx.x == 0) {
        T local = Arith<T>::from_acc(t);
        unsigned long long bits = 0;
        memcpy(&bits, &local, sizeof(T));
        mine = bits;
        *ticket = 0;
    }
    __syncthreads();
    // ---- exchange: thread p delivers to peer p, then waits for peer p's delivery
    const int par = (int)(seq & 1ull);
    if ((int)threadIdx.x < world) {
        // value, then the round's number with release semantics at system scope; the reader acquires the number before it
        // reads the value.  (One store and one load each way instead of two full system fences per slot.)
        MailSlot* dst = peers.box[threadIdx.x] + par * P2P_MAX_WORLD + rank;
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(&dst->value), "l"(mine) : "memory");
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&dst->seq), "l"(seq) : "memory");
        const MailSlot* src = peers.box[rank] + par * P2P_MAX_WORLD + threadIdx.x;
        unsigned spins = 0;
        unsigned long long seen;
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(&src->seq) : "memory");
            if (seen != seq && ++spins > (1u << 27)) __trap();           // a peer never arrived: fail loudly instead of hanging
        } while (seen != seq);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        T total = T(0);
        for (int p = 0; p < world; ++p) {                 // rank order: deterministic and identical everywhere
            const unsigned long long bits = *reinterpret_cast<const volatile unsigned long long*>(&(peers.box[rank] + par * P2P_MAX_WORLD + p)->value);
            T x; memcpy(&x, &bits, sizeof(T));
            total = Arith<T>::add(total, x);
        }
        *out = total;
        if (done_flag) {                                  // `out` and the flag may be zero-copy host memory that the host polls
            __threadfence_system();
            *reinterpret_cast<volatile unsigned long long*>(done_flag) = seq;
        }
    }
}

// out[c] = sum over ranks p (ascending) of peer_vec[p][c].  Every rank has put its own contribution into its
// exported vector buffer with earlier work on the same stream; block 0 announces that to all peers, every block
// waits for all announcements, then reads the peers' vectors directly over NVLink.
template <typename T>
__global__ void __launch_bounds__(P2P_THREADS)
vector_allreduce_kernel(T* __restrict__ out, size_t n, PeerTable vecs, size_t par_offset_bytes, PeerTable boxes, int world, int rank, unsigned long long seq) {
    const int par = (int)(seq & 1ull);
    if ((int)threadIdx.x < world) {
        if (blockIdx.x == 0) {
            __threadfence_system();
            MailSlot* dst = boxes.box[threadIdx.x] + (2 + par) * P2P_MAX_WORLD + rank;
            *reinterpret_cast<volatile unsigned long long*>(&dst->seq) = seq;
        }
        const MailSlot* src = boxes.box[rank] + (2 + par) * P2P_MAX_WORLD + threadIdx.x;
        unsigned spins = 0;
        while (*reinterpret_cast<const volatile unsigned long long*>(&src->seq) != seq) {
            if (++spins > (1u << 27)) __trap();
        }
        __threadfence_system();
    }
    __syncthreads();
    const size_t stride = (size_t)gridDim.x * P2P_THREADS;
    for (size_t c = (size_t)blockIdx.x * P2P_THREADS + threadIdx.x; c < n; c += stride) {
        T acc = T(0);
        for (int p = 0; p < world; ++p) {
            const T* v = reinterpret_cast<const T*>(reinterpret_cast<const char*>(vecs.box[p]) + par_offset_bytes);
            acc = Arith<T>::add(acc, *reinterpret_cast<const volatile T*>(v + c));
        }
        out[c] = acc;
    }
}

// Barrier between the ranks ON THE STREAM: thread p announces round `seq` to peer p and waits for peer p's announcement.  Work
// queued before it on this stream (stores into peers' memory included) has completed before any peer sees the announcement;
// work queued behind it starts after every rank has arrived.  Rows [4, 6) of the mailbox, one per parity of seq.
__global__ void __launch_bounds__(32)
peer_barrier_kernel(PeerTable boxes, int world, int rank, unsigned long long seq) {
    const int par = (int)(seq & 1ull);
    if ((int)threadIdx.x < world) {
        __threadfence_system();
        MailSlot* dst = boxes.box[threadIdx.x] + (4 + par) * P2P_MAX_WORLD + rank;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&dst->seq), "l"(seq) : "memory");
        const MailSlot* src = boxes.box[rank] + (4 + par) * P2P_MAX_WORLD + threadIdx.x;
        unsigned spins = 0;
        unsigned long long seen;
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(&src->seq) : "memory");
            if (seen != seq && ++spins > (1u << 28)) __trap();
        } while (seen != seq);
    }
}

la_status fused_entry(int dtype, bool dot, const void* a, const void* b, size_t n, void* out_dev, unsigned long long* done_flag, void* const* peer_mailboxes, int world, int rank,
                      unsigned long long seq, void* stream) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "unknown dtype %d", dtype);
    if (!out_dev || !peer_mailboxes || (n && (!a || (dot && !b)))) return la_fail(LA_ERR_ARG, "la_*_allreduce: NULL operand");
    if (world < 1 || world > P2P_MAX_WORLD || rank < 0 || rank >= world) return la_fail(LA_ERR_ARG, "la_*_allreduce: world %d / rank %d out of range (max %d)", world, rank, P2P_MAX_WORLD);
    if (seq == 0) return la_fail(LA_ERR_ARG, "la_*_allreduce: sequence numbers start at 1");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    cudaStream_t st = (cudaStream_t)stream;
    PeerTable pt;
    for (int p = 0; p < P2P_MAX_WORLD; ++p) pt.box[p] = p < world ? (MailSlot*)peer_mailboxes[p] : nullptr;
    LA_DISPATCH_DTYPE(dtype, T, {
        using A = typename Arith<T>::Acc;
        const bool packed = n > 0 && la_aligned16(a) && (!dot || la_aligned16(b));
        const size_t work = packed ? (n / Pack<T>::N + 1) : (n ? n : 1);
        size_t blocks = (work + (size_t)P2P_THREADS * P2P_UNROLL - 1) / ((size_t)P2P_THREADS * P2P_UNROLL);
        const size_t cap = (size_t)info->sm_count * 4;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        if (blocks * sizeof(A) > LA_REDUCE_SCRATCH_BYTES) blocks = LA_REDUCE_SCRATCH_BYTES / sizeof(A);
        void* scratch; LA_TRY(la_scratch(st, LA_REDUCE_SCRATCH_BYTES, &scratch));
        unsigned* ticket = (unsigned*)scratch;
        A* partials = (A*)((char*)scratch + LA_SCRATCH_ZEROED);
        if (dot) reduce_allreduce_kernel<T, true><<<(unsigned)blocks, P2P_THREADS, 0, st>>>((const T*)a, (const T*)b, n, packed, partials, ticket, (T*)out_dev, done_flag, pt, world, rank, seq);
        else     reduce_allreduce_kernel<T, false><<<(unsigned)blocks, P2P_THREADS, 0, st>>>((const T*)a, (const T*)nullptr, n, packed, partials, ticket, (T*)out_dev, done_flag, pt, world, rank, seq);
    });
    LA_LAUNCH_CHECK();
    return LA_OK;
}

}  // namespace

// comm.cu: the same launch with the result (and a completion word) wherever the caller wants it, e.g. zero-copy host memory
la_status la_fused_reduce(int dtype, bool dot, const void* a, const void* b, size_t n, void* out_dev, unsigned long long* done_flag,
                          void* const* peer_mailboxes, int world, int rank, unsigned long long seq, void* stream) {
    return fused_entry(dtype, dot, a, b, n, out_dev, done_flag, peer_mailboxes, world, rank, seq, stream);
}

la_status la_peer_barrier(void* const* peer_mailboxes, int world, int rank, unsigned long long seq, cudaStream_t st) {
    PeerTable pt;
    for (int p = 0; p < P2P_MAX_WORLD; ++p) pt.box[p] = p < world ? (MailSlot*)peer_mailboxes[p] : nullptr;
    peer_barrier_kernel<<<1, 32, 0, st>>>(pt, world, rank, seq);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

// ---------------------------------------------------------------- mailbox plumbing (cudaIpc)
extern "C" size_t la_p2p_mailbox_bytes(void) { return 6 * P2P_MAX_WORLD * sizeof(MailSlot); }   // [0,2): scalar rounds, [2,4): vector-ready flags, [4,6): stream barriers
extern "C" size_t la_p2p_handle_bytes(void) { return sizeof(cudaIpcMemHandle_t); }

// Allocates an exportable device buffer (zeroed; plain cudaMalloc so that cudaIpc can export it) and writes its
// IPC handle.  A mailbox is such a buffer of la_p2p_mailbox_bytes().
extern "C" la_status la_p2p_buffer_create(size_t bytes, void** mailbox_dev, void* ipc_handle_out) {
    if (!mailbox_dev || !ipc_handle_out || bytes == 0) return la_fail(LA_ERR_ARG, "la_p2p_buffer_create: bad argument");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    void* p = nullptr;
    LA_CUDA(cudaMalloc(&p, bytes));
    LA_CUDA(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    LA_CUDA(cudaIpcGetMemHandle(&h, p));
    memcpy(ipc_handle_out, &h, sizeof(h));
    *mailbox_dev = p;
    return LA_OK;
}
// Maps a peer's buffer (handle from la_p2p_buffer_create in the peer process) into this process.
extern "C" la_status la_p2p_buffer_open(const void* ipc_handle, void** peer_mailbox_dev) {
    if (!ipc_handle || !peer_mailbox_dev) return la_fail(LA_ERR_ARG, "la_p2p_buffer_open: NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle, sizeof(h));
    LA_CUDA(cudaIpcOpenMemHandle(peer_mailbox_dev, h, cudaIpcMemLazyEnablePeerAccess));
    return LA_OK;
}
extern "C" la_status la_p2p_buffer_close(void* peer_mailbox_dev) {
    if (peer_mailbox_dev) LA_CUDA(cudaIpcCloseMemHandle(peer_mailbox_dev));
    return LA_OK;
}
extern "C" la_status la_p2p_buffer_destroy(void* mailbox_dev) {
    if (mailbox_dev) LA_CUDA(cudaFree(mailbox_dev));
    return LA_OK;
}

// sum / dot over all ranks in one launch per rank: *out_dev (one element, on this device) receives the rank-ordered
// total.  peer_mailboxes[p] = rank p's mailbox as mapped in this process (own mailbox at [rank]).
extern "C" la_status la_sum_vec_allreduce(int dtype, const void* data, size_t n, void* out_dev, void* const* peer_mailboxes, int world, int rank,
                                          uint64_t seq, void* stream) {
    return fused_entry(dtype, false, data, nullptr, n, out_dev, nullptr, peer_mailboxes, world, rank, seq, stream);
}
extern "C" la_status la_dot_vec_vec_allreduce(int dtype, const void* a, const void* b, size_t n, void* out_dev, void* const* peer_mailboxes, int world,
                                              int rank, uint64_t seq, void* stream) {
    return fused_entry(dtype, true, a, b, n, out_dev, nullptr, peer_mailboxes, world, rank, seq, stream);
}

// Sum of every rank's vector, rank order, one launch per rank.  Each rank's contribution (n elements) must already
// sit in ITS exported buffer at byte offset (seq & 1) * half_bytes, written by earlier work on `stream`;
// peer_vectors[p] = rank p's buffer as mapped here.  out_dev must not be one of those buffers.
extern "C" la_status la_vec_allreduce(int dtype, void* out_dev, size_t n, void* const* peer_vectors, size_t half_bytes, void* const* peer_mailboxes,
                                      int world, int rank, uint64_t seq, void* stream) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_vec_allreduce: unknown dtype %d", dtype);
    if (!out_dev || !peer_vectors || !peer_mailboxes) return la_fail(LA_ERR_ARG, "la_vec_allreduce: NULL operand");
    if (world < 1 || world > P2P_MAX_WORLD || rank < 0 || rank >= world || seq == 0) return la_fail(LA_ERR_ARG, "la_vec_allreduce: bad world / rank / seq");
    if (n * la_dtype_size(dtype) > half_bytes) return la_fail(LA_ERR_ARG, "la_vec_allreduce: %zu elements do not fit the %zu-byte half buffer", n, half_bytes);
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    PeerTable vt, bt;
    for (int p = 0; p < P2P_MAX_WORLD; ++p) { vt.box[p] = p < world ? (MailSlot*)peer_vectors[p] : nullptr; bt.box[p] = p < world ? (MailSlot*)peer_mailboxes[p] : nullptr; }
    size_t blocks = (n + P2P_THREADS - 1) / P2P_THREADS;
    const size_t cap = (size_t)info->sm_count * 2;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const size_t off = (size_t)(seq & 1ull) * half_bytes;
    LA_DISPATCH_DTYPE(dtype, T, (vector_allreduce_kernel<T><<<(unsigned)blocks, P2P_THREADS, 0, (cudaStream_t)stream>>>((T*)out_dev, n, vt, off, bt, world, rank, seq)));
    LA_LAUNCH_CHECK();
    return LA_OK;
}
