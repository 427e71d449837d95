/* include/linalg_b200.h -- C ABI of liblinalg_b200.so
 *
 * The B200-native (sm_100a CUDA) replacement for the OpenCL layer of usbalbin/linear_algebra:
 * everything its Vector<T>/Matrix<T> operators reach through `get_kernels::<T>()` /
 * `cl_data::<T>()` (src/lib.rs:125-154) -- the 18 kernels of src/kernels.cl and the ocl::Buffer
 * plumbing around them.  Plain C: pointers, sizes, enums.  No torch/C++ types cross this boundary.
 *
 * Conventions
 *  - Pointers named C/A/B/data/a/b/out/in are DEVICE pointers on the current CUDA device, dense
 *    and contiguous; matrices are row-major, element (r,c) at r*cols + c (src/matrix.rs:294,
 *    src/kernels.cl:165-168).  Any alignment is accepted; 16-byte aligned pointers (what
 *    la_malloc returns) take the 128-bit vectorised paths.
 *  - Sizes are 64-bit element counts (the reference passes `as i32`, src/vector.rs:277,
 *    src/matrix.rs:255-256; 2^31 elements and more are fine here).
 *  - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Compute
 *    entry points are stream-ordered and return without waiting unless they hand a scalar back
 *    to the host (la_sum_vec, la_dot_vec_vec, la_eq_vec).  The reference blocks on every kernel
 *    (event.wait_for(), src/vector.rs:108); its drop-in wrappers call la_stream_sync() after
 *    each op to keep that behaviour.
 *  - Every function returns la_status; LA_OK == 0.  la_last_error() gives the thread-local
 *    message.  The library never aborts and has no CPU fallback: without a CUDA device every
 *    compute call fails with LA_ERR_CUDA.
 *  - Integer arithmetic wraps mod 2^n; division truncates toward zero; x/0 and MIN/-1 are
 *    unspecified values (undefined in OpenCL C).  Floating point: add/sub/mul/div are single
 *    IEEE operations (bit-exact with the reference); sum/dot/GEMV/GEMM accumulate in the
 *    element type in a fixed, deterministic order that differs from the reference's
 *    device-dependent order (tolerance: DESIGN.md).
 *  - Thread safety: any thread may call any function; scratch memory is keyed by
 *    (device, stream), so concurrent calls must use distinct streams.  la_mul_mat_mat_host
 *    calls on one device are serialised inside the library (they share its copy streams).
 *    la_set_gemm_mode and the knobs of linalg_b200_tuning.h are process-wide settings: set
 *    them before threads start calling, not while products are in flight.
 *  - Buffers and streams: la_malloc / la_free are stream-ordered pool operations on the legacy
 *    default stream.  The library orders them against every stream made by la_stream_create
 *    (la_free first lets the legacy stream wait for all work queued on those streams so far;
 *    after la_malloc those streams wait for the legacy stream), so a buffer may be freed right
 *    after queueing work that uses it and a fresh buffer may be used at once on any such
 *    stream.  Streams the caller created itself with cudaStreamNonBlocking are outside that
 *    arrangement: synchronise them before la_free of a buffer they use.
 */
#ifndef LINALG_B200_H
#define LINALG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int la_status;
enum {
    LA_OK = 0,
    LA_ERR_ARG = 1,    /* null pointer, unknown dtype/op, negative size */
    LA_ERR_SHAPE = 2,  /* shapes do not agree (the wrappers' assert_eq!, src/vector.rs:97,327,513,545) */
    LA_ERR_DTYPE = 3,  /* operation not defined for this dtype */
    LA_ERR_OOM = 4,    /* device allocation failed */
    LA_ERR_CUDA = 5,   /* CUDA runtime / driver error, incl. "no device" */
    LA_ERR_UNSUPPORTED = 6
};

/* Element types = the Rust types of `impl_type_to_str!` (src/traits.rs:30-57,76) / the OpenCL
 * names of TYPES (src/lib.rs:84-92).  Note src/traits.rs:36 maps Rust i8 to "uchar": a drop-in
 * wrapper passes LA_U8 for Vector<i8> to reproduce that; LA_I8 is OpenCL "char". */
typedef enum {
    LA_U8 = 0,  /* uchar  */
    LA_I8 = 1,  /* char   */
    LA_U16 = 2, /* ushort */
    LA_I16 = 3, /* short  */
    LA_U32 = 4, /* uint   */
    LA_I32 = 5, /* int    */
    LA_U64 = 6, /* ulong  (usize) */
    LA_I64 = 7, /* long   (isize) */
    LA_F32 = 8, /* float  */
    LA_F64 = 9  /* double (opt-in in the reference, src/lib.rs:343-347) */
} la_dtype;

typedef enum { LA_ADD = 0, LA_SUB = 1, LA_MUL = 2, LA_DIV = 3 } la_op;

/* ------------------------------------------------------------------ runtime
 * Replaces the lazy `CL_DATA` singleton, device pick and program build of src/lib.rs:125-149,
 * 310-335 (kernels here are compiled ahead of time for sm_100a; there is nothing to JIT). */
la_status la_init(int device);               /* select device, create its scratch; idempotent  */
la_status la_device_count(int* count);
la_status la_device_name(char* buf, size_t buf_len);
la_status la_sm_count(int* count);
const char* la_last_error(void);             /* thread-local, never NULL                        */
const char* la_version(void);
size_t la_dtype_size(int dtype);             /* 0 for an unknown dtype                          */

/* ------------------------------------------------------------------ buffers
 * Replace ocl::Buffer creation / read / copy: Vector::uninitialized_lock_free
 * (src/vector.rs:36-46), from_vec_lock_free (:58-69, COPY_HOST_PTR upload), to_vec (:71-78,
 * blocking read), Clone (:700-708, device-to-device copy), Vector::new (:30-34, fill). */
la_status la_malloc(void** dev_ptr, size_t bytes);                 /* 256-byte aligned; bytes==0 ok */
la_status la_free(void* dev_ptr);
la_status la_upload(void* dst_dev, const void* src_host, size_t bytes, void* stream);   /* async if src is pinned */
la_status la_download(void* dst_host, const void* src_dev, size_t bytes, void* stream); /* returns after the copy */
la_status la_copy(void* dst_dev, const void* src_dev, size_t bytes, void* stream);
/* `height` rows of `width_bytes`, rows `*_pitch` bytes apart (device to device; used to place the
 * blocks of a multi-GPU transpose). */
la_status la_copy_2d(void* dst_dev, size_t dst_pitch, const void* src_dev, size_t src_pitch, size_t width_bytes, size_t height, void* stream);
la_status la_fill(int dtype, void* dst_dev, const void* value_host, size_t n, void* stream);
/* Synthetic data generated on the device (no counterpart in the reference; bench / test utility):
 * element i = f(splitmix64(seed ^ i)): floats uniform in [-1,1) from the top 24 (f32) / 53 (f64)
 * bits, integers the top 8*sizeof(T) bits.  Reproducible on the host (linear_algebra_b200.synth). */
la_status la_synth(int dtype, void* dst_dev, uint64_t seed, size_t n, void* stream);
la_status la_host_alloc(void** host_ptr, size_t bytes);            /* pinned host memory */
la_status la_host_free(void* host_ptr);
la_status la_stream_create(void** stream);
la_status la_stream_destroy(void* stream);
la_status la_stream_sync(void* stream);      /* the reference's event.wait_for() */
/* Device-side timing on the launching stream (ocl::Event profiling has no user in the
 * reference; used by bench.py). */
la_status la_event_create(void** event);
la_status la_event_destroy(void* event);
la_status la_event_record(void* event, void* stream);
la_status la_event_elapsed_ms(void* start, void* stop, float* ms);   /* waits for `stop` */

/* ------------------------------------------------------------------ elementwise
 * {T}_add/sub/mul/div_vec_vec            src/kernels.cl:15-29   C[i] = A[i] op B[i]
 *   host: Vector::from_for_each2_lock_free src/vector.rs:96-112 (ops :366-379,403-416,482-496,571-586)
 * {T}_add/sub/mul/div_assign_vec_vec     src/kernels.cl:33-47   C[i] op= B[i]
 *   host: Vector::for_each_mut src/vector.rs:114-125 (ops :381-391,418-427,498-505,588-595)
 * {T}_mul/div_vec_scl                    src/kernels.cl:51-57   C[i] = A[i] op s
 *   host: src/vector.rs:430-456, 598-625        (op must be LA_MUL or LA_DIV)
 * {T}_mul/div_assign_vec_scl             src/kernels.cl:61-67   C[i] op= s
 *   host: src/vector.rs:458-470, 627-643
 * Matrix operators delegate to these (src/matrix.rs:134-218, 268-277).
 * `scalar_host` points at ONE element of `dtype` in host memory (passed by value to the kernel). */
la_status la_ew_vec_vec(int dtype, int op, void* C, const void* A, const void* B, size_t n, void* stream);
la_status la_ew_assign_vec_vec(int dtype, int op, void* C, const void* B, size_t n, void* stream);
la_status la_ew_vec_scl(int dtype, int op, void* C, const void* A, const void* scalar_host, size_t n, void* stream);
la_status la_ew_assign_vec_scl(int dtype, int op, void* C, const void* scalar_host, size_t n, void* stream);

/* {T}_eq_vec src/kernels.cl:71-74 + PartialEq src/vector.rs:645-675 (the 1-byte flag upload,
 * launch, and flag download become one call).  *equal_host = 1 iff A[i] == B[i] for all i
 * (`!=` semantics: NaN != NaN, +0 == -0); n == 0 gives 1.  Length/shape mismatch is decided by
 * the caller without a launch (src/vector.rs:649-651, src/matrix.rs:283-285).  Blocks. */
la_status la_eq_vec(int dtype, const void* A, const void* B, size_t n, int* equal_host, void* stream);

/* ------------------------------------------------------------------ reductions
 * {T}_sum_vec src/kernels.cl:83-102 + Vector::sum src/vector.rs:256-293 (kernel, 64-partial
 * download and host iter::sum() become one launch); {T}_dot_vec_vec src/kernels.cl:105-124 +
 * dot() src/vector.rs:326-363.  `out_host` / `out_dev` receive ONE element of `dtype`.
 * The host variants block; the _dev variants are stream-ordered (used before an all-reduce). */
la_status la_sum_vec(int dtype, const void* data, size_t n, void* out_host, void* stream);
la_status la_dot_vec_vec(int dtype, const void* a, const void* b, size_t n, void* out_host, void* stream);
la_status la_sum_vec_dev(int dtype, const void* data, size_t n, void* out_dev, void* stream);
la_status la_dot_vec_vec_dev(int dtype, const void* a, const void* b, size_t n, void* out_dev, void* stream);

/* normalize / normalized (src/vector.rs:313-323) as one stream-ordered call: dot(A, A), sqrt, 1 / length, scale,
 * without the reference's host round trip for the length.  C == A normalises in place.  f32 / f64 only (`Real`,
 * src/traits.rs:76-78).  Bit-identical to length() followed by `* (1 / length)`. */
la_status la_normalize_vec(int dtype, void* C, const void* A, size_t n, void* stream);

/* ------------------------------------------------------------------ multi-GPU building blocks (what la_comm is made of;
 * use the la_*_sharded calls below unless you manage the buffers yourself).  No counterpart in
 * the single-device reference; SURVEY.md section 8(e).  One process per GPU.  sum / dot over all ranks in ONE
 * launch per rank: the reduction kernel stores its partial into every peer's mailbox over NVLink peer memory and
 * adds the `world` partials in rank order (deterministic, same bits on every rank), instead of a separate
 * collective.  Set-up: every rank creates a mailbox, the 64-byte handles are exchanged by any means (the Python
 * mirror uses torch.distributed), every rank opens the peers' handles.  `peer_mailboxes[p]` is rank p's mailbox
 * as mapped in this process (the own one at [rank]); `seq` = 1, 2, 3, ... must advance identically on all ranks.
 * world <= 16.  A rank that never shows up makes the others trap after a bounded wait. */
size_t la_p2p_mailbox_bytes(void);
size_t la_p2p_handle_bytes(void);
la_status la_p2p_buffer_create(size_t bytes, void** dev, void* ipc_handle_out);    /* exportable, zeroed; a mailbox is one of la_p2p_mailbox_bytes() */
la_status la_p2p_buffer_open(const void* ipc_handle, void** peer_dev);
la_status la_p2p_buffer_close(void* peer_dev);
la_status la_p2p_buffer_destroy(void* dev);
la_status la_sum_vec_allreduce(int dtype, const void* data, size_t n, void* out_dev, void* const* peer_mailboxes, int world, int rank, uint64_t seq, void* stream);
la_status la_dot_vec_vec_allreduce(int dtype, const void* a, const void* b, size_t n, void* out_dev, void* const* peer_mailboxes, int world, int rank, uint64_t seq, void* stream);
/* out[c] = sum over ranks (ascending) of each rank's n-element vector, read from the peers' exported buffers over
 * NVLink in one launch per rank (used after la_mul_vec_mat on row-sharded operands).  Rank r's contribution must
 * already be at byte offset (seq & 1) * half_bytes of ITS buffer, written by earlier work on `stream`. */
la_status la_vec_allreduce(int dtype, void* out_dev, size_t n, void* const* peer_vectors, size_t half_bytes, void* const* peer_mailboxes,
                           int world, int rank, uint64_t seq, void* stream);

/* ------------------------------------------------------------------ multi-GPU: communicator and sharded operations
 * The N-device replacement for the reference's single global context (src/lib.rs:125-149, 310-335).  One process per GPU;
 * vectors and matrices are sharded in contiguous row blocks (la_partition).  An la_comm is one rank's view of the group.
 * The caller supplies ONE primitive, `exchange`: an all-gather of `bytes` bytes per rank between the processes
 * (recv holds world * bytes, in rank order; return 0 on success) -- implemented with whatever the host program already
 * has (MPI_Allgather, a torch process group, files, pipes).  It is used to set the group up (cudaIpc handles of the
 * buffers peers must reach) and for small host-side agreements; the data path itself runs over NVLink / NVSwitch peer
 * memory inside the library's kernels, with no collective library.  If peer access or cudaIpc is unavailable
 * (la_comm_info: peer_memory = 0, reason in la_last_error()) sum / dot / == / v x M fall back to the exchange callback,
 * la_mul_mat_mat_host_sharded to a plain per-rank upload of B, and la_transpose_sharded returns LA_ERR_UNSUPPORTED.
 * All la_comm_* / la_*_sharded calls are collectives: every rank makes the same calls in the same order with the same
 * global shapes.  Results are identical on every rank and deterministic (partials combined in rank order with the
 * element type's own arithmetic).  At most 16 ranks.  One thread per communicator at a time, and one stream for its
 * stream-ordered calls (v x M, transpose): the ranks meet in small kernels on that stream, not on the host.  A rank that never shows
 * up makes the others' kernels trap after a bounded wait rather than hang. */
typedef struct la_comm la_comm;
typedef int (*la_exchange_fn)(void* user, const void* send, void* recv, size_t bytes);
la_status la_comm_create(int world, int rank, la_exchange_fn exchange, void* user, la_comm** comm);   /* on the current device */
la_status la_comm_destroy(la_comm* comm);
la_status la_comm_info(const la_comm* comm, int* world, int* rank, int* peer_memory);
la_status la_comm_barrier(la_comm* comm);
/* rank's block [*lo, *hi) of n elements (rows): contiguous, the first n % world ranks hold one more */
void la_partition(size_t n, int world, int rank, size_t* lo, size_t* hi);
/* sum / dot over all ranks' shards (src/vector.rs:256-293, 326-363 for the one-device case): one launch per rank -- the
 * reduction kernel delivers its partial to every peer and adds the partials in rank order -- and the scalar arrives in
 * out_host through a zero-copy word, without a device-to-host copy. */
la_status la_sum_vec_sharded(la_comm* comm, int dtype, const void* data_local, size_t n_local, void* out_host, void* stream);
la_status la_dot_vec_vec_sharded(la_comm* comm, int dtype, const void* a_local, const void* b_local, size_t n_local, void* out_host, void* stream);
la_status la_eq_vec_sharded(la_comm* comm, int dtype, const void* A_local, const void* B_local, size_t n_local, int* equal_host, void* stream);
/* `&Vector * &Matrix` (src/vector.rs:508-539) with v and M co-sharded by rows: C (cols elements) is the full result on
 * every rank.  Stream-ordered. */
la_status la_mul_vec_mat_sharded(la_comm* comm, int dtype, void* C, const void* v_local, const void* M_local, size_t rows_local, size_t cols, void* stream);
/* Transpose of a row-sharded rows x cols matrix, row-sharded again (rank p gets rows la_partition(cols) of the transpose):
 * one kernel per rank stores the turned pieces straight into the owners' memory over NVLink. */
la_status la_transpose_sharded(la_comm* comm, int dtype, void* out_local, const void* in_local, size_t rows, size_t cols, void* stream);
/* la_mul_mat_mat_host with A and C sharded by rows (each rank passes its rows) and all of B on every rank's host: each rank
 * uploads only 1/world of B over its own PCIe link and the slices travel between the GPUs over NVLink. */
la_status la_mul_mat_mat_host_sharded(la_comm* comm, int dtype, void* C_host_local, const void* A_host_local, const void* B_host, size_t M, size_t K, size_t N);
/* (M x v needs no exchange -- la_mul_vec_transpose_mat on the local rows gives the local slice of the result -- and neither
 * do elementwise operations on co-sharded operands or A (row-sharded) x B (replicated): la_mul_mat_mat / _prepared.) */

/* ------------------------------------------------------------------ matrix x vector
 * {T}_mul_vec_mat src/kernels.cl:143-147, host `&Vector * &Matrix` src/vector.rs:508-539:
 *   C[col] = sum_j A[j] * B[j*cols + col],  A has `rows` elements, C has `cols`.
 * {T}_mul_vec_transpose_mat src/kernels.cl:150-152, host mul_transpose_mat src/vector.rs:542-569:
 *   C[r] = sum_j A[j] * B[r*cols + j],      A has `cols` elements, C has `rows`.
 * B is the rows x cols matrix. */
la_status la_mul_vec_mat(int dtype, void* C, const void* A, const void* B, size_t rows, size_t cols, void* stream);
la_status la_mul_vec_transpose_mat(int dtype, void* C, const void* A, const void* B, size_t rows, size_t cols, void* stream);

/* la_mul_vec_mat for callers that must not allocate.  No counterpart in the reference (its
 * host code allocates a cl_mem per result, src/vector.rs:516-519); added for host programs that
 * own all device memory themselves or cannot take the synchronisation of a cudaMalloc.  Same
 * arguments and results as la_mul_vec_mat.  The per-strip partial vectors live in a fixed
 * 8 MiB block inside the library image (held by every GPU that loads the library until the
 * process exits) instead of grown scratch; it calls no cudaMalloc/cudaFree or other allocating
 * routine.  Shapes whose partials exceed the block return LA_ERR_UNSUPPORTED with the byte
 * counts in la_last_error() -- roughly strips * cols * sizeof(accumulator) <= 8 MiB, e.g.
 * 32768 x 32768 f32 fits, 64 x 2^21 u8 does not.  One block per device, shared by all calls on
 * all streams: calls that overlap on the GPU race; the block keeps the previous call's contents
 * and none of it needs to start at zero.  The first call on a device checks that it is an
 * sm_10x GPU and resolves the block's device address (cudaGetSymbolAddress); with CUDA's lazy
 * module loading it is also the call that places the kernels and the 8 MiB block on the device,
 * so it may itself need device memory.  Make one small warm-up call per device and element type
 * at start-up, before the host's own allocator takes the remaining memory and before any stream
 * capture begins; later calls only launch kernels. */
la_status la_mul_vec_mat_noalloc(int dtype, void* C, const void* A, const void* B, size_t rows, size_t cols, void* stream);

/* ------------------------------------------------------------------ matrix x matrix
 * {T}_mul_mat_mat src/kernels.cl:161-169 (+ {T}_dot_product :134-140), host mul_helper
 * src/matrix.rs:240-265 (`&Matrix * &Matrix` :221-232, mul_column_row :236-238 with K == 1):
 *   C[M x N] = A[M x K] * B[K x N].  C must not alias A or B. */
la_status la_mul_mat_mat(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, void* stream);

/* la_mul_mat_mat that also says what ran (the reference has one kernel, so nothing to report there):
 * info->engine = 0 generic tiled SIMT kernel, 1 TMA-fed SIMT kernel, 2 f32 tensor engine, 3 integer byte-limb tensor
 * engine; info->padded = 1 if the operands went through zero-padded, aligned copies.  info may be NULL. */
typedef struct la_gemm_info { int engine; int padded; } la_gemm_info;
la_status la_mul_mat_mat_info(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, void* stream, la_gemm_info* info);

/* One B, many products (no counterpart in the reference, whose kernel reads B as it is; SURVEY.md section 8(e): the
 * row-block sharded product keeps B replicated on every GPU, "broadcast once").  Part of what a product does is work on B
 * alone: the tensor engines read B transposed and split into planes (f32: tf32 planes of B^T and, from K = 1024, the
 * fp16 planes of the column-scaled B^T; 8/16/32-bit integers: byte planes of B^T).  la_gemm_prepare_b does that once and keeps the result in a handle;
 * la_mul_mat_mat_prepared(C, A, handle, M) is then la_mul_mat_mat(dtype, C, A, B, M, K, N) without that part: same engine
 * selection, same kernels, bit-identical results.  Where no tensor engine applies (f64, 64-bit integers, small or ragged
 * shapes, LA_GEMM_SIMT, a B with NaN / inf / extreme exponents) the handle holds nothing extra and the product reads B
 * itself.  B must stay allocated and unchanged while the handle lives (the handle keeps the pointer, not a copy).
 * la_gemm_prepare_b allocates device memory for the planes (f32: 8, from K = 1024 6 bytes per element of B; integers: sizeof(T)) and
 * returns when they are complete; la_gemm_prepared_free waits for the device before releasing them.  A handle belongs
 * to the device it was made on and may be used from any stream and thread (it is read-only after creation). */
typedef struct la_gemm_prepared la_gemm_prepared;
la_status la_gemm_prepare_b(int dtype, const void* B, size_t K, size_t N, void* stream, la_gemm_prepared** handle);
la_status la_mul_mat_mat_prepared(void* C, const void* A, const la_gemm_prepared* handle, size_t M, void* stream);
la_status la_gemm_prepared_free(la_gemm_prepared* handle);

/* The reference user's whole sequence -- Matrix::from_vec(a) * Matrix::from_vec(b), then to_vec()
 * (src/matrix.rs:39-46, 221-265, src/vector.rs:53-78) -- as one call on HOST buffers: B is uploaded
 * once, A and C stream through in row blocks on separate copy streams so uploads, multiplies and
 * downloads overlap.  Same engines and the same gate as la_mul_mat_mat (integers: the same bits).  Blocks until C_host is complete; pinned host
 * buffers (la_host_alloc) give full PCIe rate, pageable ones work; any of the three may also be a
 * device pointer (e.g. a B that is already resident).  Calls on the same device run one at a time
 * (a second caller waits): they share one set of copy streams and one range-flag word. */
la_status la_mul_mat_mat_host(int dtype, void* C_host, const void* A_host, const void* B_host, size_t M, size_t K, size_t N);

/* f32 GEMM engine.
 *  LA_GEMM_SIMT    FP32 FMA pipe, each element accumulated k-ascending with one fused multiply-add
 *                  per k: the reference's own order (src/kernels.cl:136-138).
 *  LA_GEMM_3XTF32  tcgen05 tensor cores: every operand split x = hi + lo with an 11-bit hi; a product is hi*hi + hi*lo + lo*hi.
 *                  Below K = 1024 hi = tf32(x) and the three terms are tf32 MMAs.  From K = 1024 (K % 8 == 0) rows of A and
 *                  columns of B are scaled by powers of two (undone exactly at the end), hi = fp16(x), the residual is lifted
 *                  by 2^6 into fp16 as well, and the three terms are fp16 MMAs at twice the tf32 rate; elements 2^-23 and
 *                  more below their row's / column's maximum are left out of the planes and added afterwards in plain f32.
 *                  FP32 accumulation in chunks (32 k below K = 1024, 64 below K = 2048, 128 below 8192, 256 from
 *                  there; meets the same f32 tolerance as the SIMT
 *                  engine).  Falls back to SIMT for shapes it does not cover
 *                  (M, N >= 128, K >= 32, all multiples of 4), whenever an operand holds a NaN, an
 *                  infinity or a magnitude outside [2^-100, 2^100], and (from K = 1024) when a row of A or a column of B
 *                  sums to less than 4 times its largest magnitude -- identity, permutation, diagonal, one-hot
 *                  operands: outputs that are essentially one product, which only the exact engine computes as accurately
 *                  as the reference does (the call reads one flag back, so the
 *                  HOST blocks for the length of the pre-pass; the product itself is already queued behind
 *                  the pre-pass, so the GPU does not idle meanwhile).
 *  LA_GEMM_AUTO    (default) LA_GEMM_3XTF32 for products of at least 512^3 multiply-adds (where it overtakes the SIMT
 *                  kernels) whose K is at least LA_GEMM_AUTO_MIN_K, LA_GEMM_SIMT otherwise.  Below that K a product is a
 *                  handful of accumulation chains inside the tensor cores and its worst error over many outputs can land
 *                  slightly ABOVE that of the reference's own summation order, which the f32 gate (DESIGN.md section 4:
 *                  never less accurate than the reference order) does not allow; forcing LA_GEMM_3XTF32 there is an
 *                  explicit opt-in to at most twice the reference order's error.
 * 8/16/32-bit integer products of that size run as exact byte-limb products on the INT8 tensor cores
 * (bit-identical results, mod 2^n) unless LA_GEMM_SIMT is selected.
 * Process-wide; LA_GEMM_MODE=simt|3xtf32|auto in the environment sets the initial value. */
enum { LA_GEMM_SIMT = 0, LA_GEMM_3XTF32 = 1, LA_GEMM_AUTO = 2 };
enum { LA_GEMM_AUTO_MIN_K = 256 };
la_status la_set_gemm_mode(int mode);
la_status la_get_gemm_mode(int* mode);

/* ------------------------------------------------------------------ transpose
 * No counterpart in the reference (it only has mul_vec_transpose_mat); named by the
 * north star.  out[c*rows + r] = in[r*cols + c]; out must not alias in. */
la_status la_transpose(int dtype, void* out, const void* in, size_t rows, size_t cols, void* stream);

/* ------------------------------------------------------------------ user kernels
 * The reference's extension point: `create_kernel` (src/lib.rs:156-171) hands out a kernel from the
 * text of src/extra_kernels.cl, which `get_src` (src/lib.rs:337-354, `load_extra_src` :173-180)
 * appends to its own kernel source once per element type with {T} replaced by the OpenCL type
 * name and IS_<TYPE> defined; src/vector.rs:80-151 (generate, from_for_each2, for_each_mut,
 * map_mut, map) launches it with one work-item per element.  Here `source` is that text and
 * NVRTC compiles it for sm_100a.  It may stay in the dialect the reference documents -- `kernel
 * void {T}_name(global {T}* C, ...)`, `i` for get_global_id(0), plain helper functions,
 * get_global_size/get_local_id/barrier/local, `{T}_dot_product` (src/kernels.cl:134-140) -- or be
 * CUDA C++ (`extern "C" __global__`).  {T}: uchar char ushort short uint int ulong long float double.
 *
 * la_user_kernel_compile   compiles only (needs no device): a syntax check; *cubin_bytes may be NULL.
 * la_user_kernel_create    compiles, loads and looks up `kernel_name` (the full name, {T} already
 *                          substituted, e.g. "float_test_kernel").  A compile error returns
 *                          LA_ERR_ARG with the start of NVRTC's log in la_last_error() and the whole
 *                          log in la_user_kernel_build_log() (the reference panics with the build log).
 * la_user_kernel_launch    one work-item per index 0..global_size-1, never more (sizes that are not
 *                          a multiple of the 256-thread block run their remainder as a second,
 *                          smaller launch).  arg_values[j] points at the value of the kernel's j-th
 *                          parameter (a device pointer or a scalar), as for cuLaunchKernel.
 *                          Stream-ordered, returns without waiting.  A handle carries launch state on
 *                          the device: use it from one stream at a time (ocl::Kernel is the same). */
typedef struct la_user_kernel la_user_kernel;
la_status la_user_kernel_compile(int dtype, const char* source, size_t* cubin_bytes);
la_status la_user_kernel_create(int dtype, const char* source, const char* kernel_name, la_user_kernel** kernel);
la_status la_user_kernel_launch(la_user_kernel* kernel, size_t global_size, void** arg_values, void* stream);
la_status la_user_kernel_destroy(la_user_kernel* kernel);
const char* la_user_kernel_build_log(void);          /* NVRTC's log of this thread's last compile */
const char* la_user_kernel_type_name(int dtype);     /* what {T} becomes; NULL for an unknown dtype */

#ifdef __cplusplus
}
#endif
#endif /* LINALG_B200_H */
