/* include/linalg_b200_tuning.h -- tuning and introspection hooks of liblinalg_b200.so.
 *
 * NOT part of the drop-in boundary (include/linalg_b200.h): nothing here has a counterpart in the reference, no wrapper
 * needs it, and none of it changes results beyond what each comment states.  The hooks exist for tools/, bench.py and
 * the tests that pin every kernel variant against the oracle.  All setters write process-wide state without locking:
 * call them while no product / transpose is in flight on any thread.  Together with linalg_b200.h this header lists
 * every symbol the library exports (tests/test_abi.py checks both directions).
 */
#ifndef LINALG_B200_TUNING_H
#define LINALG_B200_TUNING_H

#include "linalg_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- SIMT GEMM (csrc/gemm.cu) */
/* Tile shape of the TMA-fed SIMT engine: 0 = default shape, used from the size where it beats the generic kernel;
 * k > 0 = shape k at any size (13 = the default shapes at any size); -1 forces the generic engine. */
la_status la_gemm_set_variant(int variant);
/* What the calling thread's last la_mul_mat_mat did (thread-local; la_mul_mat_mat_info is the explicit form):
 * engine 0 generic, 1 TMA SIMT, 2 f32 tensor engine, 3 integer byte-limb tensor engine; padded = went through
 * zero-padded copies. */
int la_gemm_last_engine(void);
int la_gemm_last_padded(void);

/* ---- f32 tensor engine (csrc/gemm_tc.cu).  Every setting below passes the same f32 gate unless stated. */
la_status la_gemm_tc_set_variant(int variant);     /* 0 CTA pairs (default); 5 / 3 one CTA per tile with / without B multicast; 1 / 2 single accumulation chain (drifts: reference only) */
la_status la_gemm_tc_set_persistent(int on);       /* persistent grid of sm_count / 2 CTA pairs; bit-identical results */
la_status la_gemm_tc_set_chunk(int k_blocks);      /* k-blocks of 32 per accumulation chain inside the tensor cores; 0 = by K */
int la_gemm_tc_last_form(void);                    /* this thread's last f32 tensor-engine product: 0 tf32 form, 1 f16-split form */
int la_gemm_tc_last_launches(void);                /* ... and the kernels it launched (pre-pass, product, fix-ups of left-out elements) */
la_status la_gemm_tc_set_f16(int mode);            /* f16-split form of the engine (row / column scaled operands, hi = fp16, lo = bf16, three 16-bit MMAs per product): -1 by K (default: from K = 1024), 0 never, 1 wherever the shapes allow (K >= 256, K % 8 == 0) */
la_status la_gemm_tc_set_a_direct(int on);         /* EXPERIMENT: A is read in place as its own hi plane (hardware truncation to tf32), lo taken against the truncated value */
la_status la_gemm_tc_set_cross(int mode);          /* tf32 form: cross terms as tf32 MMAs (-1 / 0, the default at every K) or as bf16 MMAs at twice the rate (1; ~2^-20 per product: gated at twice the reference order's error below K = 1024, and above the gate for one-product outputs such as B = identity) */
la_status la_gemm_tc_set_tile_n(int bn);           /* pair tile width 256 / 224 / 192; 0 = by wave count; bit-identical results with the K-split off */
la_status la_gemm_set_stream_tail(int mode);       /* K-split of the ragged last wave: -1 by gain model, 0 off, 1 whenever possible */
la_status la_gemm_set_raster(int tile_rows);       /* band height of the tile walk: -1 by size, 0 m-major; bit-identical results */
/* bench.py: bracket every launch of the pair kernel with events on its own stream (up to 64 kept);
 * la_gemm_kernel_time_ms waits for them, returns their sum and count, and resets. */
la_status la_gemm_kernel_timing(int on);
la_status la_gemm_kernel_time_ms(float* total_ms, int* launches);

/* ---- transpose (csrc/transpose.cu); results are bit-exact in every setting */
la_status la_transpose_set_variant(int variant);   /* 0 default policy; 1..10 LDG/STG tile kernels; 20..23 TMA loads + TMA stores; 30..34 TMA loads + st.global stores (4- / 8-byte elements) */
la_status la_transpose_set_band(int tile_rows);    /* band height of the tile walk (all kernels); -1 / 0 = along the input rows (default) */
la_status la_transpose_set_hints(int bits);        /* TMA load + store kernel: bit 0 / 1 = evict-first L2 hint on loads / stores */
la_status la_transpose_set_mode(int mode);         /* MEASUREMENT ONLY (output is then garbage): variants 20..23: 1 = loads only, 2 = stores only; default kernel: 3 = same global traffic without the turn in shared memory; 0 = transpose */

#ifdef __cplusplus
}
#endif
#endif /* LINALG_B200_TUNING_H */
