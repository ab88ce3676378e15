/*
 * pipeinfer_b200.h -- C ABI of libpipeinfer_b200.so, operator level.
 *
 * B200 (sm_100a) implementation of the arithmetic on PipeInfer's asynchronous pipelined speculation path.  Each
 * entry point names the reference interface it replaces (paths relative to the reference repository,
 * AutonomicPerfectionist/PipeInfer).  Plain C types only; "dev" pointers are CUDA device pointers on the device
 * passed to pib_init, "host" pointers are ordinary host memory.  All functions return PIB_OK (0) or a negative
 * PIB_ERR_* code; pib_last_error() gives the message.  There is no CPU fallback: without a CUDA device every entry
 * point fails with PIB_ERR_CUDA.
 *
 * The model/runtime level (llama_decode, llama_batch, llama_kv_cache_seq_*) is in pipeinfer_llama.h, the transport level
 * (ggml_mpi_*, pipeline bring-up) in pipeinfer_mpi.h, the scheduler in pipeinfer_sched.h.
 */
#ifndef PIPEINFER_B200_H
#define PIPEINFER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PIB_OK            0
#define PIB_ERR_CUDA     (-1)
#define PIB_ERR_ARG      (-2)
#define PIB_ERR_NOMEM    (-3)
#define PIB_ERR_UNSUPP   (-4)

/* ggml type ids of the weight formats on the path (ggml.h:305-333) */
#define PIB_TYPE_F32   0
#define PIB_TYPE_Q4_0  2
#define PIB_TYPE_Q8_0  8
#define PIB_TYPE_Q4_K 12
#define PIB_TYPE_Q3_K 11   /* matmul + embedding rows (Q3_K_M files: token_embd, attn_q / attn_k, ffn_gate / ffn_up); activations as for Q4_K (Q8_K) */
#define PIB_TYPE_Q5_K 13   /* matmul + embedding rows (Q4_K_S / Q5_K_M files); activations as for Q4_K (Q8_K) */
#define PIB_TYPE_Q6_K 14   /* matmul + embedding rows (output.weight / attn_v / ffn_down of Q4_K_M files, Q6_K files); activations as for Q4_K (Q8_K) */

#define PIB_MAX_BATCH 16          /* verification batch handled by the dequant-GEMV path (north_star: batch 1-16) */

typedef struct pib_weight pib_weight;   /* a weight matrix repacked into HBM (layout private to the library) */
typedef void * pib_stream;              /* cudaStream_t, or NULL for the legacy default stream */

/* ---- library ------------------------------------------------------------------------------------------------- */
/* Select the device and upload the fp16 exp/silu/gelu tables (the reference builds them in ggml_init, ggml.c:2185). */
int          pib_init(int device);
const char * pib_last_error(void);
int          pib_device_sm_count(void);
void *       pib_dev_alloc(size_t bytes);                         /* cudaMalloc; NULL on failure */
void         pib_dev_free(void * dev);
int          pib_memcpy_h2d(void * dev, const void * host, size_t bytes, pib_stream s);
int          pib_memcpy_d2h(void * host, const void * dev, size_t bytes, pib_stream s);
int          pib_stream_sync(pib_stream s);

/* ---- weights ------------------------------------------------------------------------------------------------- */
/* Upload an [N][K] matrix given in the reference's block layout (block_q4_0 / block_q8_0 / block_q4_K / block_q5_K /
 * block_q6_K rows, ggml-quants.h:10-15, 42-47, 116-122, 138-157) and repack it for the GEMV kernels.  Layout-only: no value changes.
 * Replaces: the mmap'ed tensor data that ggml_compute_forward_mul_mat reads as src0 (ggml.c:9388). */
pib_weight * pib_weight_upload(int type, const void * host_blocks, int64_t N, int64_t K);
/* Same, for blocks that are already in device memory (used by the synthetic-model generator). */
pib_weight * pib_weight_from_device(int type, const void * dev_blocks, int64_t N, int64_t K);
void         pib_weight_free(pib_weight * w);
size_t       pib_weight_bytes(const pib_weight * w);             /* algorithmic bytes: N * K * bits-per-weight / 8 */
/* Copy row `row` back out in the reference block layout (inverse of the repack; for round-trip tests). */
int          pib_weight_row_blocks(const pib_weight * w, int64_t row, void * host_blocks_out);

/* ---- K2: activation quantisation ------------------------------------------------------------------------------ */
/* Bytes of the device buffer that holds `b` quantised activation rows of length K for weights of `wtype`. */
size_t pib_act_bytes(int wtype, int64_t K, int b);          /* wtype: Q4_0 / Q8_0 / Q4_K; Q5_K and Q6_K weights use Q4_K's */
/* x_dev float [b][K] -> act_dev.  Q8_0 blocks for Q4_0/Q8_0 weights, Q8_K super-blocks for Q4_K weights, with the
 * rounding of the reference's x86 build.  Replaces: quantize_row_q8_0 (ggml-quants.c:632, AVX2 branch :701-760) and
 * quantize_row_q8_K (ggml-quants.c:2329), called from the INIT phase of mul_mat (ggml.c:9493-9509). */
int pib_quantize_act(int wtype, const float * x_dev, int b, int64_t K, void * act_dev, pib_stream s);
/* Fused row norm -> * weight -> quantise.  norm_kind 1 = RMSNorm (ggml.c:9028 + ggml_mul, llama.cpp:3590-3598).
 * xnorm_dev (optional, may be NULL) receives the normalised float rows. */
int pib_norm_quantize_act(int wtype, int norm_kind, const float * x_dev, const float * norm_w_dev, float eps, int b,
                          int64_t K, void * act_dev, float * xnorm_dev, pib_stream s);
/* Copy a quantised activation buffer back in the reference's block layout (block_q8_0 / block_q8_K rows). */
int pib_act_to_blocks(int wtype, const void * act_dev, int b, int64_t K, void * host_blocks_out);

/* ---- K1: quantised matmul at verification batch 1..16 ---------------------------------------------------------- */
/* y_dev float [b][N] = W[N][K] . act (b rows).  Integer block dot products, fp16 block scales, fp32 accumulation.
 * Replaces: ggml_compute_forward_mul_mat (ggml.c:9388) with ggml_vec_dot_q4_0_q8_0 / q8_0_q8_0 / q4_K_q8_K
 * (ggml-quants.c:2426, 3521, 5216).  `variant`: 0 = default (fastest available), 1 = warp-per-row dp4a kernel,
 * 2 = tensor-core kernel with register-prefetched weight tiles (Q4_K; the default there), 3 = persistent TMA-staged
 * tensor-core kernel (Q4_K). */
int pib_gemv(const pib_weight * w, const void * act_dev, int b, float * y_dev, int variant, pib_stream s);
/* Same product as pib_gemv(variant 0) for Q4_K weights, for callers that must not allocate device memory while they
 * run (a latency-critical serving loop that cannot take the allocator's lock or its device-wide synchronisation; a
 * process whose HBM is already fully committed; a host that forbids allocation after start-up).  No counterpart in the
 * reference.  When few output rows leave the GPU under-filled the kernel splits K over several CTAs and sums their
 * partial results through global scratch; pib_gemv takes that scratch from cudaMalloc on first use, this entry point
 * keeps it in a fixed PIB_GEMV_NOALLOC_SCRATCH_BYTES array inside the library image and calls no allocating routine.
 * Limits: split * b * N * 4 bytes of partial sums must fit in the array less 16 KiB and N <= 65536 whenever a split is
 * used (true for every matrix the current split rule splits); otherwise PIB_ERR_ARG with the sizes in pib_last_error.
 * The first call on each device looks the array's address up (cudaGetSymbolAddress): issue it before any stream
 * capture.  All calls on a device share the array: calls that overlap in time (other streams, other threads) corrupt
 * each other's sums -- serialise them.  Results are bit-identical to pib_gemv's. */
#define PIB_GEMV_NOALLOC_SCRATCH_BYTES (2u * 1024u * 1024u)
#define PIB_MAX_DEV_NOALLOC 64
int pib_gemv_q4k_noalloc(const pib_weight * w, const void * act_dev, int b, float * y_dev, pib_stream s);
/* Number of K splits pib_gemv_q4k_noalloc uses for an N x K matrix on the current device (1 = no scratch touched); lets a
 * caller -- and the tests -- see which shapes go through the built-in array.  No counterpart in the reference. */
int pib_gemv_direct_splits(int64_t N, int64_t K);
/* y = W . quantise(prologue(x)) in ONE launch, for narrow models whose passes are bound by the number of launches (Q4_K, b <= 8,
 * K <= 8192; -3 otherwise): the kernel quantises the activations itself -- what the reference does at the top of
 * ggml_compute_forward_mul_mat (ggml.c:9493-9509) -- after one of the element-wise ops that precede a matmul in build_llama:
 *   prologue 0: x                                             (quantize_row_q8_K of the attention output)
 *   prologue 1: rmsnorm(x [+ addend]) * x2, eps               (ggml_add, ggml_rms_norm ggml.c:9028, ggml_mul; sum_out, if given, receives x + addend)
 *   prologue 2: silu(x) * x2                                  (ggml_silu through the fp16 table ggml.c:1490, ggml_mul; x and x2 rows xstride apart)
 * Device pointers; x [b][xstride or K] fp32, y [b][N] fp32.  Same values as pib_quantize_act / pib_norm_quantize_act followed by pib_gemv
 * up to the order of the fp32 additions. */
int pib_gemv_fused(const pib_weight * w, int prologue, const float * x_dev, const float * x2_dev, const float * addend_dev, float * sum_out_dev,
                   int64_t xstride, float eps, int b, float * y_dev, pib_stream s);
/* Host-buffer convenience used by the parity tests and the end-to-end timing: x_host float [b][K] -> y_host [b][N];
 * includes the H2D copy of x, quantisation, GEMV and the D2H copy of y. */
int pib_mul_mat_host(const pib_weight * w, const float * x_host, int b, float * y_host, int variant);

/* The same product for a WIDE batch (prompt ingestion; 1 <= b <= 256 tokens at once) on the tcgen05 tensor cores: the Q4_K weights
 * are dequantised tile by tile to fp16, the Q8_K-quantised activations likewise, fp32 accumulation in tensor memory.  Replaces
 * ggml_compute_forward_mul_mat at large ne11 (ggml.c:9388) / ggml_cuda_op_mul_mat_cublas (ggml-cuda.cu:6562).  Tolerance-checked
 * against the oracle (fp16 operand rounding, another summation order), not bit-exact.  Host buffers in and out. */
int pib_gemm_host(const pib_weight * w, const float * x_host, int b, float * y_host);
/* Device-buffer form: act_dev = activation planes of b rows (pib_quantize_act_rows), scratch_dev = pib_gemm_scratch_bytes(K, b) bytes
 * (the fp16 copy of the activations the TMA loads from). */
size_t pib_gemm_scratch_bytes(int64_t K, int b);
int    pib_gemm(const pib_weight * w, const void * act_dev, int b, float * y_dev, void * scratch_dev, pib_stream s);
int    pib_quantize_act_rows(int wtype, const float * x_dev, int b, int64_t K, void * act_dev, pib_stream s);

/* ---- K3: embedding rows ---------------------------------------------------------------------------------------- */
/* out_dev float [n][K] = dequantised rows `ids` of W.  Replaces: ggml_compute_forward_get_rows_q (ggml.c:10161) with
 * dequantize_row_q4_0 / q8_0 / q4_K (ggml-quants.c:1074, 1168, 1917). */
int pib_get_rows(const pib_weight * w, const int32_t * ids_dev, int n, float * out_dev, pib_stream s);

#ifdef __cplusplus
}
#endif
#endif
