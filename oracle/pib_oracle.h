/*
 * pib_oracle.h -- CPU restatement of the reference arithmetic on PipeInfer's speculation path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may
 * load this library; the product (libpipeinfer_b200.so) never links, loads or calls it.
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference).  The restatement is
 * pinned two ways (tests/test_oracle_*.py): against the *compiled reference itself* (oracle/_ref/libggml_ref.so,
 * oracle/_ref/ref_decode_dump, built by oracle/Makefile from the sources where they lie) on seeded inputs, and
 * against the fixtures those produced, committed under tests/golden/.
 */
#ifndef PIB_ORACLE_H
#define PIB_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint16_t po_fp16;

/* ggml type ids (ggml.h:305-333) for the three weight formats on the path */
enum { PO_TYPE_F32 = 0, PO_TYPE_Q4_0 = 2, PO_TYPE_Q8_0 = 8, PO_TYPE_Q3_K = 11, PO_TYPE_Q4_K = 12, PO_TYPE_Q5_K = 13, PO_TYPE_Q6_K = 14 };

/* --- scalar helpers ------------------------------------------------------------------------------------------ */
float   po_fp16_to_fp32(po_fp16 h);
po_fp16 po_fp32_to_fp16(float f);
void    po_tables_init(void);                     /* ggml.c:2185-2203 */
const po_fp16 * po_table_exp_f16(void);
const po_fp16 * po_table_silu_f16(void);
const po_fp16 * po_table_gelu_f16(void);

/* --- K2: activation quantisers ------------------------------------------------------------------------------- */
void po_quantize_row_q8_0(const float * x, void * y, int k);     /* ggml-quants.c:632, AVX2 branch :701-760 */
void po_quantize_row_q8_K(const float * x, void * y, int k);     /* ggml-quants.c:2329-2363 */

/* --- K3: weight dequantisers --------------------------------------------------------------------------------- */
void po_dequantize_row_q4_0(const void * x, float * y, int k);   /* ggml-quants.c:1074 */
void po_dequantize_row_q8_0(const void * x, float * y, int k);   /* ggml-quants.c:1168 */
void po_dequantize_row_q4_K(const void * x, float * y, int k);   /* ggml-quants.c:1917 */

/* --- K1: dot products and the matmul built from them --------------------------------------------------------- */
void po_vec_dot_q4_0_q8_0(int n, float * s, const void * vx, const void * vy);  /* ggml-quants.c:2426, AVX2 :2500 */
void po_vec_dot_q8_0_q8_0(int n, float * s, const void * vx, const void * vy);  /* ggml-quants.c:3521, AVX2 :3587 */
void po_vec_dot_q4_K_q8_K(int n, float * s, const void * vx, const void * vy);  /* ggml-quants.c:5216, AVX2 :5318 */
void po_vec_dot_f16(int n, float * s, const po_fp16 * x, const po_fp16 * y);    /* ggml.c:1200 */
size_t po_row_size(int type, int64_t k);
/* y[b][N] = W[N][K] . x[b][K]  (ggml.c:9388: quantise x rows to the vec_dot type, then one vec_dot per output) */
void po_mul_mat(int type, const void * w, int64_t N, int64_t K, const float * x, int b, float * y);

/* --- K4..K9 -------------------------------------------------------------------------------------------------- */
void po_rms_norm(const float * x, float * y, int n, float eps);                 /* ggml.c:9028 */
void po_norm(const float * x, float * y, int n, float eps);                     /* ggml.c:8959 */
void po_dequantize_row_q5_K(const void * x, float * y, int k);                  /* ggml-quants.c:1917 */
void po_vec_dot_q5_K_q8_K(int n, float * s, const void * x, const void * y);    /* ggml-quants.c:5680 (AVX2 path) */
void po_dequantize_row_q3_K(const void * vx, float * y, int k);
void po_vec_dot_q3_K_q8_K(int n, float * s, const void * vx, const void * vy);
void po_dequantize_row_q6_K(const void * x, float * y, int k);                  /* ggml-quants.c:1994 */
void po_vec_dot_q6_K_q8_K(int n, float * s, const void * x, const void * y);    /* ggml-quants.c:6594 (AVX2 path :6739) */
void po_rope(float * x, int n_head, int head_dim, int n_dims, int pos, int mode,
             float freq_base, float freq_scale);                                /* ggml.c:11006 (ext_factor = 0) */
void po_soft_max(float * row, int n);                                           /* ggml.c:10551 */
void po_silu(float * x, int n);                                                 /* ggml.c:1490 (fp16 LUT) */
void po_gelu(float * x, int n);                                                 /* ggml.c:1433 (fp16 LUT) */

/* --- R6: KV-cache cell metadata ------------------------------------------------------------------------------ */
typedef struct {
    uint32_t   size;      /* n_ctx cells */
    uint32_t   head;
    uint32_t   used;
    uint32_t   n;
    int32_t    n_words;   /* 32-bit words of sequence-id bits per cell */
    int32_t *  pos;       /* [size]           -1 = empty */
    uint32_t * bits;      /* [size][n_words]  bit s set <=> cell belongs to sequence s (std::set in the reference) */
} po_kv_meta;

po_kv_meta * po_kv_new(uint32_t n_ctx, int32_t max_seq_ids);                    /* llama.cpp:1523 */
void po_kv_free(po_kv_meta * kv);
void po_kv_clear(po_kv_meta * kv);                                              /* llama.cpp:1657 */
int  po_kv_find_slot(po_kv_meta * kv, int n_tokens, const int32_t * pos, const int32_t * n_seq_id,
                     const int32_t * const * seq_id);                           /* llama.cpp:1593 */
int32_t po_kv_cell_max(const po_kv_meta * kv);                                  /* llama.cpp:1647 */
void po_kv_seq_rm(po_kv_meta * kv, int32_t seq, int32_t p0, int32_t p1);        /* llama.cpp:1666 */
void po_kv_seq_cp(po_kv_meta * kv, int32_t src, int32_t dst, int32_t p0, int32_t p1); /* llama.cpp:1699 */
void po_kv_seq_keep(po_kv_meta * kv, int32_t seq);                              /* llama.cpp:1717 */
/* mask[j][i] = 0 (visible) or 1 (masked) for token j, cell i < n_kv  (llama.cpp:5200-5240) */
void po_kv_mask(const po_kv_meta * kv, int n_tokens, int n_kv, const int32_t * pos, const int32_t * n_seq_id,
                const int32_t * const * seq_id, uint8_t * mask);

/* --- R2/R8: llama-architecture decode (build_llama llama.cpp:3872, decode :5461-5726) ------------------------ */
typedef struct {
    int32_t n_layer, n_embd, n_head, n_head_kv, n_ff, n_vocab, n_rot;
    float   rms_eps, rope_freq_base, rope_freq_scale;
    int32_t wtype;                 /* type of every 2-D tensor (files are made with `quantize --pure`) */
    const void *  tok_embd;        /* [n_vocab][n_embd] */
    const float * output_norm;     /* [n_embd] */
    const void *  output;          /* [n_vocab][n_embd] */
    const float ** attn_norm;      /* per layer */
    const void **  wq, ** wk, ** wv, ** wo;
    const float ** ffn_norm;
    const void **  ffn_gate, ** ffn_down, ** ffn_up;
    /* falcon architecture (build_falcon llama.cpp:4106): arch = 1.  wq[il] then holds the fused attn_qkv matrix
     * [n_embd + 2 n_embd_gqa][n_embd]; attn_norm2 entries are NULL in the 7B layout */
    int32_t        arch;
    float          norm_eps;
    const float ** attn_norm_b, ** attn_norm2, ** attn_norm2_b;
    const float *  output_norm_b;
    /* mixed files (quantize without --pure: Q4_K_M, Q4_K_S, Q3_K_M ...), llama architecture: per-layer types of the matrices and the
     * type of output.weight; NULL / 0 = wtype */
    const int32_t * t_wq, * t_wk, * t_wv, * t_wo, * t_gate, * t_down, * t_up;
    int32_t         t_output;
} po_model;

typedef struct {
    const po_model * model;
    po_kv_meta *     kv;
    uint32_t         n_ctx;
    po_fp16 *        k;            /* [n_layer][n_ctx][n_embd_gqa]              (llama.cpp:3565) */
    po_fp16 *        v;            /* [n_layer][n_embd_gqa][n_ctx]  transposed  (llama.cpp:3569) */
} po_context;

po_context * po_context_new(const po_model * m, uint32_t n_ctx, int32_t max_seq_ids);
void po_context_free(po_context * c);
/* returns 0 on success, -1 when no KV slot is found (this fork's llama.cpp:5562-5565); logits: [n_tokens][n_vocab] */
int po_decode(po_context * c, int n_tokens, const int32_t * token, const int32_t * pos, const int32_t * n_seq_id,
              const int32_t * const * seq_id, float * logits);

/* --- R9: greedy sampling ------------------------------------------------------------------------------------- */
int32_t po_sample_greedy(const float * logits, int n_vocab);                    /* llama.cpp:7496 */
/* sorted (descending) top-k ids and probabilities p = exp(l - max)/sum over the whole vocab (llama.cpp:7019) */
void po_sample_softmax_topk(const float * logits, int n_vocab, int k, int32_t * ids, float * logit, float * p);

/* --- (e): layer partition ------------------------------------------------------------------------------------ */
/* ranges[i] = {start, end}; restates ggml_mpi_split_range(ctx, 0, n_layer - 1, weights)  ggml-mpi.c:523-559 */
void po_split_range(int n_nodes, uint16_t start, uint16_t end, const float * weights, uint16_t * ranges);

#ifdef __cplusplus
}
#endif
#endif
