/*
 * pipeinfer_llama.h -- C ABI of libpipeinfer_b200.so, runtime level: the subset of the reference's llama.h that
 * examples/speculative/speculative.cpp, examples/main/main.cpp and common/ call on PipeInfer's path, with the same
 * names, argument meaning, struct layouts and return codes, so that those programs can be linked against this library
 * in place of llama.o + ggml*.o.  Citations are to the reference repository.
 *
 * One process drives one GPU (one pipeline stage).  Under a multi-process launch (torchrun / mpirun-style, one rank per
 * GPU) rank 0 is the pipeline head and runs the scheduler; other ranks enter llama_decode once and serve as stages
 * until shutdown, exactly like the reference's worker loop (llama.cpp:9941-9977, 9995-10007).
 */
#ifndef PIPEINFER_LLAMA_H
#define PIPEINFER_LLAMA_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

struct llama_model;
struct llama_context;
struct ggml_cgraph;                 /* opaque run handle of the async API (never dereferenced by callers) */

typedef int32_t llama_pos;
typedef int32_t llama_token;
typedef int32_t llama_seq_id;

/* llama.h:118-128 */
typedef struct llama_token_data { llama_token id; float logit; float p; } llama_token_data;
typedef struct llama_token_data_array { llama_token_data * data; size_t size; bool sorted; } llama_token_data_array;
typedef void (*llama_progress_callback)(float progress, void * ctx);

/* llama.h:142-162 (incl. PipeInfer's trailing batch_id / max_n_seq); passed BY VALUE */
typedef struct llama_batch {
    int32_t        n_tokens;
    llama_token *  token;
    float *        embd;
    llama_pos *    pos;
    int32_t *      n_seq_id;
    llama_seq_id ** seq_id;
    int8_t *       logits;
    llama_pos      all_pos_0;
    llama_pos      all_pos_1;
    llama_seq_id   all_seq_id;
    int32_t        batch_id;
    int32_t        max_n_seq;
} llama_batch;

/* llama.h:164-178; n_gpu_layers / main_gpu / tensor_split / mmap flags are accepted and ignored (always on the GPU) */
struct llama_model_params {
    int32_t       n_gpu_layers;
    int32_t       main_gpu;
    const float * tensor_split;
    llama_progress_callback progress_callback;
    void *        progress_callback_user_data;
    bool          vocab_only;
    bool          use_mmap;
    bool          use_mlock;
};

/* llama.h:180-203; thread counts and YaRN fields are accepted and ignored */
struct llama_context_params {
    uint32_t seed;
    uint32_t n_ctx;
    uint32_t n_batch;
    uint32_t n_threads;
    uint32_t n_threads_batch;
    int8_t   rope_scaling_type;
    float    rope_freq_base;
    float    rope_freq_scale;
    float    yarn_ext_factor;
    float    yarn_attn_factor;
    float    yarn_beta_fast;
    float    yarn_beta_slow;
    uint32_t yarn_orig_ctx;
    bool     mul_mat_q;
    bool     f16_kv;
    bool     logits_all;
    bool     embedding;
};

/* llama.h:389-430 */
struct llama_kv_cache_view_cell { llama_pos pos; };
struct llama_kv_cache_view {
    int32_t n_cells;
    int32_t n_max_seq;
    int32_t token_count;
    int32_t used_cells;
    int32_t max_contiguous;
    int32_t max_contiguous_idx;
    struct llama_kv_cache_view_cell * cells;
    llama_seq_id * cells_sequences;
};

/* ---- start-up / tear-down (llama.h:263-300; llama.cpp:8642, 8671, 8707) ---- */
struct llama_model_params   llama_model_default_params(void);
struct llama_context_params llama_context_default_params(void);
void llama_backend_init(bool numa);
void llama_backend_free(void);
struct llama_model *   llama_load_model_from_file(const char * path_model, struct llama_model_params params);
void                   llama_free_model(struct llama_model * model);
struct llama_context * llama_new_context_with_model(struct llama_model * model, struct llama_context_params params);
void                   llama_free(struct llama_context * ctx);
int64_t                llama_time_us(void);
void llama_set_n_threads(struct llama_context * ctx, uint32_t n_threads, uint32_t n_threads_batch);   /* no-op */

/* ---- model / vocabulary queries (llama.h:302-330, 619-670) ---- */
const struct llama_model * llama_get_model(const struct llama_context * ctx);
int   llama_n_ctx(const struct llama_context * ctx);
int   llama_n_vocab(const struct llama_model * model);
int   llama_n_ctx_train(const struct llama_model * model);
int   llama_n_embd(const struct llama_model * model);
llama_token llama_token_bos(const struct llama_model * model);
llama_token llama_token_eos(const struct llama_model * model);
llama_token llama_token_nl(const struct llama_model * model);
int   llama_add_bos_token(const struct llama_model * model);   /* tokenizer.ggml.add_bos_token: 0 / 1, -1 when absent (llama.cpp:2521-2525) */
/* Text -> ids with the model's own vocabulary kind: SentencePiece-style merging by score (llama.cpp:5970-6110) or GPT-2
 * byte-level BPE (pre-tokenizer :6249-6386, merging by rank :6117-6219); with `special`, occurrences of special tokens
 * are cut out first and emitted as their ids (:6431-6524).  Returns the count, or -needed when n_max_tokens is too
 * small (:10079-10093).  Where the reference throws out of its C API (text that is not UTF-8 for a BPE vocabulary, a
 * byte without a token) this returns INT32_MIN and pib_last_error() says why.  Host only: works on a vocab_only model
 * without a GPU. */
int   llama_tokenize(const struct llama_model * model, const char * text, int text_len, llama_token * tokens,
                     int n_max_tokens, bool add_bos, bool special);
int   llama_token_to_piece(const struct llama_model * model, llama_token token, char * buf, int length);   /* llama.cpp:10111-10166 */
/* class the BPE pre-tokenizer sees for a codepoint: 1 digit, 2 letter, 3 whitespace, 0 anything else (the reference's
 * codepoint_type, unicode.h:351-403, folded to the three classes llama.cpp:6317-6353 asks about).  For the tests. */
int   pib_tokenizer_codepoint_class(uint32_t cp);

/* ---- batches (llama.cpp:9839-9898) ---- */
struct llama_batch llama_batch_get_one(llama_token * tokens, int32_t n_tokens, llama_pos pos_0, llama_seq_id seq_id);
struct llama_batch llama_batch_init(int32_t n_tokens, int32_t embd, int32_t n_seq_max);
#ifdef __cplusplus
void llama_batch_free(struct llama_batch & batch);                 /* llama.h:585 takes a C++ reference */
#else
void llama_batch_free(struct llama_batch * batch);
#endif

/* ---- decode (llama.cpp:9995 llama_decode; :5850 / :5857 async pair; :5457 iprobe; :9981 cancel) ----
 * llama_decode: 0 on success, -1 on any failure including "no KV slot" (this fork, llama.cpp:5562-5565, 5841-5847).
 * llama_start_async_decode: enqueues the whole stage on the GPU stream and returns a run handle (NULL on failure).
 * llama_finish_async_decode: waits for the handle; 1 on success, -1 on a NULL handle (llama.cpp:5865-5867).
 * llama_mpi_iprobe: true when the OLDEST run in flight has finished (the reference: "have logits arrived?").
 * llama_cancel_run: marks batch ids cancelled; kernels of those runs that have not started yet exit at their first
 *   instruction; KV cells stay allocated and head still advances, as in the reference (llama.cpp:5700-5714). */
int llama_decode(struct llama_context * ctx, struct llama_batch batch);
#ifdef __cplusplus
struct ggml_cgraph * llama_start_async_decode(struct llama_context & lctx, struct llama_batch batch);
int                  llama_finish_async_decode(struct llama_context & lctx, struct llama_batch batch, struct ggml_cgraph * cgraph);
#else
struct ggml_cgraph * llama_start_async_decode(struct llama_context * lctx, struct llama_batch batch);
int                  llama_finish_async_decode(struct llama_context * lctx, struct llama_batch batch, struct ggml_cgraph * cgraph);
#endif
bool llama_mpi_iprobe(struct llama_context * lctx);
void llama_cancel_run(struct llama_context * ctx, int32_t * canceled, int count);
float * llama_get_logits(struct llama_context * ctx);                      /* lazy device->host copy; on the head of a
                                                                             * pipeline each row is fetched from the last stage */
float * llama_get_logits_ith(struct llama_context * ctx, int32_t i);       /* llama.cpp:10019-10025 */

/* ---- KV cache sequence ops (llama.cpp:9238-9359 wrappers over :1657-1734) ---- */
void llama_kv_cache_clear(struct llama_context * ctx);
void llama_kv_cache_seq_rm(struct llama_context * ctx, llama_seq_id seq_id, llama_pos p0, llama_pos p1);
void llama_kv_cache_seq_cp(struct llama_context * ctx, llama_seq_id seq_id_src, llama_seq_id seq_id_dst, llama_pos p0, llama_pos p1);
void llama_kv_cache_seq_keep(struct llama_context * ctx, llama_seq_id seq_id);
/* llama.cpp:9310 / :9263: seq_cp issued from the other end of the reference's MPI ring; here every op comes from the head */
void llama_kv_cache_seq_cp_back(struct llama_context * ctx, llama_seq_id seq_id_src, llama_seq_id seq_id_dst, llama_pos p0, llama_pos p1);
void llama_kv_cache_seq_cp_sync_bi(struct llama_context * ctx, llama_seq_id seq_id_src, llama_seq_id seq_id_dst, llama_pos p0, llama_pos p1);
/* llama.cpp:9348 (context shifting of `main`, not on the speculation path): refused with an error, the cache is left
 * unchanged.  The reference's MPI build aborts at the next decode instead (ggml-mpi.c:610 asserts with K_shift nodes present). */
void llama_kv_cache_seq_shift(struct llama_context * ctx, llama_seq_id seq_id, llama_pos p0, llama_pos p1, llama_pos delta);
int  llama_get_kv_cache_token_count(const struct llama_context * ctx);
int  llama_get_kv_cache_used_cells(const struct llama_context * ctx);
struct llama_kv_cache_view llama_kv_cache_view_init(const struct llama_context * ctx, int32_t n_max_seq);
void llama_kv_cache_view_free(struct llama_kv_cache_view * view);
void llama_kv_cache_view_update(const struct llama_context * ctx, struct llama_kv_cache_view * view);

/* ---- communicator helpers (llama.cpp:8901-8938).  Both heads live in one process here, so the token syncs are
 * no-ops; split_layers_weighted partitions the layers over the pipeline stages with ggml_mpi_split_range's rule. ---- */
void llama_split_comm(struct llama_context * ctx, int color);
void llama_swap_comm(struct llama_context * ctx);
void llama_sync_token(struct llama_context * ctx, llama_token * token, int root);
void llama_sync_token_data(struct llama_context * ctx, llama_token_data * data, int root);
void llama_split_layers_weighted(struct llama_context * ctx, float device_weights[], size_t num_weights);
int  llama_node_id(struct llama_context * ctx);

/* ---- samplers used under greedy decoding (llama.cpp:7019, 7496) ---- */
void        llama_sample_softmax(struct llama_context * ctx, llama_token_data_array * candidates);
llama_token llama_sample_token_greedy(struct llama_context * ctx, llama_token_data_array * candidates);
void        llama_sample_repetition_penalties(struct llama_context * ctx, llama_token_data_array * candidates,
                                              const llama_token * last_tokens, size_t penalty_last_n, float penalty_repeat,
                                              float penalty_freq, float penalty_present);                  /* llama.cpp:7286 */

/* host samplers of llama.h:720-800 so that the reference's common/sampling.cpp links unmodified (compat_ref.cu); only the
 * greedy configuration is on the path: tail-free / typical are accepted with neutral parameters, mirostat / classifier-free
 * guidance / grammar / LoRA are refused with an error */
void        llama_sample_top_k(struct llama_context * ctx, llama_token_data_array * candidates, int k, size_t min_keep);
void        llama_sample_top_p(struct llama_context * ctx, llama_token_data_array * candidates, float p, size_t min_keep);
void        llama_sample_min_p(struct llama_context * ctx, llama_token_data_array * candidates, float p, size_t min_keep);
void        llama_sample_tail_free(struct llama_context * ctx, llama_token_data_array * candidates, float z, size_t min_keep);
void        llama_sample_typical(struct llama_context * ctx, llama_token_data_array * candidates, float p, size_t min_keep);
void        llama_sample_temp(struct llama_context * ctx, llama_token_data_array * candidates, float temp);
llama_token llama_sample_token(struct llama_context * ctx, llama_token_data_array * candidates);
bool        llama_mmap_supported(void);
bool        llama_mlock_supported(void);
int         llama_max_devices(void);

/* ---- reporting (llama.h:323, 860-872; llama.cpp:10182-10240) ---- */
int          llama_vocab_type(const struct llama_model * model);           /* 0 = SPM, 1 = BPE (llama.h:70-73) */
void         llama_print_timings(struct llama_context * ctx);
void         llama_reset_timings(struct llama_context * ctx);
const char * llama_print_system_info(void);
void         llama_log_set(void (*log_callback)(int level, const char * text, void * user_data), void * user_data);

/* ---- additive extensions (no counterpart in llama.h; needed for device-side acceptance) ---- */
/* sorted top-8 {id, logit, p} of logits row i of the last finished run: 96 bytes instead of n_vocab floats */
int  llama_pib_get_top8(struct llama_context * ctx, int32_t i, llama_token_data * out8);
/* Draft-vs-target acceptance on the device (the walk of examples/speculative/speculative.cpp:439-560 under greedy sampling).
 * set_accept_table describes the verification batch that the NEXT llama_start_async_decode / llama_decode on `ctx` starts: for
 * each of n_branches (<= 8) draft branches its length len[s] (-1: inactive), and for depth i the batch row rows[s*stride+i] whose
 * logits are sampled at that depth (i <= len[s]) and the drafted token toks[s*stride+i] it is compared with (i < len[s]); `keep`
 * = the branch the walk starts on, `speculative` = 0 stops after the first sample.  The last pipeline stage then runs the walk
 * right after sampling, and after llama_finish_async_decode llama_pib_get_accepted returns the ids sampled along it (count
 * returned; -1 if the run had no table) and the branch kept: 8 + 4 n bytes cross to the host instead of 96 bytes per row. */
void    llama_pib_set_accept_table(struct llama_context * ctx, int32_t n_branches, int32_t keep, int32_t speculative, const int32_t * len,
                                   const int32_t * rows, const int32_t * toks, int32_t depth_stride);
int32_t llama_pib_get_accepted(struct llama_context * ctx, int32_t * ids, int32_t cap, int32_t * s_keep);
/* logit bias applied on the device before argmax (the reference adds it on the host, common/sampling.cpp:134-136) */
void llama_pib_set_logit_bias(struct llama_context * ctx, const llama_token * ids, const float * bias, int n);
/* random-init model of the llama architecture generated directly in HBM (no file); flavour 0 iid, 1 bigram */
struct llama_model * llama_pib_synthetic_model(int n_layer, int n_embd, int n_head, int n_head_kv, int n_ff, int n_vocab,
                                               int wtype, uint64_t seed, int flavour, float agree);
/* pipeline stages (multi-GPU, one process per GPU; see pipeinfer_mpi.h): a stage's share of a model and a context bound
 * to stage `rank` of `size`; llama_pib_stage_serve is the worker loop of ranks > 0 (reference llama.cpp:9941-9977):
 * returns 0 when the head paused the pipeline, 1 on shutdown, -1 on error */
struct llama_model * llama_pib_synthetic_model_slice(int n_layer, int n_embd, int n_head, int n_head_kv, int n_ff, int n_vocab,
                                                     int wtype, uint64_t seed, int flavour, float agree, int layer_begin,
                                                     int layer_end, int want_embd, int want_output);
struct llama_model * llama_pib_load_model_slice(const char * path_model, int layer_begin, int layer_end, int want_embd, int want_output);
struct llama_context * llama_pib_new_context_stage(struct llama_model * model, uint32_t n_ctx, int32_t max_seq_ids, int rank, int size);
int  llama_pib_n_layer(const struct llama_model * model);
int  llama_pib_stage_serve(struct llama_context * ctx);
int  llama_pib_pipe_pause(struct llama_context * ctx);
int  llama_pib_pipe_shutdown(struct llama_context * ctx);
/* 1 = this context's kernels run on a higher-priority stream than other contexts' (draft model next to a target stage) */
int  llama_pib_set_priority(struct llama_context * ctx, int high);
/* context with an explicit sequence-id capacity (the reference sizes std::set on demand) */
struct llama_context * llama_pib_new_context(struct llama_model * model, uint32_t n_ctx, int32_t max_seq_ids);
/* KV bookkeeping probes for parity tests: head / used / n, cell positions and sequence bitset rows, host mirror (0)
 * or device replica (1) */
int  llama_pib_kv_state(struct llama_context * ctx, int from_device, int32_t * scalars3, int32_t * pos, uint32_t * bits);
int  llama_pib_kv_words(const struct llama_context * ctx);
/* Ranged KV operations (no reference counterpart; the reference's scheduler loops over up to 3 000 sequence ids per operation,
 * speculative.cpp:603-612, 1252-1268).  Same bookkeeping as the literal loops:
 *   rm_range : for s in [first, last]: llama_kv_cache_seq_rm(ctx, s, p0, p1)
 *   cp_range : for d in [dst_first, dst_last]: llama_kv_cache_seq_cp(ctx, src, d, p0, p1)
 *   sweep    : for s in [first, last] (first >= 1): seq_rm(ctx, s, p0, p1); seq_cp(ctx, 0, s, p0, p1) */
void llama_pib_kv_seq_rm_range(struct llama_context * ctx, llama_seq_id first, llama_seq_id last, llama_pos p0, llama_pos p1);
void llama_pib_kv_seq_cp_range(struct llama_context * ctx, llama_seq_id src, llama_seq_id dst_first, llama_seq_id dst_last, llama_pos p0, llama_pos p1);
void llama_pib_kv_seq_sweep(struct llama_context * ctx, llama_seq_id first, llama_seq_id last, llama_pos p0, llama_pos p1);
uint64_t llama_pib_kernel_launches(const struct llama_context * ctx);
size_t   llama_pib_weight_bytes(const struct llama_model * model);
/* measurement hooks used by bench.py: host<->device bytes issued so far; CUDA-event timing of every matmul launch of
 * one layer (kind 0 qkv, 1 o, 2 gate|up, 3 down, 4 lm_head) on the launching stream; shape of a resident matrix */
void llama_pib_io_bytes(const struct llama_context * ctx, uint64_t * h2d, uint64_t * d2h);
void llama_pib_profile_layer(struct llama_context * ctx, int layer);
int  llama_pib_profile_read(struct llama_context * ctx, int32_t * kind, int32_t * b, float * ms, int cap);
int  llama_pib_matrix_shape(const struct llama_model * model, int layer, int kind, int64_t * N, int64_t * K, uint64_t * bytes);
/* debugging aids for parity work: a work buffer of the last chunk; the handle of a resident weight matrix */
int  llama_pib_debug_read(struct llama_context * ctx, int which, float * out, int n_floats);
const struct pib_weight * llama_pib_debug_weight(const struct llama_model * model, int layer, int which);

#ifdef __cplusplus
}
#endif
#endif
