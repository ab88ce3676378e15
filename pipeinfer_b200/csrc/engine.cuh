// engine.cuh -- model, context and run records of the host runtime (C++), shared by engine.cu / llama_api.cu /
// pipeline.cu / speculative.cu.
#pragma once

#include "pib_common.cuh"
#include "desc.cuh"
#include "kv_meta.cuh"

#include <deque>
#include <map>
#include <string>
#include <vector>

#define PIB_TOPK 8
#define PIB_PROMPT_MAX 256           // tokens per pass of the tensor-core prompt path (gemm_q4k_tc.cu)
struct pib_row_result {            // per logits row: sorted top-8 {id, logit, p}; entry 0 is the greedy token
    int32_t id[PIB_TOPK];
    float   logit[PIB_TOPK];
    float   p[PIB_TOPK];
};

struct pib_accept_result {         // what the device-side acceptance walk returns for one verification batch
    int32_t n_ids;                 // target tokens sampled along the walk (1 .. depth + 1)
    int32_t s_keep;                // branch the walk ended on
    int32_t ids[PIB_CHUNK + 2];
};
struct pib_accept_table {          // host form of the table in the descriptor (desc.cuh)
    int32_t n_branches = 0, keep = 0, speculative = 0;
    int32_t len[PIB_TREE_NP];
    int32_t rows[PIB_TREE_NP][PIB_TREE_DEPTH], toks[PIB_TREE_NP][PIB_TREE_DEPTH];
};

struct pib_layer {
    float *      attn_norm = nullptr;   // [n_embd]
    float *      attn_norm_b = nullptr; // falcon: LayerNorm bias
    float *      attn_norm2 = nullptr, * attn_norm2_b = nullptr;   // falcon 40B/180B layout: the norm that feeds QKV
    pib_weight * wqkv      = nullptr;   // rows: Wq (n_embd) | Wk (n_embd_gqa) | Wv (n_embd_gqa)
    pib_weight * wv        = nullptr;   // only when Wv has another weight type than Wq / Wk (Q4_K_M files): then wqkv = Wq | Wk
    pib_weight * wo        = nullptr;
    float *      ffn_norm  = nullptr;
    pib_weight * wgu       = nullptr;   // llama rows: gate (n_ff) | up (n_ff); falcon: up (n_ff) only
    pib_weight * wdown     = nullptr;
};

struct pib_model {
    // hyper-parameters (reference llm_load_hparams, llama.cpp:2203-2290)
    int   n_layer = 0, n_embd = 0, n_head = 0, n_head_kv = 0, n_ff = 0, n_vocab = 0, n_rot = 0, n_ctx_train = 0;
    float rms_eps = 1e-5f, rope_freq_base = 10000.f, rope_freq_scale = 1.f;
    int   wtype = 0;
    int   arch = 0;                      // 0 llama (build_llama, llama.cpp:3872), 1 falcon (build_falcon, :4106)
    float norm_eps = 1e-5f;              // falcon: f_norm_eps
    int   tok_bos = 1, tok_eos = 2;
    // tensors (only those of the layer range this process owns are resident)
    int          layer_begin = 0, layer_end = 0;
    pib_weight * tok_embd    = nullptr;
    float *      output_norm = nullptr;
    float *      output_norm_b = nullptr;
    pib_weight * output      = nullptr;
    std::vector<pib_layer> layers;
    // vocabulary (for llama_token_to_piece / the char-level prompt path)
    std::vector<std::string> tok_text;
    std::vector<float>       tok_score;
    std::vector<int32_t>     tok_type;
    std::string              tok_model = "llama";  // tokenizer.ggml.model: "llama" (SPM) or "gpt2" (byte-level BPE)
    std::vector<std::string> tok_merges;           // BPE merges "left right", rank = position
    int                      tok_add_bos = -1;     // tokenizer.ggml.add_bos_token: 0 / 1, -1 = not in the file
    std::string              name;
    size_t                   weight_bytes = 0;     // algorithmic bytes of all resident matrices

    int head_dim() const { return n_embd / n_head; }
    int n_embd_gqa() const { return head_dim() * n_head_kv; }
};

struct pib_run {                   // one (possibly still executing) decode: the opaque handle of the async API
    int          id        = -1;   // index in the context's run table
    bool         in_use    = false;
    int          batch_id  = 0;
    int          n_tokens  = 0;
    bool         failed    = false;
    cudaEvent_t  done      = nullptr;
    float *      logits    = nullptr;     // device [n_tokens][n_vocab]
    pib_row_result * res_dev  = nullptr;  // device [n_tokens]
    pib_row_result * res_host = nullptr;  // pinned [n_tokens]
    int          cap_tokens = 0;
    bool         has_walk  = false;       // the batch came with an acceptance table: only the walk's result crosses to the host
    bool         rows_on_host = false;
    pib_accept_result * acc_dev = nullptr, * acc_host = nullptr;
};

struct pib_stage {                 // pipeline position of this process (1-GPU: first && last)
    int  rank = 0, size = 1;
    bool first() const { return rank == 0; }
    bool last() const { return rank == size - 1; }
};

struct pib_context {
    pib_model *  model   = nullptr;
    uint32_t     n_ctx   = 0;
    int          max_seq = 1;
    int          n_words = 1;
    pib_stage    stage;
    cudaStream_t stream  = nullptr;       // everything of this context runs in order on this stream
    cudaStream_t ctrl    = nullptr;       // high-priority control stream: cancel flags overtake queued work
    int          gemv_variant = 0;

    KvMirror kv;                           // host mirror (lazy: pending ops are applied before any observation)
    KvDevice kvd;
    std::vector<KvOp> pending;             // API-level op-log not yet applied to mirror / device
    std::vector<KvOp> devops;              // applied to the mirror, waiting to ride to the device with the next decode
    __half * kcache = nullptr, * vcache = nullptr;     // [n_layer_local][n_ctx][n_embd_gqa]

    // per-chunk work buffers
    float * vsep = nullptr;               // [chunk][n_embd_gqa]: V of a layer whose Wv is a separate matrix
    float * x = nullptr, * x2 = nullptr, * qkv = nullptr, * att = nullptr, * proj = nullptr, * gu = nullptr;
    uint8_t * mask = nullptr;
    void *  act = nullptr;
    float2 * rope = nullptr;
    int      rope_n_pos = 0;
    int *    cancel_table = nullptr;       // device cancellation block: see PIB_CANCELLED (pib_common.cuh)
    int *    run_id_host  = nullptr;       // pinned {0, 1, ..., 63}: sources of the current-run writes
    int *    cancel_one   = nullptr;       // device int {1}: source of the cancel writes
    cudaEvent_t cancel_ev = nullptr;       // recorded on `ctrl` after every cancel write: the clearing memset on `stream` waits for it
    bool     cancel_ev_pending = false;

    // wide-batch (prompt) path: work buffers for up to PIB_PROMPT_MAX tokens at once, allocated on the first prompt that uses them
    struct PromptBufs { float * x = nullptr, * x2 = nullptr, * qkv = nullptr, * att = nullptr, * proj = nullptr, * gu = nullptr;
                        uint8_t * mask = nullptr; void * act = nullptr, * xhat = nullptr; bool ready = false; } pb;
    int      fuse_pro = 0;                 // 1: small llama-arch Q4_K models run norm / quantise / silu inside the consuming GEMV (PIB_FUSE_PRO, engine.cu)
    int      prompt_gemm = 1;              // 0: prompts go through the GEMV kernels in chunks of 16 (PIB_PROMPT_GEMM=0, and always in exact mode)

    // peer-store hop (pipeline.cu): this stage's exported buffer, the next stage's (mapped), the previous stage's (mapped)
    void *   hop_mine = nullptr, * hop_next = nullptr, * hop_prev = nullptr;
    size_t   hop_slot_f4 = 0;
    bool     hop_peer = false;

    // host->device staging rings (pinned host + device copies)
    uint8_t * desc_host = nullptr, * desc_dev = nullptr;
    size_t    desc_stride = 0;
    int       desc_slots = 0, desc_next = 0;
    std::vector<cudaEvent_t> desc_ev;
    KvOp *    ops_host = nullptr, * ops_dev = nullptr;
    int       ops_cap = 0, ops_next = 0;

    // logit bias (e.g. EOS -> -inf) applied on the device before argmax
    int32_t * bias_ids_dev = nullptr;
    float *   bias_val_dev = nullptr;
    int       n_bias = 0;

    pib_accept_table     tree_pending;     // set by llama_pib_set_accept_table, consumed by the next decode started
    std::vector<pib_run> runs;
    std::deque<int>      inflight;        // run ids in issue order
    int                  last_finished = -1;
    float *              logits_fetch_dev = nullptr;   // head of a pipeline: staging for one row fetched from the last stage
    std::vector<float>   logits_host;     // lazily filled rows of the last finished run
    std::vector<uint8_t> logits_row_valid;
    bool                 logits_all = true;
    uint64_t             n_kernels = 0;   // launches issued by this context (bench: gpu_launches)
    uint64_t             h2d_bytes = 0, d2h_bytes = 0;     // host<->device traffic issued by this context
    // optional per-launch timing of the matmuls of one layer (bench.py roofline): CUDA events on the launching stream
    struct ProfRec { int kind, b; cudaEvent_t e0, e1; };
    int                  prof_layer = -1;
    bool                 prof_now   = false;
    std::vector<ProfRec> prof;
    uint64_t             n_decode  = 0;
    // CUDA graphs of the layer part of single-chunk decodes, keyed by n_tokens (engine.cu: run_chunk)
    struct GraphEntry { cudaGraphExec_t exec = nullptr; int uses = 0; int n_kernels = 0; bool failed = false; };
    std::map<int, GraphEntry> graphs;
    bool                 use_graphs = true;
    uint64_t             n_graph_launches = 0;
    int64_t              t_start_us = 0;
};

// engine.cu
bool          pib_exact_mode();      // PIB_EXACT=1
pib_model *   pib_model_load_gguf(const char * path, int layer_begin, int layer_end, bool want_embd, bool want_output);
pib_model *   pib_model_synthetic(int n_layer, int n_embd, int n_head, int n_head_kv, int n_ff, int n_vocab, int wtype,
                                  uint64_t seed, int flavour, float agree, int layer_begin, int layer_end, bool want_embd,
                                  bool want_output);
void          pib_model_free(pib_model * m);
pib_context * pib_context_new(pib_model * m, uint32_t n_ctx, int max_seq, pib_stage stage);
void          pib_context_free(pib_context * c);
void          pib_kv_push(pib_context * c, const KvOp & op);          // API-level KV op (lazy, coalescing)
void          pib_kv_flush_host(pib_context * c);                     // apply pending ops to the mirror + queue for device
// start a decode of n_tokens tokens; returns run id or -1 (no KV slot / error)
int  pib_decode_start(pib_context * c, int n_tokens, const int32_t * token, const int32_t * pos, const int32_t * n_seq_id,
                      const int32_t * const * seq_id, int batch_id);
bool pib_run_ready(pib_context * c, int run);
int  pib_run_finish(pib_context * c, int run);                         // waits; makes results available; frees the slot
const pib_row_result * pib_last_results(pib_context * c);
const pib_accept_result * pib_last_accept(pib_context * c);
const float *          pib_logits_row(pib_context * c, int i);         // lazy D2H of one row of the last finished run
void pib_cancel(pib_context * c, const int32_t * batch_ids, int n);
// stages > 0: replay one decode record of the head (descriptors + resolved KV op-log) / release a finished run
int  pib_decode_replay(pib_context * c, int batch_id, int n_tokens, int n_chunks, const uint8_t * descs, size_t desc_stride,
                       const KvOp * ops, int n_ops);
void pib_stage_finish(pib_context * c, int batch_id);
int  pib_apply_devops(pib_context * c);
int  pib_kv_sync_replicas(pib_context * c);

// pipeline.cu
bool pib_pipe_ring_active();
bool pib_pipe_nccl_active();
void pib_pipe_install_hooks();
int  pib_pipe_ship_decode(int batch_id, int n_tokens, int n_chunks, const uint8_t * const * descs, size_t desc_stride,
                          const int * slots, const KvOp * ops, int n_ops);
int  pib_pipe_ship_ints(int kind, int tag, const int32_t * vals, int n);      // kind: 2 finish, 3 cancel, 5 return, 6 shutdown
int  pib_pipe_ship_bias(const int32_t * ids, const float * vals, int n);
int  pib_pipe_ship_cancel(const int32_t * batch_ids, int n);
int  pib_pipe_send_results(pib_context * c, pib_run & run);
int  pib_pipe_post_result_recv(pib_context * c, pib_run & run);
int  pib_pipe_fetch_logits(pib_context * c, int row, float * dev_staging, float * host_dst);
int  pib_stage_serve(pib_context * c);
int  pib_pipe_hop_setup(pib_context * c);
void pib_pipe_hop_free(pib_context * c);
void pib_pipe_hop_account(size_t bytes);
int  pib_hop_recv_launch(pib_context * c, int n_tokens, cudaStream_t st);
int  pib_hop_send_launch(pib_context * c, const float * a, const float * b, const float * c3, int n_tokens, cudaStream_t st);
extern int (*pib_hop_recv)(pib_context * c, float * x, int n_tokens, const int * cancel);
extern int (*pib_hop_send)(pib_context * c, const float * x, int n_tokens, const int * cancel);
void pib_set_logit_bias(pib_context * c, const int32_t * ids, const float * vals, int n);
