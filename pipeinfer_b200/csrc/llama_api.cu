// llama_api.cu -- the llama.h-compatible C ABI (include/pipeinfer_llama.h) on top of the engine.
#include "engine.cuh"
#include "tokenizer.cuh"
#include "../../include/pipeinfer_llama.h"

#include <string.h>
#include <sys/time.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <string>
#include <vector>

static int64_t now_us_api() {
    struct timeval tv;
    gettimeofday(&tv, nullptr);
    return (int64_t) tv.tv_sec * 1000000 + tv.tv_usec;
}

struct llama_model {
    pib_model * m = nullptr;
    pib_vocab *    vocab = nullptr;               // built on first use (tokenize / piece / nl / vocab_type)
    std::once_flag vocab_once;
};

struct llama_context {
    llama_model * model = nullptr;
    pib_context * c     = nullptr;
    std::vector<float> dummy_logits;
};

static void ensure_init() {
    static bool done = false;
    if (done) return;
    int dev = 0;
    const char * lr = getenv("LOCAL_RANK");
    if (lr) dev = atoi(lr);
    if (pib_init(dev) != PIB_OK) {
        fprintf(stderr, "libpipeinfer_b200: %s\n", pib_last_error());
        abort();      // the reference's hard failures are GGML_ASSERT -> abort; there is no CPU path
    }
    done = true;
}

extern "C" {

struct llama_model_params llama_model_default_params(void) {
    struct llama_model_params p;
    memset(&p, 0, sizeof(p));
    p.use_mmap = true;
    return p;
}

struct llama_context_params llama_context_default_params(void) {
    struct llama_context_params p;
    memset(&p, 0, sizeof(p));
    p.seed = 0xFFFFFFFFu; p.n_ctx = 512; p.n_batch = 512; p.n_threads = 4; p.n_threads_batch = 4; p.rope_scaling_type = -1;
    p.yarn_ext_factor = -1.0f; p.yarn_attn_factor = 1.0f; p.yarn_beta_fast = 32.0f; p.yarn_beta_slow = 1.0f;
    p.mul_mat_q = true; p.f16_kv = true;
    return p;
}

void llama_backend_init(bool numa) { (void) numa; ensure_init(); }
void llama_backend_free(void) {}

struct llama_model * llama_load_model_from_file(const char * path_model, struct llama_model_params params) {
    // vocab_only (llama.h:176): hyper-parameters and vocabulary, no tensor, no CUDA call -- the tokenizer works without a GPU
    if (!params.vocab_only) ensure_init();
    pib_model * m = params.vocab_only ? pib_model_load_gguf(path_model, 0, 0, false, false) : pib_model_load_gguf(path_model, 0, -1, true, true);
    if (!m) { fprintf(stderr, "llama_load_model_from_file: %s\n", pib_last_error()); return nullptr; }
    llama_model * lm = new llama_model();
    lm->m = m;
    return lm;
}

struct llama_model * llama_pib_synthetic_model(int n_layer, int n_embd, int n_head, int n_head_kv, int n_ff, int n_vocab,
                                               int wtype, uint64_t seed, int flavour, float agree) {
    ensure_init();
    pib_model * m = pib_model_synthetic(n_layer, n_embd, n_head, n_head_kv, n_ff, n_vocab, wtype, seed, flavour, agree, 0, -1, true, true);
    if (!m) { fprintf(stderr, "llama_pib_synthetic_model: %s\n", pib_last_error()); return nullptr; }
    llama_model * lm = new llama_model();
    lm->m = m;
    return lm;
}

// one pipeline stage's share of a model: layers [layer_begin, layer_end), the token embedding on the first stage, the
// final norm + output matrix on the last (reference: every rank loads the whole file, llama.cpp:2684-2968; here a
// stage keeps only what it computes with)
struct llama_model * llama_pib_synthetic_model_slice(int n_layer, int n_embd, int n_head, int n_head_kv, int n_ff, int n_vocab,
                                                     int wtype, uint64_t seed, int flavour, float agree, int layer_begin,
                                                     int layer_end, int want_embd, int want_output) {
    ensure_init();
    pib_model * m = pib_model_synthetic(n_layer, n_embd, n_head, n_head_kv, n_ff, n_vocab, wtype, seed, flavour, agree, layer_begin,
                                        layer_end, want_embd != 0, want_output != 0);
    if (!m) { fprintf(stderr, "llama_pib_synthetic_model_slice: %s\n", pib_last_error()); return nullptr; }
    llama_model * lm = new llama_model();
    lm->m = m;
    return lm;
}

struct llama_model * llama_pib_load_model_slice(const char * path_model, int layer_begin, int layer_end, int want_embd, int want_output) {
    ensure_init();
    pib_model * m = pib_model_load_gguf(path_model, layer_begin, layer_end, want_embd != 0, want_output != 0);
    if (!m) { fprintf(stderr, "llama_pib_load_model_slice: %s\n", pib_last_error()); return nullptr; }
    llama_model * lm = new llama_model();
    lm->m = m;
    return lm;
}

void llama_free_model(struct llama_model * model) {
    if (!model) return;
    pib_vocab_free(model->vocab);
    pib_model_free(model->m);
    delete model;
}

struct llama_context * llama_pib_new_context(struct llama_model * model, uint32_t n_ctx, int32_t max_seq_ids) {
    return llama_pib_new_context_stage(model, n_ctx, max_seq_ids, 0, 1);
}

struct llama_context * llama_pib_new_context_stage(struct llama_model * model, uint32_t n_ctx, int32_t max_seq_ids, int rank, int size) {
    if (!model) return nullptr;
    pib_stage st;
    st.rank = rank; st.size = size < 1 ? 1 : size;
    if (st.size > 1) {
        if (!pib_pipe_ring_active() || !pib_pipe_nccl_active()) { fprintf(stderr, "llama_pib_new_context_stage: open the pipeline first (pib_pipe_ring_open + pib_pipe_nccl_init)\n"); return nullptr; }
        pib_pipe_install_hooks();
    }
    pib_context * c = pib_context_new(model->m, n_ctx ? n_ctx : (uint32_t) model->m->n_ctx_train, max_seq_ids, st);
    if (!c) { fprintf(stderr, "llama_new_context_with_model: %s\n", pib_last_error()); return nullptr; }
    llama_context * lc = new llama_context();
    lc->model = model;
    lc->c     = c;
    const char * v = getenv("PIB_GEMV_VARIANT");
    if (pib_exact_mode()) c->gemv_variant = 7;          // the kernels that keep the reference's summation order
    if (v) c->gemv_variant = atoi(v);
    return lc;
}

struct llama_context * llama_new_context_with_model(struct llama_model * model, struct llama_context_params params) {
    // PipeInfer's scheduler uses sequence ids up to 1000*np + 1 (speculative.cpp:140-145); 4096 bits cover np <= 4
    const char * ms = getenv("PIB_MAX_SEQ");
    llama_context * lc = llama_pib_new_context(model, params.n_ctx, ms ? atoi(ms) : 4096);
    if (lc) lc->c->logits_all = params.logits_all;
    return lc;
}

// 1 = run this context's kernels on a higher-priority stream than other contexts' (the draft model shares the head GPU
// with the first target stage and sits on the scheduler's critical path).  Call right after creating the context.
int llama_pib_set_priority(struct llama_context * ctx, int high) {
    pib_context * c = ctx->c;
    int lo, hi;
    if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess) return -1;
    cudaStreamSynchronize(c->stream);
    for (auto & g : c->graphs) if (g.second.exec) cudaGraphExecDestroy(g.second.exec);
    c->graphs.clear();
    cudaStream_t ns = nullptr;
    const int pr = high ? (hi < lo ? hi + 1 <= lo ? hi + 1 : hi : hi) : lo;      // one step below the control stream's priority
    if (cudaStreamCreateWithPriority(&ns, cudaStreamNonBlocking, pr) != cudaSuccess) return -1;
    cudaStreamDestroy(c->stream);
    c->stream = ns;
    // A/B switch for the draft's matmul kernel only (the context the scheduler marks high-priority): its matrices are small
    // enough that the fixed cost of a launch, not bandwidth, sets the time; e.g. PIB_KEEP_PLANES=1 PIB_DRAFT_GEMV_VARIANT=1
    const char * dv = high ? getenv("PIB_DRAFT_GEMV_VARIANT") : nullptr;
    if (dv) c->gemv_variant = atoi(dv);
    return 0;
}

// ---- the rest of the llama.h surface that speculative.cpp / main.cpp / common reference (llama.h:323, 472-500, 760-800, 860-872) ----

// llama.cpp:9263 / :9310: in the reference these differ from seq_cp only in the direction the op travels along the MPI
// ring; every stage here receives every op from the head, so both are seq_cp (with the reference's src == dst shortcut)
void llama_kv_cache_seq_cp_back(struct llama_context * ctx, llama_seq_id src, llama_seq_id dst, llama_pos p0, llama_pos p1) {
    if (src == dst) return;
    llama_kv_cache_seq_cp(ctx, src, dst, p0, p1);
}
void llama_kv_cache_seq_cp_sync_bi(struct llama_context * ctx, llama_seq_id src, llama_seq_id dst, llama_pos p0, llama_pos p1) {
    if (src == dst) return;
    llama_kv_cache_seq_cp(ctx, src, dst, p0, p1);
}
// llama.cpp:1736 / :9348: position shift with re-rotation of the cached K (context shifting of `main`).  Not on the
// speculation path and not implemented: the call is refused loudly instead of leaving K rotated for the old positions.
// (The reference cannot do it on this path either: with GGML_USE_MPI -- which this fork needs to compile -- the decode that
// follows a shift aborts on GGML_ASSERT(inp0 == gf->nodes[0]), ggml-mpi.c:610, because the K_shift nodes come first.)
void llama_kv_cache_seq_shift(struct llama_context * ctx, llama_seq_id seq_id, llama_pos p0, llama_pos p1, llama_pos delta) {
    (void) ctx;
    pib_set_error("llama_kv_cache_seq_shift(seq %d, [%d, %d), %+d): K-cache re-rotation is not implemented", seq_id, p0, p1, delta);
    fprintf(stderr, "libpipeinfer_b200: %s\n", pib_last_error());
}

// llama.cpp:7286-7335
void llama_sample_repetition_penalties(struct llama_context * ctx, llama_token_data_array * candidates, const llama_token * last_tokens,
                                       size_t penalty_last_n, float penalty_repeat, float penalty_freq, float penalty_present) {
    (void) ctx;
    if (penalty_last_n == 0 || (penalty_repeat == 1.0f && penalty_freq == 0.0f && penalty_present == 0.0f)) return;
    std::map<llama_token, int> token_count;
    for (size_t i = 0; i < penalty_last_n; ++i) token_count[last_tokens[i]]++;
    for (size_t i = 0; i < candidates->size; ++i) {
        const auto it = token_count.find(candidates->data[i].id);
        if (it == token_count.end()) continue;
        const int count = it->second;
        if (candidates->data[i].logit <= 0) candidates->data[i].logit *= penalty_repeat;
        else                                candidates->data[i].logit /= penalty_repeat;
        candidates->data[i].logit -= float(count) * penalty_freq + float(count > 0) * penalty_present;
    }
    candidates->sorted = false;
}

void llama_print_timings(struct llama_context * ctx) {                                  // llama.cpp:10182-10200 (what there is to report here)
    const pib_context * c = ctx->c;
    const double wall_ms = (now_us_api() - c->t_start_us) / 1000.0;
    fprintf(stderr, "\nllama_print_timings:      decodes = %llu, kernel launches = %llu (graph launches %llu), h2d %llu B, d2h %llu B, total time = %10.2f ms\n",
            (unsigned long long) c->n_decode, (unsigned long long) c->n_kernels, (unsigned long long) c->n_graph_launches,
            (unsigned long long) c->h2d_bytes, (unsigned long long) c->d2h_bytes, wall_ms);
}
void llama_reset_timings(struct llama_context * ctx) { ctx->c->t_start_us = now_us_api(); ctx->c->n_decode = 0; }
const char * llama_print_system_info(void) {
    static char buf[160];
    int dev = 0, sms = 0, cc_major = 0, cc_minor = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&cc_minor, cudaDevAttrComputeCapabilityMinor, dev);
    snprintf(buf, sizeof(buf), "CUDA = 1 | sm = %d%d | SMs = %d | PDL = %d | pipeline stages = %d | ", cc_major, cc_minor, sms,
             pib_pdl_enabled() ? 1 : 0, pib_pipe_ring_active() ? 2 : 1);
    return buf;
}
void llama_log_set(void (*log_callback)(int level, const char * text, void * user_data), void * user_data) { (void) log_callback; (void) user_data; }

int llama_pib_stage_serve(struct llama_context * ctx) { return pib_stage_serve(ctx->c); }
// head: let every stage drain its stream and return from llama_pib_stage_serve (kind 5), or end for good (kind 6)
int llama_pib_pipe_pause(struct llama_context * ctx) {
    if (ctx->c->stage.size < 2) return 0;
    if (pib_kv_sync_replicas(ctx->c) != PIB_OK) return -1;
    const int rc = pib_pipe_ship_ints(5, 0, nullptr, 0);
    cudaStreamSynchronize(ctx->c->stream);
    return rc;
}
int llama_pib_pipe_shutdown(struct llama_context * ctx) {
    if (ctx->c->stage.size < 2) return 0;
    return pib_pipe_ship_ints(6, 0, nullptr, 0);
}

void llama_free(struct llama_context * ctx) {
    if (!ctx) return;
    pib_context_free(ctx->c);
    delete ctx;
}

int64_t llama_time_us(void) {
    struct timeval tv;
    gettimeofday(&tv, nullptr);
    return (int64_t) tv.tv_sec * 1000000 + tv.tv_usec;
}

void llama_set_n_threads(struct llama_context * ctx, uint32_t a, uint32_t b) { (void) ctx; (void) a; (void) b; }

const struct llama_model * llama_get_model(const struct llama_context * ctx) { return ctx->model; }
int llama_n_ctx(const struct llama_context * ctx) { return (int) ctx->c->n_ctx; }
int llama_n_vocab(const struct llama_model * model) { return model->m->n_vocab; }
int llama_pib_n_layer(const struct llama_model * model) { return model->m->n_layer; }
int llama_n_ctx_train(const struct llama_model * model) { return model->m->n_ctx_train; }
int llama_n_embd(const struct llama_model * model) { return model->m->n_embd; }
llama_token llama_token_bos(const struct llama_model * model) { return model->m->tok_bos; }
llama_token llama_token_eos(const struct llama_model * model) { return model->m->tok_eos; }
size_t llama_pib_weight_bytes(const struct llama_model * model) { return model->m->weight_bytes; }

// ---- tokenizer (tokenizer.cu): SPM and byte-level BPE, special tokens cut out first when `special` ------------------
static const pib_vocab * vocab_of(const llama_model * model) {
    llama_model * lm = const_cast<llama_model *>(model);
    std::call_once(lm->vocab_once, [lm] { lm->vocab = pib_vocab_build(lm->m); });
    return lm->vocab;
}

int llama_vocab_type(const struct llama_model * model) { return pib_vocab_type(vocab_of(model)); }      // llama.h:70-73
llama_token llama_token_nl(const struct llama_model * model) { return pib_vocab_linefeed(vocab_of(model)); }
int llama_add_bos_token(const struct llama_model * model) { return model->m->tok_add_bos; }

// returns the number of tokens, -needed when n_max_tokens is too small (llama.cpp:10079-10093); INT32_MIN when the text
// cannot be tokenized at all (invalid UTF-8 for a BPE vocabulary, a byte without a token) -- the reference throws there
int llama_tokenize(const struct llama_model * model, const char * text, int text_len, llama_token * tokens, int n_max_tokens,
                   bool add_bos, bool special) {
    std::vector<int32_t> out;
    if (pib_vocab_tokenize(vocab_of(model), std::string(text, (size_t) (text_len > 0 ? text_len : 0)), add_bos, special, out) != 0) {
        fprintf(stderr, "llama_tokenize: %s\n", pib_last_error());
        return PIB_TOKENIZE_INVALID;
    }
    if ((int) out.size() > n_max_tokens) return -(int) out.size();
    for (size_t i = 0; i < out.size(); i++) tokens[i] = out[i];
    return (int) out.size();
}

int llama_token_to_piece(const struct llama_model * model, llama_token token, char * buf, int length) {
    std::string r;
    if (!pib_vocab_piece(vocab_of(model), token, r)) return 0;
    if ((int) r.size() > length) return -(int) r.size();
    memcpy(buf, r.data(), r.size());
    return (int) r.size();
}

// ---- batches ----------------------------------------------------------------------------------------------------------
struct llama_batch llama_batch_get_one(llama_token * tokens, int32_t n_tokens, llama_pos pos_0, llama_seq_id seq_id) {
    struct llama_batch b;
    memset(&b, 0, sizeof(b));
    b.n_tokens = n_tokens; b.token = tokens; b.all_pos_0 = pos_0; b.all_pos_1 = 1; b.all_seq_id = seq_id; b.batch_id = 0; b.max_n_seq = 1;
    return b;
}

struct llama_batch llama_batch_init(int32_t n_tokens, int32_t embd, int32_t n_seq_max) {
    struct llama_batch b;
    memset(&b, 0, sizeof(b));
    b.max_n_seq = n_seq_max;
    if (embd) b.embd = (float *) malloc(sizeof(float) * n_tokens * embd);
    else      b.token = (llama_token *) malloc(sizeof(llama_token) * n_tokens);
    b.pos      = (llama_pos *) malloc(sizeof(llama_pos) * n_tokens);
    b.n_seq_id = (int32_t *) malloc(sizeof(int32_t) * n_tokens);
    b.seq_id   = (llama_seq_id **) malloc(sizeof(llama_seq_id *) * n_tokens);
    for (int i = 0; i < n_tokens; ++i) b.seq_id[i] = (llama_seq_id *) malloc(sizeof(llama_seq_id) * n_seq_max);
    b.logits   = (int8_t *) malloc(sizeof(int8_t) * n_tokens);
    return b;
}

void llama_batch_free(struct llama_batch & batch) {
    // like the reference (llama.cpp:9877-9898) this frees seq_id[i] only for i < n_tokens
    free(batch.token); free(batch.embd); free(batch.pos); free(batch.n_seq_id);
    if (batch.seq_id) { for (int i = 0; i < batch.n_tokens; ++i) free(batch.seq_id[i]); free(batch.seq_id); }
    free(batch.logits);
    batch.token = nullptr; batch.embd = nullptr; batch.pos = nullptr; batch.n_seq_id = nullptr; batch.seq_id = nullptr; batch.logits = nullptr;
}

// ---- decode -----------------------------------------------------------------------------------------------------------
static int start_from_batch(llama_context * ctx, const llama_batch & batch) {
    if (batch.embd && !batch.token) { pib_set_error("decode: embedding input is not supported"); return -1; }
    const int n = batch.n_tokens;
    if (n <= 0) return -1;
    std::vector<llama_pos> pos;
    std::vector<int32_t>   nseq;
    std::vector<llama_seq_id> seq1;
    std::vector<const llama_seq_id *> seqp;
    const llama_pos * p = batch.pos;
    if (!p) { pos.resize(n); for (int i = 0; i < n; i++) pos[i] = batch.all_pos_0 + i * batch.all_pos_1; p = pos.data(); }
    const int32_t * ns = batch.n_seq_id;
    const llama_seq_id * const * sp = (const llama_seq_id * const *) batch.seq_id;
    if (!batch.seq_id) {
        nseq.assign(n, 1); seq1.assign(n, batch.all_seq_id); seqp.resize(n);
        for (int i = 0; i < n; i++) seqp[i] = &seq1[i];
        ns = nseq.data(); sp = seqp.data();
    }
    return pib_decode_start(ctx->c, n, batch.token, p, ns, sp, batch.batch_id);
}

int llama_decode(struct llama_context * ctx, struct llama_batch batch) {
    const int run = start_from_batch(ctx, batch);
    if (run < 0) return -1;
    return pib_run_finish(ctx->c, run) == 0 ? 0 : -1;
}

struct ggml_cgraph * llama_start_async_decode(struct llama_context & lctx, struct llama_batch batch) {
    const int run = start_from_batch(&lctx, batch);
    return run < 0 ? nullptr : reinterpret_cast<struct ggml_cgraph *>((intptr_t) (run + 1));
}

int llama_finish_async_decode(struct llama_context & lctx, struct llama_batch batch, struct ggml_cgraph * cgraph) {
    (void) batch;
    if (!cgraph) return -1;
    return pib_run_finish(lctx.c, (int) ((intptr_t) cgraph - 1)) == 0 ? 1 : -1;
}

bool llama_mpi_iprobe(struct llama_context * lctx) {
    pib_context * c = lctx->c;
    return !c->inflight.empty() && pib_run_ready(c, c->inflight.front());
}

void llama_cancel_run(struct llama_context * ctx, int32_t * canceled, int count) { pib_cancel(ctx->c, canceled, count); }

float * llama_get_logits_ith(struct llama_context * ctx, int32_t i) { return const_cast<float *>(pib_logits_row(ctx->c, i)); }
float * llama_get_logits(struct llama_context * ctx) {
    pib_context * c = ctx->c;
    if (c->last_finished < 0) return nullptr;
    const int n = c->runs[c->last_finished].n_tokens;
    for (int i = 0; i < n; i++) pib_logits_row(c, i);
    return c->logits_host.data();
}

int llama_pib_get_top8(struct llama_context * ctx, int32_t i, llama_token_data * out8) {
    pib_context * c = ctx->c;
    const pib_row_result * r = pib_last_results(c);
    if (!r || i < 0 || i >= c->runs[c->last_finished].n_tokens) return -1;
    for (int k = 0; k < PIB_TOPK; k++) { out8[k].id = r[i].id[k]; out8[k].logit = r[i].logit[k]; out8[k].p = r[i].p[k]; }
    return 0;
}

// acceptance on the device (north_star item 4): the table of the NEXT decode started on this context
void llama_pib_set_accept_table(struct llama_context * ctx, int32_t n_branches, int32_t keep, int32_t speculative, const int32_t * len,
                                const int32_t * rows, const int32_t * toks, int32_t depth_stride) {
    pib_accept_table & T = ctx->c->tree_pending;
    T.n_branches = 0;
    if (n_branches < 1 || n_branches > PIB_TREE_NP || depth_stride < 1) return;       // too wide for the device table: the caller walks the rows itself
    memset(T.rows, 0, sizeof(T.rows)); memset(T.toks, 0, sizeof(T.toks));
    for (int s = 0; s < PIB_TREE_NP; s++) T.len[s] = -1;
    for (int s = 0; s < n_branches; s++) {
        if (len[s] > PIB_TREE_DEPTH - 1) return;
        T.len[s] = len[s];
        for (int i = 0; i < PIB_TREE_DEPTH && i < depth_stride; i++) { T.rows[s][i] = rows[s * depth_stride + i]; T.toks[s][i] = toks[s * depth_stride + i]; }
    }
    T.keep = keep; T.speculative = speculative; T.n_branches = n_branches;
}

// after llama_finish_async_decode of a run that carried a table: the ids sampled along the walk, and the branch kept
int32_t llama_pib_get_accepted(struct llama_context * ctx, int32_t * ids, int32_t cap, int32_t * s_keep) {
    const pib_accept_result * r = pib_last_accept(ctx->c);
    if (!r) return -1;
    for (int i = 0; i < r->n_ids && i < cap; i++) ids[i] = r->ids[i];
    if (s_keep) *s_keep = r->s_keep;
    return r->n_ids;
}

void llama_pib_set_logit_bias(struct llama_context * ctx, const llama_token * ids, const float * bias, int n) {
    pib_set_logit_bias(ctx->c, ids, bias, n);
}

uint64_t llama_pib_kernel_launches(const struct llama_context * ctx) { return ctx->c->n_kernels; }

// ---- KV cache ---------------------------------------------------------------------------------------------------------
void llama_kv_cache_clear(struct llama_context * ctx) { pib_kv_push(ctx->c, { KVOP_CLEAR, 0, 0, 0, 0, 0 }); }
void llama_kv_cache_seq_rm(struct llama_context * ctx, llama_seq_id seq_id, llama_pos p0, llama_pos p1) {
    pib_kv_push(ctx->c, { KVOP_RM, seq_id, 0, p0, p1, 0 });
}
void llama_kv_cache_seq_cp(struct llama_context * ctx, llama_seq_id src, llama_seq_id dst, llama_pos p0, llama_pos p1) {
    if (src == dst) return;                    // the reference's wrapper returns before touching the cache (llama.cpp:9303-9305): head keeps its value
    pib_kv_push(ctx->c, { KVOP_CP, src, dst, p0, p1, 0 });
}
void llama_kv_cache_seq_keep(struct llama_context * ctx, llama_seq_id seq_id) { pib_kv_push(ctx->c, { KVOP_KEEP, seq_id, 0, 0, 0, 0 }); }
// ranged forms (extension): exactly what the per-id loops of the scheduler collapse to in the op-log, without the 12 000 calls
void llama_pib_kv_seq_rm_range(struct llama_context * ctx, llama_seq_id first, llama_seq_id last, llama_pos p0, llama_pos p1) {
    kvlog_push_ranged(ctx->c->pending, KVOP_RM, first, last, 0, p0, p1);
}
void llama_pib_kv_seq_cp_range(struct llama_context * ctx, llama_seq_id src, llama_seq_id dst_first, llama_seq_id dst_last, llama_pos p0, llama_pos p1) {
    kvlog_push_ranged(ctx->c->pending, KVOP_CP, src, dst_first, dst_last, p0, p1);
}
void llama_pib_kv_seq_sweep(struct llama_context * ctx, llama_seq_id first, llama_seq_id last, llama_pos p0, llama_pos p1) {
    kvlog_push_ranged(ctx->c->pending, KVOP_SWEEP, first, last, 0, p0, p1);
}

int llama_get_kv_cache_used_cells(const struct llama_context * ctx) {
    pib_kv_flush_host(ctx->c);
    return (int) ctx->c->kv.used;
}

int llama_get_kv_cache_token_count(const struct llama_context * ctx) {
    pib_kv_flush_host(ctx->c);
    int n = 0;
    for (uint32_t w : ctx->c->kv.bits) n += __builtin_popcount(w);
    return n;
}

struct llama_kv_cache_view llama_kv_cache_view_init(const struct llama_context * ctx, int32_t n_max_seq) {
    (void) ctx;
    struct llama_kv_cache_view v;
    memset(&v, 0, sizeof(v));
    v.n_max_seq = n_max_seq;
    return v;
}

void llama_kv_cache_view_free(struct llama_kv_cache_view * view) {
    free(view->cells); free(view->cells_sequences);
    view->cells = nullptr; view->cells_sequences = nullptr;
}

// same fields and fill rules as the reference (llama.cpp:9160-9222)
void llama_kv_cache_view_update(const struct llama_context * ctx, struct llama_kv_cache_view * view) {
    pib_context * c = ctx->c;
    pib_kv_flush_host(c);
    const KvMirror & kv = c->kv;
    if ((uint32_t) view->n_cells < kv.size || !view->cells) {
        view->n_cells = (int32_t) kv.size;
        view->cells = (struct llama_kv_cache_view_cell *) realloc(view->cells, sizeof(struct llama_kv_cache_view_cell) * view->n_cells);
        view->cells_sequences = (llama_seq_id *) realloc(view->cells_sequences, sizeof(llama_seq_id) * view->n_max_seq * view->n_cells);
    }
    int32_t used = 0, tokens = 0, cur = -1, best_idx = -1;
    uint32_t best = 0;
    for (int32_t i = 0; i < (int32_t) kv.size; i++) {
        llama_seq_id * cs = view->cells_sequences + (size_t) i * view->n_max_seq;
        view->cells[i].pos = kv.pos[i];
        int k = 0, cnt = 0;
        for (int w = 0; w < kv.n_words; w++) {
            uint32_t bits = kv.bits[(size_t) i * kv.n_words + w];
            cnt += __builtin_popcount(bits);
            while (bits) { const int b = __builtin_ctz(bits); bits &= bits - 1; if (k < view->n_max_seq) cs[k++] = w * 32 + b; }
        }
        tokens += cnt;
        if (cnt > 0) {
            if (cur >= 0 && (uint32_t) (i - cur) > best) { best = (uint32_t) (i - cur); best_idx = cur; }
            cur = -1;
            used++;
        } else if (cur < 0) cur = i;
        for (; k < view->n_max_seq; k++) cs[k] = -1;
    }
    if (cur >= 0 && kv.size - (uint32_t) cur > best) { best_idx = cur; best = kv.size - (uint32_t) cur; }
    view->max_contiguous = (int32_t) best; view->max_contiguous_idx = best_idx; view->token_count = tokens; view->used_cells = used;
}

int llama_pib_kv_words(const struct llama_context * ctx) { return ctx->c->n_words; }

int llama_pib_kv_state(struct llama_context * ctx, int from_device, int32_t * scalars3, int32_t * pos, uint32_t * bits) {
    pib_context * c = ctx->c;
    pib_kv_flush_host(c);
    const size_t nb = (size_t) c->n_ctx * c->n_words;
    if (!from_device) {
        scalars3[0] = (int32_t) c->kv.head; scalars3[1] = (int32_t) c->kv.used; scalars3[2] = (int32_t) c->kv.n;
        memcpy(pos, c->kv.pos.data(), sizeof(int32_t) * c->n_ctx);
        memcpy(bits, c->kv.bits.data(), sizeof(uint32_t) * nb);
        return 0;
    }
    // ops that were flushed to the mirror but not yet shipped ride with the next decode; push them now (to every stage)
    if (pib_kv_sync_replicas(c) != PIB_OK) return -1;
    int32_t s[4];
    if (cudaMemcpyAsync(s, c->kvd.scalars, 16, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaMemcpyAsync(pos, c->kvd.pos, sizeof(int32_t) * c->n_ctx, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaMemcpyAsync(bits, c->kvd.bits, sizeof(uint32_t) * nb, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) return -1;
    scalars3[0] = s[0]; scalars3[1] = s[1]; scalars3[2] = s[2];
    return 0;
}

// ---- communicator helpers: single-process heads => no-ops ---------------------------------------------------------------
void llama_split_comm(struct llama_context * ctx, int color) { (void) ctx; (void) color; }
void llama_swap_comm(struct llama_context * ctx) { (void) ctx; }
void llama_sync_token(struct llama_context * ctx, llama_token * token, int root) { (void) ctx; (void) token; (void) root; }
void llama_sync_token_data(struct llama_context * ctx, llama_token_data * data, int root) { (void) ctx; (void) data; (void) root; }
void llama_split_layers_weighted(struct llama_context * ctx, float device_weights[], size_t num_weights) {
    (void) ctx; (void) device_weights; (void) num_weights;
}
int llama_node_id(struct llama_context * ctx) { return ctx->c->stage.rank; }

// ---- host samplers kept for API completeness (the scheduler uses the device results) -----------------------------------
void llama_sample_softmax(struct llama_context * ctx, llama_token_data_array * cand) {
    (void) ctx;
    if (!cand->sorted) {
        std::sort(cand->data, cand->data + cand->size, [](const llama_token_data & a, const llama_token_data & b) { return a.logit > b.logit; });
        cand->sorted = true;
    }
    const float max_l = cand->data[0].logit;
    float cum = 0.0f;
    for (size_t i = 0; i < cand->size; ++i) { cand->data[i].p = expf(cand->data[i].logit - max_l); cum += cand->data[i].p; }
    for (size_t i = 0; i < cand->size; ++i) cand->data[i].p /= cum;
}

llama_token llama_sample_token_greedy(struct llama_context * ctx, llama_token_data_array * cand) {
    (void) ctx;
    size_t best = 0;
    for (size_t i = 1; i < cand->size; i++) if (cand->data[best].logit < cand->data[i].logit) best = i;
    return cand->data[best].id;
}

}   // extern "C"

// debugging aid for parity work: read back a work buffer of the last chunk (0: residual stream x, 1: ffn_inp x2,
// 2: attention output, 3: last projection output)
extern "C" int llama_pib_debug_read(struct llama_context * ctx, int which, float * out, int n_floats) {
    pib_context * c = ctx->c;
    const float * src = which == 0 ? c->x : which == 1 ? c->x2 : which == 2 ? c->att : c->proj;
    if (cudaStreamSynchronize(c->stream) != cudaSuccess) return -1;
    return cudaMemcpy(out, src, sizeof(float) * (size_t) n_floats, cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -1;
}

// debugging aid: handle of a resident weight matrix (layer < 0: 0 = token_embd, 1 = output; else 0 qkv, 1 o, 2 gate|up, 3 down)
extern "C" const pib_weight * llama_pib_debug_weight(const struct llama_model * model, int layer, int which) {
    const pib_model * m = model->m;
    if (layer < 0) return which == 0 ? m->tok_embd : m->output;
    const pib_layer & L = m->layers[layer];
    return which == 0 ? L.wqkv : which == 1 ? L.wo : which == 2 ? L.wgu : L.wdown;
}

// ---- measurement hooks (bench.py) -----------------------------------------------------------------------------------
extern "C" void llama_pib_io_bytes(const struct llama_context * ctx, uint64_t * h2d, uint64_t * d2h) {
    *h2d = ctx->c->h2d_bytes; *d2h = ctx->c->d2h_bytes;
}
// time every matmul launch of `layer` (and the lm_head) with CUDA events on the launching stream; -1 = off
extern "C" void llama_pib_profile_layer(struct llama_context * ctx, int layer) { ctx->c->prof_layer = layer; }
// fetch and clear the records: kind (0 qkv, 1 o, 2 gate|up, 3 down, 4 lm_head), batch, milliseconds; returns the count
extern "C" int llama_pib_profile_read(struct llama_context * ctx, int32_t * kind, int32_t * b, float * ms, int cap) {
    pib_context * c = ctx->c;
    cudaStreamSynchronize(c->stream);
    int n = 0;
    for (auto & r : c->prof) {
        float t = 0.f;
        cudaEventElapsedTime(&t, r.e0, r.e1);
        if (n < cap) { kind[n] = r.kind; b[n] = r.b; ms[n] = t; n++; }
        cudaEventDestroy(r.e0); cudaEventDestroy(r.e1);
    }
    c->prof.clear();
    return n;
}
// shape of a resident matrix: N, K, algorithmic bytes (kind as above; layer ignored for the lm_head)
extern "C" int llama_pib_matrix_shape(const struct llama_model * model, int layer, int kind, int64_t * N, int64_t * K, uint64_t * bytes) {
    const pib_model * m = model->m;
    const pib_weight * w = kind == 4 ? m->output : (layer < 0 || layer >= m->n_layer) ? nullptr :
                           kind == 0 ? m->layers[layer].wqkv : kind == 1 ? m->layers[layer].wo : kind == 2 ? m->layers[layer].wgu : m->layers[layer].wdown;
    if (!w) return -1;
    *N = w->N; *K = w->K; *bytes = w->bytes;
    return 0;
}
