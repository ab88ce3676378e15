// tests/sched_sim/sim_ref_surface.cpp -- TEST INFRASTRUCTURE: the rest of the llama.h surface, so that the REFERENCE's own
// examples/speculative/speculative.cpp + common/*.cpp (compiled from /root/reference, unmodified) link against the
// virtual-clock engine model of sim_llama.cpp instead of llama.o / ggml*.o.  oracle/Makefile builds it as
// oracle/_ref/speculative_sim.  Purpose: drive the reference's scheduler and this repository's scheduler with the same
// clock, model and KV bookkeeping and compare what they ask of the target (tests/test_scheduler_sim.py).
//
// Configuration comes from the environment: SIM_CFG="key=value,..." over the fields of sim_config, SIM_TRACE=<file>,
// SIM_OUT=<file> (counters as JSON, written by llama_backend_free).  A model path containing "target" is the target.
// Tokens: text <-> ids is byte + 3 (BOS = 1 in front); pieces print as " <id>".
#include "sim_llama.cpp"

#include <string>

static std::vector<llama_context *> g_contexts;
static std::vector<float> g_logits;

static void apply_env_config() {
    sim_config cfg = DEFAULTS;
    const char * e = getenv("SIM_CFG");
    std::string s = e ? e : "";
    size_t p = 0;
    while (p < s.size()) {
        size_t q = s.find(',', p);
        if (q == std::string::npos) q = s.size();
        const std::string kv = s.substr(p, q - p);
        const size_t eq = kv.find('=');
        if (eq != std::string::npos) {
            const std::string k = kv.substr(0, eq);
            const double v = atof(kv.c_str() + eq + 1);
#define F(name) if (k == #name) cfg.name = (decltype(cfg.name)) v;
            F(n_stages) F(stage_us) F(stage_tok_us) F(hop_us) F(ret_us) F(skip_us) F(draft_us) F(draft_tok_us) F(enqueue_us) F(call_us)
            F(agree) F(p_top) F(p_second) F(jitter) F(share_gpu0) F(n_vocab) F(seed) F(spec_issue_extra_us) F(nonspec_issue_extra_us) F(head_frac) F(tail_frac)
#undef F
        }
        p = q + 1;
    }
    sim_configure(&cfg);
}

extern "C" {

const char * pib_last_error(void) { return g_err; }

void llama_backend_init(bool) { apply_env_config(); }
void llama_backend_free(void) {
    if (getenv("SIM_OUT")) {
        FILE * f = fopen(getenv("SIM_OUT"), "w");
        if (f) {
            fprintf(f, "{\"now_us\": %.0f, \"bad_context_rows\": %ld, \"read_cancelled_rows\": %ld", g_now, g_bad_context, g_read_cancelled);
            for (llama_context * c : g_contexts)
                fprintf(f, ", \"%s\": {\"decodes\": %ld, \"refused\": %ld, \"cancelled_runs\": %ld, \"cells_used\": %u}", c->is_target ? "target" : "draft",
                        c->n_decodes, c->n_failed, c->n_cancelled_runs, c->kv.used);
            fprintf(f, "}\n");
            fclose(f);
        }
    }
    if (g_trace) { fclose(g_trace); g_trace = nullptr; }
}

struct llama_model_params llama_model_default_params(void) { struct llama_model_params p; memset(&p, 0, sizeof(p)); p.use_mmap = true; return p; }
struct llama_context_params llama_context_default_params(void) {
    struct llama_context_params p;
    memset(&p, 0, sizeof(p));
    p.seed = 0xFFFFFFFFu; p.n_ctx = 512; p.n_batch = 512; p.n_threads = 1; p.n_threads_batch = 1; p.rope_scaling_type = -1; p.yarn_ext_factor = -1.0f;
    p.yarn_attn_factor = 1.0f; p.yarn_beta_fast = 32.0f; p.yarn_beta_slow = 1.0f; p.mul_mat_q = true; p.f16_kv = true;
    return p;
}
struct llama_model * llama_load_model_from_file(const char * path, struct llama_model_params) {
    llama_model * m = new llama_model();
    m->is_target = strstr(path, "target") != nullptr;
    return m;
}
void llama_free_model(struct llama_model * m) { delete m; }
struct llama_context * llama_new_context_with_model(struct llama_model * m, struct llama_context_params p) {
    const int max_seq = getenv("SIM_MAX_SEQ") ? atoi(getenv("SIM_MAX_SEQ")) : 3002;
    llama_context * c = sim_new_context(p.n_ctx, max_seq, m->is_target);
    c->model = *m;
    g_contexts.push_back(c);
    return c;
}
void llama_free(struct llama_context *) {}          // counters are read at llama_backend_free

int llama_n_ctx(const struct llama_context * c) { return (int) c->kv.size; }
int llama_n_vocab(const struct llama_model *) { return G.n_vocab; }
int llama_n_embd(const struct llama_model *) { return 64; }
int llama_pib_n_layer(const struct llama_model *) { return 80; }
llama_token llama_token_bos(const struct llama_model *) { return 1; }
llama_token llama_token_nl(const struct llama_model *) { return 13; }
int llama_add_bos_token(const struct llama_model *) { return -1; }
int llama_vocab_type(const struct llama_model *) { return 0; }
int llama_tokenize(const struct llama_model *, const char * text, int len, llama_token * out, int cap, bool add_bos, bool) {
    const int n = len + (add_bos ? 1 : 0);
    if (n > cap) return -n;
    int k = 0;
    if (add_bos) out[k++] = 1;
    for (int i = 0; i < len; i++) out[k++] = 3 + (unsigned char) text[i];
    return k;
}
int llama_token_to_piece(const struct llama_model *, llama_token t, char * buf, int cap) {
    char tmp[16];
    const int n = snprintf(tmp, sizeof(tmp), " %d", t);
    if (n > cap) return -n;
    memcpy(buf, tmp, n);
    return n;
}

struct llama_batch llama_batch_init(int32_t n_tokens, int32_t embd, int32_t n_seq_max) {
    struct llama_batch b;
    memset(&b, 0, sizeof(b));
    b.max_n_seq = n_seq_max;
    (void) embd;
    b.token    = (llama_token *) malloc(sizeof(llama_token) * n_tokens);
    b.pos      = (llama_pos *) malloc(sizeof(llama_pos) * n_tokens);
    b.n_seq_id = (int32_t *) malloc(sizeof(int32_t) * n_tokens);
    b.seq_id   = (llama_seq_id **) malloc(sizeof(llama_seq_id *) * n_tokens);
    for (int i = 0; i < n_tokens; ++i) b.seq_id[i] = (llama_seq_id *) malloc(sizeof(llama_seq_id) * n_seq_max);
    b.logits   = (int8_t *) malloc(sizeof(int8_t) * n_tokens);
    return b;
}

// n_vocab floats for the reference's sampler: log(p) of the row's candidates, nothing else has any mass
float * llama_get_logits_ith(struct llama_context * c, int32_t i) {
    g_now += G.call_us;
    g_logits.assign(G.n_vocab, -1e9f);
    if (i < 0 || i >= (int) c->rows.size()) return g_logits.data();
    if (!c->rows[i].valid) g_read_cancelled++;
    if (!c->rows[i].whole && getenv("SIM_DEBUG")) fprintf(stderr, "sim: t=%.0f %s row %d read with a broken context (batch %d, seq %d, pos %d, saw %d cells, first gap/duplicate at %d)\n", g_now, c->is_target ? "target" : "draft", i, c->rows[i].batch_id, c->rows[i].seq0, c->rows[i].pos, c->rows[i].n_seen, c->rows[i].first_gap);
    for (int k = 7; k >= 0; k--) g_logits[c->rows[i].top[k].id] = c->is_target ? (float) (8 - k) : logf(std::max(c->rows[i].top[k].p, 1e-30f));
    return g_logits.data();
}

void llama_sample_softmax(struct llama_context *, llama_token_data_array * a) {
    if (!a->sorted) { std::sort(a->data, a->data + a->size, [](const llama_token_data & x, const llama_token_data & y) { return x.logit > y.logit; }); a->sorted = true; }
    const float mx = a->data[0].logit;
    float sum = 0;
    for (size_t i = 0; i < a->size; i++) { a->data[i].p = expf(a->data[i].logit - mx); sum += a->data[i].p; }
    for (size_t i = 0; i < a->size; i++) a->data[i].p /= sum;
}
llama_token llama_sample_token_greedy(struct llama_context *, llama_token_data_array * a) {
    return std::max_element(a->data, a->data + a->size, [](const llama_token_data & x, const llama_token_data & y) { return x.logit < y.logit; })->id;
}
void llama_sample_repetition_penalties(struct llama_context *, llama_token_data_array *, const llama_token *, size_t n, float, float, float) {
    if (n) { fprintf(stderr, "sim: repetition penalties are not modelled (--repeat-last-n 0)\n"); abort(); }
}

struct llama_kv_cache_view llama_kv_cache_view_init(const struct llama_context *, int32_t n_max_seq) {
    struct llama_kv_cache_view v;
    memset(&v, 0, sizeof(v));
    v.n_max_seq = n_max_seq;
    return v;
}

// examples/speculative_orig times itself in "cycles"
int64_t ggml_cycles(void) { return (int64_t) g_now; }
int64_t ggml_cycles_per_ms(void) { return 1000; }
void llama_batch_free(struct llama_batch & b) {
    free(b.token); free(b.pos); free(b.n_seq_id);
    if (b.seq_id) { for (int i = 0; i < b.n_tokens; ++i) free(b.seq_id[i]); free(b.seq_id); }
    free(b.logits);
}

// one process, no communicators
int  llama_node_id(struct llama_context *) { return 0; }
void llama_split_comm(struct llama_context *, int) {}
void llama_swap_comm(struct llama_context *) {}
void llama_sync_token(struct llama_context *, llama_token *, int) {}
void llama_sync_token_data(struct llama_context *, llama_token_data *, int) {}
void llama_split_layers_weighted(struct llama_context *, float *, size_t) {}
void llama_set_n_threads(struct llama_context *, uint32_t, uint32_t) {}
const char * llama_print_system_info(void) { return "sim"; }
void llama_print_timings(struct llama_context *) {}
void llama_reset_timings(struct llama_context *) {}

}   // extern "C"
