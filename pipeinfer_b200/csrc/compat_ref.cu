// compat_ref.cu -- the remaining symbols the reference's own examples/speculative/speculative.cpp + common/*.cpp link
// against (llama.h sampling section :700-800, grammar :810-850, misc :263-300; ggml.h time / cpu-feature probes), so that
// the UNMODIFIED reference program can be linked with this library in place of llama.o + ggml*.o (INTEGRATION.md §1;
// oracle/Makefile builds it as oracle/_ref/speculative_b200 and tests/test_gpu_cli.py runs it).
//
// Samplers are host code over llama_token_data_array exactly as in the reference; the ones on the path's greedy
// configuration (softmax, greedy, repetition penalties) live in llama_api.cu.  Grammar sampling, LoRA, mirostat,
// classifier-free guidance are outside the path: they are accepted when their
// parameters are neutral (as the reference's own early returns do) and refused loudly otherwise.
#include "engine.cuh"
#include "../../include/pipeinfer_llama.h"

#include <string.h>
#include <sys/time.h>

#include <algorithm>
#include <random>
#include <vector>

static void not_on_path(const char * what) {
    pib_set_error("%s is not implemented (outside the speculation path: greedy sampling only)", what);
    fprintf(stderr, "libpipeinfer_b200: %s\n", pib_last_error());
}

extern "C" {

// ---- ggml.h ----
int64_t ggml_time_us(void) { return llama_time_us(); }
void    ggml_print_backtrace(void) {}
int ggml_cpu_has_avx(void) { return 0; }
int ggml_cpu_has_avx2(void) { return 0; }
int ggml_cpu_has_avx512(void) { return 0; }
int ggml_cpu_has_avx512_vbmi(void) { return 0; }
int ggml_cpu_has_avx512_vnni(void) { return 0; }
int ggml_cpu_has_fma(void) { return 0; }
int ggml_cpu_has_neon(void) { return 0; }
int ggml_cpu_has_arm_fma(void) { return 0; }
int ggml_cpu_has_f16c(void) { return 0; }
int ggml_cpu_has_fp16_va(void) { return 0; }
int ggml_cpu_has_wasm_simd(void) { return 0; }
int ggml_cpu_has_blas(void) { return 0; }
int ggml_cpu_has_cublas(void) { return 1; }
int ggml_cpu_has_clblast(void) { return 0; }
int ggml_cpu_has_gpublas(void) { return 1; }
int ggml_cpu_has_sse3(void) { return 0; }
int ggml_cpu_has_ssse3(void) { return 0; }
int ggml_cpu_has_vsx(void) { return 0; }

// ---- llama.h misc ----
bool llama_mmap_supported(void) { return true; }
bool llama_mlock_supported(void) { return false; }
int  llama_max_devices(void) { return 16; }
int  llama_model_apply_lora_from_file(const struct llama_model * model, const char * path_lora, float scale, const char * path_base_model,
                                      int n_threads) {
    (void) model; (void) path_lora; (void) scale; (void) path_base_model; (void) n_threads;
    not_on_path("llama_model_apply_lora_from_file");
    return 1;
}

// ---- what examples/main/main.cpp needs on top (llama.h:249-260, 350, 524-540, 692, 860-872) ----
int64_t ggml_time_ms(void) { return llama_time_us() / 1000; }
void    llama_set_rng_seed(struct llama_context * ctx, uint32_t seed) { (void) ctx; (void) seed; }      // greedy path: no RNG in use
int     llama_model_desc(const struct llama_model * model, char * buf, size_t buf_size) {
    return snprintf(buf, buf_size, "pipeinfer_b200 model, %d layers, n_embd %d, n_vocab %d", llama_pib_n_layer(model), llama_n_embd(model), llama_n_vocab(model));
}
struct llama_timings { double t_start_ms, t_end_ms, t_load_ms, t_sample_ms, t_p_eval_ms, t_eval_ms; int32_t n_sample, n_p_eval, n_eval; };
struct llama_timings llama_get_timings(struct llama_context * ctx) {
    (void) ctx;
    struct llama_timings t;
    memset(&t, 0, sizeof(t));
    t.t_end_ms = (double) llama_time_us() / 1000.0;
    return t;
}
void llama_dump_timing_info_yaml(FILE * stream, const struct llama_context * ctx) { (void) ctx; fprintf(stream, "\n# timings: see llama_print_timings\n"); }
// session files (prompt caches) are outside the path: loading reports "no session", saving reports failure
bool llama_load_session_file(struct llama_context * ctx, const char * path_session, llama_token * tokens_out, size_t n_token_capacity,
                             size_t * n_token_count_out) {
    (void) ctx; (void) path_session; (void) tokens_out; (void) n_token_capacity;
    if (n_token_count_out) *n_token_count_out = 0;
    not_on_path("llama_load_session_file");
    return false;
}
bool llama_save_session_file(struct llama_context * ctx, const char * path_session, const llama_token * tokens, size_t n_token_count) {
    (void) ctx; (void) path_session; (void) tokens; (void) n_token_count;
    not_on_path("llama_save_session_file");
    return false;
}

// ---- samplers (llama.cpp:7019-7500) ----
// The three truncation samplers below exist so that the reference's unmodified common/sampling.cpp links against this library;
// greedy decoding (the path of SURVEY.md section 8) never calls them.  Behaviour as llama.cpp:7048-7130, own formulation.
static inline bool by_logit_desc(const llama_token_data & a, const llama_token_data & b) { return a.logit > b.logit; }

void llama_sample_top_k(struct llama_context *, llama_token_data_array * cand, int k, size_t min_keep) {
    const size_t want = std::min(cand->size, (size_t) std::max<long long>(k, (long long) min_keep));
    if (!cand->sorted) {
        // the same library algorithms as the reference (heap-based partial sort unless everything is kept): with equal logits
        // the order among them is whatever those algorithms leave, and callers compare ids
        llama_token_data * first = cand->data, * last = cand->data + cand->size;
        if (want < cand->size) std::partial_sort(first, first + want, last, by_logit_desc);
        else                   std::sort(first, last, by_logit_desc);
        cand->sorted = true;
    }
    cand->size = want;
}

void llama_sample_top_p(struct llama_context * ctx, llama_token_data_array * cand, float p, size_t min_keep) {
    if (!(p < 1.0f)) return;
    llama_sample_softmax(ctx, cand);                   // sorted, probabilities filled in
    // shortest prefix whose mass reaches p (fp32 running sum, left to right), but never fewer than min_keep entries
    float  mass = 0.0f;
    size_t n    = 0;
    while (n < cand->size) {
        mass += cand->data[n++].p;
        if (mass >= p && n >= min_keep) break;
    }
    cand->size = n;
}

void llama_sample_min_p(struct llama_context * ctx, llama_token_data_array * cand, float p, size_t min_keep) {
    if (cand->size == 0 || !(p > 0.0f)) return;
    llama_sample_softmax(ctx, cand);
    const float floor_p = p * cand->data[0].p;         // relative to the most likely token
    size_t n = 1;
    while (n < cand->size && !(cand->data[n].p < floor_p && n >= min_keep)) n++;
    cand->size = n;
}
// tail-free sampling: cut the candidates where the cumulated, normalised |second difference| of the sorted probabilities passes z
void llama_sample_tail_free(struct llama_context * ctx, llama_token_data_array * candidates, float z, size_t min_keep) { // :7131
    if (z >= 1.0f || candidates->size <= 2) return;
    llama_sample_softmax(ctx, candidates);
    const llama_token_data * c = candidates->data;
    const size_t n2 = candidates->size - 2;
    std::vector<float> curv(n2);
    float total = 0.0f;                                    // summed left to right in fp32, like std::accumulate there
    for (size_t i = 0; i < n2; ++i) {
        const float d0 = c[i].p - c[i + 1].p, d1 = c[i + 1].p - c[i + 2].p;
        curv[i] = std::fabs(d0 - d1);
        total += curv[i];
    }
    for (float & v : curv) v = total > 1e-6f ? v / total : 1.0f / (float) n2;
    float run = 0.0f;
    for (size_t i = 0; i < n2; ++i) {
        run += curv[i];
        if (run > z && i >= min_keep) { candidates->size = i; return; }
    }
}
// locally typical sampling: keep the candidates whose surprise is closest to the entropy until their mass passes p
void llama_sample_typical(struct llama_context * ctx, llama_token_data_array * candidates, float p, size_t min_keep) {  // :7190
    if (p >= 1.0f) return;
    llama_sample_softmax(ctx, candidates);
    const size_t n = candidates->size;
    float entropy = 0.0f;
    for (size_t i = 0; i < n; ++i) entropy += -candidates->data[i].p * logf(candidates->data[i].p);
    std::vector<float> dist(n);
    for (size_t i = 0; i < n; ++i) dist[i] = fabsf(-logf(candidates->data[i].p) - entropy);
    std::vector<size_t> order(n);
    for (size_t i = 0; i < n; ++i) order[i] = i;
    std::sort(order.begin(), order.end(), [&](size_t a, size_t b) { return dist[a] < dist[b]; });
    float  mass = 0.0f;
    size_t keep = n;
    for (size_t i = 0; i < n; ++i) {
        mass += candidates->data[order[i]].p;
        if (mass > p && i >= min_keep - 1) { keep = i + 1; break; }      // min_keep == 0 wraps and keeps everything, as there
    }
    std::vector<llama_token_data> kept(keep);
    for (size_t i = 0; i < keep; ++i) kept[i] = candidates->data[order[i]];
    std::copy(kept.begin(), kept.end(), candidates->data);
    candidates->size = keep;
}
void llama_sample_temp(struct llama_context * ctx, llama_token_data_array * candidates, float temp) {                   // :7259
    (void) ctx;
    for (size_t i = 0; i < candidates->size; ++i) candidates->data[i].logit /= temp;
}
void llama_sample_classifier_free_guidance(struct llama_context * ctx, llama_token_data_array * candidates,
                                           struct llama_context * guidance_ctx, float scale) {
    (void) ctx; (void) candidates; (void) guidance_ctx; (void) scale;
    not_on_path("llama_sample_classifier_free_guidance");
}
llama_token llama_sample_token(struct llama_context * ctx, llama_token_data_array * candidates) {                       // :7511
    (void) ctx;
    static std::mt19937 rng(0xFFFFFFFFu);                   // the reference seeds ctx->rng from params.seed; one stream here
    llama_sample_softmax(nullptr, candidates);
    std::vector<float> probs;
    probs.reserve(candidates->size);
    for (size_t i = 0; i < candidates->size; ++i) probs.push_back(candidates->data[i].p);
    std::discrete_distribution<> dist(probs.begin(), probs.end());
    return candidates->data[dist(rng)].id;
}
llama_token llama_sample_token_mirostat(struct llama_context * ctx, llama_token_data_array * candidates, float tau, float eta, int m, float * mu) {
    (void) tau; (void) eta; (void) m; (void) mu;
    not_on_path("llama_sample_token_mirostat");
    return llama_sample_token_greedy(ctx, candidates);
}
llama_token llama_sample_token_mirostat_v2(struct llama_context * ctx, llama_token_data_array * candidates, float tau, float eta, float * mu) {
    (void) tau; (void) eta; (void) mu;
    not_on_path("llama_sample_token_mirostat_v2");
    return llama_sample_token_greedy(ctx, candidates);
}

// ---- grammar (llama.h:810-850): only an absent grammar is supported ----
struct llama_grammar;
struct llama_grammar * llama_grammar_init(const void ** rules, size_t n_rules, size_t start_rule_index) {
    (void) rules; (void) n_rules; (void) start_rule_index;
    not_on_path("llama_grammar_init");
    return nullptr;
}
void llama_grammar_free(struct llama_grammar * grammar) { (void) grammar; }
struct llama_grammar * llama_grammar_copy(const struct llama_grammar * grammar) { (void) grammar; return nullptr; }
void llama_sample_grammar(struct llama_context * ctx, llama_token_data_array * candidates, const struct llama_grammar * grammar) {
    (void) ctx; (void) candidates; (void) grammar;
}
void llama_grammar_accept_token(struct llama_context * ctx, struct llama_grammar * grammar, llama_token token) {
    (void) ctx; (void) grammar; (void) token;
}

}   // extern "C"
