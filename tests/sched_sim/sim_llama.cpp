// tests/sched_sim/sim_llama.cpp -- TEST INFRASTRUCTURE: a model of the engine under the scheduler, on a virtual clock.
//
// The scheduler (pipeinfer_b200/csrc/scheduler.cu) is written against the llama.h surface only.  This file implements
// that surface without a GPU so the scheduler object file itself (the one that ships in libpipeinfer_b200.so) can be
// driven on the CPU:
//   * time is virtual: llama_time_us() returns a clock that only the calls below advance (host cost per call, waiting
//     for a run); a target decode occupies the S stages of a pipeline one after the other, runs queue behind each other
//     per stage, a cancelled run stops at the next kernel (it takes almost no time from the cancel on); the
//     draft decodes synchronously and, if asked, steals GPU time from stage 0 (same GPU, higher priority);
//   * KV bookkeeping is the real one: the same op-log front end and host mirror the engine uses (kv_meta.cu), so "no
//     slot" failures happen where they would;
//   * the "model" is a hash of the context a row can SEE through the KV metadata (cells that share a sequence id with
//     the row and hold a position <= the row's, the predicate of llama.cpp:5200-5240): next = H(tokens at positions
//     0..pos).  If the scheduler's sequence juggling ever shows a row a wrong, incomplete or duplicated context, that
//     row's prediction differs from the true greedy continuation and the generated text differs from sim_truth().
//     The draft predicts the same token with probability `agree`, else another one.
// Nothing here is product code and nothing in the product calls it.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <deque>
#include <vector>

#include <cuda_runtime.h>      // types only (cudaStream_t in kv_meta.cuh); no CUDA call is made

#include "../../pipeinfer_b200/csrc/kv_meta.cuh"
#include "../../include/pipeinfer_llama.h"

// ---- symbols the two product objects expect from the rest of the library ------------------------------------------------
static char g_err[512];
void pib_set_error(const char * fmt, ...) { va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap); }
bool pib_pdl_enabled() { return false; }
void pib_kvtest_ship_decode(pib_kvtest *, int, const int32_t *, const uint32_t *) {}
extern "C" const char * sim_last_error(void) { return g_err; }

// ---- configuration and global state ---------------------------------------------------------------------------------------
struct sim_config {
    int32_t n_stages;            // pipeline depth of the target
    double  stage_us, stage_tok_us;   // a stage takes stage_us + stage_tok_us * n_tokens
    double  hop_us;              // activations to the next stage
    double  ret_us;              // sampled rows back to the head
    double  skip_us;             // a stage of a cancelled run
    double  draft_us, draft_tok_us;   // one synchronous draft decode
    double  enqueue_us;          // host time to issue a target decode
    double  call_us;             // host time of any other API call
    double  agree;               // probability that the draft's first candidate is the target's token
    double  p_top, p_second;     // probabilities the draft reports for its first / second candidate
    double  jitter;              // +- relative noise on every duration (deterministic in seed)
    double  share_gpu0;          // 0..1: how much of a draft decode's duration stage-0 work of the target loses (same GPU, the draft has priority)
    int32_t n_vocab;
    uint64_t seed;
    double  spec_issue_extra_us;     // extra host time to issue a speculative target run / a non-speculative one (the state of the
    double  nonspec_issue_extra_us;  // library BEFORE the host KV-mirror fix: ~900 and ~500 us)
    double  head_frac, tail_frac;    // share of the target's layers on the first / last stage (0 = all stages equal); stage_us is the MEAN
};
// defaults: fitted to the 8-GPU runs of the 70B / 1.1B pair in profiles/ (plain 89, classic 182, PipeInfer 206 / 204 / 184 / 163
// tok/s at --draft 3 / 2 / 5 / 8): the model lands within 10 % of each
static const sim_config DEFAULTS = { 8, 1290, 60, 20, 30, 150, 710, 0, 150, 1.5, 0.8, 0.9, 0.05, 0, 0.0, 32000, 1, 0, 0, 0, 0 };
static sim_config G = DEFAULTS;
static double g_now = 0;
static uint64_t g_rng = 88172645463325252ull;
static std::vector<std::pair<double, double>> g_draft_busy;      // GPU-0 intervals taken by the draft
static long g_bad_context = 0, g_read_cancelled = 0;

// optional log of what the scheduler asked of the TARGET context (env SIM_TRACE=<file>): issue order, batch contents, finishes
// and cancels -- the logical schedule, comparable between two schedulers driven with the same clock
static FILE * g_trace = nullptr;
static void trace_open() {
    if (g_trace) { fclose(g_trace); g_trace = nullptr; }
    if (getenv("SIM_TRACE")) g_trace = fopen(getenv("SIM_TRACE"), "w");
}

static uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33; return x; }
static double noise() {
    if (G.jitter <= 0) return 1.0;
    g_rng ^= g_rng << 13; g_rng ^= g_rng >> 7; g_rng ^= g_rng << 17;
    return 1.0 + G.jitter * (2.0 * ((g_rng >> 11) * (1.0 / 9007199254740992.0)) - 1.0);
}

// the target's greedy token after a context with hash h
static llama_token target_next(uint64_t h) { return 3 + (llama_token) (mix(h ^ G.seed) % (uint64_t) (G.n_vocab - 3)); }
static uint64_t push_tok(uint64_t h, llama_token t) { return mix(h * 0x9E3779B97F4A7C15ull + (uint64_t) (uint32_t) t + 1); }

struct sim_row { llama_token_data top[8]; bool valid; bool whole; int32_t pos, n_seen, first_gap, seq0, batch_id; std::vector<int64_t> owners; };
struct sim_run {
    int32_t batch_id = 0, n_tokens = 0;
    uint32_t cell0 = 0;             // its KV cells are [cell0, cell0 + n_tokens)
    int64_t ticket = 0;
    double  t_issue = 0, t_cancel = -1, t_done = 0;
    bool    skipped = false;
    std::vector<sim_row> rows;
    std::vector<double>  dur;       // per stage, drawn at issue time
};

struct llama_model { bool is_target = false; };
struct llama_context {
    llama_model           model;
    bool                  is_target = false;
    KvMirror              kv;
    std::vector<KvOp>     pending, devops;
    std::vector<int32_t>  cell_tok;
    std::vector<int64_t>  cell_owner;        // ticket of the decode that wrote the cell
    int64_t               tickets = 0;
    double                draft_free = 0;
    std::vector<int64_t>  skipped;           // tickets of runs that were cancelled before they were through: their K/V is garbage    // the draft's own stream: its decodes do not overlap
    std::deque<sim_run>   flight;            // oldest at the front
    std::vector<double>   stage_free;        // when each stage finished its last retired run
    std::vector<sim_row>  rows;              // rows of the last finished decode
    long n_decodes = 0, n_failed = 0, n_cancelled_runs = 0;
};

// ---- the model: what a row predicts from the context the KV metadata lets it see ------------------------------------------
static void predict_rows(llama_context * c, const llama_batch & b, std::vector<sim_row> & rows) {
    const KvMirror & kv = c->kv;
    const int W = kv.n_words;
    rows.assign(b.n_tokens, sim_row());
    std::vector<uint32_t> want(W);
    std::vector<std::pair<int32_t, int32_t>> seen;      // (position, token)
    std::vector<int64_t> owners;
    for (int i = 0; i < b.n_tokens; i++) {
        std::fill(want.begin(), want.end(), 0u);
        for (int j = 0; j < b.n_seq_id[i]; j++) want[b.seq_id[i][j] >> 5] |= 1u << (b.seq_id[i][j] & 31);
        seen.clear();
        owners.clear();
        for (uint32_t cell = 0; cell < kv.size; cell++) {
            if (kv.pos[cell] < 0 || kv.pos[cell] > b.pos[i]) continue;
            const uint32_t * bits = &kv.bits[(size_t) cell * W];
            bool shares = false;
            for (int w = 0; w < W && !shares; w++) shares = (bits[w] & want[w]) != 0;
            if (shares) { seen.push_back({ kv.pos[cell], c->cell_tok[cell] }); owners.push_back(c->cell_owner[cell]); }
        }
        std::sort(seen.begin(), seen.end());
        uint64_t h = 0x1234;
        bool     whole = (int) seen.size() == b.pos[i] + 1;
        int32_t first_gap = -1;
        for (size_t k = 0; k < seen.size(); k++) {
            if (seen[k].first != (int32_t) k) { whole = false; if (first_gap < 0) first_gap = (int32_t) k; }
            h = push_tok(h, seen[k].second);
        }
        if (!whole) { g_bad_context++; h = mix(h ^ 0xBADC0DEull); }      // a hole or a duplicate: the prediction is garbage
        const llama_token truth = target_next(h);
        llama_token other = target_next(h ^ 0x5555AAAA5555ull);
        if (other == truth) other = 3 + (truth - 3 + 1) % (G.n_vocab - 3);
        sim_row & r = rows[i];
        r.valid = true; r.whole = whole; r.pos = b.pos[i]; r.n_seen = (int32_t) seen.size(); r.first_gap = first_gap; r.seq0 = b.seq_id[i][0]; r.batch_id = b.batch_id;
        std::sort(owners.begin(), owners.end());
        owners.erase(std::unique(owners.begin(), owners.end()), owners.end());
        r.owners = owners;
        const bool agrees = c->is_target || (mix(h ^ 0x0DDBA11ull ^ G.seed) % 1000000) < (uint64_t) (G.agree * 1e6);
        for (int k = 0; k < 8; k++) {
            r.top[k].id = k == 0 ? (agrees ? truth : other) : k == 1 ? (agrees ? other : truth) : 3 + (llama_token) ((uint64_t) (truth + 7919 * k) % (uint64_t) (G.n_vocab - 3));
            r.top[k].p  = (float) (c->is_target ? (k == 0 ? 1.0 : 0.0) : k == 0 ? G.p_top : k == 1 ? G.p_second : (1.0 - G.p_top - G.p_second) / 6.0);
            r.top[k].logit = logf(std::max(r.top[k].p, 1e-30f));
        }
    }
}

// ---- the pipeline's timeline ---------------------------------------------------------------------------------------------
static double run_on_gpu0(double start, double need) {      // stage-0 work that yields (partly) to the draft's intervals
    if (G.share_gpu0 <= 0) return start + need;
    const double rate = std::max(0.0, 1.0 - G.share_gpu0);    // the target's speed while a draft decode runs on the same GPU
    double t = start;
    for (const auto & iv : g_draft_busy) {
        if (iv.second <= t) continue;
        if (iv.first >= t + need) break;
        if (iv.first > t) { need -= iv.first - t; t = iv.first; }
        const double span = iv.second - t;
        if (rate > 0 && need <= span * rate) return t + need / rate;
        need -= span * rate;
        t = iv.second;
    }
    return t + need;
}

// (re)computes t_done of every run in flight; with `retire` also advances stage_free past the oldest run
static void resolve(llama_context * c, bool retire) {
    std::vector<double> free_at = c->stage_free;
    bool first = true;
    for (sim_run & r : c->flight) {
        double t = r.t_issue;
        for (int s = 0; s < G.n_stages; s++) {
            const double start = std::max(t, free_at[s]);
            // cancellation is kernel-granular in the engine (every kernel of the run tests the flag first): a stage that has not
            // started is skipped, a stage in progress stops at the next kernel; either way the run's KV cells are not all written
            double d = r.dur[s];
            if (r.t_cancel >= 0 && start + d > r.t_cancel) {
                d = std::max(0.0, r.t_cancel - start) + G.skip_us;
                r.skipped = true;
            }
            const double end   = s == 0 ? run_on_gpu0(start, d) : start + d;
            free_at[s] = end;
            t = end + (s + 1 < G.n_stages ? G.hop_us : G.ret_us);
        }
        r.t_done = t;
        if (first && retire) c->stage_free = free_at;
        first = false;
    }
}

static void fill_bits(const KvMirror & kv, const llama_batch & b, std::vector<uint32_t> & tb) {
    tb.assign((size_t) b.n_tokens * kv.n_words, 0u);
    for (int i = 0; i < b.n_tokens; i++)
        for (int j = 0; j < b.n_seq_id[i]; j++) tb[(size_t) i * kv.n_words + (b.seq_id[i][j] >> 5)] |= 1u << (b.seq_id[i][j] & 31);
}

// issue: KV ops so far take effect, a slot is found, the rows are predicted from what is visible NOW (the engine replays
// the metadata in stream order, so a decode sees the bookkeeping as of its issue)
static bool issue(llama_context * c, const llama_batch & b, sim_run & run) {
    c->n_decodes++;
    kvlog_flush(c->kv, c->pending, c->devops);
    c->devops.clear();
    std::vector<uint32_t> tb;
    fill_bits(c->kv, b, tb);
    if (!kvlog_decode_begin(c->kv, c->devops, b.n_tokens, b.pos, tb.data())) { c->n_failed++; return false; }
    run.cell0 = c->kv.head; run.ticket = ++c->tickets;
    for (int i = 0; i < b.n_tokens; i++) { c->cell_tok[c->kv.head + i] = b.token[i]; c->cell_owner[c->kv.head + i] = run.ticket; }
    predict_rows(c, b, run.rows);
    kvlog_decode_end(c->kv, b.n_tokens);
    c->devops.clear();
    run.batch_id = b.batch_id; run.n_tokens = b.n_tokens;
    if (g_trace && c->is_target) {
        fprintf(g_trace, "D %d %d", b.batch_id, b.n_tokens);
        for (int i = 0; i < b.n_tokens; i++) {
            uint64_t hs = 0;
            for (int j = 0; j < b.n_seq_id[i]; j++) hs = mix(hs ^ (uint64_t) (b.seq_id[i][j] + 1));
            fprintf(g_trace, " %d@%d/%d:%d:%04x", b.token[i], b.pos[i], b.n_seq_id[i], b.seq_id[i][0], (unsigned) (hs & 0xffff));
        }
        fprintf(g_trace, " # t=%.0f cell=%u\n", g_now, run.cell0);
    }
    return true;
}

// ---- the llama.h surface the scheduler uses --------------------------------------------------------------------------------
extern "C" {

int64_t llama_time_us(void) { return (int64_t) g_now; }
const struct llama_model * llama_get_model(const struct llama_context * ctx) { return &ctx->model; }
llama_token llama_token_eos(const struct llama_model *) { return 2; }       // never predicted: tokens start at 3
int  llama_pib_kv_words(const struct llama_context * ctx) { return ctx->kv.n_words; }
void llama_pib_set_logit_bias(struct llama_context *, const llama_token *, const float *, int) { g_now += G.call_us; }

void llama_kv_cache_seq_rm(struct llama_context * ctx, llama_seq_id seq, llama_pos p0, llama_pos p1) { kvlog_push(ctx->pending, { KVOP_RM, seq, 0, p0, p1, 0 }); g_now += G.call_us * 0.02; }
void llama_kv_cache_seq_cp(struct llama_context * ctx, llama_seq_id src, llama_seq_id dst, llama_pos p0, llama_pos p1) { if (src == dst) return; kvlog_push(ctx->pending, { KVOP_CP, src, dst, p0, p1, 0 }); g_now += G.call_us * 0.02; }
void llama_kv_cache_seq_keep(struct llama_context * ctx, llama_seq_id seq) { kvlog_push(ctx->pending, { KVOP_KEEP, seq, 0, 0, 0, 0 }); g_now += G.call_us * 0.02; }

void llama_pib_kv_seq_rm_range(struct llama_context * ctx, llama_seq_id first, llama_seq_id last, llama_pos p0, llama_pos p1) { kvlog_push_ranged(ctx->pending, KVOP_RM, first, last, 0, p0, p1); g_now += G.call_us * 0.02; }
void llama_pib_kv_seq_cp_range(struct llama_context * ctx, llama_seq_id src, llama_seq_id d0, llama_seq_id d1, llama_pos p0, llama_pos p1) { kvlog_push_ranged(ctx->pending, KVOP_CP, src, d0, d1, p0, p1); g_now += G.call_us * 0.02; }
void llama_pib_kv_seq_sweep(struct llama_context * ctx, llama_seq_id first, llama_seq_id last, llama_pos p0, llama_pos p1) { kvlog_push_ranged(ctx->pending, KVOP_SWEEP, first, last, 0, p0, p1); g_now += G.call_us * 0.02; }

struct ggml_cgraph * llama_start_async_decode(struct llama_context & c, struct llama_batch b) {
    g_now += G.enqueue_us * noise();
    if (c.is_target) g_now += (b.n_tokens == 1 && b.n_seq_id[0] == 1 && b.seq_id[0][0] == 0) ? G.nonspec_issue_extra_us : G.spec_issue_extra_us;
    sim_run run;
    const uint32_t head_before = c.kv.head;
    (void) head_before;
    if (!issue(&c, b, run)) return nullptr;
    run.t_issue = g_now;
    if (c.is_target) {
        run.dur.resize(G.n_stages);
        for (int s = 0; s < G.n_stages; s++) {
            double w = 1.0;                                   // this stage's share of the layers, relative to an even split
            if (G.n_stages > 1 && G.head_frac > 0) {
                const double tail = G.tail_frac > 0 ? G.tail_frac : (1.0 - G.head_frac) / (G.n_stages - 1);
                const double mid  = G.n_stages > 2 ? (1.0 - G.head_frac - tail) / (G.n_stages - 2) : 0.0;
                w = (s == 0 ? G.head_frac : s == G.n_stages - 1 ? tail : mid) * G.n_stages;
            }
            run.dur[s] = (G.stage_us + G.stage_tok_us * b.n_tokens) * w * noise();
        }
    } else {                                                    // the draft runs at once, on GPU 0, and the caller waits
        const double d = (G.draft_us + G.draft_tok_us * b.n_tokens) * noise(), t0 = std::max(g_now, c.draft_free);
        g_draft_busy.push_back({ t0, t0 + d });
        run.t_done = c.draft_free = t0 + d;
    }
    c.flight.push_back(std::move(run));
    return (struct ggml_cgraph *) (uintptr_t) (0x1000 + c.flight.back().batch_id);
}

int llama_finish_async_decode(struct llama_context & c, struct llama_batch b, struct ggml_cgraph * handle) {
    if (!handle || c.flight.empty()) return -1;
    if (c.is_target) resolve(&c, true);
    sim_run & r = c.flight.front();                             // runs finish in issue order
    if (r.batch_id != b.batch_id) { pib_set_error("sim: finish of batch %d but the oldest run is %d", b.batch_id, r.batch_id); return -1; }
    g_now = std::max(g_now, r.t_done) + G.call_us;
    if (g_trace && c.is_target) fprintf(g_trace, "F %d%s # t=%.0f\n", r.batch_id, r.t_cancel >= 0 ? " cancelled" : "", g_now);
    c.rows = std::move(r.rows);
    if (r.t_cancel >= 0) for (sim_row & row : c.rows) row.valid = false;
    if (r.skipped) c.skipped.push_back(r.ticket);
    // a row that attended to cells of a run which never computed them (cancelled ahead of it in the stream) is garbage
    for (sim_row & row : c.rows)
        for (int64_t o : row.owners)
            if (std::binary_search(c.skipped.begin(), c.skipped.end(), o) && row.whole) {
                row.whole = false; row.first_gap = -2; g_bad_context++;
                for (int k = 0; k < 8; k++) row.top[k].id = 3 + (llama_token) (mix((uint64_t) row.top[k].id * 31 + (uint64_t) o) % (uint64_t) (G.n_vocab - 3));
            }
    c.flight.pop_front();
    while (!g_draft_busy.empty() && c.is_target && g_draft_busy.front().second < c.stage_free[0] - 1e6) g_draft_busy.erase(g_draft_busy.begin());
    return 1;
}

int llama_decode(struct llama_context * ctx, struct llama_batch b) {
    struct ggml_cgraph * h = llama_start_async_decode(*ctx, b);
    if (!h) return -1;
    return llama_finish_async_decode(*ctx, b, h) == 1 ? 0 : -1;
}

bool llama_mpi_iprobe(struct llama_context * c) {
    g_now += std::max(G.call_us, 0.2);        // polling takes time even when the other calls are modelled as free: a scheduler that spins on it must see the clock move
    if (c->flight.empty()) return false;
    if (c->is_target) resolve(c, false);
    return c->flight.front().t_done <= g_now;
}

void llama_cancel_run(struct llama_context * c, int32_t * ids, int count) {
    g_now += G.call_us;
    if (c->is_target) resolve(c, false);
    if (g_trace && c->is_target) { fprintf(g_trace, "C"); for (int i = 0; i < count; i++) fprintf(g_trace, " %d", ids[i]); fprintf(g_trace, " # t=%.0f\n", g_now); }
    for (sim_run & r : c->flight)
        for (int i = 0; i < count; i++)
            if (r.batch_id == ids[i] && r.t_cancel < 0) {
                r.t_cancel = g_now;
                c->n_cancelled_runs++;
                // a run that is cancelled before it is through leaves garbage K/V behind in the layers it skipped: whoever
                // still sees these cells afterwards predicts garbage
                if (r.t_done > g_now)
                    for (int k = 0; k < r.n_tokens; k++)
                        if (c->cell_owner[r.cell0 + k] == r.ticket) c->cell_tok[r.cell0 + k] = (int32_t) (-100 - (r.ticket * 64 + k) % 1000000);
            }
}

// the virtual engine has no device to walk on: the scheduler falls back to walking the sampled rows
void    llama_pib_set_accept_table(struct llama_context *, int32_t, int32_t, int32_t, const int32_t *, const int32_t *, const int32_t *, int32_t) {}
int32_t llama_pib_get_accepted(struct llama_context *, int32_t *, int32_t, int32_t *) { return -1; }

int llama_pib_get_top8(struct llama_context * c, int32_t i, llama_token_data * out8) {
    g_now += G.call_us;
    if (i < 0 || i >= (int) c->rows.size()) return -1;
    if (!c->rows[i].valid) g_read_cancelled++;
    if (!c->rows[i].whole && getenv("SIM_DEBUG")) fprintf(stderr, "sim: t=%.0f %s row %d read with a broken context (batch %d, seq %d, pos %d, saw %d cells, first gap/duplicate at %d)\n", g_now, c->is_target ? "target" : "draft", i, c->rows[i].batch_id, c->rows[i].seq0, c->rows[i].pos, c->rows[i].n_seen, c->rows[i].first_gap);
    memcpy(out8, c->rows[i].top, sizeof(c->rows[i].top));
    return 0;
}

// ---- the harness's own entry points ------------------------------------------------------------------------------------------
void sim_configure(const sim_config * cfg) { trace_open(); G = *cfg; g_now = 0; g_rng = 88172645463325252ull ^ mix(cfg->seed); g_draft_busy.clear(); g_bad_context = 0; g_read_cancelled = 0; }
void sim_default_config(sim_config * cfg) { *cfg = DEFAULTS; }
double sim_now_us(void) { return g_now; }

struct llama_context * sim_new_context(uint32_t n_ctx, int32_t max_seq, int is_target) {
    llama_context * c = new llama_context();
    c->is_target = is_target != 0;
    c->kv.init(n_ctx, max_seq);
    c->cell_tok.assign(n_ctx, -1);
    c->cell_owner.assign(n_ctx, 0);
    c->stage_free.assign(G.n_stages > 0 ? G.n_stages : 1, 0.0);
    return c;
}
void sim_free_context(struct llama_context * c) { delete c; }

// counters: decodes issued, decodes refused (no KV slot), runs cancelled, rows predicted from a broken context, rows of a
// cancelled run that the scheduler read, cells in use
void sim_stats(const struct llama_context * c, int64_t * out6) {
    if (g_trace) fflush(g_trace);
    out6[0] = c->n_decodes; out6[1] = c->n_failed; out6[2] = c->n_cancelled_runs; out6[3] = g_bad_context; out6[4] = g_read_cancelled; out6[5] = c->kv.used;
}

// the true greedy continuation of a prompt under the model above
void sim_truth(const llama_token * prompt, int n_prompt, llama_token * out, int n_out) {
    uint64_t h = 0x1234;
    for (int i = 0; i < n_prompt; i++) h = push_tok(h, prompt[i]);
    for (int i = 0; i < n_out; i++) { out[i] = target_next(h); h = push_tok(h, out[i]); }
}

}   // extern "C"
