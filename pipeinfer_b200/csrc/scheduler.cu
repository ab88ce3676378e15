// scheduler.cu -- PipeInfer's continuous asynchronous speculation scheduler (host, integer logic) on top of the
// llama.h-compatible ABI of this library.
//
// Behavioural restatement of the reference's examples/speculative/speculative.cpp (S1-S6 of SURVEY.md §8a), re-organised as
// one class; every step cites the reference lines whose behaviour it reproduces.  What differs by design:
//   * both pipeline heads live in this process, so the MPI_Bcast lock-step of the two head ranks (llama_sync_token /
//     llama_sync_token_data / llama_swap_comm, speculative.cpp:323-326, 454, 973-983) disappears;
//   * sampling reads 8 x {id, logit, p} per row produced on the device (sampling.cu) instead of sorting n_vocab floats
//     on the host (common/sampling.cpp:102-216); greedy only (--temp <= 0, --repeat-last-n 0: the configuration of the
//     authors' launch scripts, examples/speculative/scripts/*.sh);
//   * the target runs asynchronously on its stream (or pipeline), the draft synchronously on its own stream.
#include "engine.cuh"
#include "../../include/pipeinfer_llama.h"
#include "../../include/pipeinfer_sched.h"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <deque>
#include <vector>

namespace {

struct Draft {                       // speculative.cpp:16-28 seq_draft
    bool active = false, drafting = false, skip = false;
    int  i_batch_dft = 0;
    std::vector<int>         i_batch_tgt;
    std::vector<llama_token> tokens, prefix_tokens;
};

struct Batch {                       // an owned llama_batch (tokens, positions, per-token sequence-id lists)
    std::vector<llama_token>               token;
    std::vector<llama_pos>                 pos;
    std::vector<std::vector<llama_seq_id>> seq;
    int32_t                                batch_id = 0;
    std::vector<int32_t>        nseq_;
    std::vector<llama_seq_id *> ptr_;
    std::vector<int8_t>         lg_;
    void clear() { token.clear(); pos.clear(); seq.clear(); }
    int  size() const { return (int) token.size(); }
    void add(llama_token t, llama_pos p, std::vector<llama_seq_id> s) { token.push_back(t); pos.push_back(p); seq.push_back(std::move(s)); }
    llama_batch view(int max_n_seq) {
        const int n = size();
        nseq_.resize(n); ptr_.resize(n); lg_.assign(n, 1);
        for (int i = 0; i < n; i++) { nseq_[i] = (int32_t) seq[i].size(); ptr_[i] = seq[i].data(); }
        llama_batch b;
        b.n_tokens = n; b.token = token.data(); b.embd = nullptr; b.pos = pos.data(); b.n_seq_id = nseq_.data(); b.seq_id = ptr_.data();
        b.logits = lg_.data(); b.all_pos_0 = 0; b.all_pos_1 = 0; b.all_seq_id = 0; b.batch_id = batch_id; b.max_n_seq = max_n_seq;
        return b;
    }
};

struct Run {                         // speculative.cpp:30-43 seq_async_run
    struct ggml_cgraph * handle = nullptr;
    Batch              batch;
    std::vector<Draft> drafts;
    int  n_past_tgt = 0, prefix_n_past_tgt = 0, n_past_dft = 0, s_keep = 0, seq_offset = 0, n_past_max = 0;
    bool speculative = false, canceled = false;
};

class Scheduler {
  public:
    Scheduler(llama_context * tgt, llama_context * dft, const pib_spec_params & p) : ctx_tgt(tgt), ctx_dft(dft), P(p) {
        np      = std::max(1, P.n_parallel);
        max_seq = 1000 * np + 1;                                           // :140-145
        for (int i = 0; i < 1000; i++) free_offsets.push_back(i * np + 1);
        drafts.assign(np, Draft());
    }

    int run(const llama_token * prompt, int n_prompt, llama_token * out, int out_cap, pib_spec_result * res, double * itl_us, int itl_cap);

  private:
    llama_context * ctx_tgt, * ctx_dft;
    pib_spec_params P;
    int np, max_seq;
    std::deque<int>    free_offsets;
    std::vector<Draft> drafts;
    std::deque<Run>    tgt_runs;          // push_front new, back() oldest  (:1262 / :331)
    std::vector<llama_token> generated;
    int32_t batch_id = 0;
    int n_predict = 0, n_drafted = 0, n_accept = 0, n_rejected = 0, n_cancel = 0, n_spec_runs = 0;
    int n_past_tgt = 0, n_past_dft = 0, spec_past_tgt = 0, spec_past_dft = 0;
    bool first_run = true, failed = false;
    llama_token id = 0;
    Run current;                          // the run whose logits are being verified (only the scalar fields are used)
    Batch batch_tgt, batch_dft;

    // optional event trace (env PIB_TRACE=<file>): time in us since the start of the generation, event, a, b, c
    struct Ev { int64_t t; const char * what; int a, b, c; };
    std::vector<Ev> trace;
    bool            tracing = getenv("PIB_TRACE") != nullptr;
    // PIB_REFERENCE_HAZARDS=1 restores the two behaviours of the reference that can make the text differ from the greedy
    // text (see check_for_cancel and speculation_loop); for A/B runs only
    bool            reference_hazards = getenv("PIB_REFERENCE_HAZARDS") != nullptr && atoi(getenv("PIB_REFERENCE_HAZARDS")) != 0;
    // Scheduler policy (PIB_SCHED_CHAIN; default 2 since round 2: measured on the GPU, 98 -> 146 accepted tokens/s on one B200 with
    // the 70B / 1.1B pair, text identical; PIB_SCHED_CHAIN=0 is the reference's policy, event for event -- tests/test_scheduler_sim.py).
    // 1: keep the chain of speculative runs alive across a verification.  The
    // reference restarts drafting from the last verified token after EVERY run it verifies (speculative.cpp:392-394), so when
    // the draft was right the runs already in flight are drafted and verified a second time and later dropped as stale.  With
    // this policy drafting resumes at the tip of the chain whenever the newest speculative run in flight is still consistent
    // with what has been generated; a rejection restarts the chain as before.  The text is the greedy text either way.
    // PIB_SCHED_CHAIN=2 additionally skips, when the chain is resumed, the non-speculative run of the token just sampled (on
    // both models): the next link in flight starts with exactly that token, so its first row yields the same logits and its
    // KV cell is committed along with that link's accepted tokens.
    // PIB_SCHED_CHAIN=3 additionally merges, when the chain has to be restarted, the target's non-speculative run of the sampled
    // token into the first speculative run (the token becomes row 0 of that run, as in classic speculation): one pass of the
    // target instead of two, which is what matters when there is no pipeline to hide the second one (one or two GPUs).
    int             chain_policy = getenv("PIB_SCHED_CHAIN") ? atoi(getenv("PIB_SCHED_CHAIN")) : 2;
    // acceptance walked on the device that holds the logits (PIB_DEVICE_WALK=0: the host walks the sampled rows, 96 bytes each)
    bool            device_walk = !(getenv("PIB_DEVICE_WALK") && atoi(getenv("PIB_DEVICE_WALK")) == 0);
    int             chain_start_tgt = 0, chain_start_dft = 0;      // n_past at which the current chain began: prefix of all its links
    llama_token_data tip_top8[8];                                   // the draft's candidates after the tip of branch 0
    bool            have_tip = false;
    struct Tip { std::vector<Draft> drafts; int spec_past_tgt = 0, spec_past_dft = 0, seq_offset = 0; llama_token id = 0; bool first_run = true; } tip;
    int             n_chain_resumed = 0;
    void save_tip() { tip.drafts = drafts; tip.spec_past_tgt = spec_past_tgt; tip.spec_past_dft = spec_past_dft; tip.seq_offset = current.seq_offset; tip.id = id; tip.first_run = first_run; }
    bool chain_alive() const;
    bool next_link_starts_with(llama_token tok) const;
    bool resume_chain(int verified_offset, bool verified_spec);
    bool merge_first = false;                              // policy 3: the next first link carries the sampled token itself
    int  pending_pos_tgt = -1, pending_pos_dft = -1;       // policy 2: position of a sampled token that lives only in the next link's partition
    int64_t         t_trace0 = 0;
    void ev(const char * what, int a = 0, int b = 0, int c = 0) { if (tracing) trace.push_back({ llama_time_us() - t_trace0, what, a, b, c }); }
    void dump_trace() {
        if (!tracing) return;
        FILE * f = fopen(getenv("PIB_TRACE"), "a");
        if (!f) return;
        fprintf(f, "# generation\n");
        for (const Ev & e : trace) fprintf(f, "%lld %s %d %d %d\n", (long long) e.t, e.what, e.a, e.b, e.c);
        fclose(f);
    }

    static llama_token greedy(llama_context * c, int row) { llama_token_data t[8]; llama_pib_get_top8(c, row, t); return t[0].id; }

    void check_for_cancel();
    void begin_async_run(llama_context * ctx, std::deque<Run> & q, const std::vector<Draft> & dr, int & n_past, int n_past_dft_arg,
                         bool is_spec, Batch & batch, int n_past_max, int prefix_n_past, int seq_offset);
    void begin_non_spec_run(llama_context * ctx, std::deque<Run> & q, int & n_past, int n_past_dft_arg);
    bool start_spec_run(int prefix_n_past, int n_past_dft_arg, int orig_offset, float p_adjust);
    void speculation_loop();
    float p_adjust_of(int iter) const { return iter * P.p_recovery - std::max(n_rejected * P.p_decay, 0.0f); }   // :846-848
};

// :1277-1359 -- a run is cancelled when it is entirely stale (n_past_max < n_past_tgt) or when prefix_tokens ++ tokens of
// every non-empty branch diverges from what has been generated since the run's prefix position
void Scheduler::check_for_cancel() {
    std::vector<int32_t> ids;
    for (Run & run : tgt_runs) {
        if (run.canceled) continue;
        bool correct_prefix = true;
        if (run.speculative && n_past_tgt >= run.prefix_n_past_tgt) {
            for (int d = np - 1; d >= 0; d--) {
                if (run.drafts[d].tokens.empty()) continue;
                correct_prefix = true;
                std::vector<llama_token> cat = run.drafts[d].prefix_tokens;
                cat.insert(cat.end(), run.drafts[d].tokens.begin(), run.drafts[d].tokens.end());
                size_t di = 0, gi = (size_t) run.prefix_n_past_tgt;
                while (gi < generated.size() && di < cat.size() && generated.size() > (size_t) run.prefix_n_past_tgt) {
                    if (generated[gi] != cat[di]) { correct_prefix = false; break; }
                    di++;
                    gi = (size_t) run.prefix_n_past_tgt + di;
                }
                if (correct_prefix) run.s_keep = d;
            }
        }
        if (run.n_past_max < n_past_tgt || !correct_prefix) {
            // A run that is merely stale (everything it covers has been verified by other runs, its own draft tokens were
            // right) is dropped from the bookkeeping but NOT stopped on the GPUs: a run chained behind it attends to its
            // KV cells, and cells of a run that was stopped before it got through every stage hold garbage.  Only runs with
            // a wrong prefix are stopped -- whatever is chained behind those has the same wrong prefix.  The reference
            // stops both kinds (speculative.cpp:1344-1356); tests/test_scheduler_sim.py shows the text that costs.
            if (run.speculative && (!correct_prefix || reference_hazards)) ids.push_back(run.batch.batch_id);
            run.canceled = true;
        }
    }
    if (!ids.empty()) { n_cancel += (int) ids.size(); llama_cancel_run(ctx_tgt, ids.data(), (int) ids.size()); ev("cancel", (int) ids.size(), ids[0]); }
}

// :1182-1275
void Scheduler::begin_async_run(llama_context * ctx, std::deque<Run> & q, const std::vector<Draft> & dr, int & n_past, int n_past_dft_arg,
                                bool is_spec, Batch & batch, int n_past_max, int prefix_n_past, int seq_offset) {
    batch_id++;
    Run run;
    run.seq_offset = seq_offset;
    run.batch = batch;
    run.batch.batch_id = batch_id;
    run.drafts = dr;
    run.n_past_tgt = n_past; run.prefix_n_past_tgt = prefix_n_past; run.n_past_max = n_past_max; run.n_past_dft = n_past_dft_arg;
    run.speculative = is_spec;
    if (!is_spec) llama_pib_kv_seq_rm_range(ctx, 0, max_seq, n_past, n_past + 1);     // :1252-1255, for i in 0..max_seq: seq_rm(i, ...)
    else          { for (int i = 0; i < np; i++) llama_kv_cache_seq_rm(ctx, i + seq_offset, n_past, n_past + 1); } // :1256-1260
    if (ctx == ctx_tgt && device_walk && np <= 8) {
        // the acceptance table of this batch, indexed as the verification walk will see the drafts (first token of every branch
        // erased, :431): the last stage's device walks it right after sampling and only the ids sampled on the way come back
        int32_t len[8], rows[8][17], toks[8][17];
        bool fits = true;
        memset(rows, 0, sizeof(rows)); memset(toks, 0, sizeof(toks));
        for (int s = 0; s < np && fits; s++) {
            const int nt = (int) dr[s].tokens.size() - 1, nr = (int) dr[s].i_batch_tgt.size();
            len[s] = dr[s].active ? nt : -1;                                   // an inactive branch is never compared, but may be the one sampled from
            if (nt > 16 || nr > 17) { fits = false; break; }
            for (int i = 0; i < nr; i++) rows[s][i] = dr[s].i_batch_tgt[i];
            for (int i = 0; i < nt; i++) toks[s][i] = dr[s].tokens[i + 1];
        }
        if (fits) llama_pib_set_accept_table(ctx, np, run.s_keep, is_spec ? 1 : 0, len, &rows[0][0], &toks[0][0], 17);
    }
    run.handle = llama_start_async_decode(*ctx, run.batch.view(max_seq + 1));                                      // :1261
    if (ctx == ctx_tgt) ev(is_spec ? "start_spec" : "start_nonspec", batch_id, (int) run.batch.size(), n_past);
    if (!run.handle) failed = true;     // no KV slot left: the reference would spin forever here
    q.push_front(std::move(run));
    if (!is_spec) llama_pib_kv_seq_cp_range(ctx, 0, 1, max_seq, n_past, n_past + 1);  // :1264-1268, for i in 1..max_seq: seq_cp(0, i, ...)
}

// :850-879
void Scheduler::begin_non_spec_run(llama_context * ctx, std::deque<Run> & q, int & n_past, int n_past_dft_arg) {
    std::vector<Draft> ns(np);
    for (int s = 0; s < np; s++) {
        ns[s].i_batch_tgt = { 0 };
        ns[s].i_batch_dft = drafts[s].i_batch_dft;
        ns[s].tokens      = { id };
        ns[s].active = drafts[s].active; ns[s].drafting = drafts[s].drafting; ns[s].skip = drafts[s].skip;
    }
    Batch b;
    b.add(id, n_past, { 0 });
    begin_async_run(ctx, q, ns, n_past, n_past_dft_arg, false, b, n_past + 1, n_past, 0);
    n_past++;
}

// :881-1180 -- one speculative run: seed the partition's KV from sequence 0 (first run after a verification) or from the
// previous partition (chained runs), tree-draft up to n_draft tokens with synchronous draft decodes, launch the target
// verification batch asynchronously.  Returns true when speculation should stop.
bool Scheduler::start_spec_run(int prefix_n_past, int n_past_dft_arg, int orig_offset, float p_adjust) {
    const int off = current.seq_offset;
    for (int i = 0; i < np; i++) {                                                                       // :891-905
        llama_kv_cache_seq_rm(ctx_tgt, i + off, first_run ? prefix_n_past : prefix_n_past - 1, -1);
        llama_kv_cache_seq_rm(ctx_dft, i + off, first_run ? n_past_dft_arg : n_past_dft_arg - 1, -1);
        llama_kv_cache_seq_cp(ctx_tgt, first_run ? 0 : orig_offset, i + off, first_run ? prefix_n_past : prefix_n_past - 1, spec_past_tgt + 1);
        llama_kv_cache_seq_cp(ctx_dft, first_run ? 0 : orig_offset, i + off, first_run ? n_past_dft_arg : n_past_dft_arg - 1, spec_past_dft + 1);
    }
    batch_tgt.clear();
    for (int s = 0; s < np; s++) {                                                                       // :910-922
        drafts[s].active = false;
        if (!first_run) drafts[s].prefix_tokens.insert(drafts[s].prefix_tokens.end(), drafts[s].tokens.begin(), drafts[s].tokens.end());
        else drafts[s].prefix_tokens.clear();
        drafts[s].tokens.clear();
        drafts[s].i_batch_tgt.clear();
    }
    drafts[0].tokens.push_back(id);                                                                      // erased after drafting
    const int merged = merge_first && first_run ? 1 : 0;                                                 // policy 3: `id` is row 0 of this run
    if (merged) { drafts[0].i_batch_tgt.push_back(0); batch_tgt.add(id, spec_past_tgt, { off }); }
    batch_dft.clear();
    int n_seq_cur = 0, max_ran_seq = 0, n_past_cur = spec_past_dft;
    for (int s = 0; s < np; s++) { drafts[s].skip = true; drafts[s].active = false; drafts[s].drafting = false; }
    drafts[0].active = true; drafts[0].drafting = true; drafts[0].skip = false; drafts[0].i_batch_dft = 0;

    for (int i = 0; i < P.n_draft; ++i) {                                                                // :957-1104
        batch_dft.clear();
        for (int s = 0; s <= max_ran_seq; ++s) {
            if (!drafts[s].drafting || drafts[s].skip) continue;
            llama_token_data cur_p[8];
            if (chain_policy && !first_run && have_tip && i == 0 && s == 0) memcpy(cur_p, tip_top8, sizeof(cur_p));
            else llama_pib_get_top8(ctx_dft, drafts[s].i_batch_dft, cur_p);                              // :975-983
            if (cur_p[0].p < P.p_accept + p_adjust) { drafts[s].drafting = false; continue; }           // :998-1003
            std::vector<int> sa(1, s);
            for (int f = 1; f < 8; ++f) {                                                                // :1009-1050 branch split
                if (n_seq_cur < np - 1 && cur_p[f].p > P.p_split + p_adjust) {
                    n_seq_cur++;
                    llama_kv_cache_seq_rm(ctx_dft, n_seq_cur + off, n_past_dft_arg, n_past_cur);
                    llama_kv_cache_seq_cp(ctx_dft, s + off, n_seq_cur + off, n_past_dft_arg, n_past_cur);
                    for (int t = 0; t < batch_tgt.size(); ++t)                                           // earlier tokens join the new branch
                        for (size_t p = 0; p < batch_tgt.seq[t].size(); ++p)
                            if (batch_tgt.seq[t][p] == s + off) { batch_tgt.seq[t].push_back(n_seq_cur + off); break; }
                    drafts[n_seq_cur].active = true; drafts[n_seq_cur].drafting = true; drafts[n_seq_cur].skip = false;
                    drafts[n_seq_cur].tokens = drafts[s].tokens;
                    drafts[n_seq_cur].i_batch_tgt = drafts[s].i_batch_tgt;
                    sa.push_back(n_seq_cur);
                } else break;
            }
            for (int is = 0; is < (int) sa.size(); ++is) {                                               // :1054-1081
                const llama_token tok = cur_p[is].id;
                const int sq = sa[is];
                drafts[sq].tokens.push_back(tok);
                drafts[sq].i_batch_tgt.push_back(batch_tgt.size());
                batch_tgt.add(tok, spec_past_tgt + merged + i, { sq + off });
                drafts[sq].i_batch_dft = batch_dft.size();
                batch_dft.add(tok, n_past_cur, { sq + off });
                if (batch_tgt.size() > P.n_draft + merged) drafts[sq].drafting = false;
            }
        }
        if (batch_dft.size() == 0) break;
        llama_decode(ctx_dft, batch_dft.view(max_seq + 1));                                              // :1090 synchronous draft step
        ev("draft_step", (int) batch_dft.size());
        ++n_past_cur;
        ++n_drafted;
        max_ran_seq = n_seq_cur;
        batch_dft.clear();
        if (batch_tgt.size() > P.n_draft + merged) break;
    }
    if (batch_dft.size() != 0) llama_decode(ctx_dft, batch_dft.view(max_seq + 1));                       // :1107-1113

    if (batch_tgt.size() == merged) {                                                                    // :1129-1135 (nothing drafted)
        free_offsets.push_back(off);
        n_rejected++;
        return true;
    }
    for (int s = 0; s < np && !merged; ++s) {                                                            // :1149-1157
        if (!drafts[s].active) continue;
        drafts[s].tokens.erase(drafts[s].tokens.begin());                                                // (a merged run keeps `id` in front: it is in the batch)
    }
    if (first_run) ++n_drafted;
    n_spec_runs++;
    begin_async_run(ctx_tgt, tgt_runs, drafts, spec_past_tgt, n_past_dft_arg, true, batch_tgt,
                    spec_past_tgt + (int) drafts[0].tokens.size(), prefix_n_past, off);                 // :1163
    spec_past_tgt += (int) drafts[0].tokens.size();
    spec_past_dft += (int) drafts[0].tokens.size() - merged;
    id = drafts[0].tokens.back();
    first_run = false;
    merge_first = false;
    if (chain_policy) have_tip = llama_pib_get_top8(ctx_dft, drafts[0].i_batch_dft, tip_top8) == 0;   // what the next link starts from
    return false;
}

// after a verification: is the newest speculative run in flight still consistent with the text (check_for_cancel has just
// run) and does it reach beyond what is verified?  Then go on drafting from its end instead of from the verified token.
bool Scheduler::chain_alive() const {
    if (!chain_policy || tip.first_run || !have_tip) return false;
    const Run * newest = nullptr;
    for (const Run & r : tgt_runs) if (r.speculative) { newest = &r; break; }      // front() is the newest
    return newest && !newest->canceled && newest->seq_offset == tip.seq_offset && newest->n_past_max > n_past_tgt;
}

// policy 2's precondition, checked explicitly: the oldest speculative run in flight that has not been cancelled begins at the
// position the sampled token will take and its kept branch begins with that token
bool Scheduler::next_link_starts_with(llama_token tok) const {
    for (auto it = tgt_runs.rbegin(); it != tgt_runs.rend(); ++it) {                // back() is the oldest
        if (!it->speculative || it->canceled) continue;
        const Draft & d = it->drafts[it->s_keep];
        return it->n_past_tgt == n_past_tgt && !d.tokens.empty() && d.tokens[0] == tok;
    }
    return false;
}

bool Scheduler::resume_chain(int verified_offset, bool verified_spec) {
    if (!chain_alive()) return false;
    if (verified_spec) free_offsets.push_back(verified_offset);                   // its partition is not reused by a restart
    drafts = tip.drafts;
    spec_past_tgt = tip.spec_past_tgt; spec_past_dft = tip.spec_past_dft;
    id = tip.id;
    current.seq_offset = tip.seq_offset;
    first_run = false;
    n_chain_resumed++;
    ev("resume_chain", spec_past_tgt, n_past_tgt);
    return true;
}

// :765-844
void Scheduler::speculation_loop() {
    bool is_waiting = llama_mpi_iprobe(ctx_tgt);
    if (is_waiting) free_offsets.push_back(current.seq_offset);
    int   iter = 0;
    float p_adjust;
    while (!is_waiting && P.p_accept + (p_adjust = p_adjust_of(iter)) < 1.0f) {
        if (P.max_spec_runs > 0 && (int) tgt_runs.size() >= P.max_spec_runs) break;   // optional bound (0 = reference behaviour)
        const int orig_offset = current.seq_offset;
        bool should_run = true;
        do {                                                                                             // :791-807
            should_run = true;
            for (const Run & r : tgt_runs) {
                if (r.seq_offset == current.seq_offset && r.speculative) {
                    should_run = false;
                    if (!free_offsets.empty()) { current.seq_offset = free_offsets.front(); free_offsets.pop_front(); }
                    break;
                }
            }
        } while (!should_run && !free_offsets.empty());
        if (!should_run) break;
        // A chained run seeds its partition from the partition of the run before it (start_spec_run).  If that run has been
        // cancelled and retired meanwhile, the partition handed out here is that same one: the reference then clears it and
        // copies it onto itself (speculative.cpp:891-905 with orig_offset == seq_offset), which leaves the new run without
        // the positions from prefix - 1 on.  Stop speculating instead; the next verified run restarts the chain.
        if (!first_run && current.seq_offset == orig_offset && !reference_hazards) break;
        auto it = std::find(free_offsets.begin(), free_offsets.end(), current.seq_offset);
        if (it != free_offsets.end()) free_offsets.erase(it);
        if (first_run || !chain_policy) { chain_start_tgt = n_past_tgt - (merge_first ? 1 : 0); chain_start_dft = n_past_dft; }
        if (start_spec_run(chain_start_tgt, chain_start_dft, orig_offset, p_adjust)) break;
        is_waiting = llama_mpi_iprobe(ctx_tgt);
        first_run  = false;
        iter++;
    }
}

int Scheduler::run(const llama_token * prompt, int n_prompt, llama_token * out, int out_cap, pib_spec_result * res, double * itl_us, int itl_cap) {
    if (n_prompt < 2) { pib_set_error("speculative: the prompt needs at least 2 tokens"); return -1; }
    const int64_t t_enc_start = llama_time_us();
    std::vector<llama_seq_id> all_ids(max_seq + 1);
    for (int i = 0; i <= max_seq; i++) all_ids[i] = i;
    {   // :230-246 prompt: every token belongs to every sequence; target decodes n-1 tokens, then the last one
        Batch bt, bd;
        for (int i = 0; i < n_prompt - 1; i++) { bt.add(prompt[i], i, all_ids); bd.add(prompt[i], i, all_ids); }
        if (llama_decode(ctx_tgt, bt.view(max_seq + 1)) != 0) return -1;
        bt.clear();
        bt.add(prompt[n_prompt - 1], n_prompt - 1, all_ids);
        bd.add(prompt[n_prompt - 1], n_prompt - 1, all_ids);
        if (llama_decode(ctx_tgt, bt.view(max_seq + 1)) != 0) return -1;
        if (llama_decode(ctx_dft, bd.view(max_seq + 1)) != 0) return -1;
    }
    const int64_t t_enc_end = llama_time_us();
    n_past_tgt = n_past_dft = n_prompt;
    generated.assign(prompt, prompt + n_prompt);
    drafts[0].i_batch_tgt = { 0 };                                                                       // :284-286
    current = Run();
    current.n_past_tgt = n_past_tgt - 1; current.n_past_max = n_past_tgt; current.n_past_dft = n_past_dft - 1;
    current.seq_offset = free_offsets.front();
    spec_past_tgt = n_past_tgt; spec_past_dft = n_past_dft;
    std::deque<Run> dft_runs;
    bool has_eos = false, first_token = false, has_run_first_token = false;
    int64_t ttft = llama_time_us(), itt_start = 0;
    std::vector<int64_t> itl;
    const int64_t t_dec_start = llama_time_us();
    t_trace0 = t_dec_start;
    first_run = true;
    int n_out = 0;

    while (!failed) {
        int i_dft = 0, s_keep = 0;
        const bool is_waiting = llama_mpi_iprobe(ctx_tgt);                                               // :323
        if (!tgt_runs.empty() && is_waiting) {                                                           // :328-395
            check_for_cancel();
            Run & run = tgt_runs.back();
            if (run.canceled) {
                llama_finish_async_decode(*ctx_tgt, run.batch.view(max_seq + 1), run.handle);           // drain (:366)
                ev("finish_cancelled", run.batch.batch_id);
                if (run.speculative) free_offsets.push_back(run.seq_offset);
                tgt_runs.pop_back();
                continue;
            }
            if (chain_policy) save_tip();
            drafts = run.drafts;
            current.speculative = run.speculative; current.n_past_max = run.n_past_max; current.n_past_tgt = run.n_past_tgt;
            current.n_past_dft = run.n_past_dft; current.seq_offset = run.seq_offset;
            s_keep = run.s_keep;
            for (int s = 0; s < np; ++s) if (drafts[s].active) drafts[s].tokens.erase(drafts[s].tokens.begin());
            llama_finish_async_decode(*ctx_tgt, run.batch.view(max_seq + 1), run.handle);
            ev("finish", run.batch.batch_id, run.speculative ? 1 : 0, (int) run.batch.size());
            tgt_runs.pop_back();
            spec_past_tgt = n_past_tgt; spec_past_dft = n_past_dft;
            first_run = true;
        } else if (!tgt_runs.empty()) {                                                                  // :397-403 speculate while waiting
            speculation_loop();
            continue;
        }

        // ---- verification walk (:439-560) ----
        int32_t walk_ids[18];
        int     walk_n = -1, walk_pos = 0;
        bool any_match = false;
        int  keep0 = s_keep;
        bool keeps_nonempty = true;
        while (keeps_nonempty) {
            // :448 -- from the device's walk of this batch when it carried a table (same loop, same order), else from the row
            if (walk_n < 0) walk_n = device_walk ? llama_pib_get_accepted(ctx_tgt, walk_ids, 18, nullptr) : 0;
            if (walk_n > 0 && walk_pos < walk_n) id = walk_ids[walk_pos++];
            else id = greedy(ctx_tgt, drafts[keep0].i_batch_tgt[i_dft]);
            if (current.n_past_tgt + i_dft == n_past_tgt - 1) {                                          // a token nobody has seen yet (:465)
                any_match = true;
                ++n_predict;
                if (current.speculative) n_accept++;
                if (has_run_first_token) {
                    if (first_token) { ttft = llama_time_us() - ttft; first_token = false; }
                    else itl.push_back(llama_time_us() - itt_start);
                    itt_start = llama_time_us();
                }
                generated.push_back(id);
                if (n_out < out_cap) out[n_out] = id;
                n_out++;
            }
            if (id == llama_token_eos(llama_get_model(ctx_tgt)) && !P.ignore_eos) has_eos = true;
            if (!current.speculative) break;
            bool matches = false;                                                                        // :518-535
            keeps_nonempty = false;
            for (int s = 0; s < np; ++s) {
                if (!drafts[s].active) continue;
                if (i_dft < (int) drafts[s].tokens.size() && id == drafts[s].tokens[i_dft]) {
                    if (!matches) { keep0 = s; s_keep = s; }
                    matches = true;
                    keeps_nonempty = true;
                } else drafts[s].active = false;
            }
            if (matches) {                                                                               // :537-552
                if (current.n_past_tgt + i_dft == n_past_tgt - 1) { ++n_accept; ++n_past_tgt; ++n_past_dft; }
                ++i_dft;
            }
        }
        if (!any_match) {                                                                                // :569-577
            ev("no_new_token", current.n_past_tgt, i_dft);
            if (resume_chain(current.seq_offset, current.speculative)) continue;
            if (current.speculative) free_offsets.push_back(current.seq_offset);
            continue;
        }
        n_rejected = 0;
        ev("accepted", n_past_tgt, i_dft, current.speculative ? 1 : 0);
        check_for_cancel();                                                                              // :581

        if (current.speculative) {                                                                       // :585-628 KV commit
            // policy 2: the token sampled by the previous verification sits in this run's first cell only
            const int c0_tgt = pending_pos_tgt == current.n_past_tgt ? current.n_past_tgt : current.n_past_tgt + 1;
            const int c0_dft = pending_pos_dft == current.n_past_dft ? current.n_past_dft : current.n_past_dft + 1;
            pending_pos_tgt = pending_pos_dft = -1;
            for (int i = 0; i < np; i++) {
                llama_kv_cache_seq_rm(ctx_tgt, i + current.seq_offset, n_past_tgt, -1);
                llama_kv_cache_seq_rm(ctx_dft, i + current.seq_offset, n_past_dft, -1);
            }
            llama_kv_cache_seq_rm(ctx_tgt, 0, c0_tgt, n_past_tgt);
            llama_kv_cache_seq_rm(ctx_dft, 0, c0_dft, n_past_dft);
            llama_kv_cache_seq_cp(ctx_tgt, s_keep + current.seq_offset, 0, c0_tgt, n_past_tgt);
            llama_kv_cache_seq_cp(ctx_dft, s_keep + current.seq_offset, 0, c0_dft, n_past_dft);
            // :603-612, for i in 1..max_seq: seq_rm(i, range); seq_cp(0, i, range) on both contexts -- one ranged record each
            llama_pib_kv_seq_sweep(ctx_tgt, 1, max_seq, c0_tgt, n_past_tgt);
            llama_pib_kv_seq_sweep(ctx_dft, 1, max_seq, c0_dft, n_past_dft);
        }
        if (chain_policy >= 2 && chain_alive() && next_link_starts_with(id)) {
            // the next link in flight begins with `id` at this very position: no run of its own for it, on either model
            pending_pos_tgt = n_past_tgt++;
            pending_pos_dft = n_past_dft++;
            if (!has_run_first_token) { has_run_first_token = true; first_token = true; }
            ev("skip_nonspec", n_past_tgt);
            resume_chain(current.seq_offset, current.speculative);
            speculation_loop();
            if (n_predict > P.n_predict || has_eos || failed) break;
            continue;
        }
        const bool try_merge = chain_policy >= 3;
        if (!try_merge) begin_non_spec_run(ctx_tgt, tgt_runs, n_past_tgt, n_past_dft);                   // :630 (async)
        else { pending_pos_tgt = n_past_tgt++; merge_first = true; }                                     // `id` rides in the next first link
        begin_non_spec_run(ctx_dft, dft_runs, n_past_dft, n_past_dft);                                   // :633 (finished right away)
        if (!has_run_first_token) { has_run_first_token = true; first_token = true; }
        {
            Run & dr = dft_runs.back();
            llama_finish_async_decode(*ctx_dft, dr.batch.view(max_seq + 1), dr.handle);
            dft_runs.pop_back();
        }
        spec_past_tgt = n_past_tgt - (try_merge ? 1 : 0); spec_past_dft = n_past_dft;
        if (try_merge) pending_pos_dft = n_past_dft;       // the merged run also verifies its FIRST drafted token: the draft's cell for it is committed from the partition
        if (!try_merge && resume_chain(current.seq_offset, current.speculative)) {
            speculation_loop();
            if (n_predict > P.n_predict || has_eos || failed) break;
            continue;
        }
        bool can_speculate = true;
        if (!current.speculative) {                                                                      // :653-662
            if (free_offsets.empty()) can_speculate = false;
            else { current.seq_offset = free_offsets.front(); free_offsets.pop_front(); }
        }
        if (can_speculate) speculation_loop();                                                           // :666
        if (merge_first) {
            // no first link was issued (the oldest run is already waiting, no partition or run slot free, nothing drafted):
            // the sampled token gets its own run after all, as in the reference -- there must be a run in flight that yields the next token
            merge_first = false; pending_pos_tgt = pending_pos_dft = -1; n_past_tgt--;
            begin_non_spec_run(ctx_tgt, tgt_runs, n_past_tgt, n_past_dft);
            spec_past_tgt = n_past_tgt;
        }
        if (!can_speculate) continue;
        if (n_predict > P.n_predict || has_eos || failed) break;                                         // :672
    }
    const int64_t t_dec_end = llama_time_us();
    ev("end", n_predict);
    dump_trace();
    if (chain_policy > 0) {
        // the text is complete: what is still in flight is only waited for, so it is stopped first (a stopped run costs its launches,
        // not its weights).  Not counted in n_cancelled: the reference's statistic is about rejected speculation.  Under the reference
        // policy (PIB_SCHED_CHAIN=0) the schedule stays the reference's event for event: it waits for them.
        std::vector<int32_t> ids;
        for (auto & r : tgt_runs) if (!r.canceled) ids.push_back(r.batch.batch_id);
        if (!ids.empty()) llama_cancel_run(ctx_tgt, ids.data(), (int) ids.size());
    }
    for (int i = (int) tgt_runs.size() - 1; i >= 0; i--) llama_finish_async_decode(*ctx_tgt, tgt_runs[i].batch.view(max_seq + 1), tgt_runs[i].handle);
    tgt_runs.clear();
    if (res) {
        res->n_predict = n_predict; res->n_drafted = n_drafted; res->n_accept = n_accept; res->n_cancelled = n_cancel; res->n_spec_runs = n_spec_runs;
        res->t_encode_s = (t_enc_end - t_enc_start) / 1e6; res->t_decode_s = (t_dec_end - t_dec_start) / 1e6;
        res->decode_tok_per_s = res->t_decode_s > 0 ? n_predict / res->t_decode_s : 0;                  // :697
        res->ttft_s = ttft / 1e6;
        res->n_itl = (int) itl.size();
        double sum = 0;
        for (int64_t v : itl) sum += (double) v;
        res->itl_mean_s = itl.empty() ? 0 : sum / itl.size() / 1e6;                                      // the reference reports the mean (:683-688)
        std::vector<int64_t> srt = itl;
        std::sort(srt.begin(), srt.end());
        res->itl_p50_s = srt.empty() ? 0 : srt[srt.size() / 2] / 1e6;
    }
    for (int i = 0; i < (int) itl.size() && i < itl_cap; i++) itl_us[i] = (double) itl[i];
    return std::min(n_out, out_cap);          // never more than was written (a run may accept past n_predict)
}

}   // namespace

extern "C" {

pib_spec_params pib_spec_default_params(void) {
    pib_spec_params p;
    p.n_draft = 16; p.n_parallel = 1; p.p_accept = 0.5f; p.p_split = 0.1f; p.p_recovery = 0.0f; p.p_decay = 0.0f;   // common.h:54-61
    p.n_predict = 128; p.ignore_eos = 1; p.max_spec_runs = 0;
    return p;
}

int pib_speculative_generate(struct llama_context * ctx_tgt, struct llama_context * ctx_dft, const llama_token * prompt, int n_prompt,
                             const pib_spec_params * params, llama_token * out_tokens, int out_cap, pib_spec_result * result,
                             double * itl_us, int itl_cap) {
    if (!ctx_tgt || !ctx_dft || !prompt || !params) { pib_set_error("pib_speculative_generate: bad arguments"); return -1; }
    const int need = 1000 * std::max(1, params->n_parallel) + 2;
    if (llama_pib_kv_words(ctx_tgt) * 32 < need || llama_pib_kv_words(ctx_dft) * 32 < need) {
        pib_set_error("pib_speculative_generate: contexts need a sequence-id capacity of %d (llama_pib_new_context)", need);
        return -1;
    }
    if (params->ignore_eos) {          // common.cpp:1061-1063: logit_bias[eos] = -inf, applied on the device before argmax
        const llama_token e_t = llama_token_eos(llama_get_model(ctx_tgt)), e_d = llama_token_eos(llama_get_model(ctx_dft));
        const float ninf = -INFINITY;
        llama_pib_set_logit_bias(ctx_tgt, &e_t, &ninf, 1);
        llama_pib_set_logit_bias(ctx_dft, &e_d, &ninf, 1);
    }
    Scheduler s(ctx_tgt, ctx_dft, *params);
    return s.run(prompt, n_prompt, out_tokens, out_cap, result, itl_us, itl_cap);
}

// plain (non-speculative) greedy decode: the reference's examples/main under the same launch (the "Sequential" bar of the
// paper, scripts/dolphin_sequential_8_nodes.sh) -- the denominator of the 3x target
int pib_plain_generate(struct llama_context * ctx, const llama_token * prompt, int n_prompt, int n_predict, llama_token * out_tokens,
                       pib_spec_result * result, double * itl_us, int itl_cap) {
    if (!ctx || !prompt || n_prompt < 1) { pib_set_error("pib_plain_generate: bad arguments"); return -1; }
    const int64_t t0 = llama_time_us();
    Batch b;
    for (int i = 0; i < n_prompt; i++) b.add(prompt[i], i, { 0 });
    if (llama_decode(ctx, b.view(1)) != 0) return -1;
    const int64_t t1 = llama_time_us();
    llama_token_data t8[8];
    llama_pib_get_top8(ctx, n_prompt - 1, t8);
    llama_token id = t8[0].id;
    int n_past = n_prompt, n_out = 0;
    std::vector<int64_t> itl;
    int64_t last = llama_time_us(), ttft = 0;
    while (n_out < n_predict) {
        out_tokens[n_out++] = id;
        const int64_t now = llama_time_us();
        if (n_out == 1) ttft = now - t1; else itl.push_back(now - last);
        last = now;
        if (n_out >= n_predict) break;
        b.clear();
        b.add(id, n_past++, { 0 });
        if (llama_decode(ctx, b.view(1)) != 0) return -1;
        llama_pib_get_top8(ctx, 0, t8);
        id = t8[0].id;
    }
    const int64_t t2 = llama_time_us();
    if (result) {
        memset(result, 0, sizeof(*result));
        result->n_predict = n_out; result->t_encode_s = (t1 - t0) / 1e6; result->t_decode_s = (t2 - t1) / 1e6;
        result->decode_tok_per_s = n_out / std::max(1e-9, result->t_decode_s);
        result->ttft_s = ttft / 1e6; result->n_itl = (int) itl.size();
        double sum = 0; for (int64_t v : itl) sum += (double) v;
        result->itl_mean_s = itl.empty() ? 0 : sum / itl.size() / 1e6;
        std::vector<int64_t> srt = itl; std::sort(srt.begin(), srt.end());
        result->itl_p50_s = srt.empty() ? 0 : srt[srt.size() / 2] / 1e6;
    }
    for (int i = 0; i < (int) itl.size() && i < itl_cap; i++) itl_us[i] = (double) itl[i];
    return n_out;
}

// classic synchronous speculative decoding: the reference's examples/speculative_orig (the "Speculative" bar of plot.py).
// Draft up to n_draft tokens as a tree over n_parallel sequences, verify them with ONE target decode, keep the matching
// prefix, repeat (speculative_orig/speculative.cpp:222-505; cited lines are of that file).  Greedy only, like the rest.
int pib_speculative_orig_generate(struct llama_context * ctx_tgt, struct llama_context * ctx_dft, const llama_token * prompt, int n_prompt,
                                  const pib_spec_params * params, llama_token * out_tokens, int out_cap, pib_spec_result * result,
                                  double * itl_us, int itl_cap) {
    if (!ctx_tgt || !ctx_dft || !prompt || !params || n_prompt < 2) { pib_set_error("pib_speculative_orig_generate: bad arguments"); return -1; }
    const pib_spec_params & P = *params;
    const int n_seq_dft = std::max(1, P.n_parallel), n_draft = P.n_draft;
    if (llama_pib_kv_words(ctx_tgt) * 32 < n_seq_dft + 1 || llama_pib_kv_words(ctx_dft) * 32 < n_seq_dft + 1) {
        pib_set_error("pib_speculative_orig_generate: contexts need a sequence-id capacity of %d", n_seq_dft + 1);
        return -1;
    }
    if (P.ignore_eos) {
        const llama_token e_t = llama_token_eos(llama_get_model(ctx_tgt)), e_d = llama_token_eos(llama_get_model(ctx_dft));
        const float ninf = -INFINITY;
        llama_pib_set_logit_bias(ctx_tgt, &e_t, &ninf, 1);
        llama_pib_set_logit_bias(ctx_dft, &e_d, &ninf, 1);
    }
    const int M = n_seq_dft + 1;
    auto top8 = [](llama_context * c, int row, llama_token_data * t) { llama_pib_get_top8(c, row, t); };
    const int64_t t_enc_start = llama_time_us();
    std::vector<llama_seq_id> seq_ids(M);
    for (int i = 0; i < M; i++) seq_ids[i] = i;
    Batch batch_dft, batch_tgt;
    {   // :172-184 prompt: every token in every sequence; the target decodes n-1 tokens, then the last one
        for (int i = 0; i < n_prompt - 1; i++) { batch_dft.add(prompt[i], i, seq_ids); batch_tgt.add(prompt[i], i, seq_ids); }
        if (llama_decode(ctx_tgt, batch_tgt.view(M)) != 0) return -1;
        batch_tgt.clear();
        batch_dft.add(prompt[n_prompt - 1], n_prompt - 1, seq_ids);
        batch_tgt.add(prompt[n_prompt - 1], n_prompt - 1, seq_ids);
        if (llama_decode(ctx_tgt, batch_tgt.view(M)) != 0) return -1;
        if (llama_decode(ctx_dft, batch_dft.view(M)) != 0) return -1;
    }
    const int64_t t_enc_end = llama_time_us();
    int n_predict = 0, n_drafted = 0, n_accept = 0, n_past_tgt = n_prompt, n_past_dft = n_prompt, n_out = 0, n_runs = 0;
    bool has_eos = false, first_token = false, has_run_first_token = false;
    std::vector<Draft> drafts(n_seq_dft);
    drafts[0].i_batch_tgt = { 0 };                                                                        // :222-224
    int64_t ttft = llama_time_us(), itt_start = 0;
    std::vector<int64_t> itl;
    const int64_t t_dec_start = llama_time_us();
    llama_token_data t8[8];
    while (true) {
        int i_dft = 0, s_keep = 0;
        while (true) {                                                                                    // :247-339 verification walk
            top8(ctx_tgt, drafts[s_keep].i_batch_tgt[i_dft], t8);
            const llama_token id = t8[0].id;
            if (has_run_first_token) {
                if (first_token) { ttft = llama_time_us() - ttft; first_token = false; }
                else itl.push_back(llama_time_us() - itt_start);
                itt_start = llama_time_us();
            }
            if (n_out < out_cap) out_tokens[n_out] = id;
            n_out++;
            if (id == llama_token_eos(llama_get_model(ctx_tgt)) && !P.ignore_eos) has_eos = true;
            ++n_predict;
            bool matches = false;
            for (int s = 0; s < n_seq_dft; ++s) {
                if (!drafts[s].active) continue;
                if (i_dft < (int) drafts[s].tokens.size() && id == drafts[s].tokens[i_dft]) { s_keep = s; matches = true; }
                else drafts[s].active = false;
            }
            if (matches) { ++n_accept; ++n_past_tgt; ++n_past_dft; ++i_dft; continue; }
            // :305-338 keep the winning branch, feed the new token to the draft
            llama_kv_cache_seq_keep(ctx_dft, s_keep);
            llama_kv_cache_seq_cp(ctx_dft, s_keep, 0, -1, -1);
            llama_kv_cache_seq_keep(ctx_dft, 0);
            llama_kv_cache_seq_rm(ctx_tgt, s_keep, n_past_tgt, -1);
            llama_kv_cache_seq_keep(ctx_tgt, s_keep);
            llama_kv_cache_seq_cp(ctx_tgt, s_keep, 0, -1, -1);
            llama_kv_cache_seq_keep(ctx_tgt, 0);
            for (int s = 0; s < n_seq_dft; ++s) { drafts[s].active = false; drafts[s].tokens.clear(); drafts[s].i_batch_tgt.clear(); }
            drafts[0].tokens.push_back(id);
            drafts[0].i_batch_tgt.push_back(0);
            batch_dft.clear();
            batch_dft.add(id, n_past_dft, { 0 });
            llama_kv_cache_seq_rm(ctx_dft, -1, n_past_dft, -1);
            if (llama_decode(ctx_dft, batch_dft.view(M)) != 0) return -1;
            ++n_past_dft;
            break;
        }
        if (n_predict > P.n_predict || has_eos) break;                                                    // :341-343
        int n_seq_cur = 1, n_past_cur = n_past_dft;
        for (int s = 0; s < n_seq_dft; ++s) { drafts[s].active = false; drafts[s].drafting = false; }
        drafts[0].active = true; drafts[0].drafting = true; drafts[0].i_batch_dft = 0;
        batch_tgt.clear();
        batch_tgt.add(drafts[0].tokens[0], n_past_tgt, { 0 });
        for (int i = 0; i < n_draft; ++i) {                                                               // :362-475 tree drafting
            batch_dft.clear();
            for (int s = 0; s < n_seq_dft; ++s) drafts[s].skip = false;
            for (int s = 0; s < n_seq_dft; ++s) {
                if (!drafts[s].drafting || drafts[s].skip) continue;
                top8(ctx_dft, drafts[s].i_batch_dft, t8);
                if (t8[0].p < P.p_accept) { drafts[s].drafting = false; continue; }                      // :389-393
                std::vector<int> sa(1, s);
                for (int f = 1; f < 8; ++f) {                                                             // :398-441 branch splitting
                    if (n_seq_cur < n_seq_dft && t8[f].p > P.p_split) {
                        llama_kv_cache_seq_rm(ctx_dft, n_seq_cur, -1, -1);
                        llama_kv_cache_seq_cp(ctx_dft, s, n_seq_cur, -1, -1);
                        for (int t = 0; t < batch_tgt.size(); ++t)
                            for (size_t p = 0; p < batch_tgt.seq[t].size(); ++p)
                                if (batch_tgt.seq[t][p] == s) { batch_tgt.seq[t].push_back(n_seq_cur); break; }
                        drafts[n_seq_cur].active = true; drafts[n_seq_cur].drafting = true; drafts[n_seq_cur].skip = true;
                        drafts[n_seq_cur].tokens = drafts[s].tokens;
                        drafts[n_seq_cur].i_batch_dft = drafts[s].i_batch_dft;
                        drafts[n_seq_cur].i_batch_tgt = drafts[s].i_batch_tgt;
                        sa.push_back(n_seq_cur);
                        n_seq_cur++;
                    } else break;
                }
                for (int is = 0; is < (int) sa.size(); ++is) {                                            // :444-465
                    const llama_token id = t8[is].id;
                    const int sq = sa[is];
                    drafts[sq].tokens.push_back(id);
                    drafts[sq].i_batch_tgt.push_back(batch_tgt.size());
                    batch_tgt.add(id, n_past_tgt + i + 1, { sq });
                    drafts[sq].i_batch_dft = batch_dft.size();
                    batch_dft.add(id, n_past_cur, { sq });
                    if (batch_tgt.size() > n_draft) drafts[sq].drafting = false;
                }
            }
            if (batch_dft.size() == 0) break;
            if (llama_decode(ctx_dft, batch_dft.view(M)) != 0) return -1;
            ++n_past_cur;
            ++n_drafted;
            if (batch_tgt.size() > n_draft) break;
        }
        llama_kv_cache_seq_keep(ctx_tgt, 0);                                                              // :478-495 verify
        for (int s = 1; s < n_seq_dft; ++s) llama_kv_cache_seq_cp(ctx_tgt, 0, s, -1, -1);
        if (!has_run_first_token) { has_run_first_token = true; first_token = true; }
        if (llama_decode(ctx_tgt, batch_tgt.view(M)) != 0) return -1;
        ++n_past_tgt;
        ++n_runs;
        for (int s = 0; s < n_seq_dft; ++s) if (drafts[s].active) drafts[s].tokens.erase(drafts[s].tokens.begin());   // :498-504
    }
    const int64_t t_dec_end = llama_time_us();
    if (result) {
        memset(result, 0, sizeof(*result));
        result->n_predict = n_predict; result->n_drafted = n_drafted; result->n_accept = n_accept; result->n_spec_runs = n_runs;
        result->t_encode_s = (t_enc_end - t_enc_start) / 1e6; result->t_decode_s = (t_dec_end - t_dec_start) / 1e6;
        result->decode_tok_per_s = n_predict / std::max(1e-9, result->t_decode_s);
        result->ttft_s = ttft / 1e6; result->n_itl = (int) itl.size();
        double sum = 0; for (int64_t v : itl) sum += (double) v;
        result->itl_mean_s = itl.empty() ? 0 : sum / itl.size() / 1e6;
        std::vector<int64_t> srt = itl; std::sort(srt.begin(), srt.end());
        result->itl_p50_s = srt.empty() ? 0 : srt[srt.size() / 2] / 1e6;
    }
    for (int i = 0; i < (int) itl.size() && i < itl_cap; i++) itl_us[i] = (double) itl[i];
    return std::min(n_out, out_cap);
}

}   // extern "C"
