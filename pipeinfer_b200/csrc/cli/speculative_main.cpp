// speculative -- command-line front end with the flags of the reference's examples/speculative (speculative.cpp:60-230,
// common/common.cpp gpt_params_parse): same meaning, same stdout (the generated text only, :488-496), same summary block
// on stderr (:694-718) and the same results.csv row (encode_speed,decode_speed,itt,ttft).
//
//   speculative -m target.gguf -md draft.gguf -p "prompt" | -f prompt.txt  [-n 128] [-c 1024] [--draft 16] [-np 1]
//               [-pa 0.5] [-ps 0.1] [-pr 0] [-pd 0] [--ignore-eos] [--mpi-layer-split 0.1,0.15,.../1.0]
//               [--temp -1] [--repeat-last-n 0] [-t N] [-tb N]           (greedy only; thread counts are accepted)
//
// One GPU: run it directly.  N GPUs: launch one process per GPU with any launcher that sets RANK / WORLD_SIZE / LOCAL_RANK /
// MASTER_PORT (torchrun --no-python, mpirun with a wrapper ...): rank 0 is the head (scheduler + draft + first layers), the
// other ranks serve their layers until the head shuts the pipeline down -- the reference's mpirun topology without the
// separate draft node (the draft model shares GPU 0, on its own stream).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../../include/pipeinfer_b200.h"
#include "../../../include/pipeinfer_llama.h"
#include "../../../include/pipeinfer_mpi.h"
#include "../../../include/pipeinfer_sched.h"

static void usage() {
    fprintf(stderr, "usage: speculative -m TARGET.gguf -md DRAFT.gguf (-p PROMPT | -f FILE) [-n N] [-c CTX] [--draft N] [-np N] [-pa P] [-ps P]\n"
                    "                   [-pr P] [-pd P] [--ignore-eos] [--mpi-layer-split w0,w1,...[/draft]] [--max-spec-runs N]\n");
}

int main(int argc, char ** argv) {
    std::string model, draft, prompt, split;
    int n_predict = 128, n_ctx = 512, max_spec_runs = 0;
    pib_spec_params P = pib_spec_default_params();
    P.ignore_eos = 0;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&]() -> const char * { if (i + 1 >= argc) { usage(); exit(1); } return argv[++i]; };
        if (a == "-m" || a == "--model") model = next();
        else if (a == "-md" || a == "--model-draft") draft = next();
        else if (a == "-p" || a == "--prompt") prompt = next();
        else if (a == "-f" || a == "--file") { std::ifstream f(next()); std::stringstream ss; ss << f.rdbuf(); prompt = ss.str(); if (!prompt.empty() && prompt.back() == '\n') prompt.pop_back(); }
        else if (a == "-n" || a == "--n-predict") n_predict = atoi(next());
        else if (a == "-c" || a == "--ctx-size") n_ctx = atoi(next());
        else if (a == "--draft") P.n_draft = atoi(next());
        else if (a == "-np" || a == "--parallel") P.n_parallel = atoi(next());
        else if (a == "-pa" || a == "--p-accept") P.p_accept = (float) atof(next());
        else if (a == "-ps" || a == "--p-split") P.p_split = (float) atof(next());
        else if (a == "-pr" || a == "--p-recovery") P.p_recovery = (float) atof(next());
        else if (a == "-pd" || a == "--p-decay") P.p_decay = (float) atof(next());
        else if (a == "--ignore-eos") P.ignore_eos = 1;
        else if (a == "--mpi-layer-split") split = next();
        else if (a == "--max-spec-runs") max_spec_runs = atoi(next());
        else if (a == "--temp" || a == "--repeat-last-n" || a == "-t" || a == "-tb" || a == "--threads" || a == "--threads-batch" ||
                 a == "-s" || a == "--seed" || a == "-b" || a == "--batch-size" || a == "-ngl" || a == "-ngld") next();   // accepted, no effect
        else if (a == "-h" || a == "--help") { usage(); return 0; }
        else { fprintf(stderr, "speculative: unknown argument %s\n", a.c_str()); usage(); return 1; }
    }
    if (model.empty() || draft.empty()) { usage(); return 1; }
    P.n_predict = n_predict; P.max_spec_runs = max_spec_runs;

    llama_backend_init(false);
    // stdout carries the generated text and nothing else: whatever the communication library prints while the pipeline is
    // brought up (NCCL's version banner) goes to stderr
    fflush(stdout);
    const int saved_stdout = dup(1);
    dup2(2, 1);
    const int world = pib_pipe_open_env();
    fflush(stdout);
    dup2(saved_stdout, 1);
    close(saved_stdout);
    if (world < 0) { fprintf(stderr, "speculative: %s\n", pib_last_error()); return 1; }
    const int rank = pib_pipe_rank();
    const int max_seq = 1000 * (P.n_parallel < 1 ? 1 : P.n_parallel) + 2;              // speculative.cpp:140-145

    llama_model *   m_tgt = nullptr;
    llama_context * c_tgt = nullptr;
    if (world > 1) {
        // --mpi-layer-split w0,w1,...[/draft weights]: the part in front of '/' weighs the target's stages (speculative.cpp:97-107)
        std::vector<float> w;
        std::string tw = split.substr(0, split.find('/'));
        std::stringstream ss(tw);
        for (std::string tok; std::getline(ss, tok, ',');) if (!tok.empty()) w.push_back((float) atof(tok.c_str()));
        while ((int) w.size() < world) w.push_back(w.empty() ? 1.0f / world : w.back());
        // the layer count comes from the file: load the head's vocabulary-only view first is not needed -- every rank reads the
        // header through its own slice load below, so ask the library for the split after a cheap probe load of layer 0
        llama_model * probe = llama_pib_load_model_slice(model.c_str(), 0, 1, 0, 0);
        if (!probe) return 1;
        const int n_layer = llama_pib_n_layer(probe);
        llama_free_model(probe);
        std::vector<uint16_t> r((size_t) 2 * world);
        pib_split_range(world, 0, n_layer - 1, w.data(), r.data());
        const int b = r[2 * rank], e = rank == world - 1 ? n_layer : r[2 * rank + 1];
        m_tgt = llama_pib_load_model_slice(model.c_str(), b, e, rank == 0, rank == world - 1);
        if (!m_tgt) return 1;
        c_tgt = llama_pib_new_context_stage(m_tgt, (uint32_t) n_ctx, max_seq, rank, world);
        if (!c_tgt) return 1;
        if (rank != 0) {                                                            // a pipeline stage: llama.cpp:9941-9977
            int rc;
            while ((rc = llama_pib_stage_serve(c_tgt)) == 0) {}
            if (rc < 0) fprintf(stderr, "speculative (stage %d): %s\n", rank, pib_last_error());
            llama_free(c_tgt); llama_free_model(m_tgt);
            pib_pipe_nccl_free(); pib_pipe_ring_close();
            return rc < 0 ? 1 : 0;
        }
    } else {
        m_tgt = llama_load_model_from_file(model.c_str(), llama_model_default_params());
        if (!m_tgt) return 1;
        c_tgt = llama_pib_new_context(m_tgt, (uint32_t) n_ctx, max_seq);
        if (!c_tgt) return 1;
    }
    llama_model * m_dft = llama_load_model_from_file(draft.c_str(), llama_model_default_params());
    if (!m_dft) return 1;
    llama_context * c_dft = llama_pib_new_context(m_dft, (uint32_t) n_ctx, max_seq);
    if (!c_dft) return 1;
    llama_pib_set_priority(c_dft, 1);
    if (llama_n_vocab(m_tgt) != llama_n_vocab(m_dft)) { fprintf(stderr, "speculative: draft and target vocabularies differ\n"); return 1; }   // speculative.cpp:186-200

    // BOS in front unless the file says otherwise; without a say, SPM vocabularies get one (common.cpp llama_should_add_bos_token)
    auto wants_bos = [](const llama_model * m) { const int a = llama_add_bos_token(m); return a != -1 ? a != 0 : llama_vocab_type(m) == 0; };
    if (wants_bos(m_tgt) != wants_bos(m_dft)) { fprintf(stderr, "speculative: draft and target disagree about a leading BOS token\n"); return 1; }   // speculative.cpp:181-192
    std::vector<llama_token> inp(prompt.size() + 8);
    const int n_inp = llama_tokenize(m_tgt, prompt.c_str(), (int) prompt.size(), inp.data(), (int) inp.size(), wants_bos(m_tgt), true);
    if (n_inp < 2) { fprintf(stderr, "speculative: the prompt needs at least 2 tokens\n"); return 1; }
    inp.resize(n_inp);
    if (n_inp > n_ctx - 4) { fprintf(stderr, "speculative: the prompt exceeds the context size\n"); return 1; }   // :212-216
    fprintf(stderr, "\n\n");
    char piece[256];
    for (llama_token t : inp) { const int n = llama_token_to_piece(m_tgt, t, piece, sizeof(piece)); if (n > 0) fprintf(stderr, "%.*s", n, piece); }
    fflush(stderr);

    std::vector<llama_token> out((size_t) n_predict + 64);
    std::vector<double>      itl((size_t) n_predict + 64);
    pib_spec_result          res;
    const int n_out = pib_speculative_generate(c_tgt, c_dft, inp.data(), n_inp, &P, out.data(), (int) out.size(), &res, itl.data(), (int) itl.size());
    if (n_out < 0) { fprintf(stderr, "speculative: %s\n", pib_last_error()); return 1; }
    for (int i = 0; i < n_out && i < (int) out.size(); i++) {
        const int n = llama_token_to_piece(m_tgt, out[i], piece, sizeof(piece));     // < 0: the piece does not fit (never with 256 bytes)
        if (n > 0) fwrite(piece, 1, (size_t) n, stdout);
    }
    fflush(stdout);

    const double encode_speed = res.t_encode_s > 0 ? n_inp / res.t_encode_s : 0.0;
    {
        std::ofstream results("results.csv", std::ios_base::app);
        results << encode_speed << ',' << res.decode_tok_per_s << ',' << res.itl_mean_s << ',' << res.ttft_s << '\n';
    }
    fprintf(stderr, "\n\n");
    fprintf(stderr, "OVERALL SYSTEM MEASUREMENTS:\n");
    fprintf(stderr, "encoded %4d tokens in %8.3f seconds, speed: %8.3f t/s\n", n_inp, res.t_encode_s, encode_speed);
    fprintf(stderr, "decoded %4d tokens in %8.3f seconds, speed: %8.3f t/s\n", res.n_predict, res.t_decode_s, res.decode_tok_per_s);
    fprintf(stderr, "Average inter-token latency: %f seconds\n", res.itl_mean_s);
    fprintf(stderr, "Time-to-first-token: %f seconds\n", res.ttft_s);
    fprintf(stderr, "\nTARGET PIPELINE MEASUREMENTS\n");
    fprintf(stderr, "n_draft   = %d\n", P.n_draft);
    fprintf(stderr, "n_predict = %d\n", res.n_predict);
    fprintf(stderr, "n_drafted = %d\n", res.n_drafted);
    fprintf(stderr, "n_accept  = %d\n", res.n_accept);
    fprintf(stderr, "accept    = %.3f%%\n", res.n_drafted ? 100.0f * res.n_accept / res.n_drafted : 0.0f);

    if (world > 1) llama_pib_pipe_shutdown(c_tgt);
    llama_free(c_dft); llama_free_model(m_dft);
    llama_free(c_tgt); llama_free_model(m_tgt);
    if (world > 1) { pib_pipe_nccl_free(); pib_pipe_ring_close(); }
    llama_backend_free();
    return 0;
}
