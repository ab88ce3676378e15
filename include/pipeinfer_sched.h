/*
 * pipeinfer_sched.h -- C ABI of libpipeinfer_b200.so, scheduler level: PipeInfer's continuous asynchronous speculation
 * loop and the plain pipelined decode it is compared with.
 *
 * The reference ships these as programs, not as functions: examples/speculative/speculative.cpp (main, :78-763) and
 * examples/main/main.cpp; the `speculative` CLI of this repository (pipeinfer_b200/csrc/cli) parses the same flags
 * (-m -md --draft -np -pa -ps -pr -pd --mpi-layer-split -n -c ...) and calls the functions below.  bench.py and the
 * tests call them through ctypes.
 */
#ifndef PIPEINFER_SCHED_H
#define PIPEINFER_SCHED_H

#include "pipeinfer_llama.h"

#ifdef __cplusplus
extern "C" {
#endif

/* gpt_params fields the scheduler reads (common/common.h:50-61; flags common/common.cpp:415-469) */
typedef struct pib_spec_params {
    int32_t n_draft;        /* --draft      tokens drafted per speculative run                         */
    int32_t n_parallel;     /* -np          draft tree branches (sequence partition width)             */
    float   p_accept;       /* -pa          stop drafting below this draft probability                 */
    float   p_split;        /* -ps          branch when a runner-up candidate exceeds this probability */
    float   p_recovery;     /* -pr          per-iteration increase of the thresholds while speculating  */
    float   p_decay;        /* -pd          decrease after rejected speculation                         */
    int32_t n_predict;      /* -n           stop after more than this many new tokens (speculative.cpp:672) */
    int32_t ignore_eos;     /* --ignore-eos EOS logit = -inf (applied on the device)                   */
    int32_t max_spec_runs;  /* extension: bound on runs in flight, 0 = unbounded as in the reference   */
} pib_spec_params;

/* the figures the reference prints / appends to results.csv (speculative.cpp:681-718) */
typedef struct pib_spec_result {
    int32_t n_predict, n_drafted, n_accept, n_cancelled, n_spec_runs, n_itl;
    double  t_encode_s, t_decode_s;
    double  decode_tok_per_s;          /* n_predict / (t_dec_end - t_dec_start)   speculative.cpp:697 */
    double  ttft_s, itl_mean_s, itl_p50_s;
} pib_spec_result;

pib_spec_params pib_spec_default_params(void);

/* Greedy generation with continuous asynchronous speculation.  Both contexts must have been created with a
 * sequence-id capacity of 1000 * n_parallel + 2 (llama_pib_new_context / PIB_MAX_SEQ).  Writes the newly generated
 * token ids to out_tokens (at most out_cap) and returns how many were generated, or -1 on error.
 * itl_us (optional) receives the inter-token latency samples in microseconds.
 * Replaces: main() of examples/speculative/speculative.cpp. */
int pib_speculative_generate(struct llama_context * ctx_tgt, struct llama_context * ctx_dft, const llama_token * prompt,
                             int n_prompt, const pib_spec_params * params, llama_token * out_tokens, int out_cap,
                             pib_spec_result * result, double * itl_us, int itl_cap);

/* Plain greedy decode, one token per pass through the pipeline.  Replaces: the generation loop of examples/main/main.cpp
 * under the reference's "sequential" launch (examples/speculative/scripts/dolphin_sequential_8_nodes.sh). */
int pib_plain_generate(struct llama_context * ctx, const llama_token * prompt, int n_prompt, int n_predict,
                       llama_token * out_tokens, pib_spec_result * result, double * itl_us, int itl_cap);

/* Classic synchronous speculative decoding: draft a tree of up to n_draft tokens over n_parallel sequences, verify it with
 * one target decode, keep the matching prefix.  Replaces: the loop of examples/speculative_orig/speculative.cpp:222-505
 * (the "Speculative" bar next to "Sequential" and PipeInfer in plot.py).  Same parameters and result as above; the
 * contexts need a sequence-id capacity of n_parallel + 1. */
int pib_speculative_orig_generate(struct llama_context * ctx_tgt, struct llama_context * ctx_dft, const llama_token * prompt,
                                  int n_prompt, const pib_spec_params * params, llama_token * out_tokens, int out_cap,
                                  pib_spec_result * result, double * itl_us, int itl_cap);

#ifdef __cplusplus
}
#endif
#endif
