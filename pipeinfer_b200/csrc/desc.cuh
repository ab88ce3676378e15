// desc.cuh -- the per-micro-batch descriptor that travels host -> device (and stage -> stage over NCCL).
//
// It replaces the seven MPI messages of the reference's ggml_mpi_eval_init (reference ggml-mpi.c:236-347: batch_id,
// n_tokens, logits mask, tokens, n_seq_id, pos, flattened seq ids) with one fixed-layout record per chunk of at most
// PIB_CHUNK tokens.  Sequence-id lists become bitset rows (a prompt token that carries all 1000*np+2 ids costs n_words
// words instead of 3002 ints).
#pragma once

#include <stdint.h>

#define PIB_CHUNK 16                 // tokens per device pass (the GEMV batch limit, PIB_MAX_BATCH)

// int32 header fields
#define PIB_DH_NTOK     0            // tokens in this chunk
#define PIB_DH_KVHEAD   1            // first KV cell of this chunk
#define PIB_DH_BATCHID  2            // llama_batch.batch_id (reference llama.h:160) -> index into the cancel table
#define PIB_DH_NKV      3            // attended range n (informational; kernels read the device scalar)
#define PIB_DH_TREE_NP  4            // branches of the acceptance table below (0: none), PIB_DH_TREE_KEEP the branch the walk starts on,
#define PIB_DH_TREE_KEEP 5           //   PIB_DH_TREE_SPEC 1 for a speculative run (0: sample row 0 and stop)
#define PIB_DH_TREE_SPEC 6
#define PIB_DH_WORDS    16

#define PIB_DESC_HDR_OFF   0
#define PIB_DESC_TOK_OFF   (PIB_DH_WORDS * 4)                       // int32 token[PIB_CHUNK]
#define PIB_DESC_POS_OFF   (PIB_DESC_TOK_OFF + PIB_CHUNK * 4)       // int32 pos[PIB_CHUNK]
// acceptance table of a verification batch (speculative.cpp:439-560 walks it on the host; here the last stage's device does):
// per draft branch s its length (-1: inactive), the batch row of its i-th position and the token drafted there
#define PIB_TREE_NP        8
#define PIB_TREE_DEPTH     (PIB_CHUNK + 1)
#define PIB_DESC_TREE_OFF  (PIB_DESC_POS_OFF + PIB_CHUNK * 4)       // int32 len[NP], rows[NP][DEPTH], toks[NP][DEPTH]
#define PIB_DESC_TREE_BYTES ((PIB_TREE_NP + 2 * PIB_TREE_NP * PIB_TREE_DEPTH) * 4)
#define PIB_DESC_BITS_OFF  (PIB_DESC_TREE_OFF + PIB_DESC_TREE_BYTES)  // uint32 tokbits[PIB_CHUNK][n_words]

static inline size_t pib_desc_stride(int n_words) {
    return ((size_t) PIB_DESC_BITS_OFF + (size_t) PIB_CHUNK * n_words * 4 + 255) & ~(size_t) 255;
}
