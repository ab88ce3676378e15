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
#define PIB_DH_WORDS    16

#define PIB_DESC_HDR_OFF   0
#define PIB_DESC_TOK_OFF   (PIB_DH_WORDS * 4)                       // int32 token[PIB_CHUNK]
#define PIB_DESC_POS_OFF   (PIB_DESC_TOK_OFF + PIB_CHUNK * 4)       // int32 pos[PIB_CHUNK]
#define PIB_DESC_BITS_OFF  (PIB_DESC_POS_OFF + PIB_CHUNK * 4)       // uint32 tokbits[PIB_CHUNK][n_words]

static inline size_t pib_desc_stride(int n_words) {
    return ((size_t) PIB_DESC_BITS_OFF + (size_t) PIB_CHUNK * n_words * 4 + 255) & ~(size_t) 255;
}
