// tokenizer.cuh -- host-side vocabulary: text <-> token ids (SentencePiece-style and GPT-2 byte-level BPE).
// Replaces, for the callers on the path (prompt in, pieces out), the reference's llama_vocab + llm_tokenizer_spm /
// llm_tokenizer_bpe + tokenizer_st_partition (llama.cpp:1345-1390, 5970-6601) and llama_token_to_piece (:10111-10166).
// No CUDA call anywhere in here: a vocab_only model tokenizes on a machine without a GPU.
#pragma once

#include <stdint.h>

#include <string>
#include <vector>

struct pib_model;
struct pib_vocab;

enum { PIB_VOCAB_SPM = 0, PIB_VOCAB_BPE = 1 };          // llama.h:70-73

#define PIB_TOKENIZE_INVALID INT32_MIN                  // text is not UTF-8 / a byte has no token (the reference throws)

pib_vocab * pib_vocab_build(const pib_model * m);       // reads m->tok_*; fills the model's linefeed id through *nl
void        pib_vocab_free(pib_vocab * v);
int         pib_vocab_type(const pib_vocab * v);
int         pib_vocab_linefeed(const pib_vocab * v);
int         pib_vocab_n_special(const pib_vocab * v);   // size of the special-token cache (llama.cpp:2538-2625)
// appends to out; returns 0, or PIB_TOKENIZE_INVALID with pib_last_error() set
int         pib_vocab_tokenize(const pib_vocab * v, const std::string & text, bool add_bos, bool special, std::vector<int32_t> & out);
// false when the token prints nothing (control, unused, out of range)
bool        pib_vocab_piece(const pib_vocab * v, int32_t token, std::string & out);
