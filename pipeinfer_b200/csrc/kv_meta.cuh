// kv_meta.cuh -- KV-cache cell bookkeeping: host mirror + device replica kept identical by replaying one op-log.
//
// Reference semantics (llama.cpp:1292-1760 of the reference): every cell has a position (-1 = empty) and a *set* of
// sequence ids; `head` is the next-slot search hint, `used` the number of non-empty cells, `n` the attended range.
// PipeInfer's scheduler uses up to 1000*np + 2 sequence ids (speculative.cpp:140-145), so a set becomes a bitset row.
//
//  * KvMirror (host) is authoritative for everything the API returns synchronously: llama_decode's "no slot" failure,
//    llama_kv_cache_view_*, used-cell counts.  It also resolves find_slot so the device never searches.
//  * The device replica (pos[], bits[][], scalars) is what kernels read: the multi-sequence attention mask is computed
//    from it on the device and seq_rm / seq_cp / seq_keep -- the invalidation of cancelled speculative branches -- run
//    there as one metadata kernel that replays the op-log (kv_meta.cu).  A pipeline stage on another GPU applies the
//    same log in the same order, which is how the reference keeps ranks identical (llama.cpp:9245-9346).
#pragma once

#include <stdint.h>

#include <vector>

enum {
    KVOP_CLEAR = 1,     // llama_kv_cache_clear
    KVOP_RM    = 2,     // a..b = sequence range (a = -1: all), p0, p1   llama_kv_cache_seq_rm   (PipeInfer predicate)
    KVOP_CP    = 3,     // a=src, b..pad = destination range, p0, p1     llama_kv_cache_seq_cp   (head = 0)
    KVOP_KEEP  = 4,     // a=seq                                     llama_kv_cache_seq_keep
    KVOP_ALLOC = 5,     // a=head, b=n_tokens, p0=n (attended range), p1=descriptor slot: cells [head, head+n_tokens)
                        //   take pos / seq bits of the batch         llama_kv_cache_find_slot (resolved on the host)
    KVOP_ADVANCE = 6,   // a=n_tokens, b=first cell of the decode: head = b + a, wrap to 0 (a multi-chunk decode issues one ALLOC
                        //   per chunk, each moving head to its own first cell)        llama.cpp:5708-5713
    KVOP_SWEEP = 7,     // a=first seq, b=last seq, p0, p1: for i in [a, b]: RM(i, p0, p1); CP(0, i, p0, p1)
                        //   the per-sequence commit loop of speculative.cpp:603-612 collapsed into one op
    KVOP_SETHEAD = 8,   // a=head: the pre-decode heuristic (llama.cpp:5548-5550) resolved on the host
};

struct KvOp {
    int32_t op, a, b, p0, p1, pad;
};

struct KvMirror {
    uint32_t              size = 0, head = 0, used = 0, n = 0;
    int32_t               n_words = 0;
    std::vector<int32_t>  pos;
    std::vector<uint32_t> bits;     // [size][n_words]

    void init(uint32_t n_ctx, int32_t max_seq_ids);
    bool has(uint32_t cell, int32_t seq) const {
        return seq >= 0 && seq < n_words * 32 && ((bits[(size_t) cell * n_words + (seq >> 5)] >> (seq & 31)) & 1u);
    }
    bool empty(uint32_t cell) const;
    void clear();
    void seq_rm(int32_t first, int32_t last, int32_t p0, int32_t p1);
    void seq_cp(int32_t src, int32_t dfirst, int32_t dlast, int32_t p0, int32_t p1);
    void seq_keep(int32_t seq);
    void sweep(int32_t first, int32_t last, int32_t p0, int32_t p1);
    // find_slot: returns false when there is no run of n_tokens free cells; on success head points at the slot and the
    // cells are filled from pos / per-token bitset rows (tokbits: [n_tokens][n_words])
    bool    find_slot(int n_tokens, const int32_t * tpos, const uint32_t * tokbits);
    int32_t cell_max() const;
    void    apply(const KvOp & op, const int32_t * tpos, const uint32_t * tokbits);
};

// CPU-only test harness around the mirror and the op-log front end (kv_meta.cu; pipeline.cu for the replicated variant)
struct pib_kvtest {
    KvMirror          kv;
    std::vector<KvOp> pending, devops;
    bool              ship = false;     // publish every decode's record on the control ring (head of a pipeline)
};

// device replica
struct KvDevice {
    int32_t *  pos     = nullptr;   // [size]
    uint32_t * bits    = nullptr;   // [size][n_words]
    int32_t *  scalars = nullptr;   // {head, used, n, size}
    uint32_t   size    = 0;
    int32_t    n_words = 0;
};

int kv_device_init(KvDevice * kv, uint32_t n_ctx, int32_t n_words);
void kv_device_free(KvDevice * kv);
// replay n_ops ops (device array) ; ALLOC ops read the batch descriptor at desc_base + slot * desc_stride:
//   int32 pos[PIB_DESC_MAX_TOK] then uint32 tokbits[PIB_DESC_MAX_TOK][n_words]   (see context.cuh)
int kv_device_apply(const KvDevice * kv, const KvOp * ops_dev, int n_ops, const uint8_t * desc_base, size_t desc_stride,
                    cudaStream_t st);
// mask[t][i] = 1 when cell i is NOT visible to token t (reference llama.cpp:5200-5240), for i < n (read from scalars)
int kv_device_mask(const KvDevice * kv, const int32_t * tpos_dev, const uint32_t * tokbits_dev, int n_tokens,
                   uint8_t * mask_dev, cudaStream_t st);

// op-log front end (kv_meta.cu)
void kvlog_push(std::vector<KvOp> & pending, const KvOp & op);
void kvlog_push_ranged(std::vector<KvOp> & pending, int kind, int32_t a, int32_t b, int32_t c, int32_t p0, int32_t p1);
void kvlog_flush(KvMirror & kv, std::vector<KvOp> & pending, std::vector<KvOp> & devops);
bool kvlog_decode_begin(KvMirror & kv, std::vector<KvOp> & devops, int n_tokens, const int32_t * pos, const uint32_t * tokbits);
void kvlog_decode_end(KvMirror & kv, int n_tokens);
