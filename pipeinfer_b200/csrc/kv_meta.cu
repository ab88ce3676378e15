// kv_meta.cu -- host mirror of the KV-cell bookkeeping and the device-side metadata kernel that replays the op-log.
// Semantics and reference lines: see kv_meta.cuh.  Bit-exact against the oracle / the reference's llama_kv_cache_view
// (tests/test_kv_meta.py on the CPU for the mirror, tests/test_gpu_engine.py for the device replica).
#include "pib_common.cuh"
#include "kv_meta.cuh"
#include "desc.cuh"

#include <limits.h>
#include <string.h>

// ---------------------------------------------------------------------------------------------------------------------
// host mirror
void KvMirror::init(uint32_t n_ctx, int32_t max_seq_ids) {
    size    = n_ctx;
    n_words = (max_seq_ids + 31) / 32;
    if (n_words < 1) n_words = 1;
    pos.assign(n_ctx, -1);
    bits.assign((size_t) n_ctx * n_words, 0u);
    head = used = n = 0;
}

bool KvMirror::empty(uint32_t cell) const {
    const uint32_t * r = &bits[(size_t) cell * n_words];
    for (int w = 0; w < n_words; w++) if (r[w]) return false;
    return true;
}

void KvMirror::clear() {
    std::fill(pos.begin(), pos.end(), -1);
    std::fill(bits.begin(), bits.end(), 0u);
    head = 0;
    used = 0;
}

static inline void host_range(uint32_t * r, int W, int first, int last, bool set) {
    for (int s = first; s <= last;) {
        const int      w = s >> 5, lo = s & 31;
        const int      hi = (last >> 5) == w ? (last & 31) : 31;
        const uint32_t m  = (hi == 31 ? 0xffffffffu : ((1u << (hi + 1)) - 1u)) & ~((1u << lo) - 1u);
        if (w < W) { if (set) r[w] |= m; else r[w] &= ~m; }
        s = (w + 1) << 5;
    }
}

// seq_rm for every sequence in [first, last] (first < 0: all sequences), evaluated per cell in one pass.  Identical to
// the literal loop: within a run of seq_rm calls cells only lose bits, so "became empty at some point" == "is empty at
// the end", `used` drops once per emptied cell, and head = min(head, first empty cell in range).
void KvMirror::seq_rm(int32_t first, int32_t last, int32_t p0, int32_t p1) {
    uint32_t new_head = size;
    if (p0 < 0) p0 = 0;
    if (p1 < 0) p1 = INT_MAX;
    for (uint32_t i = 0; i < size; ++i) {
        const int32_t p = pos[i];
        if ((p >= p0 || p < 0) && p < p1) {
            // an empty cell (pos < 0 <=> no sequence bit: every op keeps that invariant) has nothing to lose; it only counts
            // for the head rule below.  Skipping its row is what makes the ~12 k-op commit of an accepted run cheap.
            if (p < 0) { if (new_head == size) new_head = i; continue; }
            uint32_t * r = &bits[(size_t) i * n_words];
            if (first < 0) memset(r, 0, sizeof(uint32_t) * n_words);
            else host_range(r, n_words, first, last, false);
            if (empty(i)) {
                if (p >= 0) used--;
                pos[i] = -1;
                if (new_head == size) new_head = i;
            }
        }
    }
    if (new_head != size && new_head < head) head = new_head;
}

// seq_cp(src -> d) for every d in [dfirst, dlast] in one pass (head = 0 as after any seq_cp)
void KvMirror::seq_cp(int32_t src, int32_t dfirst, int32_t dlast, int32_t p0, int32_t p1) {
    if (p0 < 0) p0 = 0;
    if (p1 < 0) p1 = INT_MAX;
    head = 0;
    if (dfirst < 0) return;
    for (uint32_t i = 0; i < size; ++i)
        if (has(i, src) && pos[i] >= p0 && pos[i] < p1) host_range(&bits[(size_t) i * n_words], n_words, dfirst, dlast, true);
}

void KvMirror::seq_keep(int32_t seq) {
    uint32_t new_head = size;
    for (uint32_t i = 0; i < size; ++i) {
        uint32_t * r = &bits[(size_t) i * n_words];
        if (!has(i, seq)) {
            if (pos[i] >= 0) used--;
            pos[i] = -1;
            memset(r, 0, sizeof(uint32_t) * n_words);
            if (new_head == size) new_head = i;
        } else {
            memset(r, 0, sizeof(uint32_t) * n_words);
            r[seq >> 5] = 1u << (seq & 31);
        }
    }
    if (new_head != size && new_head < head) head = new_head;
}

// for i in [first, last]: seq_rm(i, p0, p1); seq_cp(0, i, p0, p1) -- evaluated per cell in one pass.
// Per cell in the position range: cells holding sequence 0 end up with all of [first, last] set (each removed bit is
// re-added by the copy); other cells lose [first, last] and are reset when that empties them.  Sequence 0 itself is
// never touched (first >= 1), `used` is order independent, and `head` ends at 0 because the last op is a seq_cp.
void KvMirror::sweep(int32_t first, int32_t last, int32_t p0, int32_t p1) {
    if (first > last) return;
    if (first < 1) {                 // not the pattern this op was made for: fall back to the literal loop
        for (int32_t i = first; i <= last; i++) { seq_rm(i, i, p0, p1); seq_cp(0, i, i, p0, p1); }
        return;
    }
    int32_t q0 = p0 < 0 ? 0 : p0, q1 = p1 < 0 ? INT_MAX : p1;
    for (uint32_t i = 0; i < size; ++i) {
        const int32_t p  = pos[i];
        uint32_t *    r  = &bits[(size_t) i * n_words];
        const bool rm_in = (p >= q0 || p < 0) && p < q1;
        const bool cp_in = p >= q0 && p < q1;
        if (p < 0 || !rm_in) continue;        // empty cells have no bits to lose, cells outside the range are untouched (cp_in implies rm_in)
        const bool has0  = r[0] & 1u;
        for (int32_t s = first; s <= last; ) {
            const int w = s >> 5, lo = s & 31;
            const int hi = (last >> 5) == w ? (last & 31) : 31;
            const uint32_t m = (hi == 31 ? 0xffffffffu : ((1u << (hi + 1)) - 1u)) & ~((1u << lo) - 1u);
            if (w < n_words) {
                if (has0 && cp_in) r[w] |= m;          // removed (if in rm range) and copied back
                else if (rm_in)    r[w] &= ~m;
            }
            s = (w + 1) << 5;
        }
        if (rm_in && empty(i)) {
            if (p >= 0) used--;
            pos[i] = -1;
        }
    }
    head = 0;
}

bool KvMirror::find_slot(int n_tokens_i, const int32_t * tpos, const uint32_t * tokbits) {
    const uint32_t n_ctx = size, n_tokens = (uint32_t) n_tokens_i;
    if (n_tokens > n_ctx) return false;
    uint32_t n_tested = 0;
    while (true) {
        if (head + n_tokens > n_ctx) {
            n_tested += n_ctx - head;
            head = 0;
            continue;
        }
        bool found = true;
        for (uint32_t i = 0; i < n_tokens; i++) {
            if (pos[head + i] >= 0) {
                found = false;
                head += i + 1;
                n_tested += i + 1;
                break;
            }
        }
        if (found) break;
        if (n_tested >= n_ctx) return false;
    }
    for (uint32_t i = 0; i < n_tokens; i++) {
        pos[head + i] = tpos[i];
        uint32_t *       r = &bits[(size_t) (head + i) * n_words];
        const uint32_t * t = tokbits + (size_t) i * n_words;
        for (int w = 0; w < n_words; w++) r[w] |= t[w];
    }
    used += n_tokens;
    return true;
}

// host replay of one device-level op-log record (what kv_apply_kernel does on the replica)
void KvMirror::apply(const KvOp & op, const int32_t * tpos, const uint32_t * tokbits) {
    switch (op.op) {
        case KVOP_CLEAR:   clear(); break;
        case KVOP_RM:      seq_rm(op.a, op.b, op.p0, op.p1); break;
        case KVOP_CP:      seq_cp(op.a, op.b, op.pad, op.p0, op.p1); break;
        case KVOP_KEEP:    seq_keep(op.a); break;
        case KVOP_SWEEP:   sweep(op.a, op.b, op.p0, op.p1); break;
        case KVOP_SETHEAD: head = (uint32_t) op.a; break;
        case KVOP_ALLOC:
            head = (uint32_t) op.a;
            for (int i = 0; i < op.b; i++) {
                pos[head + i] = tpos[i];
                uint32_t * r = &bits[(size_t) (head + i) * n_words];
                for (int w = 0; w < n_words; w++) r[w] |= tokbits[(size_t) i * n_words + w];
            }
            used += (uint32_t) op.b;
            n = (uint32_t) op.p0;
            break;
        case KVOP_ADVANCE: head = (uint32_t) (op.b + op.a); if (head >= size) head = 0; break;
        default: break;
    }
}

int32_t KvMirror::cell_max() const {
    for (uint32_t i = size - 1; i > 0; --i)
        if (pos[i] >= 0 && !empty(i)) return (int32_t) (i + 1);
    return 0;
}

// ---------------------------------------------------------------------------------------------------------------------
// device replica
#define KVT 1024

__device__ __forceinline__ bool row_empty(const uint32_t * r, int W) {
    for (int w = 0; w < W; w++) if (r[w]) return false;
    return true;
}

// set or clear the bits [first, last] of a cell row
__device__ __forceinline__ void dev_range(uint32_t * r, int W, int first, int last, bool set) {
    for (int s = first; s <= last;) {
        const int      w = s >> 5, lo = s & 31;
        const int      hi = (last >> 5) == w ? (last & 31) : 31;
        const uint32_t m  = (hi == 31 ? 0xffffffffu : ((1u << (hi + 1)) - 1u)) & ~((1u << lo) - 1u);
        if (w < W) { if (set) r[w] |= m; else r[w] &= ~m; }
        s = (w + 1) << 5;
    }
}

// One CTA replays the log in order; every op is a parallel sweep over the cells followed by a block-wide reduction of
// (first emptied cell, number of cells that became empty) where the op needs them.
__global__ void __launch_bounds__(KVT) kv_apply_kernel(int32_t * __restrict__ pos, uint32_t * __restrict__ bits,
                                                       int32_t * __restrict__ scalars, int size, int W,
                                                       const KvOp * __restrict__ ops, int n_ops,
                                                       const uint8_t * __restrict__ desc_base, size_t desc_stride) {
    __shared__ int s_min[KVT / 32];
    __shared__ int s_cnt[KVT / 32];
    __shared__ int s_head, s_used, s_n;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { s_head = scalars[0]; s_used = scalars[1]; s_n = scalars[2]; }
    __syncthreads();
    for (int k = 0; k < n_ops; k++) {
        const KvOp op       = ops[k];
        int        my_first = size;     // first cell (by index) this thread reset
        int        my_cnt   = 0;        // used cells this thread reset
        const int  q0 = op.p0 < 0 ? 0 : op.p0, q1 = op.p1 < 0 ? INT_MAX : op.p1;
        if (op.op == KVOP_CLEAR) {
            for (int i = tid; i < size; i += KVT) {
                pos[i] = -1;
                for (int w = 0; w < W; w++) bits[(size_t) i * W + w] = 0u;
            }
        } else if (op.op == KVOP_RM) {
            const int first = op.a, last = op.b;
            for (int i = tid; i < size; i += KVT) {
                const int p = pos[i];
                if ((p >= q0 || p < 0) && p < q1) {
                    uint32_t * r = bits + (size_t) i * W;
                    if (first < 0) { for (int w = 0; w < W; w++) r[w] = 0u; }
                    else dev_range(r, W, first, last, false);
                    if (row_empty(r, W)) {
                        if (p >= 0) my_cnt++;
                        pos[i] = -1;
                        if (i < my_first) my_first = i;
                    }
                }
            }
        } else if (op.op == KVOP_CP) {
            const int src = op.a, dfirst = op.b, dlast = op.pad;
            if (dfirst >= 0 && src >= 0 && src < W * 32) {
                for (int i = tid; i < size; i += KVT) {
                    const int  p = pos[i];
                    uint32_t * r = bits + (size_t) i * W;
                    if (((r[src >> 5] >> (src & 31)) & 1u) && p >= q0 && p < q1) dev_range(r, W, dfirst, dlast, true);
                }
            }
        } else if (op.op == KVOP_KEEP) {
            const int seq = op.a;
            for (int i = tid; i < size; i += KVT) {
                uint32_t * r   = bits + (size_t) i * W;
                const bool has = seq >= 0 && seq < W * 32 && ((r[seq >> 5] >> (seq & 31)) & 1u);
                for (int w = 0; w < W; w++) r[w] = 0u;
                if (!has) {
                    if (pos[i] >= 0) my_cnt++;
                    pos[i] = -1;
                    if (i < my_first) my_first = i;
                } else {
                    r[seq >> 5] = 1u << (seq & 31);
                }
            }
        } else if (op.op == KVOP_SWEEP) {
            const int first = op.a, last = op.b;
            for (int i = tid; i < size; i += KVT) {
                const int  p     = pos[i];
                uint32_t * r     = bits + (size_t) i * W;
                const bool rm_in = (p >= q0 || p < 0) && p < q1;
                const bool cp_in = p >= q0 && p < q1;
                const bool has0  = r[0] & 1u;
                for (int s = first; s <= last;) {
                    const int      w = s >> 5, lo = s & 31;
                    const int      hi = (last >> 5) == w ? (last & 31) : 31;
                    const uint32_t m  = (hi == 31 ? 0xffffffffu : ((1u << (hi + 1)) - 1u)) & ~((1u << lo) - 1u);
                    if (w < W) {
                        if (has0 && cp_in) r[w] |= m;
                        else if (rm_in)    r[w] &= ~m;
                    }
                    s = (w + 1) << 5;
                }
                if (rm_in && row_empty(r, W)) {
                    if (p >= 0) my_cnt++;
                    pos[i] = -1;
                }
            }
        } else if (op.op == KVOP_ALLOC) {
            const uint8_t *  d    = desc_base + (size_t) op.p1 * desc_stride;
            const int32_t *  tpos = reinterpret_cast<const int32_t *>(d + PIB_DESC_POS_OFF);
            const uint32_t * tb   = reinterpret_cast<const uint32_t *>(d + PIB_DESC_BITS_OFF);
            for (int j = tid; j < op.b * W; j += KVT) {
                const int t = j / W, w = j % W;
                bits[(size_t) (op.a + t) * W + w] |= tb[(size_t) t * W + w];
                if (w == 0) pos[op.a + t] = tpos[t];
            }
        }
        // block reductions only where the op's scalars depend on the sweep
        if (op.op == KVOP_RM || op.op == KVOP_KEEP || op.op == KVOP_SWEEP) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                my_first = min(my_first, __shfl_xor_sync(0xffffffffu, my_first, o));
                my_cnt += __shfl_xor_sync(0xffffffffu, my_cnt, o);
            }
            if (lane == 0) { s_min[warp] = my_first; s_cnt[warp] = my_cnt; }
            __syncthreads();
            if (tid == 0) {
                int mn = size, cnt = 0;
                for (int w = 0; w < KVT / 32; w++) { mn = min(mn, s_min[w]); cnt += s_cnt[w]; }
                s_used -= cnt;
                if (op.op == KVOP_SWEEP) s_head = 0;
                else if (mn != size && mn < s_head) s_head = mn;
            }
        } else if (tid == 0) {
            if (op.op == KVOP_CLEAR) { s_head = 0; s_used = 0; }
            else if (op.op == KVOP_CP) s_head = 0;
            else if (op.op == KVOP_SETHEAD) s_head = op.a;
            else if (op.op == KVOP_ALLOC) { s_head = op.a; s_used += op.b; s_n = op.p0; }
            else if (op.op == KVOP_ADVANCE) { s_head = op.b + op.a; if (s_head >= size) s_head = 0; }
        }
        __syncthreads();   // also orders this op's global writes before the next op's reads (same CTA)
    }
    if (tid == 0) { scalars[0] = s_head; scalars[1] = s_used; scalars[2] = s_n; scalars[3] = size; }
}

__global__ void kv_mask_kernel(const int32_t * __restrict__ pos, const uint32_t * __restrict__ bits,
                               const int32_t * __restrict__ scalars, int W, const int32_t * __restrict__ tpos,
                               const uint32_t * __restrict__ tokbits, int n_ctx, uint8_t * __restrict__ mask) {
    pib_pdl_wait();
    pib_pdl_trigger();
    const int t  = blockIdx.y;
    const int n  = scalars[2];
    const int tp = tpos[t];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t * r = bits + (size_t) i * W;
        const uint32_t * q = tokbits + (size_t) t * W;
        uint32_t         any = 0;
        for (int w = 0; w < W; w++) any |= r[w] & q[w];
        mask[(size_t) t * n_ctx + i] = (uint8_t) (!any || pos[i] > tp);
    }
}

int kv_device_init(KvDevice * kv, uint32_t n_ctx, int32_t n_words) {
    kv->size    = n_ctx;
    kv->n_words = n_words;
    PIB_CHECK(cudaMalloc(&kv->pos, sizeof(int32_t) * n_ctx));
    PIB_CHECK(cudaMalloc(&kv->bits, sizeof(uint32_t) * (size_t) n_ctx * n_words));
    PIB_CHECK(cudaMalloc(&kv->scalars, 64));
    PIB_CHECK(cudaMemset(kv->pos, 0xff, sizeof(int32_t) * n_ctx));
    PIB_CHECK(cudaMemset(kv->bits, 0, sizeof(uint32_t) * (size_t) n_ctx * n_words));
    PIB_CHECK(cudaMemset(kv->scalars, 0, 64));
    return PIB_OK;
}

void kv_device_free(KvDevice * kv) {
    if (kv->pos) cudaFree(kv->pos);
    if (kv->bits) cudaFree(kv->bits);
    if (kv->scalars) cudaFree(kv->scalars);
    kv->pos = nullptr; kv->bits = nullptr; kv->scalars = nullptr;
}

int kv_device_apply(const KvDevice * kv, const KvOp * ops_dev, int n_ops, const uint8_t * desc_base, size_t desc_stride,
                    cudaStream_t st) {
    if (n_ops <= 0) return PIB_OK;
    kv_apply_kernel<<<1, KVT, 0, st>>>(kv->pos, kv->bits, kv->scalars, (int) kv->size, kv->n_words, ops_dev, n_ops,
                                       desc_base, desc_stride);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

int kv_device_mask(const KvDevice * kv, const int32_t * tpos_dev, const uint32_t * tokbits_dev, int n_tokens,
                   uint8_t * mask_dev, cudaStream_t st) {
    dim3 grid((kv->size + 255) / 256, n_tokens);
    pib_launch(kv_mask_kernel, grid, dim3(256), 0, st, kv->pos, kv->bits, kv->scalars, kv->n_words, tpos_dev, tokbits_dev,
                                          (int) kv->size, mask_dev);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// op-log front end shared by the engine and the CPU-only test hooks.
// kvlog_push: the public llama_kv_cache_seq_* calls only append; RM(i, p0, p1) directly followed by CP(0 -> i, p0, p1)
// for consecutive i -- the scheduler's commit loop, speculative.cpp:603-612, ~2*max_seq calls -- collapses into one
// KVOP_SWEEP with identical results.
void kvlog_push(std::vector<KvOp> & q, const KvOp & op_in) {
    KvOp op = op_in;
    if (op.op == KVOP_RM && op.a >= 0) op.b = op.a;          // single sequence == range [a, a]
    if (op.op == KVOP_CP) op.pad = op.b;                     // single destination == range [b, b]
    if (!q.empty()) {
        KvOp & last = q.back();
        const bool same = last.p0 == op.p0 && last.p1 == op.p1;
        if (op.op == KVOP_CP && op.a == 0 && op.b >= 1 && same && last.op == KVOP_RM && last.a == op.b && last.b == op.b) {
            // RM(i) directly followed by CP(0 -> i): start or extend a sweep
            if (q.size() >= 2) {
                KvOp & prev = q[q.size() - 2];
                if (prev.op == KVOP_SWEEP && prev.b + 1 == op.b && prev.p0 == op.p0 && prev.p1 == op.p1) {
                    prev.b = op.b;
                    q.pop_back();
                    return;
                }
            }
            last.op = KVOP_SWEEP;
            return;
        }
        // RM(i), RM(i+1), ...  /  CP(s -> i), CP(s -> i+1), ...  with one position range: one ranged op
        if (op.op == KVOP_RM && op.a >= 0 && last.op == KVOP_RM && last.a >= 0 && same && last.b + 1 == op.a) { last.b = op.b; return; }
        if (op.op == KVOP_CP && last.op == KVOP_CP && same && last.a == op.a && last.pad + 1 == op.b &&
            (op.a < last.b || op.a > op.b)) { last.pad = op.b; return; }
    }
    q.push_back(op);
}

// the ranged records the coalescing above arrives at, pushed directly: what the scheduler's per-id loops (3 000 ids, 12 000 calls
// per accepted run) collapse to anyway.  kind = KVOP_RM: sequences [a, b]; KVOP_CP: a -> [b, c] (a outside [b, c]); KVOP_SWEEP: [a, b], a >= 1.
void kvlog_push_ranged(std::vector<KvOp> & q, int kind, int32_t a, int32_t b, int32_t c, int32_t p0, int32_t p1) {
    if (a > b && kind != KVOP_CP) return;
    if (kind == KVOP_CP && b > c) return;
    if (kind == KVOP_CP && a >= b && a <= c) {               // the source inside the destinations: the literal loop skips src == dst
        if (a > b) kvlog_push_ranged(q, kind, a, b, a - 1, p0, p1);
        if (a < c) kvlog_push_ranged(q, kind, a, a + 1, c, p0, p1);
        return;
    }
    if (kind == KVOP_SWEEP && a < 1) {
        // the closed form of the sweep (host mirror and device kernel alike) assumes sequence 0 is the untouched source; ids
        // below 1 take the literal per-id records instead, so that the mirror and the device replica cannot diverge
        // (negative ids name no sequence; for id 0 the literal loop is seq_rm(0) followed by a copy onto itself, a no-op)
        if (b >= 0) q.push_back({ KVOP_RM, 0, 0, p0, p1, 0 });
        if (b < 1) return;
        a = 1;
    }
    q.push_back({ kind, a, b, p0, p1, kind == KVOP_CP ? c : 0 });
}

void kvlog_flush(KvMirror & kv, std::vector<KvOp> & pending, std::vector<KvOp> & devops) {
    for (const KvOp & op : pending) {
        switch (op.op) {
            case KVOP_CLEAR: kv.clear(); break;
            case KVOP_RM:    kv.seq_rm(op.a, op.b, op.p0, op.p1); break;
            case KVOP_CP:    kv.seq_cp(op.a, op.b, op.pad, op.p0, op.p1); break;
            case KVOP_KEEP:  kv.seq_keep(op.a); break;
            case KVOP_SWEEP: kv.sweep(op.a, op.b, op.p0, op.p1); break;
        }
        devops.push_back(op);
    }
    pending.clear();
}

// bookkeeping of one decode before the compute is enqueued (reference llama.cpp:5548-5571): the head-reset heuristic,
// find_slot, n = min(n_ctx, max(32, cell_max)).  Returns false when there is no slot (the mirror's head then stays
// where the failed search left it, like the reference's).
bool kvlog_decode_begin(KvMirror & kv, std::vector<KvOp> & devops, int n_tokens, const int32_t * pos, const uint32_t * tokbits) {
    if (kv.head > kv.used + 2 * (uint32_t) n_tokens) { kv.head = 0; devops.push_back({ KVOP_SETHEAD, 0, 0, 0, 0, 0 }); }
    if (!kv.find_slot(n_tokens, pos, tokbits)) {
        devops.push_back({ KVOP_SETHEAD, (int32_t) kv.head, 0, 0, 0, 0 });
        return false;
    }
    const int32_t cm = kv.cell_max();
    const int32_t n  = cm > 32 ? cm : 32;
    kv.n = (uint32_t) (n < (int32_t) kv.size ? n : (int32_t) kv.size);
    return true;
}

// after the compute has been enqueued (llama.cpp:5708-5713)
void kvlog_decode_end(KvMirror & kv, int n_tokens) {
    kv.head += (uint32_t) n_tokens;
    if (kv.head >= kv.size) kv.head = 0;
}

// ---- CPU-only test hooks (no CUDA call): drive the host mirror through the same front end the engine uses -------------
void pib_kvtest_ship_decode(pib_kvtest * t, int n_tokens, const int32_t * pos, const uint32_t * tokbits);   // pipeline.cu

extern "C" void * pib_kvtest_new(uint32_t n_ctx, int32_t max_seq) {
    pib_kvtest * t = new pib_kvtest();
    t->kv.init(n_ctx, max_seq);
    return t;
}
extern "C" void pib_kvtest_free(void * h) { delete (pib_kvtest *) h; }
extern "C" void pib_kvtest_op(void * h, int op, int a, int b, int p0, int p1) { kvlog_push(((pib_kvtest *) h)->pending, { op, a, b, p0, p1, 0 }); }
extern "C" int pib_kvtest_pending(void * h) { return (int) ((pib_kvtest *) h)->pending.size(); }
extern "C" void pib_kvtest_op_ranged(void * h, int kind, int a, int b, int c, int p0, int p1) { kvlog_push_ranged(((pib_kvtest *) h)->pending, kind, a, b, c, p0, p1); }
// the pending (not yet flushed) records, 6 int32 each: op, a, b, p0, p1, pad
extern "C" int pib_kvtest_dump_pending(void * h, int32_t * out, int cap_records) {
    const std::vector<KvOp> & q = ((pib_kvtest *) h)->pending;
    const int n = (int) q.size() < cap_records ? (int) q.size() : cap_records;
    memcpy(out, q.data(), sizeof(KvOp) * (size_t) n);
    return (int) q.size();
}
extern "C" int pib_kvtest_decode(void * h, int n_tokens, const int32_t * pos, const int32_t * n_seq_id, const int32_t * const * seq_id) {
    pib_kvtest * t = (pib_kvtest *) h;
    kvlog_flush(t->kv, t->pending, t->devops);
    const int W = t->kv.n_words;
    std::vector<uint32_t> tb((size_t) n_tokens * W, 0u);
    for (int i = 0; i < n_tokens; i++)
        for (int j = 0; j < n_seq_id[i]; j++) tb[(size_t) i * W + (seq_id[i][j] >> 5)] |= 1u << (seq_id[i][j] & 31);
    if (!kvlog_decode_begin(t->kv, t->devops, n_tokens, pos, tb.data())) return -1;
    if (t->ship) pib_kvtest_ship_decode(t, n_tokens, pos, tb.data());
    kvlog_decode_end(t->kv, n_tokens);
    return 0;
}
extern "C" void pib_kvtest_state(void * h, int32_t * scalars3, int32_t * pos, uint32_t * bits) {
    pib_kvtest * t = (pib_kvtest *) h;
    kvlog_flush(t->kv, t->pending, t->devops);
    scalars3[0] = (int32_t) t->kv.head; scalars3[1] = (int32_t) t->kv.used; scalars3[2] = (int32_t) t->kv.n;
    memcpy(pos, t->kv.pos.data(), sizeof(int32_t) * t->kv.size);
    memcpy(bits, t->kv.bits.data(), sizeof(uint32_t) * t->kv.bits.size());
}
