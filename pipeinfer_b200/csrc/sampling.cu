// sampling.cu -- device-side token acceptance inputs (K10 / R9): per-row argmax and sorted top-8 probabilities.
//
// The reference pulls n_vocab floats per sampled row to the host and runs std::sort over the whole vocabulary there
// (common/sampling.cpp:131, llama.cpp:7019-7046 llama_sample_softmax, :7496 llama_sample_token_greedy).  Here the
// logits stay in HBM; per row the host receives 8 x {id, logit, p} (the llama_token_data the scheduler consumes at
// speculative.cpp:977-1010), entry 0 being the greedy token:
//   id    first index of the maximum (std::max_element semantics)
//   p_i   expf(l_i - max) / sum_j expf(l_j - max); the sum is accumulated in double (the reference adds floats in
//         sorted order), so p agrees to ~1e-7 relative -- the ids are exact.
// logit_bias entries (e.g. EOS = -inf under --ignore-eos, common.cpp:1061-1063 / sampling.cpp:134-136) are applied
// in the kernel before anything else, in place, as the reference does.
#include "engine.cuh"

#define SMP_THREADS 1024
#define SMP_TOPK    8

__global__ void __launch_bounds__(SMP_THREADS) sample_topk_kernel(float * __restrict__ logits, int n_vocab,
                                                                  const int32_t * __restrict__ bias_ids, const float * __restrict__ bias_val,
                                                                  int n_bias, pib_row_result * __restrict__ out,
                                                                  const uint8_t * __restrict__ desc, const int * __restrict__ cancel) {
    (void) desc;
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    __shared__ float  s_val[SMP_THREADS / 32];
    __shared__ int    s_idx[SMP_THREADS / 32];
    __shared__ double s_sum[SMP_THREADS / 32];
    __shared__ float  s_best;
    __shared__ int    s_besti;
    float *   l    = logits + (size_t) blockIdx.x * n_vocab;
    const int tid  = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < n_bias; k += SMP_THREADS) l[bias_ids[k]] += bias_val[k];
    __syncthreads();
    // one pass: every thread keeps the 8 largest of its strided elements, sorted (value descending, index ascending: the
    // indices a thread sees only grow, so strict comparisons keep the earlier one of two equal values in front)
    float tv[SMP_TOPK];
    int   ti[SMP_TOPK];
#pragma unroll
    for (int j = 0; j < SMP_TOPK; j++) { tv[j] = -INFINITY; ti[j] = 0x7fffffff; }
#pragma unroll 4
    for (int i = tid; i < n_vocab; i += SMP_THREADS) {
        const float v = l[i];
        if (v > tv[SMP_TOPK - 1] || ti[SMP_TOPK - 1] == 0x7fffffff) {
            tv[SMP_TOPK - 1] = v; ti[SMP_TOPK - 1] = i;
#pragma unroll
            for (int j = SMP_TOPK - 1; j > 0; j--) {
                if (tv[j] > tv[j - 1] || ti[j - 1] == 0x7fffffff) {
                    const float fv = tv[j]; tv[j] = tv[j - 1]; tv[j - 1] = fv;
                    const int   fi = ti[j]; ti[j] = ti[j - 1]; ti[j - 1] = fi;
                }
            }
        }
    }
    // eight rounds over the heads of the per-thread lists (std::max_element semantics: first index of the maximum)
    float max_l = 0.f;
    for (int k = 0; k < SMP_TOPK; k++) {
        float bv = tv[0];
        int   bi = ti[0];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int   oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            bv = s_val[lane]; bi = s_idx[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int   oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (oi != 0x7fffffff && (bi == 0x7fffffff || ov > bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
            }
            if (lane == 0) {
                s_best = bv; s_besti = bi;
                out[blockIdx.x].id[k]    = bi == 0x7fffffff ? 0 : bi;
                out[blockIdx.x].logit[k] = bv;
            }
        }
        __syncthreads();
        if (ti[0] == s_besti && s_besti != 0x7fffffff) {          // the winner pops its head
#pragma unroll
            for (int j = 0; j < SMP_TOPK - 1; j++) { tv[j] = tv[j + 1]; ti[j] = ti[j + 1]; }
            tv[SMP_TOPK - 1] = -INFINITY; ti[SMP_TOPK - 1] = 0x7fffffff;
        }
        if (k == 0) max_l = s_best;
    }
    double sum = 0.0;
    for (int i = tid; i < n_vocab; i += SMP_THREADS) sum += (double) expf(l[i] - max_l);
    sum = warp_sum_d(sum);
    if (lane == 0) s_sum[warp] = sum;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < SMP_THREADS / 32; w++) t += s_sum[w];
        for (int k = 0; k < SMP_TOPK; k++) out[blockIdx.x].p[k] = (float) ((double) expf(out[blockIdx.x].logit[k] - max_l) / t);
    }
}

int pib_sample_topk(float * logits_dev, int n_rows, int n_vocab, const int32_t * bias_ids_dev, const float * bias_val_dev,
                    int n_bias, void * out_dev, const uint8_t * desc, const int * cancel, cudaStream_t st) {
    pib_launch(sample_topk_kernel, dim3(n_rows), dim3(SMP_THREADS), 0, st, logits_dev, n_vocab, bias_ids_dev, bias_val_dev, n_bias,
                                                       (pib_row_result *) out_dev, desc, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

// ---- acceptance walk (reference speculative.cpp:439-560, greedy) ----------------------------------------------------------------
// The verification loop of the scheduler, on the device that holds the sampled rows: start on branch `keep` at depth 0; sample the
// target's token at that branch's row; every active branch whose drafted token at this depth differs is dropped, the first that
// agrees becomes the branch to follow; stop at the first depth where none agrees.  Only the ids sampled along the way (and the
// branch kept) cross to the host: 8 + 4 * n_ids bytes instead of 96 bytes per row.
__global__ void accept_walk_kernel(const pib_row_result * __restrict__ rows, const uint8_t * __restrict__ desc, pib_accept_result * __restrict__ out,
                                   const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (threadIdx.x != 0) return;
    const int32_t * hdr = reinterpret_cast<const int32_t *>(desc + PIB_DESC_HDR_OFF);
    const int np = hdr[PIB_DH_TREE_NP];
    if (np <= 0 || PIB_CANCELLED(cancel)) { out->n_ids = 0; out->s_keep = 0; return; }
    const int32_t * len = reinterpret_cast<const int32_t *>(desc + PIB_DESC_TREE_OFF);
    const int32_t * trow = len + PIB_TREE_NP, * ttok = trow + PIB_TREE_NP * PIB_TREE_DEPTH;
    int  keep = hdr[PIB_DH_TREE_KEEP], n = 0, i = 0;
    bool act[PIB_TREE_NP];
    for (int s = 0; s < PIB_TREE_NP; s++) act[s] = s < np && len[s] >= 0;
    while (n < PIB_CHUNK + 2 && i < PIB_TREE_DEPTH) {
        const int id = rows[trow[keep * PIB_TREE_DEPTH + i]].id[0];
        out->ids[n++] = id;
        if (!hdr[PIB_DH_TREE_SPEC]) break;
        bool matches = false;
        for (int s = 0; s < np; s++) {
            if (!act[s]) continue;
            if (i < len[s] && id == ttok[s * PIB_TREE_DEPTH + i]) { if (!matches) keep = s; matches = true; }
            else act[s] = false;
        }
        if (!matches) break;
        i++;
    }
    out->n_ids = n;
    out->s_keep = keep;
}

int pib_accept_walk(const void * rows_dev, const uint8_t * desc, void * out_dev, const int * cancel, cudaStream_t st) {
    pib_launch(accept_walk_kernel, dim3(1), dim3(32), 0, st, (const pib_row_result *) rows_dev, desc, (pib_accept_result *) out_dev, cancel);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}
