// attention.cu -- RoPE + KV store (K5, K6) and multi-sequence masked attention over the KV cache (K7).
//
// Reference op chain (llama.cpp:3913-3944 build_llama, :3546 llm_build_kv_store, :3698 llm_build_kqv; CPU kernels
// ggml.c:11006 rope, :6528 dup f32->f16, :1200 vec_dot_f16, :9937 scale, :10551 soft_max):
//   Q,K <- rope(pos)             theta_i = pos * base^(-2i/d) by repeated multiply, cosf/sinf of the host libm
//   cache K (f16), cache V (f16)  at cells kv_head + t
//   S = (K . fp16(Q)) * 1/sqrt(d)   fp32 accumulate of exact fp16 products
//   S += mask (0 / -inf)            cell visible iff it shares a sequence id with the token and cell.pos <= token.pos
//   P = softmax(S)                  exp through the fp16 table on fp16(S - max), sum in double, P = val * (float)(1/sum)
//   O = V . fp16(P)
// Everything that is rounded in the reference is rounded identically here (fp16 Q, fp16 K/V, fp16 exp table, fp16 P);
// the double sum of fp16 values is exact in any order, so only the fp32 dot-product association differs.
//
// cos/sin come from a table built on the host with the reference's recurrence (rope_table_build), so RoPE itself is
// bit-exact.  The V cache is kept [cell][n_embd_gqa] (not transposed as in the reference, llama.cpp:3569): the layout
// is private and this one is coalesced for both the store and the P.V pass.
#include "pib_common.cuh"
#include "desc.cuh"

#include <math.h>

#include <algorithm>
#include <vector>

// ---- RoPE table -----------------------------------------------------------------------------------------------------
// tab[pos][i] = {cos(theta), sin(theta)}, theta as in reference rope_f32 (ggml.c:11062, 11111-11123) with ext_factor 0:
// theta_base = pos; per pair: theta = freq_scale * theta_base; theta_base *= theta_scale.  mode 2 (NeoX) pre-multiplies
// theta_base by freq_scale once more (ggml.c:11137).
void rope_table_build(std::vector<float2> & tab, int n_pos, int n_dims, int mode, float freq_base, float freq_scale) {
    const float theta_scale = powf(freq_base, -2.0f / n_dims);
    tab.resize((size_t) n_pos * (n_dims / 2));
    for (int p = 0; p < n_pos; p++) {
        float theta_base = (float) p;
        if (mode & 2) theta_base *= freq_scale;
        for (int i = 0; i < n_dims / 2; i++) {
            const float theta = freq_scale * theta_base;
            tab[(size_t) p * (n_dims / 2) + i] = make_float2(cosf(theta), sinf(theta));
            theta_base *= theta_scale;
        }
    }
}

// ---- attention ------------------------------------------------------------------------------------------------------
// grid (n_head, n_tokens), 256 threads; one CTA per (query head, token).  Every fp32 sum is formed in the reference's order
// (ggml_vec_dot_f16 for both K.Q and V.P), so the output is bit-identical to the CPU's: phase 1 gives every thread whole cells
// (a K row in 128-bit loads, the 32 lane accumulators of the reference in registers); phase 3 gives every thread a pair of
// output dims and a share of the 32 accumulators, each a sequential chain over its cells.  K/V of a KV head are re-read by its
// G query heads through L2 -- a few MB per layer at these context lengths.  Masked cells are never loaded.
#define ATQ_THREADS 256
#define ATQ_VCAP    256          /* V rows staged in shared memory (cells 0 .. ATQ_VCAP - 1); later cells are read from global memory */
// RoPE, KV store and attention in one launch.  A CTA rotates its own query head itself, rotates / converts the K and V rows of
// ALL tokens of the batch for its KV head into shared memory (<= 16 rows: the batch's cells are the contiguous range starting at
// the descriptor's kv head, and a token of a tree batch attends to earlier tokens of the same batch), and the first query head
// of each KV group also writes its token's K / V row to the cache.  Cells of the batch are therefore never read from global
// memory, cells outside it never written: no ordering between CTAs is needed.  Values and roundings are the reference's
// (rope_f32 ggml.c:11006, dup f32 -> f16 :6528).
//
// The kernel is a chain of dependent memory round trips, not of arithmetic, so everything that does not depend on the QKV
// projection is fetched BEFORE the wait on it: the descriptor words and positions, the RoPE entries, the mask bytes and K rows of
// the first 128 cells (registers), and the V rows of up to ATQ_VCAP cells (cp.async into shared memory).  All of it was final
// before this launch's predecessor started: cache rows of earlier decodes, and the mask / scalars of the pass's first kernels.
template <int HD>
__global__ void __launch_bounds__(ATQ_THREADS, 2) attention_qh_kernel(const float * __restrict__ qkv, const float * __restrict__ vsep,
                                                                   const uint8_t * __restrict__ desc, const float2 * __restrict__ rope, int rope_n_pos, int mode,
                                                                   __half * __restrict__ kcache, __half * __restrict__ vcache, const uint8_t * __restrict__ mask,
                                                                   const int32_t * __restrict__ scalars, const uint16_t * __restrict__ exp_lut,
                                                                   int E, int GQA, int n_ctx, float kq_scale, float * __restrict__ out,
                                                                   const int * __restrict__ cancel, int gdiv, int vcap, int early) {
    // early == 0: the launch before this one may still be writing cache rows this one reads (the attention of the previous chunk
    // of a wide prompt pass, engine.cu run_prompt_pass): nothing is fetched ahead of the wait then
    if (!early) pib_pdl_wait();
    extern __shared__ float smem[];
    __shared__ __align__(16) float sq[HD];
    __shared__ float  s_redf[ATQ_THREADS / 32];
    __shared__ double s_redd[ATQ_THREADS / 32];
    __shared__ __align__(16) __half kb[PIB_CHUNK][HD], vb[PIB_CHUNK][HD];      // this KV head's rows of the batch's tokens
    const int head = blockIdx.x, kvh = head / gdiv, t = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int half = HD / 2, CH = HD * 2 / 16;                             // pairs per head; 16-byte chunks per cache row
    float *         s    = smem;                       // [n] scores -> probabilities
    float *         red  = smem + n_ctx;               // [32 accumulators][HD]
    __half *        vs   = reinterpret_cast<__half *>(smem + n_ctx + 32 * HD);  // [vcap][HD] staged V rows
    const uint8_t * mrow = mask + (size_t) t * n_ctx;
    const int32_t * hdr  = reinterpret_cast<const int32_t *>(desc + PIB_DESC_HDR_OFF);
    const int32_t * tpos = reinterpret_cast<const int32_t *>(desc + PIB_DESC_POS_OFF);
    const int       rstride = E + (vsep ? 1 : 2) * GQA;

    // ---- before the wait ---------------------------------------------------------------------------------------------
    const int n = min(scalars[2], n_ctx), ntok = hdr[PIB_DH_NTOK], cell0 = hdr[PIB_DH_KVHEAD];
    const __half * vhead = vcache + (size_t) kvh * HD;
    const __half * khead = kcache + (size_t) kvh * HD;
    {
        const int      nst = min(n, vcap);
        const uint32_t vs_a = (uint32_t) __cvta_generic_to_shared(vs);
        for (int i = tid; i < nst * CH; i += ATQ_THREADS) {
            const int c = i / CH, ch = i - c * CH;
            if (c >= cell0 && c < cell0 + ntok) continue;                      // the batch's own rows come from vb
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(vs_a + (uint32_t) (c * HD * 2 + ch * 16)), "l"(vhead + (size_t) c * GQA + ch * 8) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    // phase 1 works on cell tid / 2 (+ 128 per round), two lanes per cell: round 0's mask byte and K half-row go to registers now
    const int hsel = tid & 1;
    uint4     kreg[HD / 16];
    bool      live0;
    {
        const int  i = tid >> 1, ic = max(min(i, n - 1), 0), bi = ic - cell0;
        const bool inb = bi >= 0 && bi < ntok;
        const uint4 * kr = reinterpret_cast<const uint4 *>(khead + (size_t) ic * GQA);
#pragma unroll
        for (int q = 0; q < HD / 16; q++) kreg[q] = inb ? make_uint4(0, 0, 0, 0) : kr[(q >> 1) * 4 + 2 * (q & 1) + hsel];
        live0 = i < n && !mrow[ic];
    }
    // RoPE entries of this thread's first two rotation items (covers batches of up to 2 * 256 / half - 1 tokens)
    float2 cs_pre[2];
#pragma unroll
    for (int it = 0; it < 2; it++) {
        const int i = tid + it * ATQ_THREADS, r = i / half, j = i - r * half;
        cs_pre[it] = make_float2(1.f, 0.f);
        if (r <= ntok) cs_pre[it] = rope[(size_t) min(tpos[r < ntok ? r : t], rope_n_pos - 1) * half + j];
    }

    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) { asm volatile("cp.async.wait_all;" ::: "memory"); return; }

    // ---- rotate: pair j of a head = elements (2j, 2j+1), or (j, j + HD/2) in NeoX mode; unfused multiplies as the reference (-std=c11)
    {
        int it = 0;
        for (int i = tid; i < (ntok + 1) * half; i += ATQ_THREADS, it++) {
            const int     r = i / half, j = i - r * half;              // r < ntok: K row of batch token r; r == ntok: this CTA's query
            const int     tok = r < ntok ? r : t;
            const float * src = qkv + (size_t) tok * rstride + (r < ntok ? E + (size_t) kvh * HD : (size_t) head * HD);
            const float2  cs  = it == 0 ? cs_pre[0] : it == 1 ? cs_pre[1] : rope[(size_t) min(tpos[tok], rope_n_pos - 1) * half + j];
            const int     i0 = (mode & 2) ? j : 2 * j, i1 = (mode & 2) ? j + half : 2 * j + 1;
            const float   x0 = src[i0], x1 = src[i1];
            const __half  y0 = __float2half_rn(__fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y)));
            const __half  y1 = __float2half_rn(__fadd_rn(__fmul_rn(x0, cs.y), __fmul_rn(x1, cs.x)));
            if (r < ntok) { kb[r][i0] = y0; kb[r][i1] = y1; }
            else          { sq[i0] = __half2float(y0); sq[i1] = __half2float(y1); }
        }
        for (int i = tid; i < ntok * HD; i += ATQ_THREADS) {
            const int r = i / HD, d = i - r * HD;
            const float * v = vsep ? vsep + (size_t) r * GQA : qkv + (size_t) r * rstride + E + GQA;
            vb[r][d] = __float2half_rn(v[(size_t) kvh * HD + d]);
        }
    }
    __syncthreads();
    if (head == kvh * gdiv) {                           // first query head of the group: this token's K / V row goes to the cache
        const size_t co = (size_t) (cell0 + t) * GQA + (size_t) kvh * HD;
        if (tid < HD / 8) reinterpret_cast<uint4 *>(kcache + co)[tid] = reinterpret_cast<const uint4 *>(kb[t])[tid];
        else if (tid < HD / 4) reinterpret_cast<uint4 *>(vcache + co)[tid - HD / 8] = reinterpret_cast<const uint4 *>(vb[t])[tid - HD / 8];
    }

    // phase 1: scores, in the reference's order (ggml_vec_dot_f16, ggml.c:1200-1235 with the AVX2 reduce macro :772-790):
    // 32 fp32 accumulators, element e of the head goes to accumulator e mod 32, then ((0+2)+(1+3)) over the four 8-lane
    // vectors, low half + high half, two horizontal adds.
    // A cell is shared by two neighbouring lanes: the even one forms vectors 0 and 2 of the four, the odd one 1 and 3, and the
    // first addition of the reduce -- (0 + 2) and (1 + 3) -- happens before the halves meet through one shuffle per lane.
    float lmax = -INFINITY;
    for (int i0 = 0; i0 < n; i0 += ATQ_THREADS / 2) {
        const int  i  = i0 + (tid >> 1), ic = max(min(i, n - 1), 0), bi = ic - cell0;
        const bool inb = bi >= 0 && bi < ntok;
        bool live = live0;
        if (i0 > 0) {
            // later rounds (n > 128): the row's address does not depend on the mask byte, both loads go out together
            const uint4 * kr = reinterpret_cast<const uint4 *>(khead + (size_t) ic * GQA);
#pragma unroll
            for (int q = 0; q < HD / 16; q++) kreg[q] = inb ? make_uint4(0, 0, 0, 0) : kr[(q >> 1) * 4 + 2 * (q & 1) + hsel];
            live = i < n && !mrow[ic];
        }
        if (inb) {
            const uint4 * kr = reinterpret_cast<const uint4 *>(kb[bi]);
#pragma unroll
            for (int q = 0; q < HD / 16; q++) kreg[q] = kr[(q >> 1) * 4 + 2 * (q & 1) + hsel];
        }
        float r[8];
#pragma unroll
        for (int l = 0; l < 8; l++) r[l] = 0.f;
        if (live) {
            float a[2][8];
#pragma unroll
            for (int j = 0; j < 2; j++)
#pragma unroll
                for (int l = 0; l < 8; l++) a[j][l] = 0.f;
#pragma unroll
            for (int st = 0; st < HD / 32; st++)
#pragma unroll
                for (int jj = 0; jj < 2; jj++) {
                    const int     j  = 2 * jj + hsel;
                    const uint4   k8 = kreg[st * 2 + jj];
                    const __half2 h0 = *reinterpret_cast<const __half2 *>(&k8.x), h1 = *reinterpret_cast<const __half2 *>(&k8.y);
                    const __half2 h2 = *reinterpret_cast<const __half2 *>(&k8.z), h3 = *reinterpret_cast<const __half2 *>(&k8.w);
                    const float4  qa = *reinterpret_cast<const float4 *>(sq + st * 32 + j * 8), qb = *reinterpret_cast<const float4 *>(sq + st * 32 + j * 8 + 4);
                    a[jj][0] = fmaf(__low2float(h0), qa.x, a[jj][0]); a[jj][1] = fmaf(__high2float(h0), qa.y, a[jj][1]);
                    a[jj][2] = fmaf(__low2float(h1), qa.z, a[jj][2]); a[jj][3] = fmaf(__high2float(h1), qa.w, a[jj][3]);
                    a[jj][4] = fmaf(__low2float(h2), qb.x, a[jj][4]); a[jj][5] = fmaf(__high2float(h2), qb.y, a[jj][5]);
                    a[jj][6] = fmaf(__low2float(h3), qb.z, a[jj][6]); a[jj][7] = fmaf(__high2float(h3), qb.w, a[jj][7]);
                }
#pragma unroll
            for (int l = 0; l < 8; l++) r[l] = __fadd_rn(a[0][l], a[1][l]);             // even lane: 0 + 2, odd lane: 1 + 3
        }
        // both lanes of a pair take the same branch (same cell), so the shuffles below are convergent
#pragma unroll
        for (int l = 0; l < 8; l++) {
            const float o = __shfl_xor_sync(0xffffffffu, r[l], 1);
            r[l] = hsel ? __fadd_rn(o, r[l]) : __fadd_rn(r[l], o);                      // (0 + 2) + (1 + 3)
        }
        float sc = -INFINITY;
        if (live) {
            const float t0 = __fadd_rn(r[0], r[4]), t1 = __fadd_rn(r[1], r[5]), t2 = __fadd_rn(r[2], r[6]), t3 = __fadd_rn(r[3], r[7]);
            sc = __fmul_rn(__fadd_rn(__fadd_rn(t0, t1), __fadd_rn(t2, t3)), kq_scale);
        }
        if (i < n && !hsel) s[i] = sc;
        lmax = fmaxf(lmax, sc);
    }
    lmax = warp_max(lmax);
    if (lane == 0) s_redf[warp] = lmax;
    __syncthreads();
    float mx = s_redf[0];
#pragma unroll
    for (int w = 1; w < ATQ_THREADS / 32; w++) mx = fmaxf(mx, s_redf[w]);

    // phase 2: exp through the fp16 table, exact double sum, fp16-rounded probabilities
    double sum = 0.0;
    for (int i = tid; i < n; i += ATQ_THREADS) {
        const float x = s[i];
        float       v = 0.0f;
        if (x != -INFINITY) { v = pib_h2f(exp_lut[pib_f2h(__fsub_rn(x, mx))]); sum += (double) v; }
        s[i] = v;
    }
    sum = warp_sum_d(sum);
    if (lane == 0) s_redd[warp] = sum;
    asm volatile("cp.async.wait_all;" ::: "memory");            // this thread's V chunks have landed; the barrier publishes them
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int w = 0; w < ATQ_THREADS / 32; w++) tot += s_redd[w];   // every thread forms the same total, in warp order
    const float inv = (float) (1.0 / tot);
    for (int i = tid; i < n; i += ATQ_THREADS) s[i] = __half2float(__float2half_rn(__fmul_rn(s[i], inv)));
    __syncthreads();

    // phase 3: O = V . P in the reference's order (the same ggml_vec_dot_f16, now over the cells): cell c < np = n & ~31 goes
    // to accumulator c mod 32, each accumulator a sequential fma chain over its cells; then the same tree per output element;
    // the n - np cells left over are added one by one in double.  A thread owns a pair of output dims and APG of the 32
    // accumulators; the 32 x HD partial sums meet in shared memory.
    constexpr int TPR = HD / 2, NGR = ATQ_THREADS / TPR, APG = 32 / NGR;      // hd 128: 4 groups x 8 accumulators; hd 64: 8 x 4
    const int grp = tid / TPR, dp = tid % TPR;
    const int np  = n & ~31;
    auto vrow = [&](int c) -> const __half * {          // row of cell c for this KV head: batch rows, staged rows, or the cache
        const int bi = c - cell0;
        return (bi >= 0 && bi < ntok) ? vb[bi] : c < vcap ? vs + (size_t) c * HD : vhead + (size_t) c * GQA;
    };
    float o0[APG], o1[APG];
#pragma unroll
    for (int u = 0; u < APG; u++) { o0[u] = 0.f; o1[u] = 0.f; }
#pragma unroll 2
    for (int base = 0; base < np; base += 32) {
        __half2 vv[APG];
        float   pp[APG];
#pragma unroll
        for (int u = 0; u < APG; u++) {
            const int c = base + grp * APG + u;
            pp[u] = s[c];
            vv[u] = pp[u] == 0.f ? __half2()            // p == 0: the fma leaves the accumulator as it is (and the row may never have been written)
                                 : *reinterpret_cast<const __half2 *>(vrow(c) + dp * 2);
        }
#pragma unroll
        for (int u = 0; u < APG; u++) { o0[u] = fmaf(__low2float(vv[u]), pp[u], o0[u]); o1[u] = fmaf(__high2float(vv[u]), pp[u], o1[u]); }
    }
#pragma unroll
    for (int u = 0; u < APG; u++) { red[(grp * APG + u) * HD + dp * 2] = o0[u]; red[(grp * APG + u) * HD + dp * 2 + 1] = o1[u]; }
    __syncthreads();
    if (tid < HD) {
        float r[8];
#pragma unroll
        for (int l = 0; l < 8; l++)
            r[l] = __fadd_rn(__fadd_rn(red[(0 * 8 + l) * HD + tid], red[(2 * 8 + l) * HD + tid]), __fadd_rn(red[(1 * 8 + l) * HD + tid], red[(3 * 8 + l) * HD + tid]));
        const float t0 = __fadd_rn(r[0], r[4]), t1 = __fadd_rn(r[1], r[5]), t2 = __fadd_rn(r[2], r[6]), t3 = __fadd_rn(r[3], r[7]);
        double sumf = (double) __fadd_rn(__fadd_rn(t0, t1), __fadd_rn(t2, t3));
        for (int c = np; c < n; c++) {
            const float p = s[c];
            if (p != 0.f) sumf += (double) __fmul_rn(__half2float(vrow(c)[tid]), p);
        }
        out[(size_t) t * E + (size_t) head * HD + tid] = (float) sumf;
    }
}

template <int HD>
static int launch_att_qh(const float * qkv, const float * vsep, const uint8_t * desc, const float2 * rope, int rope_n_pos, int mode, __half * kc, __half * vc,
                         const uint8_t * mask, const int32_t * scalars, int E, int GQA, int n_ctx, int n_tokens, int n_head, int gdiv, float * out,
                         const int * cancel, cudaStream_t st, int early) {
    const int    vcap = std::min(n_ctx, ATQ_VCAP);
    const size_t smem = ((size_t) n_ctx + 32 * HD) * sizeof(float) + (size_t) vcap * HD * sizeof(__half);   // scores + 32 x HD partial sums + staged V rows
    static size_t configured = 0;
    if (smem > configured) {
        // several CTAs per SM (a verification batch of 3+ tokens is more CTAs than SMs): ask for the large shared-memory carve-out
        PIB_CHECK(cudaFuncSetAttribute(attention_qh_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) std::max(smem, (size_t) 32 * 1024)));
        PIB_CHECK(cudaFuncSetAttribute(attention_qh_kernel<HD>, cudaFuncAttributePreferredSharedMemoryCarveout, (int) cudaSharedmemCarveoutMaxShared));
        configured = smem;
    }
    pib_launch(attention_qh_kernel<HD>, dim3(n_head, n_tokens), dim3(ATQ_THREADS), smem, st, qkv, vsep, desc, rope, rope_n_pos, mode, kc, vc, mask, scalars,
               pib_get_luts().exp_f16, E, GQA, n_ctx, 1.0f / sqrtf((float) HD), out, cancel, gdiv, vcap, early);
    PIB_CHECK(cudaGetLastError());
    return PIB_OK;
}

bool pib_attention_supported(int hd, int n_head, int n_head_kv, int n_ctx) {
    if (hd != 64 && hd != 128) return false;
    if (n_head % n_head_kv) return false;
    return ((size_t) n_ctx + 32 * hd) * sizeof(float) + (size_t) std::min(n_ctx, ATQ_VCAP) * hd * 2 <= 200 * 1024;      // the scores of one (head, token) live in shared memory
}

// RoPE (mode 0: adjacent pairs, 2: NeoX) + KV store + attention of one chunk: qkv = the fused projection's output [n_tokens][E + 2 GQA]
// (or [E + GQA] with V in vsep)
int pib_attention(const float * qkv, const float * vsep, const uint8_t * desc, const float2 * rope, int rope_n_pos, int mode, __half * kc, __half * vc,
                  const uint8_t * mask, const int32_t * scalars, int n_tokens, int n_head, int n_head_kv, int hd, int n_ctx, float * out,
                  const int * cancel, cudaStream_t st, bool cache_final) {
    const int E = n_head * hd, GQA = n_head_kv * hd, G = n_head / n_head_kv;
    if (hd == 128) return launch_att_qh<128>(qkv, vsep, desc, rope, rope_n_pos, mode, kc, vc, mask, scalars, E, GQA, n_ctx, n_tokens, n_head, G, out, cancel, st, cache_final ? 1 : 0);
    if (hd == 64)  return launch_att_qh<64>(qkv, vsep, desc, rope, rope_n_pos, mode, kc, vc, mask, scalars, E, GQA, n_ctx, n_tokens, n_head, G, out, cancel, st, cache_final ? 1 : 0);
    pib_set_error("attention: unsupported head_dim %d", hd);
    return PIB_ERR_UNSUPP;
}
